#!/usr/bin/env python
"""bench.py -- log-mel audio-seconds/sec of the B200 front-end (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path over one batch of the headline workload (BASELINE.json
configs[1]): 64 synthetic 10 s clips at 22.05 kHz = 640 audio-seconds = 55,168 frames per GPU.

* `value`  device-timed whole-job throughput, waveforms resident in HBM.  Blocks of K steps are repeated until
           the timed region is >= 0.5 s (`repeats`, `timed_region_s`): a 20-step run is 3 ms and cannot stand as a
           measurement.
* `e2e`    the same metric through the public module (`TorchMelSpectrogram.forward`) from pinned HOST buffers with
           the result read back to the host, copies inside the timed region; `e2e_pcm16` is the same loop fed with
           what a 16-bit WAV holds (int16, half the H2D bytes).  Both at every N.
* `sweep`  BASELINE config 4, strong-scaled over the N ranks: a fixed synthetic corpus (seeded per UTTERANCE, so the
           data do not depend on the rank layout), length-sorted batches, LPT sharding, statistics fused into the
           log-mel kernel, one fp64 NCCL all-reduce; rank 0 also runs the whole corpus alone and reports the largest
           relative difference between the sharded and the single-rank statistics.
* `configs`, `ref_gpu`, `roofline`, `cpu_baseline`, `ingest`, `vq`: rank 0, N = 1.

`--impl reference` times the reference's own CPU call sequence (torchaudio MelSpectrogram -> log(clamp) ->
/ mel_norms, the oracle's `callsite_logmel`) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B_CLIPS, CLIP_LEN, SR = 64, 220500, 22050
FLOP_PER_FRAME = 2.185e6          # algorithmic FLOP per frame (DFT-as-GEMM, unpruned, single pass), V1
FLOP_PER_FRAME_V2 = 2.0 * 1024 * 1026 + 3 * 513 + 2 * 513 * 100 + 513 + 200   # SURVEY 8d "For V2": 100 mels, + sqrt
BYTES_PER_FRAME = 1344.0          # 1024 B in + 320 B out
# tensor FLOP k_logmel2 actually issues per frame: 2 parities x sp sub-passes x [33 cos + 32 sin k-steps] x 3 TF32
# products x 2 * nb * 8  (V1: sp 1, nb 192; V2: sp 2, nb 128)
EXEC_FLOP_PER_FRAME = 2 * (32 + 32) * 3 * 2 * 192 * 8      # the centre tap is added by the epilogue: 32 cosine k-steps
EXEC_FLOP_PER_FRAME_V2 = 2 * 2 * (32 + 32) * 3 * 2 * 128 * 8
WORKLOAD = "64x10s_22050Hz_uniform[-1,1)_TorchMelSpectrogram_V1"
ROTATE = 4                        # distinct input batches: 4 x 56.4 MB = 226 MB > 126 MB L2
MIN_REGION_S = 0.5
SWEEP_UTTS = 40000                # config 4 corpus in the bench (>= 20 batches of 256 clips per rank at N = 8)
SWEEP_CLIPS = 256


def frames_of(L: int) -> int:
    return 1 + L // 256


def config_dict(world: int) -> dict:
    return {"workload": WORKLOAD, "clips_per_step_per_gpu": B_CLIPS, "clip_seconds": 10,
            "frames_per_step_per_gpu": B_CLIPS * frames_of(CLIP_LEN),
            "sharding": f"dp{world}: one batch per rank, no data-path collective (the `sweep` object is the sharded path)",
            "l2": f"inputs rotate over {ROTATE} distinct batches (226 MB > 126 MB L2)"}


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.thread = [], None, None
        self.gpu_index = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "50", "-i", str(self.gpu_index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()
        time.sleep(0.12)                       # the first sample lands before the timed region starts

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t_begin: float = 0.0, t_end: float = float("inf")) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.08)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        for ts, r in self.rows:
            if not (t_begin <= ts <= t_end + 0.06):
                continue
            f = [c.strip() for c in r.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def peaks() -> dict:
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_burst": d["bf16_tflops"], "bf16_sustained": d["bf16_tflops_sustained"],
                "src": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "src": "fallback"}


def wait_for_clocks(dev, max_wait_s: float = 8.0) -> dict:
    """After a sustained 1 kW matmul the board holds the SM clock near 1.1 GHz for a second or two even when idle
    (tools/gpu_c5_check.py: config 5 took 1.05 ms right after the TF32 measurement, 0.65 ms three seconds later), so
    whatever is timed next would be measured at the previous workload's clock.  Idle until the clock is back."""
    import torch
    torch.cuda.synchronize(dev)
    t0 = time.time()
    mhz = None
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(dev.index or 0)
        top = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
        while time.time() - t0 < max_wait_s:
            mhz = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
            if mhz >= 0.97 * top:
                break
            time.sleep(0.25)
    except Exception:
        time.sleep(3.0)
    return {"waited_s": round(time.time() - t0, 2), "sm_mhz": mhz}


def measure_tf32_peak(dev, sustain_s: float = 2.0, part: str = "both") -> dict:
    """torch.matmul fp32 with allow_tf32, 8192^3 -- best of 10 (burst) and back to back for
    `sustain_s` seconds (sustained), CUDA events; the same method as MEASURED_PEAKS.json's bf16 rows.
    part = "burst" (20 ms of work) | "sustained" | "both"."""
    import torch
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        a = torch.randn(n, n, device=dev)
        b = torch.randn(n, n, device=dev)
        c = torch.empty(n, n, device=dev)
        flop = 2.0 * n ** 3
        for _ in range(3):
            torch.matmul(a, b, out=c)
        torch.cuda.synchronize(dev)
        best = float("inf")
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); torch.matmul(a, b, out=c); e1.record()
            torch.cuda.synchronize(dev)
            best = min(best, e0.elapsed_time(e1))
        if part == "burst":
            return {"burst": flop / (best * 1e-3) / 1e12}
        per = max(1, int(sustain_s / (best * 1e-3)))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(per):
            torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize(dev)
        sustained = flop * per / (e0.elapsed_time(e1) * 1e-3) / 1e12
        return {"burst": flop / (best * 1e-3) / 1e12, "sustained": sustained, "sustain_s": e0.elapsed_time(e1) * 1e-3,
                "how": "torch.matmul fp32, allow_tf32=True, 8192^3 (2*N^3 flop): best of 10 / back to back"}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def cpu_reference_throughput(steps: int, warmup: int, batch):
    """The reference's CPU path on the host cores: audio-s/s over `steps` passes of `batch`."""
    import torch
    from oracle import mel_oracle as orc
    norms = orc.default_mel_norms(80)
    for _ in range(max(1, warmup)):
        orc.callsite_logmel(batch, orc.V1, norms)
    t0 = time.perf_counter()
    for _ in range(steps):
        orc.callsite_logmel(batch, orc.V1, norms)
    dt = time.perf_counter() - t0
    audio_s = batch.shape[0] * batch.shape[1] / SR
    return audio_s * steps / dt, dt / steps, torch.get_num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from oracle import mel_oracle as orc
    torch.set_num_threads(os.cpu_count() or 1)      # torchrun pins OMP_NUM_THREADS=1: use every host core instead
    x = torch.rand(B_CLIPS, CLIP_LEN, generator=torch.Generator().manual_seed(1234)) * 2 - 1
    # a step is a bounded sample of the 64 x 10 s batch: as many of its clips as keep K + W steps within ~90 s
    norms = orc.default_mel_norms(80)
    orc.callsite_logmel(x, orc.V1, norms)
    t0 = time.perf_counter()
    orc.callsite_logmel(x, orc.V1, norms)
    t_full = time.perf_counter() - t0
    per_step = 90.0 / max(1, args.steps + args.warmup)
    clips = max(1, min(B_CLIPS, int(B_CLIPS * per_step / t_full)))
    val, sec_per_step, cores = cpu_reference_throughput(args.steps, args.warmup, x[:clips])
    line = {
        "impl": "reference", "metric": "log-mel audio-seconds/sec", "value": val, "unit": "audio-s/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec_per_step * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args.gpus),
        "frames_per_s": val * SR / 256.0,
        "cpu_baseline": {"value": val, "unit": "audio-s/s", "cores": cores, "kind": "port",
                         "sample": f"{args.steps} steps x {clips} of the batch's 64 10 s clips through the reference's call "
                                   "sequence (torchaudio MelSpectrogram/torch.stft MKL -> log(clamp) -> /mel_norms) on the "
                                   f"host; the full batch takes {t_full * 1e3:.0f} ms"},
        "e2e": {"value": val, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------------
# synthetic corpus of config 4: utterance u's samples depend on u alone (counter hash), never on the rank layout
# ----------------------------------------------------------------------------------------------------------------
def synth_utterances(utt_ids, max_len, dev):
    """[B, max_len] fp32 in [-1, 1) on a 16-byte row pitch (read in place): x[b, i] = hash(utt_ids[b], i)."""
    import torch
    B = len(utt_ids)
    pitch = (max_len + 3) // 4 * 4
    out = torch.empty(B, pitch, dtype=torch.float32, device=dev)
    ids = torch.as_tensor(utt_ids, dtype=torch.int64, device=dev)[:, None]
    step = 1 << 20                                      # columns per chunk: bounds the int64 temporaries (~2 GB each)
    for c0 in range(0, pitch, step):
        c1 = min(pitch, c0 + step)
        i = torch.arange(c0, c1, dtype=torch.int64, device=dev)[None, :]
        h = i * 6364136223846793005 + ids * 1442695040888963407 + 0x2545F491
        h = (h ^ (h >> 29)) * (-4658895280553007687)    # 0xBF58476D1CE4E5B9 as int64 (wraps)
        h = h ^ (h >> 32)
        out[:, c0:c1] = (h & 0xFFFFFF).to(torch.float32) * (1.0 / 8388608.0) - 1.0
    return out[:, :max_len]


def run_sweep(plan, dev, world, rank, dist, barrier):
    """BASELINE config 4 (strong scaling): returns the `sweep` object on rank 0."""
    import numpy as np
    import torch
    from xtts_b200.sweep import MelSweep, norms_from_stats
    rng = np.random.default_rng(4321)
    lens = [int(round(SR * u)) for u in rng.uniform(1.0, 30.0, size=SWEEP_UTTS)]
    sweep = MelSweep(plan, lens, max_clips=SWEEP_CLIPS, rank=rank, world_size=world)
    mine = sweep.my_buckets()
    resident = {}
    for b in mine:                                      # this rank's share of the corpus, resident in HBM
        resident[b.indices] = (synth_utterances(b.indices, b.max_len, dev),
                               torch.tensor([lens[i] for i in b.indices], device=dev))
    fetch = lambda idx, max_len: resident[tuple(idx)]
    full = list(sweep.owned[rank])
    sweep.owned[rank] = full[:3]
    sweep.run_stats(fetch)                              # warm-up on three batches (incl. the all-reduce)
    sweep.owned[rank] = full
    times, ar_us, acc = [], [], None
    for _ in range(3):
        barrier()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        a = torch.zeros(2 * plan.n_mels + 1, dtype=torch.float64, device=dev)
        for b in mine:
            wav, l = resident[b.indices]
            plan.forward(wav, l, t_max=plan.frames(b.max_len), stats_acc=a)
        e1.record()
        if world > 1:
            dist.all_reduce(a, op=dist.ReduceOp.SUM)
        e2.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e2)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times.append(float(t.item()))
        ar_us.append(e1.elapsed_time(e2) * 1e3)
        acc = a
    plan.status()
    fmax = torch.tensor([float(sum(b.frames for b in mine))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(fmax, op=dist.ReduceOp.MAX)
    # the whole corpus on rank 0 alone (generated batch by batch, not timed): sharded == single-rank statistics?
    rel = None
    if rank == 0:
        single = torch.zeros_like(acc)
        for b in sweep.buckets:
            if b.indices in resident:
                wav, l = resident[b.indices]
            else:
                wav = synth_utterances(b.indices, b.max_len, dev)
                l = torch.tensor([lens[i] for i in b.indices], device=dev)
            plan.forward(wav, l, t_max=plan.frames(b.max_len), stats_acc=single)
        plan.status()
        rel = float(((acc[:-1] - single[:-1]).abs() / single[:-1].abs()).max())
        assert float(acc[-1]) == float(single[-1]) == float(sum(frames_of(L) for L in lens))
    if world > 1:
        barrier()
    if rank != 0:
        return None
    ms = min(times)
    audio = sum(lens) / SR
    frames = sum(frames_of(L) for L in lens)
    mean, _ = norms_from_stats(acc)
    return {"workload": f"config 4: {SWEEP_UTTS} synthetic utterances of 1-30 s (seeded per utterance), length-sorted "
                        f"batches of {SWEEP_CLIPS} clips, LPT-sharded over the ranks, statistics fused into the log-mel "
                        "kernel, one fp64 all-reduce (NCCL)",
            "scaling": "strong", "audio_s_per_s": audio / (ms * 1e-3), "frames_per_s": frames / (ms * 1e-3),
            "ms": ms, "ms_all_passes": times, "allreduce_us": statistics.median(ar_us),
            "batches_total": len(sweep.buckets), "batches_per_rank": len(mine),
            "imbalance": float(fmax.item()) * world / frames,
            "audio_seconds": audio, "frames": frames,
            "resident_input_GB_rank0": sum(w.numel() for w, _ in resident.values()) * 4 / 1e9,
            "mel_norms_max_rel_diff_vs_single_rank": rel, "mel_norms_head": [round(float(v), 6) for v in mean[:4]],
            "launches_per_batch": 2,
            "timing": "CUDA events around this rank's batches + the all-reduce, max over ranks, best of 3 passes"}


def timed(torch, fn, iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def run_configs(dev, tf32_peak, hbm_gbs):
    """The other BASELINE configs + the V2 parametrisation + the reference's own GPU path, one GPU, inputs resident."""
    import numpy as np
    import torch
    from oracle import mel_oracle as orc
    from xtts_b200 import MelSpectrogramFeatures, TorchMelSpectrogram
    out = {}
    g = torch.Generator(device=dev).manual_seed(99)
    v1 = TorchMelSpectrogram(mel_norm_file=None)
    v1.mel_norms = torch.from_numpy(orc.default_mel_norms(80))
    plan1 = v1._plan_for(dev)

    def entry(ms, frames, audio_s, v2=False, kernel_ms=None):
        alg = FLOP_PER_FRAME_V2 if v2 else FLOP_PER_FRAME
        exe = EXEC_FLOP_PER_FRAME_V2 if v2 else EXEC_FLOP_PER_FRAME
        k = kernel_ms if kernel_ms is not None else ms
        return {"ms": ms, "kernel_ms": kernel_ms, "audio_s_per_s": audio_s / (ms * 1e-3), "frames_per_s": frames / (ms * 1e-3),
                "roofline_frac": alg * frames / (k * 1e-3) / 1e12 / tf32_peak,
                "executed_frac": exe * frames / (k * 1e-3) / 1e12 / tf32_peak,
                "hbm_frac": (1024.0 + (400.0 if v2 else 320.0)) * frames / (k * 1e-3) / 1e9 / hbm_gbs}

    def kernel_alone(plan, fn, n):
        plan.profile_enable(n)
        for _ in range(n):
            fn()
        ms = plan.profile_collect(n)
        plan.profile_enable(0)
        return statistics.median(ms) if ms else None

    # C1: one 10 s clip
    x = torch.rand(1, CLIP_LEN, device=dev, generator=g) * 2 - 1
    out["C1_1x10s"] = entry(timed(torch, lambda: v1(x), 200), frames_of(CLIP_LEN), 10.0,
                            kernel_ms=kernel_alone(plan1, lambda: v1(x), 50))
    # C5: 32 x 120 s
    x = torch.rand(32, 2646000, device=dev, generator=g) * 2 - 1
    out["C5_32x120s"] = entry(timed(torch, lambda: v1(x), 20), 32 * frames_of(2646000), 32 * 120.0,
                              kernel_ms=kernel_alone(plan1, lambda: v1(x), 10))
    del x
    # V2 dense: 64 x 10 s at 24 kHz through MelSpectrogramFeatures (what save_mel_to_disk.py and the loaders call)
    v2 = MelSpectrogramFeatures()
    plan2 = v2._plan_for(dev)
    xs = [torch.rand(64, 240000, device=dev, generator=g) * 2 - 1 for _ in range(ROTATE)]
    it = iter(range(10 ** 9))
    fn = lambda: v2(xs[next(it) % ROTATE])
    out["V2_64x10s_24kHz"] = entry(timed(torch, fn, 200), 64 * frames_of(240000), 640.0, v2=True,
                                   kernel_ms=kernel_alone(plan2, fn, 100))
    del xs
    # C3 through V2: 256 clips of 1..30 s at 24 kHz, length-sorted ragged batches of 64 clips, masks + frame counts
    rng = np.random.default_rng(1234)
    lens = sorted(int(round(24000 * u)) for u in rng.uniform(1.0, 30.0, size=256))
    batches = []
    for k in range(0, 256, 64):
        ll = lens[k:k + 64]
        pitch = (max(ll) + 3) // 4 * 4
        batches.append((torch.rand(64, pitch, device=dev, generator=g)[:, :max(ll)] * 2 - 1, torch.tensor(ll, device=dev)))

    def c3():
        for wav, l in batches:
            plan2.forward(wav, l, want_mask=True, t_max=plan2.frames(int(wav.shape[1])))
    f3 = sum(frames_of(L) for L in lens)
    out["C3_v2_256_ragged_1-30s"] = entry(timed(torch, c3, 20), f3, sum(lens) / 24000.0, v2=True)
    out["C3_v2_256_ragged_1-30s"]["note"] = "4 length-sorted batches of 64 clips, per-clip reflect padding, masks + frame counts"
    del batches
    # the reference's own GPU path on the same B200: torchaudio (cuFFT + cuBLAS + elementwise), utils.py:509-522 with inp on CUDA
    x = torch.rand(B_CLIPS, CLIP_LEN, device=dev, generator=g) * 2 - 1
    tr = orc._callsite_transform(orc.V1).to(dev)
    nd = torch.from_numpy(orc.default_mel_norms(80)).to(dev)
    ref_fn = lambda: torch.log(torch.clamp(tr(x), min=1e-5)) / nd[None, :, None]
    ms_ref = timed(torch, ref_fn, 30)
    ms_ours = timed(torch, lambda: v1(x), 100)
    err = float((v1(x) - ref_fn()).abs().max())
    tr.to("cpu")                                        # the cached transform serves the CPU baseline too
    ref_gpu = {"workload": WORKLOAD, "ms": ms_ref, "audio_s_per_s": 640.0 / (ms_ref * 1e-3), "ours_ms_same_batch": ms_ours,
               "speedup": ms_ref / ms_ours, "max_abs_diff_ours_vs_ref_gpu": err,
               "how": "the reference's call sequence (torchaudio MelSpectrogram -> log(clamp) -> / mel_norms) with the input on "
                      "the B200: cuFFT + cuBLAS + elementwise kernels, CUDA events, same batch resident in HBM"}
    return out, ref_gpu


def run_ingest(dev, mod, pk, audio_step):
    # SURVEY 8f N3, reported beside the headline (not part of `value`): 16-bit stereo PCM at 44.1 kHz resident
    # in HBM -> load_audio_pcm16 (fused decode + 2:1 sinc decimator + clip) -> the same mel call
    import torch
    from xtts_b200 import load_audio_pcm16
    pcm = torch.randint(-32768, 32767, (B_CLIPS, 2 * CLIP_LEN, 2), dtype=torch.int16, device=dev)
    flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device=dev)
    ms_r, ms_p = [], []
    for i in range(3 + 20):
        flush.fill_(0.0)
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        wav = load_audio_pcm16(pcm, 2 * SR, SR)
        e1.record()
        mod(wav)
        e2.record()
        torch.cuda.synchronize(dev)
        if i >= 3:
            ms_r.append(e0.elapsed_time(e1))
            ms_p.append(e0.elapsed_time(e2))
    r_ms, p_ms = statistics.median(ms_r), statistics.median(ms_p)
    r_bytes = B_CLIPS * (2 * CLIP_LEN * 2 * 2 + CLIP_LEN * 4)
    return {"workload": "64x10s 16-bit stereo PCM @ 44100 Hz -> 22050 Hz -> log-mel, resident in HBM",
            "pipeline_audio_s_per_s": audio_step / (p_ms * 1e-3), "pipeline_ms": p_ms,
            "resample_kernel": "k_decimate<2,28,int16>", "resample_ms": r_ms,
            "resample_GB/s": r_bytes / (r_ms * 1e-3) / 1e9,
            "resample_hbm_frac": r_bytes / (r_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
            "l2": "256 MB flush write before every iteration"}


def run_vq(dev, tf32_peak):
    # SURVEY 8f N2, reported beside the headline (not part of `value`): the DVAE codebook search behind the mel
    # path at this batch's size -- 64 clips x 216 encoder rows of dim 512 against 8192 codes, rows resident
    import torch
    from xtts_b200.vq import VQPlan
    g = torch.Generator(device=dev).manual_seed(4242)
    emb = torch.randn(512, 8192, device=dev, generator=g)
    vplan = VQPlan(emb, dev)
    rows = B_CLIPS * 216
    zs = [emb.T[torch.randint(0, 8192, (rows,), device=dev, generator=g)]
          + 0.5 * torch.randn(rows, 512, device=dev, generator=g) for _ in range(4)]
    ms_v = []
    for i in range(3 + 20):
        e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
        e0.record()
        vind = vplan.nearest(zs[i & 3])[0]
        e1.record()
        torch.cuda.synchronize(dev)
        if i >= 3:
            ms_v.append(e0.elapsed_time(e1))
    vplan.status()
    v_ms = statistics.median(ms_v)
    v_flop = 2.0 * rows * 512 * 8192
    zl = zs[(3 + 20 - 1) & 3]                      # the batch of the last timed call
    dist_ref = zl.pow(2).sum(1, keepdim=True) - 2 * zl @ emb + emb.pow(2).sum(0, keepdim=True)
    return {"workload": "64x216 rows of dim 512 -> nearest of 8192 codes (Quantize.forward, xtts_dvae.py:87-90), resident in HBM",
            "kernels": {"k_vq_tile": 1, "k_vq_tc": 1, "k_vq_finish": 1}, "ms": v_ms,
            "rows_per_s": rows / (v_ms * 1e-3),
            "algorithmic_tflops": v_flop / (v_ms * 1e-3) / 1e12,
            "frac_of_tf32_peak": v_flop / (v_ms * 1e-3) / 1e12 / tf32_peak,
            "executed_frac": 3.0 * v_flop / (v_ms * 1e-3) / 1e12 / tf32_peak,
            "agreement_with_library_sequence_on_gpu": float(((-dist_ref).max(1)[1] == vind).float().mean()),
            "l2": "inputs rotate over 4 batches (113 MB of rows); the 33 MB of codebook tiles stay in L2"}


def run_ours(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the log-mel path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from xtts_b200.hostmem import bind_to_gpu_numa, unbind
    all_cpus = os.sched_getaffinity(0)
    numa = bind_to_gpu_numa(local)          # before any pinned allocation: the staging buffers live next to the GPU
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from oracle import mel_oracle as orc
    from xtts_b200 import TorchMelSpectrogram, decode_pcm16

    norms = torch.from_numpy(orc.default_mel_norms(80))
    mod = TorchMelSpectrogram(mel_norm_file=None)
    mod.mel_norms = norms
    plan = mod._plan_for(dev)

    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    inputs = [torch.rand(B_CLIPS, CLIP_LEN, device=dev, generator=gen) * 2 - 1 for _ in range(ROTATE)]
    frames_step = B_CLIPS * frames_of(CLIP_LEN)
    audio_step = B_CLIPS * CLIP_LEN / SR
    K, W = args.steps, max(3, args.warmup)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def max_over_ranks(v: float) -> float:
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- device-resident throughput (`value`) ----------------
    for i in range(W):
        plan.forward(inputs[i % ROTATE])
    plan.status()
    # how many K-step blocks make the timed region >= MIN_REGION_S (the same count on every rank)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    probe = min(K, 50)
    e0.record()
    for i in range(probe):
        plan.forward(inputs[i % ROTATE])
    e1.record()
    torch.cuda.synchronize(dev)
    est_step_ms = max_over_ranks(e0.elapsed_time(e1) / probe)
    R = max(1, int(MIN_REGION_S * 1e3 / (est_step_ms * K) + 0.999))
    KP = min(K * R, 2048)                  # kernel-only CUDA-event pairs: the first KP timed steps
    plan.profile_enable(KP)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    t_begin = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    e0.record()
    for i in range(K * R):
        plan.forward(inputs[i % ROTATE])
        launches += plan.last_launches
    e1.record()
    barrier()
    t_end = time.perf_counter()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop(t_begin, t_end) if rank == 0 else None
    kern_ms = plan.profile_collect(KP)
    plan.profile_enable(0)
    plan.status()
    ms_step = ms_total / (K * R)
    value = audio_step * world / (ms_step * 1e-3)

    # ---------------- end to end through the public module from pinned host memory ----------------
    T = frames_of(CLIP_LEN)
    host_out = [torch.empty(B_CLIPS, T, 80, dtype=torch.float32).pin_memory() for _ in range(2)]
    copy_s, out_s = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    main_s = torch.cuda.current_stream(dev)

    def e2e_loop(n, host_in, dev_in, fn):
        # double-buffered: H2D of step i+1 and D2H of step i-1 overlap the kernel of step i
        ready = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]
        with torch.cuda.stream(copy_s):
            dev_in[0].copy_(host_in[0], non_blocking=True)
            ready[0].record(copy_s)
        for i in range(n):
            s = i & 1
            if i + 1 < n:
                with torch.cuda.stream(copy_s):
                    if i >= 1:
                        copy_s.wait_event(consumed[s ^ 1])
                    dev_in[s ^ 1].copy_(host_in[s ^ 1], non_blocking=True)
                    ready[s ^ 1].record(copy_s)
            main_s.wait_event(ready[s])
            mel = fn(dev_in[s])                        # the call a user makes
            consumed[s].record(main_s)
            done = torch.cuda.Event()
            done.record(main_s)
            with torch.cuda.stream(out_s):
                out_s.wait_event(done)
                host_out[s].copy_(mel.transpose(1, 2), non_blocking=True)
                mel.record_stream(out_s)
        out_s.synchronize()

    def e2e_measure(host_in, dev_in, fn):
        e2e_loop(W, host_in, dev_in, fn)
        barrier()
        t0 = time.perf_counter()
        e2e_loop(20, host_in, dev_in, fn)
        torch.cuda.synchronize(dev)
        est = max_over_ranks((time.perf_counter() - t0) / 20)
        n = min(4000, max(K if K <= 400 else 400, 200, int(MIN_REGION_S / est + 0.999)))
        barrier()
        t0 = time.perf_counter()
        e2e_loop(n, host_in, dev_in, fn)
        torch.cuda.synchronize(dev)
        barrier()
        sec = max_over_ranks(time.perf_counter() - t0)
        return audio_step * n * world / sec, sec / n * 1e3, n

    def copy_gbs(dst, src):
        ms = timed(torch, lambda: dst.copy_(src, non_blocking=True), 20)
        return src.numel() * src.element_size() / (ms * 1e-3) / 1e9

    host_f32 = [torch.empty(B_CLIPS, CLIP_LEN, dtype=torch.float32).pin_memory() for _ in range(2)]
    for h, d in zip(host_f32, inputs):
        h.copy_(d)
    dev_f32 = [torch.empty(B_CLIPS, CLIP_LEN, device=dev) for _ in range(2)]
    e2e_val, e2e_ms, e2e_n = e2e_measure(host_f32, dev_f32, lambda w: mod(w))
    # sanity: the last e2e result equals the device-resident result
    chk = float((host_out[(e2e_n - 1) & 1].to(dev) - plan.forward(dev_f32[(e2e_n - 1) & 1])[0]).abs().max())
    # pinned-memory bandwidth of this host (what bounds e2e): one direction at a time, 56 MB transfers, all ranks at once
    barrier()
    h2d_gbs = copy_gbs(dev_f32[0], host_f32[0])
    barrier()
    d2h_gbs = copy_gbs(host_f32[1], dev_f32[1])
    # both directions at once (what the e2e loop does: H2D of step i+1 while step i-1 drains), rank 0's share
    barrier()
    def duplex():
        with torch.cuda.stream(copy_s):
            dev_f32[0].copy_(host_f32[0], non_blocking=True)
        with torch.cuda.stream(out_s):
            host_f32[1].copy_(dev_f32[1], non_blocking=True)
    for _ in range(3):
        duplex()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for _ in range(20):
        duplex()
    torch.cuda.synchronize(dev)
    duplex_gbs = 20 * B_CLIPS * CLIP_LEN * 4 / (time.perf_counter() - t0) / 1e9
    del host_f32, dev_f32
    # the same loop fed with what a 16-bit WAV at the target rate holds: pinned int16 -> H2D -> decode_pcm16 (k_pcm16)
    # -> the same mel call -> D2H
    host_pcm = [torch.randint(-32768, 32767, (B_CLIPS, CLIP_LEN), dtype=torch.int16).pin_memory() for _ in range(2)]
    dev_pcm = [torch.empty(B_CLIPS, CLIP_LEN, dtype=torch.int16, device=dev) for _ in range(2)]
    pcm_fn = lambda p: mod(decode_pcm16(p[:, :, None]))
    pcm_val, pcm_ms, pcm_n = e2e_measure(host_pcm, dev_pcm, pcm_fn)
    chk_pcm = float((host_out[(pcm_n - 1) & 1].to(dev)
                     - plan.forward(dev_pcm[(pcm_n - 1) & 1].float() / 32768.0)[0]).abs().max())
    del host_pcm, dev_pcm

    # ---------------- config 4: the sharded sweep (every N) ----------------
    sweep_line = None if args.no_sweep else run_sweep(plan, dev, world, rank, dist, barrier)
    mod.mel_norms = norms
    plan = mod._plan_for(dev)                  # the sweep cleared the plan's norms: the module re-uploads its own

    if rank == 0:
        pk = peaks()
        # The burst figure costs 20 ms of matmul; the sustained one is a multi-second 1 kW load that leaves the SM clock
        # near 1.1 GHz for a while, so it runs AFTER the side measurements unless the headline's regime needs it first
        # (and then the clock is given time to come back).
        region_s = ms_total * 1e-3
        regime = "sustained" if region_s >= 2.0 else "burst"
        sustain_s = 4.0 if world == 1 else 1.5
        tf32 = measure_tf32_peak(dev, part="burst")
        clock_wait = None
        if regime == "sustained":
            tf32 = measure_tf32_peak(dev, sustain_s)
            clock_wait = wait_for_clocks(dev)
        tf32_peak = tf32[regime]
        k_ms = statistics.mean(kern_ms) if kern_ms else None
        achieved = FLOP_PER_FRAME * frames_step / (k_ms * 1e-3) / 1e12 if k_ms else None
        traffic, traffic_src = None, None
        tpath = os.path.join(ROOT, "profiles", "k_logmel_ncu_metrics.json")
        if os.path.exists(tpath):
            try:
                tj = json.load(open(tpath))
                traffic = tj.get("dram_bytes_per_launch")
                traffic_src = {"file": "profiles/k_logmel_ncu_metrics.json", "kernel": tj.get("kernel"),
                               "capture": tj.get("source"), "commit": tj.get("commit")}
            except Exception:
                traffic = None
        cpu_line, configs, ref_gpu, ingest_line, vq_line = None, None, None, None, None
        if world == 1 and not args.no_extras:
            configs, ref_gpu = run_configs(dev, tf32_peak, pk["hbm_gbs"])
            ingest_line = run_ingest(dev, mod, pk, audio_step)
            vq_line = run_vq(dev, tf32_peak)
        if regime == "burst":
            tf32 = measure_tf32_peak(dev, sustain_s)
            clock_wait = wait_for_clocks(dev)         # the CPU baseline below does not care; recorded for the reader
        if world == 1 and not args.no_cpu_baseline:
            xs = inputs[0].cpu()
            unbind(all_cpus)                    # the CPU leg gets every core again, not just the GPU's NUMA node
            torch.set_num_threads(os.cpu_count() or 1)
            val, sec, cores = cpu_reference_throughput(12, 1, xs)
            cpu_line = {"value": val, "unit": "audio-s/s", "cores": cores, "kind": "port",
                        "sample": "12 passes of the same 64x10 s batch through the reference's call sequence "
                                  "(oracle.callsite_logmel: torchaudio MelSpectrogram on CPU/MKL)"}
        line = {
            "metric": "log-mel audio-seconds/sec", "value": value, "unit": "audio-s/s", "n_gpus": world,
            "steps": K, "warmup": W, "repeats": R, "timed_region_s": region_s, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "tf32x3(f32 accumulate)", "data": "synthetic",
            "config": config_dict(world),
            "frames_per_s": value * SR / 256.0,
            "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": "audio-s/s", "h2d_bytes_per_step": B_CLIPS * CLIP_LEN * 4,
                    "d2h_bytes_per_step": B_CLIPS * T * 80 * 4, "ms_per_step": e2e_ms, "steps": e2e_n,
                    "how": "TorchMelSpectrogram.forward on pinned host batches, double-buffered H2D/D2H streams, wall clock, "
                           "max over ranks",
                    "numa": numa, "max_abs_diff_vs_resident": chk,
                    "host_limiter": {"pinned_h2d_GB/s": h2d_gbs, "pinned_d2h_GB/s": d2h_gbs,
                                     "duplex_GB/s_each_way": duplex_gbs,
                                     "copy_bound_audio_s_per_s_per_gpu": audio_step / (B_CLIPS * CLIP_LEN * 4 / (h2d_gbs * 1e9)),
                                     "note": "rank 0's rate while every rank copies at once, one direction at a time; the "
                                             "step's H2D copy alone bounds e2e"}},
            "e2e_pcm16": {"value": pcm_val, "unit": "audio-s/s", "h2d_bytes_per_step": B_CLIPS * CLIP_LEN * 2,
                          "d2h_bytes_per_step": B_CLIPS * T * 80 * 4, "ms_per_step": pcm_ms, "steps": pcm_n,
                          "how": "the e2e loop fed with pinned int16 mono PCM at 22050 Hz (what a WAV holds): H2D -> "
                                 "decode_pcm16 -> TorchMelSpectrogram.forward -> D2H",
                          "copy_bound_audio_s_per_s_per_gpu": audio_step / (B_CLIPS * CLIP_LEN * 2 / (h2d_gbs * 1e9)),
                          "copy_bound_duplex": audio_step / (B_CLIPS * CLIP_LEN * 2 / (duplex_gbs * 1e9)),
                          "max_abs_diff_vs_resident": chk_pcm},
            "gpu_launches": launches,
            "kernels_per_step": {"k_prep(+edge buffers)": 1, "k_logmel2": 1},
            "roofline": {"bound": "tensor", "kernel": "k_logmel2", "achieved": achieved, "peak": tf32_peak,
                         "unit": "TFLOP/s", "frac": (achieved / tf32_peak) if achieved else None, "traffic": traffic,
                         "traffic_src": traffic_src,
                         "kernel_ms": k_ms, "kernel_share_of_step": (k_ms / ms_step) if k_ms else None,
                         "tf32_peak_measured": tf32, "clock_recovery_after_sustained_matmul": clock_wait,
                         "peak_src": f"TF32 peak measured in this run (torch.matmul fp32, allow_tf32, 8192^3); the {regime} "
                                     f"figure, because the timed region lasted {region_s:.2f} s (>= 2 s would take the sustained one)",
                         "bf16_half_for_comparison": {"burst": pk["bf16_burst"] / 2, "sustained": pk["bf16_sustained"] / 2,
                                                      "src": pk["src"]},
                         "algorithmic_flop_per_frame": FLOP_PER_FRAME,
                         "executed_flop_per_frame": EXEC_FLOP_PER_FRAME,
                         "executed_frac": (EXEC_FLOP_PER_FRAME * frames_step / (k_ms * 1e-3) / 1e12 / tf32_peak) if k_ms else None,
                         "note": "achieved = ALGORITHMIC flops (SURVEY 8d) / kernel time; frac > 1 is possible because the "
                                 "double fold issues 0.55x the algorithmic flops; executed_frac is the tensor pipe's own load",
                         "hbm_frac": (BYTES_PER_FRAME * frames_step / (k_ms * 1e-3) / 1e9 / pk["hbm_gbs"]) if k_ms else None},
            "sweep": sweep_line,
            "configs": configs,
            "ref_gpu": ref_gpu,
            "cpu_baseline": cpu_line,
            "ingest": ingest_line,
            "vq": vq_line,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip configs / ref_gpu / ingest / vq (they run at N = 1 only)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
