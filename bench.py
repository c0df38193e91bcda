#!/usr/bin/env python
"""bench.py -- log-mel audio-seconds/sec of the B200 front-end (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path over one batch of the headline workload (BASELINE.json
configs[1]): 64 synthetic 10 s clips at 22.05 kHz = 640 audio-seconds = 55,168 frames per GPU.
`value` is device-timed whole-job throughput with the waveforms resident in HBM; `e2e` is the same
metric through the public module (`TorchMelSpectrogram.forward`) from pinned HOST buffers with the
result read back to the host, copies inside the timed region.  `--impl reference` times the
reference's own CPU call sequence (torchaudio MelSpectrogram -> log(clamp) -> / mel_norms, the
oracle's `callsite_logmel`) on the host cores.
"""
from __future__ import annotations

import argparse
import ctypes
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
FLOP_PER_FRAME = 2.185e6          # algorithmic FLOP per frame (DFT-as-GEMM, unpruned, single pass)
BYTES_PER_FRAME = 1344.0          # 1024 B in + 320 B out
# tensor FLOP k_logmel2 actually issues per frame (V1): 2 parity passes x [33 cos + 32 sin k-steps] x 3 TF32 products
# x 2*192*8 -- the double fold halves K, bin pruning drops N from 2*513 to 2*384, the 3-term split triples the rest
EXEC_FLOP_PER_FRAME = 2 * (33 + 32) * 3 * 2 * 192 * 8
WORKLOAD = "64x10s_22050Hz_uniform[-1,1)_TorchMelSpectrogram_V1"
ROTATE = 4                        # distinct input batches: 4 x 56.4 MB = 226 MB > 126 MB L2


def frames_of(L: int) -> int:
    return 1 + L // 256


def config_dict(world: int) -> dict:
    return {"workload": WORKLOAD, "clips_per_step_per_gpu": B_CLIPS, "clip_seconds": 10,
            "frames_per_step_per_gpu": B_CLIPS * frames_of(CLIP_LEN),
            "sharding": f"dp{world}: one batch per rank, no data-path collective",
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
                                          "-lms", "100", "-i", str(self.gpu_index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        for r in self.rows:
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
    from xtts_b200 import TorchMelSpectrogram, _cabi
    lib = _cabi.lib()

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

    # ---------------- device-resident throughput (`value`) ----------------
    for i in range(W):
        plan.forward(inputs[i % ROTATE])
    plan.status()
    KP = min(K, 1024)                      # kernel-only CUDA-event pairs: first KP timed steps
    _cabi.check(lib.ttts_mel_profile_enable(KP))
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    e0.record()
    for i in range(K):
        plan.forward(inputs[i % ROTATE])
        launches += plan.last_launches
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    kbuf = (ctypes.c_float * KP)()
    nk = lib.ttts_mel_profile_collect(kbuf, KP)
    _cabi.check(lib.ttts_mel_profile_enable(0))
    kern_ms = [kbuf[i] for i in range(nk)]
    plan.status()
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = audio_step * K * world / (ms_total * 1e-3)

    # ---------------- end to end through the public module from pinned host memory ----------------
    host_in = [torch.empty(B_CLIPS, CLIP_LEN, dtype=torch.float32).pin_memory() for _ in range(2)]
    for h, d in zip(host_in, inputs):
        h.copy_(d)
    T = frames_of(CLIP_LEN)
    host_out = [torch.empty(B_CLIPS, T, 80, dtype=torch.float32).pin_memory() for _ in range(2)]
    dev_in = [torch.empty(B_CLIPS, CLIP_LEN, device=dev) for _ in range(2)]
    copy_s, out_s = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    main_s = torch.cuda.current_stream(dev)

    def e2e_loop(n, host_in=host_in, dev_in=dev_in, fn=lambda w: mod(w)):
        # double-buffered: H2D of step i+1 and D2H of step i-1 overlap the kernel of step i
        ready = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]
        drained = [torch.cuda.Event() for _ in range(2)]
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
                if i >= 2:
                    pass
                host_out[s].copy_(mel.transpose(1, 2), non_blocking=True)
                mel.record_stream(out_s)
                drained[s].record(out_s)
        out_s.synchronize()

    e2e_loop(W)
    barrier()
    t0 = time.perf_counter()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record(main_s)
    e2e_loop(K)
    torch.cuda.synchronize(dev)
    g1.record(main_s)
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    e2e_val = audio_step * K * world / e2e_s
    # sanity: e2e result equals the device-resident result
    chk = float((host_out[(K - 1) & 1].to(dev) - plan.forward(dev_in[(K - 1) & 1])[0]).abs().max())

    # the same loop fed with what a 16-bit WAV at the target rate holds (int16, half the H2D bytes), N=1 only:
    # pinned int16 -> H2D -> decode_pcm16 (k_pcm16) -> the same mel call -> D2H.  Reported under "ingest".
    e2e_pcm_val = None
    if world == 1:
        from xtts_b200 import decode_pcm16
        host_pcm = [torch.randint(-32768, 32767, (B_CLIPS, CLIP_LEN), dtype=torch.int16).pin_memory() for _ in range(2)]
        dev_pcm = [torch.empty(B_CLIPS, CLIP_LEN, dtype=torch.int16, device=dev) for _ in range(2)]
        pcm_fn = lambda p: mod(decode_pcm16(p[:, :, None]))
        Kp = max(50, K // 10)
        e2e_loop(W, host_pcm, dev_pcm, pcm_fn)
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        e2e_loop(Kp, host_pcm, dev_pcm, pcm_fn)
        torch.cuda.synchronize(dev)
        e2e_pcm_val = audio_step * Kp / (time.perf_counter() - t0)
        chk_pcm = float((host_out[(Kp - 1) & 1].to(dev)
                         - plan.forward(dev_pcm[(Kp - 1) & 1].float() / 32768.0)[0]).abs().max())
        del host_pcm, dev_pcm

    if rank == 0:
        pk = peaks()
        tf32_peak = pk["bf16_sustained"] / 2.0
        k_ms = statistics.mean(kern_ms) if kern_ms else None
        achieved = FLOP_PER_FRAME * frames_step / (k_ms * 1e-3) / 1e12 if k_ms else None
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "k_logmel_ncu_metrics.json")
        if os.path.exists(tpath):
            try:
                traffic = json.load(open(tpath)).get("dram_bytes_per_launch")
            except Exception:
                traffic = None
        cpu_line = None
        if world == 1 and not args.no_cpu_baseline:
            xs = inputs[0].cpu()
            unbind(all_cpus)                    # the CPU leg gets every core again, not just the GPU's NUMA node
            torch.set_num_threads(os.cpu_count() or 1)
            val, sec, cores = cpu_reference_throughput(12, 1, xs)
            cpu_line = {"value": val, "unit": "audio-s/s", "cores": cores, "kind": "port",
                        "sample": "12 passes of the same 64x10 s batch through the reference's call sequence "
                                  "(oracle.callsite_logmel: torchaudio MelSpectrogram on CPU/MKL)"}
        ingest_line = None
        if world == 1:
            # SURVEY 8f N3, reported beside the headline (not part of `value`): 16-bit stereo PCM at 44.1 kHz resident
            # in HBM -> load_audio_pcm16 (fused decode + 2:1 sinc decimator + clip) -> the same mel call
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
            ingest_line = {"workload": "64x10s 16-bit stereo PCM @ 44100 Hz -> 22050 Hz -> log-mel, resident in HBM",
                           "pipeline_audio_s_per_s": audio_step / (p_ms * 1e-3), "pipeline_ms": p_ms,
                           "resample_kernel": "k_decimate<2,28,int16>", "resample_ms": r_ms,
                           "resample_GB/s": r_bytes / (r_ms * 1e-3) / 1e9,
                           "resample_hbm_frac": r_bytes / (r_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
                           "l2": "256 MB flush write before every iteration",
                           "e2e_pcm16": {"value": e2e_pcm_val, "unit": "audio-s/s",
                                         "h2d_bytes_per_step": B_CLIPS * CLIP_LEN * 2,
                                         "d2h_bytes_per_step": B_CLIPS * T * 80 * 4,
                                         "how": "the e2e loop fed with pinned int16 mono PCM at 22050 Hz: H2D -> decode_pcm16 "
                                                "-> TorchMelSpectrogram.forward -> D2H",
                                         "max_abs_diff_vs_resident": chk_pcm}}
            del pcm, flush
        vq_line = None
        if world == 1:
            # SURVEY 8f N2, reported beside the headline (not part of `value`): the DVAE codebook search behind the mel
            # path at this batch's size -- 64 clips x 216 encoder rows of dim 512 against 8192 codes, rows resident
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
            vq_line = {"workload": "64x216 rows of dim 512 -> nearest of 8192 codes (Quantize.forward, xtts_dvae.py:87-90), resident in HBM",
                       "kernels": {"k_vq_tile": 1, "k_vq_tc": 1, "k_vq_finish": 1}, "ms": v_ms,
                       "rows_per_s": rows / (v_ms * 1e-3),
                       "algorithmic_tflops": v_flop / (v_ms * 1e-3) / 1e12,
                       "frac_of_tf32_peak": v_flop / (v_ms * 1e-3) / 1e12 / tf32_peak,
                       "executed_frac": 3.0 * v_flop / (v_ms * 1e-3) / 1e12 / tf32_peak,
                       "agreement_with_library_sequence_on_gpu": float(((-dist_ref).max(1)[1] == vind).float().mean()),
                       "l2": "inputs rotate over 4 batches (113 MB of rows); the 33 MB of codebook tiles stay in L2"}
            del zs, zl, emb, dist_ref
        line = {
            "metric": "log-mel audio-seconds/sec", "value": value, "unit": "audio-s/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "tf32x3(f32 accumulate)", "data": "synthetic",
            "config": config_dict(world),
            "frames_per_s": value * SR / 256.0,
            "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": "audio-s/s", "h2d_bytes_per_step": B_CLIPS * CLIP_LEN * 4,
                    "d2h_bytes_per_step": B_CLIPS * T * 80 * 4, "ms_per_step": e2e_s / K * 1e3,
                    "how": "TorchMelSpectrogram.forward on pinned host batches, double-buffered H2D/D2H streams, wall clock",
                    "numa": numa,
                    "max_abs_diff_vs_resident": chk},
            "gpu_launches": launches,
            "kernels_per_step": {"k_prep(+edge buffers)": 1, "k_logmel2": 1},
            "roofline": {"bound": "tensor", "kernel": "k_logmel2", "achieved": achieved, "peak": tf32_peak,
                         "unit": "TFLOP/s", "frac": (achieved / tf32_peak) if achieved else None, "traffic": traffic,
                         "kernel_ms": k_ms, "kernel_share_of_step": (k_ms / (ms_total / K)) if k_ms else None,
                         "peak_src": f"{pk['src']}: bf16_tflops_sustained/2 (TF32 issues at half the bf16 rate)",
                         "algorithmic_flop_per_frame": FLOP_PER_FRAME,
                         "executed_flop_per_frame": EXEC_FLOP_PER_FRAME,
                         "executed_frac": (EXEC_FLOP_PER_FRAME * frames_step / (k_ms * 1e-3) / 1e12 / tf32_peak) if k_ms else None,
                         "note": "achieved = ALGORITHMIC flops (SURVEY 8d) / kernel time; frac > 1 is possible because the "
                                 "double fold issues 0.55x the algorithmic flops; executed_frac is the tensor pipe's own load",
                         "hbm_frac": (BYTES_PER_FRAME * frames_step / (k_ms * 1e-3) / 1e9 / pk["hbm_gbs"]) if k_ms else None},
            "cpu_baseline": cpu_line,
            "ingest": ingest_line,
            "vq": vq_line,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
