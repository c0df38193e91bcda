"""Throughput of every BASELINE.json config on one GPU (device-resident inputs, CUDA events)."""
import os, sys, json, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from xtts_b200 import MelPlan, TorchMelSpectrogram, MelSpectrogramFeatures, _cabi
from xtts_b200.sweep import MelSweep

dev = torch.device("cuda:0")
lib = _cabi.lib()


def timed(fn, iters, warm=3):
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


out = {}
v1 = TorchMelSpectrogram(mel_norm_file=None)
v1.mel_norms = torch.linspace(-6, -1, 80)
plan = v1._plan_for(dev)
g = torch.Generator(device=dev).manual_seed(1234)
for name, B, L, it in (("C1 1x10s", 1, 220500, 200), ("C2 64x10s", 64, 220500, 100), ("C5 32x120s", 32, 2646000, 20)):
    x = torch.rand(B, L, device=dev, generator=g) * 2 - 1
    ms = timed(lambda: v1(x), it)
    out[name] = dict(ms=ms, audio_s_per_s=B * L / 22050 / (ms * 1e-3), frames_per_s=B * (1 + L // 256) / (ms * 1e-3))
    del x

# C3: 256 ragged clips through the bucketed sweep (masks + n_frames + stats)
rng = np.random.default_rng(1234)
lens = [int(round(22050 * u)) for u in rng.uniform(1.0, 30.0, size=256)]
for bsec, mc in ((2.0, 64), (30.0, 64), (30.0, 256)):
    sweep = MelSweep(plan, lens, bucket_seconds=bsec, max_clips=mc)
    batches = {}
    for b in sweep.buckets:
        wav = torch.rand(len(b.indices), b.max_len, device=dev, generator=g) * 2 - 1
        batches[b.indices] = (wav, torch.tensor([lens[i] for i in b.indices], device=dev))
    fetch = lambda idx, max_len: batches[tuple(idx)]
    ms = timed(lambda: sweep.run_stats(fetch), 10)
    out[f"C3 256 ragged 1-30s, {bsec:g} s buckets of <= {mc} (masks+stats)"] = dict(ms=ms, audio_s_per_s=sum(lens) / 22050 / (ms * 1e-3),
        frames_per_s=sum(1 + L // 256 for L in lens) / (ms * 1e-3), buckets=len(sweep.buckets),
        padded_over_real=sum(b.max_len * len(b.indices) for b in sweep.buckets) / sum(lens))
    del batches
plan.set_norms(torch.linspace(-6, -1, 80))

# V2 (what save_mel_to_disk.py and the loaders call): 64 x 10 s at 24 kHz
v2 = MelSpectrogramFeatures()
x = torch.rand(64, 240000, device=dev, generator=g) * 2 - 1
ms = timed(lambda: v2(x), 100)
out["V2 64x10s 24kHz 100 mel"] = dict(ms=ms, audio_s_per_s=64 * 10 / (ms * 1e-3), frames_per_s=64 * (1 + 240000 // 256) / (ms * 1e-3))
# the reference's own GPU path (torchaudio: cuFFT + cuBLAS + elementwise kernels) on the same B200
from oracle import mel_oracle as orc
norms_d = torch.linspace(-6, -1, 80, device=dev)
for name, B, L, it in (("REF-GPU torchaudio C2 64x10s", 64, 220500, 30), ("REF-GPU torchaudio C5 32x120s", 32, 2646000, 5)):
    x = torch.rand(B, L, device=dev, generator=g) * 2 - 1
    tr = orc._callsite_transform(orc.V1).to(dev)
    fn = lambda: torch.log(torch.clamp(tr(x), min=1e-5)) / norms_d[None, :, None]
    ms = timed(fn, it)
    out[name] = dict(ms=ms, audio_s_per_s=B * L / 22050 / (ms * 1e-3), frames_per_s=B * (1 + L // 256) / (ms * 1e-3))
    del x
for k, v in out.items():
    print(f"{k:45s} {v['ms']:9.3f} ms  {v['audio_s_per_s']:.3e} audio-s/s  {v['frames_per_s']:.3e} frames/s")
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "configs.json"), "w"), indent=1)
