"""Timing of the general-shape log-mel path (GeneralMelPlan: reflect pad + tcgen05 contraction + ttts_stft_logmel) against
the reference's own call sequence on the same GPU (torchaudio: cuFFT + cuBLAS + elementwise), 64 x 10 s resident."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import mel_oracle as orc
from xtts_b200 import TorchMelSpectrogram

dev = torch.device("cuda:0")


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
for name, (n_fft, hop, win, n_mels, sr) in {"1024/256 (fused kernel)": (1024, 256, 1024, 80, 22050),
                                             "2048/512": (2048, 512, 2048, 80, 22050),
                                             "512/128": (512, 128, 512, 80, 22050),
                                             "1024/300/800": (1024, 300, 800, 80, 22050),
                                             "400/160 @16k": (400, 160, 400, 64, 16000)}.items():
    m = TorchMelSpectrogram(filter_length=n_fft, hop_length=hop, win_length=win, n_mel_channels=n_mels, mel_fmax=sr // 2 if sr == 16000 else 8000,
                            sampling_rate=sr, mel_norm_file=None)
    cfg = orc.MelConfig(sample_rate=sr, n_fft=n_fft, hop=hop, win_length=win, n_mels=n_mels, f_max=float(sr // 2 if sr == 16000 else 8000))
    x = torch.rand(64, 10 * sr, device=dev) * 2 - 1
    tr = orc._callsite_transform(cfg).to(dev)
    ref = lambda: torch.log(torch.clamp(tr(x), min=1e-5))
    ms_ours, ms_ref = timed(lambda: m(x), 50), timed(ref, 20)
    err = float((m(x) - ref()).abs().max())
    out[name] = {"ours_ms": round(ms_ours, 4), "torchaudio_gpu_ms": round(ms_ref, 4), "speedup": round(ms_ref / ms_ours, 2),
                 "max_abs_diff": err, "frames": int(64 * orc.frame_count(10 * sr, cfg))}
    tr.to("cpu")
    print(name, out[name], flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "general_shapes_bench.json"), "w"), indent=1)
