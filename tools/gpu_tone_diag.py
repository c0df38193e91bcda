"""High-dynamic-range inputs (clean tones): |ours - fp64| against the dual-oracle bound of SURVEY 8c (ii), per kernel."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from oracle import mel_oracle as orc
from xtts_b200 import MelPlan
dev = torch.device("cuda:0")
sr, L = 22050, 66150
t = np.arange(L, dtype=np.float64) / sr
rng = np.random.default_rng(5)
sig = {"tone440": 0.8 * np.sin(2 * np.pi * 440.0 * t),
       "tone+1e-3noise": 0.5 * np.sin(2 * np.pi * 1234.5 * t) + 1e-3 * rng.standard_normal(L),
       "two tones": 0.3 * np.sin(2 * np.pi * 3000.0 * t) + 0.3 * np.sin(2 * np.pi * 3100.0 * t)}
kw = dict(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0, f_max=8000, power=2.0, norm="slaney")
for name, s in sig.items():
    x = s.astype(np.float32)[None]
    ref32 = orc.callsite_logmel(x, orc.V1, None).numpy().astype(np.float64)
    truth = np.asarray(orc.restated_logmel(x, orc.V1, None), dtype=np.float64)
    bound = np.maximum(1e-4, 1.5 * np.abs(ref32 - truth))
    for algo in ("1", "2"):
        os.environ["TTTS_MEL_ALGO"] = algo
        plan = MelPlan(device=dev, **kw)
        for impl in ("tcgen05", "simt"):
            y = plan.forward(torch.from_numpy(x).to(dev), impl=impl)[0].transpose(1, 2).cpu().numpy().astype(np.float64)
            err = np.abs(y - truth)
            bad = err > bound
            i = np.unravel_index(np.argmax(err - bound), err.shape)
            print(f"{name:16s} algo {algo} {impl:7s}: max err {err.max():.2e} ref32 err max {np.abs(ref32-truth).max():.2e} "
                  f"violations {bad.sum()}/{bad.size} worst: err {err[i]:.2e} bound {bound[i]:.2e} truth {truth[i]:.3f} (mel {i[1]})")
