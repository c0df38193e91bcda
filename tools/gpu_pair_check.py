"""CTA-pair (cta_group::2) variant of k_logmel2 against the single-CTA variant: bit-level agreement and timing."""
import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from xtts_b200 import MelPlan, _cabi
dev = torch.device("cuda:0")
lib = _cabi.lab()          # TTTS_MEL_PAIR is read by the lab build only
kw = dict(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0, f_max=8000, power=2.0, norm="slaney")
plans = {}
for pair in ("0", "1"):
    os.environ["TTTS_MEL_PAIR"] = pair
    plans[pair] = MelPlan(device=dev, lab=True, **kw)
x = torch.rand(64, 220500, device=dev) * 2 - 1
xs = torch.rand(7, 50000, device=dev) * 2 - 1
lens = torch.tensor([50000, 513, 8192, 30001, 12345, 49999, 2600], device=dev)
for name, fn in (("dense 64x10s", lambda p: p.forward(x)[0]), ("ragged 7 clips", lambda p: p.forward(xs, lens)[0]),
                 ("one clip", lambda p: p.forward(xs[:1])[0])):
    a, b = fn(plans["0"]), fn(plans["1"])
    torch.cuda.synchronize()
    print(f"{name}: max |pair - single| = {float((a - b).abs().max()):.3e}  finite {bool(torch.isfinite(b).all())}", flush=True)
for pair in ("0", "1", "0", "1"):
    p = plans[pair]
    for _ in range(5):
        p.forward(x)
    torch.cuda.synchronize()
    p.profile_enable(30)
    for _ in range(30):
        p.forward(x)
    ms = sorted(p.profile_collect(30))
    n = len(ms)
    print(f"pair={pair}: k_logmel2 median {ms[n//2]*1e3:.1f} us  min {ms[0]*1e3:.1f} us", flush=True)
    p.profile_enable(0)
