"""One 64x10 s forward (after 2 warm-ups) for an ncu capture of the fused kernel:
ncu --set full --import-source on --clock-control none -k regex:k_logmel -s 2 -c 1 -o gpurun_out/x python tools/ncu_one.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from xtts_b200 import MelPlan
dev = torch.device("cuda:0")
plan = MelPlan(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0, f_max=8000,
               power=2.0, norm="slaney", device=dev)
x = torch.rand(64, 220500, device=dev) * 2 - 1
for _ in range(3):
    y = plan.forward(x)
torch.cuda.synchronize()
print("ok", float(y[0].abs().mean()) if isinstance(y, tuple) else float(y.abs().mean()))
