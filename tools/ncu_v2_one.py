"""One 64 x 10 s forward of the V2 extractor (MelSpectrogramFeatures: 24 kHz, 100 mel, 512 live bins, power 1) for ncu:
ncu --set full --import-source on --clock-control none -k regex:k_logmel2 -s 2 -c 1 -o gpurun_out/prof_v2 python tools/ncu_v2_one.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from xtts_b200 import MelSpectrogramFeatures
v2 = MelSpectrogramFeatures()
x = torch.rand(64, 240000, device="cuda") * 2 - 1
for _ in range(3):
    y = v2(x)
torch.cuda.synchronize()
print("ok", tuple(y.shape), float(y.mean()))
