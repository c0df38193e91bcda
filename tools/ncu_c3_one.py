"""BASELINE config 3 as one ragged batch (256 clips, 1-30 s) for an ncu launch list: which kernel takes the time?
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv python tools/ncu_c3_one.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from xtts_b200 import MelPlan
dev = torch.device("cuda:0")
plan = MelPlan(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0, f_max=8000,
               power=2.0, norm="slaney", device=dev)
rng = np.random.default_rng(1234)
lens = sorted(int(round(22050 * u)) for u in rng.uniform(1.0, 30.0, size=256))
if len(sys.argv) > 1 and sys.argv[1] == "unsorted":
    rng.shuffle(lens)
L = torch.tensor(lens, device=dev)
width = max(lens) if "aligned" not in sys.argv else (max(lens) + 3) // 4 * 4      # "aligned": rows on a 16-byte pitch -> read in place
x = (torch.rand(256, width, device=dev) * 2 - 1)[:, :max(lens)]
for _ in range(3):
    y = plan.forward(x, lengths=L)
torch.cuda.synchronize()
ts = []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); y = plan.forward(x, lengths=L); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
print("pitch", x.stride(0), "forward ms", sorted(ts)[5], "audio-s", sum(lens) / 22050, "frames", sum(1 + l // 256 for l in lens))
