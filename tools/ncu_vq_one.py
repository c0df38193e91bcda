"""Three full-size codebook searches (64 x 216 rows, dim 512, 8192 codes) for ncu captures of k_vq_tc:
ncu --set full --import-source on --clock-control none -k regex:k_vq_tc -s 2 -c 1 -o gpurun_out/x python tools/ncu_vq_one.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from xtts_b200.vq import VQPlan
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(41)
embed = torch.randn(512, 8192, generator=g)
plan = VQPlan(embed, dev)
pick = torch.randint(0, 8192, (64 * 216,), generator=g)
z = (embed.T[pick] + 0.5 * torch.randn(64 * 216, 512, generator=g)).to(dev)
for _ in range(3):
    ind = plan.nearest(z)[0]
torch.cuda.synchronize()
plan.status()
print("ok", float((ind.cpu() == pick).float().mean()))
