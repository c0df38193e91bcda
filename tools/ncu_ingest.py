"""One launch of each ingestion kernel for an ncu capture:  ncu --set full -k regex:'k_resample|k_decimate' python tools/ncu_ingest.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xtts_b200 import Resampler  # noqa: E402

o, n = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (16000, 22050)
x = torch.rand((64, o * 10), device="cuda") * 2 - 1
rs = Resampler(o, n)
for _ in range(3):
    y = rs(x)
torch.cuda.synchronize()
print(y.shape)
