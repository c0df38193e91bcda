"""Leg-removal timings of k_resample_tc (TTTS_RESAMPLE_EXP: 1 no converters, 2 no taps loads, 4 no epilogue, 32 no MMAs,
128 converters without global loads, 256 converters without TMEM stores).
Results are INVALID while a flag is set; this only attributes time."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xtts_b200 import Resampler
flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda")
for o, n in ((44100, 24000), (24000, 22050), (16000, 22050)):
    x = torch.rand((64, o * 10), device="cuda") * 2 - 1
    rs = Resampler(o, n)
    line = f"{rs.orig}:{rs.new}"
    for exp in (0, 38, 38 | 128, 38 | 256, 1, 2, 4, 32, 1 | 2, 1 | 4, 2 | 4, 1 | 2 | 4, 1 | 2 | 32, 2 | 4 | 32, 1 | 2 | 4 | 32):
        os.environ["TTTS_RESAMPLE_EXP"] = str(exp)
        ms = []
        for i in range(8):
            flush.fill_(0.0)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); rs(x); b.record(); torch.cuda.synchronize()
            if i >= 3: ms.append(a.elapsed_time(b))
        line += f"  exp{exp}={np.median(ms)*1e3:.0f}us"
    os.environ["TTTS_RESAMPLE_EXP"] = "0"
    print(line, flush=True)
