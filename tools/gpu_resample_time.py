"""Median time of the resampler for a few ratios (fp32 and fused int16 input), 64 x 10 s, L2 flushed."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xtts_b200 import Resampler
flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda")
def med(fn, reps=10):
    """Per-call time of `reps` back-to-back calls behind a long flush kernel (the host runs ahead of the GPU, so
    launch overhead is hidden; working sets are > L2)."""
    ms = []
    for i in range(8):
        for _ in range(4):
            flush.fill_(0.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record(); torch.cuda.synchronize()
        if i >= 2: ms.append(a.elapsed_time(b) / reps)
    return np.median(ms) * 1e3
line = ""
RATES = [tuple(int(v) for v in a.split(':')) for a in sys.argv[1:]] or [(44100, 24000), (24000, 22050), (16000, 22050), (48000, 22050)]
for o, n in RATES:
    x = torch.rand((64, o * 10), device="cuda") * 2 - 1
    pcm = torch.randint(-32768, 32767, (64, o * 10, 2), dtype=torch.int16, device="cuda")
    rs = Resampler(o, n)
    line += f"  {rs.orig}:{rs.new} f32 {med(lambda: rs(x)):.0f} i16 {med(lambda: rs.from_pcm16(pcm)):.0f}"
print("us:" + line, flush=True)
