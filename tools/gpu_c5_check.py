"""Why does config 5 (32 x 120 s) take 0.68 ms alone and 1.1 ms inside bench.py?  Times it fresh, after a sustained
matmul (the TF32 peak measurement that precedes run_configs), after a pause, and with a big allocation behind it."""
import statistics
import sys
import time

import torch

sys.path.insert(0, ".")
import pynvml
from oracle import mel_oracle as orc
from xtts_b200 import TorchMelSpectrogram

pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
dev = torch.device("cuda:0")


def clk():
    return pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0


def timed(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    mid = clk()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters, mid


v1 = TorchMelSpectrogram(mel_norm_file=None)
v1.mel_norms = torch.from_numpy(orc.default_mel_norms(80))
plan = v1._plan_for(dev)
g = torch.Generator(device=dev).manual_seed(99)
x = torch.rand(32, 2646000, device=dev, generator=g) * 2 - 1
x1 = torch.rand(1, 220500, device=dev, generator=g) * 2 - 1


def kernel_alone(fn, n):
    plan.profile_enable(n)
    for _ in range(n):
        fn()
    ms = plan.profile_collect(n)
    plan.profile_enable(0)
    return statistics.median(ms)


def report(tag):
    ms, c = timed(lambda: v1(x), 20)
    k = kernel_alone(lambda: v1(x), 10)
    ms1, _ = timed(lambda: v1(x1), 200)
    k1 = kernel_alone(lambda: v1(x1), 50)
    print(f"{tag:40s} C5 ms {ms:.3f} kernel {k:.3f}  clocks/power mid-loop {c} | C1 ms {ms1:.4f} kernel {k1:.4f}", flush=True)


report("fresh")
report("again")
time.sleep(1.0)
report("after 1 s idle")
a = torch.randn(8192, 8192, device=dev)
torch.backends.cuda.matmul.allow_tf32 = True
t0 = time.time()
while time.time() - t0 < 4.0:
    for _ in range(20):
        a @ a
    torch.cuda.synchronize()
print("burn done", clk())
report("right after a 4 s matmul burn")
report("again")
time.sleep(3.0)
report("after 3 s idle")
junk = torch.empty(60 << 30, dtype=torch.uint8, device=dev)
junk.fill_(1)
report("with 60 GB allocated")
del junk
torch.cuda.empty_cache()
report("after empty_cache")
out = torch.empty(32, plan.frames(2646000), 80, device=dev)
for it in (20, 200):
    ms, c = timed(lambda: v1(x), it)
    print(f"iters {it}: {ms:.3f}", c)
