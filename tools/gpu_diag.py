"""Stage-level diagnosis on a GPU box: simt vs oracle, tcgen05 vs oracle, DFT dump vs fp64."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from oracle import mel_oracle as orc
from xtts_b200 import MelPlan, TorchMelSpectrogram

dev = torch.device("cuda:0")
print(torch.cuda.get_device_name(0), flush=True)
norms = orc.default_mel_norms(80)
plan = MelPlan(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0, f_max=8000,
               power=2.0, norm="slaney", device=dev, lab=True)
print("plan: first_bin", plan.first_bin, "num_bins", plan.num_bins, "live", plan.live_bins, flush=True)

def run(B, L, impl, seed=1):
    x = orc.synth("uniform", B, L, seed)
    xd = torch.from_numpy(x).to(dev)
    out, nf, _ = plan.forward(xd, impl=impl)
    torch.cuda.synchronize()
    plan.status()
    ref = orc.callsite_logmel(x, orc.V1, None).numpy().transpose(0, 2, 1)
    o = out.cpu().numpy()
    err = np.abs(o - ref)
    print(f"[{impl}] B={B} L={L} T={o.shape[1]} max_err={err.max():.3e} mean_err={err.mean():.3e} nan={np.isnan(o).sum()}", flush=True)
    if err.max() > 1e-4:
        bad = np.argwhere(err > 1e-4)
        print("   first bad (b,t,m):", bad[:5].tolist(), " n_bad", len(bad), "of", err.size)
        print("   ours", o[tuple(bad[0])], "ref", ref[tuple(bad[0])])
    return x, o, ref

for impl in ("simt", "tcgen05"):
    try:
        run(2, 4000, impl)
        run(3, 22050, impl)
    except Exception as e:
        print(f"[{impl}] FAILED: {e!r}", flush=True)
        if impl == "tcgen05":
            raise

# DFT stage dump vs fp64
x = orc.synth("uniform", 1, 40000, 5)
xd = torch.from_numpy(x).to(dev)
out, nf, _, reim = plan.forward(xd, debug_reim=True)
torch.cuda.synchronize(); plan.status()
st = orc.restated_stages(x[0], orc.V1)
T = st.re.shape[0]
nbt, k0, live = plan.num_bins, plan.first_bin, plan.live_bins
re = reim[:T, :live].cpu().numpy(); im = reim[:T, nbt:nbt + live].cpu().numpy()
scale = np.abs(st.re[:, k0:k0 + live]).max()
print("DFT re err (rel to max)", np.abs(re - st.re[:, k0:k0 + live]).max() / scale,
      " im err", np.abs(im - (-st.im[:, k0:k0 + live])).max() / scale, flush=True)

# timing of the headline config
x = torch.rand(64, 220500, device=dev, lab=True) * 2 - 1
for impl in ("tcgen05", "simt"):
    for _ in range(3):
        plan.forward(x, impl=impl)
    torch.cuda.synchronize()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    n = 10 if impl == "tcgen05" else 2
    t0.record()
    for _ in range(n):
        plan.forward(x, impl=impl)
    t1.record(); torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / n
    print(f"[{impl}] 64x10s: {ms:.3f} ms/call -> {640.0 / (ms * 1e-3):.3e} audio-s/s, {55168 / (ms * 1e-3):.3e} frames/s", flush=True)
plan.status()
a = plan.forward(x, impl="tcgen05")[0]; b = plan.forward(x, impl="simt")[0]
torch.cuda.synchronize()
print("tcgen05 vs simt on 64x10s: max abs diff", float((a - b).abs().max()), flush=True)
