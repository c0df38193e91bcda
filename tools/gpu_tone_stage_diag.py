"""Where does the ~1e-4 log-mel error at a tone's peak come from?  Stage by stage against the fp64 restatement, at the
worst element of the committed tone golden (tests/golden/v1_signals.npz, x_tone)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import mel_oracle as orc
from xtts_b200 import MelPlan
dev = torch.device("cuda:0")
g = np.load(os.path.join(ROOT, "tests/golden/v1_signals.npz"))
x = g["x_tone"]
norms = g["norms"]
plan = MelPlan(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0, f_max=8000,
               power=2.0, norm="slaney", device=dev, lab=True)
plan.set_norms(torch.from_numpy(norms))
out, nf, _, reim = plan.forward(torch.from_numpy(x).to(dev), debug_reim=True)
plan.status()
T = int(nf[0]); k0, live, nbt = plan.first_bin, plan.live_bins, plan.num_bins
y = out.cpu().numpy().transpose(0, 2, 1).astype(np.float64)
truth = orc.restated_logmel(x, orc.V1, norms)
ref = g["y_tone"].astype(np.float64)
err = np.abs(y - truth)
b, m, t = np.unravel_index(err.argmax(), err.shape)
print("worst element", (b, m, t), "ours", y[b, m, t], "truth", truth[b, m, t], "ref", ref[b, m, t], "ours err", err[b, m, t], "ref err", abs(ref[b, m, t] - truth[b, m, t]))
st = orc.restated_stages(x[b], orc.V1)
re = reim[b * T:(b + 1) * T, :live].cpu().numpy().astype(np.float64)
im = reim[b * T:(b + 1) * T, nbt:nbt + live].cpu().numpy().astype(np.float64)
fb = orc.melscale_fbanks64(orc.V1)                                # [bins, mels]
bins = np.nonzero(fb[:, m])[0]
print("filter", m, "bins", bins[0], "..", bins[-1])
P_true = st.re[t, bins] ** 2 + st.im[t, bins] ** 2
P_ours = re[t, bins - k0] ** 2 + im[t, bins - k0] ** 2
print("power true", P_true)
print("rel err of our power per bin", (P_ours - P_true) / P_true)
print("rel err re", (re[t, bins - k0] - st.re[t, bins]) / np.abs(st.re[t, bins] + 1j * st.im[t, bins]))
print("rel err im", (-im[t, bins - k0] - st.im[t, bins]) / np.abs(st.re[t, bins] + 1j * st.im[t, bins]))
mel_true = float(fb[bins, m] @ P_true)
mel_ours64 = float(fb[bins, m] @ P_ours)
fb32 = torch.from_numpy(fb.astype(np.float32))
print("mel true", mel_true, "mel from our re/im in fp64", mel_ours64, "rel", (mel_ours64 - mel_true) / mel_true)
print("log true", np.log(mel_true) / norms[m], "log of ours64", np.log(mel_ours64) / norms[m], "kernel", y[b, m, t], "truth table", truth[b, m, t])
fbp = plan.fb.double().numpy()
print("filterbank (product fp32) vs fp64 restatement, rel diff at these bins", (fbp[bins, m] - fb[bins, m]) / fb[bins, m])
mel_fb32 = float(fbp[bins, m] @ P_true)
print("log with the fp32 filterbank and true power", np.log(mel_fb32) / norms[m], "diff to truth", np.log(mel_fb32) / norms[m] - truth[b, m, t])
