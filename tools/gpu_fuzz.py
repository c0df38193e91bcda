"""Randomised parity sweep: ragged batches of random size and lengths through V1 / V2 ("center", "same") / a general
transform size against the reference's call per clip (oracle.callsite_logmel_ragged).  Frame counts and masks bit-exact,
values to the tests' tolerances.  usage: python tools/gpu_fuzz.py [cases] [seed]"""
import os, sys, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import mel_oracle as orc
from xtts_b200 import MelSpectrogramFeatures, TorchMelSpectrogram

dev = torch.device("cuda:0")
cases = int(sys.argv[1]) if len(sys.argv) > 1 else 120
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 2026)
norms = orc.default_mel_norms(80)
v1 = TorchMelSpectrogram(mel_norm_file=None); v1.mel_norms = torch.from_numpy(norms)
g1 = TorchMelSpectrogram(filter_length=512, hop_length=128, win_length=512, mel_norm_file=None)
mods = {
    "v1": (v1, orc.V1, norms, 513),
    "v2_center": (MelSpectrogramFeatures(), orc.V2, None, 513),
    "v2_same": (MelSpectrogramFeatures(padding="same"), orc.V2_SAME, None, 1024),
    "general_512_128": (g1, orc.MelConfig(n_fft=512, hop=128, win_length=512), None, 257),
}
worst = {k: 0.0 for k in mods}
bad = []
notes = []
for c in range(cases):
    name = list(mods)[c % len(mods)]
    mod, cfg, nr, lmin = mods[name]
    B = int(rng.integers(1, 24))
    kind = ["uniform", "gauss"][int(rng.integers(0, 2))]
    lens = [int(v) for v in np.where(rng.random(B) < 0.3, rng.integers(lmin, lmin + 2100, B), rng.integers(lmin, 70000, B))]
    Lmax = max(lens) + int(rng.integers(0, 3))           # row pitch not always a multiple of 4: both staging paths
    x = orc.synth(kind, B, Lmax, 10_000 + c)
    scale = float(10.0 ** rng.uniform(-3, 0))
    x = (x * scale).astype(np.float32)
    on_dev = bool(rng.integers(0, 2))
    lt = torch.tensor(lens, device=dev) if on_dev else torch.tensor(lens)
    y, nf, mask = mod(torch.from_numpy(x).to(dev), lengths=lt)
    ry, rn, rmask = orc.callsite_logmel_ragged(x, lens, cfg, nr)
    T = ry.shape[2]
    y = y[:, :, :T].cpu().numpy(); mask = mask[:, :T].cpu().numpy()
    ok_shape = nf.cpu().tolist() == rn.tolist() and np.array_equal(mask, rmask)
    d = np.abs(y - ry.numpy())
    if cfg.power == 1.0:                                   # V2's stated bar: 1e-4, or 2e-5 absolute on the mel magnitude
        bar = np.maximum(1e-4, 2e-5 / np.exp(ry.numpy()))
        over = int((d > bar).sum())
    elif name.startswith("general"):                       # one-bin filters (n_fft 512 with 80 mels): same kind of bar on the power,
        ref_mel = np.exp(ry.numpy() * (nr[None, :, None] if nr is not None else 1.0))   # |error of the mel power| <= 5e-8 for |x| <= 1
        over = int((d > np.maximum(1e-4, 5e-8 / ref_mel)).sum())
    else:
        over = int((d > 1e-4).sum())
    worst[name] = max(worst[name], float(d.max()))
    dual = None
    if over and cfg.power == 2.0:
        # where the fp32 reference itself is off the fp64 truth (a one-bin filter whose bin happens to be tiny:
        # cancellation), the bar is max(1e-4, 1.5 |ref - truth|) against the fp64 restatement
        b = int(np.unravel_index(int(d.argmax()), d.shape)[0])
        nfb = int(rn[b])
        truth = orc.restated_logmel(x[b:b + 1, :lens[b]], cfg, nr)[0][:, :nfb]
        refb, oursb = ry.numpy()[b][:, :nfb].astype(np.float64), y[b][:, :nfb].astype(np.float64)
        bound = np.maximum(1e-4, 1.5 * np.abs(refb - truth))
        dual = float((np.abs(oursb - truth) / bound).max())
        if dual <= 1.0:
            over = 0
            notes.append({"case": c, "name": name, "max_abs_diff_vs_ref": float(d.max()), "dual_oracle_ratio": dual})
    if not ok_shape or over:
        i = np.unravel_index(int(d.argmax()), d.shape)
        rv = float(ry.numpy()[i]) * (float(nr[i[1]]) if nr is not None else 1.0)          # un-normalised log-mel there
        bad.append({"case": c, "name": name, "B": B, "lens": lens[:6], "shape_ok": ok_shape, "over": over, "max": float(d.max()),
                    "at": [int(v) for v in i], "ref_logmel_unnormalised": rv, "ref_mel": float(np.exp(rv)), "scale": scale})
print(json.dumps({"cases": cases, "worst_abs_diff": worst, "passed_by_the_dual_oracle_rule": notes, "failures": bad}, indent=1))
sys.exit(1 if bad else 0)
