"""Error distribution of the tcgen05 path on the headline batch (64 x 10 s): vs the fp32 reference call
sequence (oracle, CPU), vs the fp64 restatement on a sample, and vs the CUDA-core kernel."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from oracle import mel_oracle as orc
from xtts_b200 import MelPlan

dev = torch.device("cuda:0")
plan = MelPlan(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0, f_max=8000,
               power=2.0, norm="slaney", device=dev, lab=True)
res = []
for seed in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    for kind in ("uniform", "gauss"):
        x = orc.synth(kind, 64, 220500, 1000 + seed)
        xd = torch.from_numpy(x).to(dev)
        a = plan.forward(xd, impl="tcgen05")[0].cpu().numpy().transpose(0, 2, 1)
        s = plan.forward(xd, impl="simt")[0].cpu().numpy().transpose(0, 2, 1)
        plan.status()
        ref = orc.callsite_logmel(x, orc.V1, None).numpy()
        truth = orc.restated_logmel(x[:4], orc.V1, None)
        e = np.abs(a - ref); es = np.abs(s - ref)
        row = dict(kind=kind, seed=seed, n=int(e.size), max=float(e.max()), p999999=float(np.quantile(e, 0.999999)),
                   p9999=float(np.quantile(e, 0.9999)), mean=float(e.mean()), over_1e4=int((e > 1e-4).sum()),
                   simt_max=float(es.max()), simt_over_1e4=int((es > 1e-4).sum()),
                   ref_vs_fp64_max=float(np.abs(ref[:4] - truth).max()), ours_vs_fp64_max=float(np.abs(a[:4] - truth).max()),
                   simt_vs_fp64_max=float(np.abs(s[:4] - truth).max()))
        if row["over_1e4"]:
            idx = np.argwhere(e > 1e-4)[:3]
            row["examples"] = [dict(idx=i.tolist(), ours=float(a[tuple(i)]), ref=float(ref[tuple(i)])) for i in idx]
        res.append(row)
        print(json.dumps(row), flush=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "accuracy.json"), "w"), indent=1)
