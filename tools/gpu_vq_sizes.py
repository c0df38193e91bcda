"""Codebook search (dim 512, 8192 codes) at the row counts of the BASELINE configs behind the DVAE encoder (4 mel frames
per row): one 10 s clip, 8 clips, config 2 (64 x 10 s), config 5 (32 x 120 s), config 3's largest bucket (256 x 30 s).
    python tools/gpu_vq_sizes.py > gpurun_out/vq_sizes.json"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from xtts_b200.vq import VQPlan

dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(7)
emb = torch.randn(512, 8192, device=dev, generator=g)
plan = VQPlan(emb, dev)
out = {"dim": 512, "n_embed": 8192, "cases": []}
for name, clips, tp in (("1x10s", 1, 216), ("8x10s", 8, 216), ("64x10s (config 2)", 64, 216), ("32x120s (config 5)", 32, 2584),
                        ("256x30s (config 3, longest bucket)", 256, 646)):
    rows = clips * tp
    z = emb.T[torch.randint(0, 8192, (rows,), device=dev, generator=g)] + 0.5 * torch.randn(rows, 512, device=dev, generator=g)
    zc = z.view(clips, tp, 512).permute(0, 2, 1).contiguous()    # the encoder's layout [B, dim, T']
    for _ in range(3):
        plan.nearest(z)
    torch.cuda.synchronize()
    res = {}
    for label, fn in (("rows", lambda: plan.nearest(z)[0]), ("bct", lambda: plan.nearest_bct(zc)),
                      ("library_sequence", lambda: (-(z.pow(2).sum(1, keepdim=True) - 2 * z @ emb + emb.pow(2).sum(0, keepdim=True))).max(1)[1])):
        if label == "library_sequence" and rows > 100000:
            res[label + "_ms"] = None                              # the [rows, 8192] distance table alone is 5.4 GB
            continue
        ts = []
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); r = fn(); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        res[label + "_ms"] = sorted(ts)[len(ts) // 2]
    res["equal"] = bool(torch.equal(plan.nearest(z)[0], plan.nearest_bct(zc).reshape(-1)))
    res["tflops_algorithmic"] = 2.0 * rows * 512 * 8192 / (res["rows_ms"] * 1e-3) / 1e12
    out["cases"].append({"workload": name, "rows": rows, **res})
plan.status()
print(json.dumps(out))
