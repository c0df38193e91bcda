"""Time the codebook search (k_vq_tile + k_vq_tc + k_vq_finish) at BASELINE config 2's size behind the DVAE encoder:
64 clips x 10 s -> 64 x 216 rows of dim 512 against 8192 codes (ttts/vqvae/config.json:19,22), next to the
reference's own call sequence on the same GPU (cuBLAS sgemm + elementwise + max: xtts_dvae.py:88-90).

    python tools/gpu_vq_bench.py [--rows N] [--iters K] [--legs] [--out gpurun_out/vq_bench.json]

Inputs rotate over 4 batches (4 x 28 MB of rows + 33 MB of codebook tiles: the rows come from HBM, the codebook
stays in L2, as in a sweep).  CUDA events on the launching stream.  --legs adds the timing-only experiments of
TTTS_VQ_EXP (results invalid: pipeline legs removed).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import vq_oracle as vo  # noqa: E402  (inputs + the acceptance check; never the thing timed)
from xtts_b200.vq import Quantize  # noqa: E402


def timed(fn, iters, warmup=3):
    for _ in range(warmup):
        fn(0)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn(i)
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
    return {"median_ms": ts[len(ts) // 2], "best_ms": ts[0]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=64 * 216)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--legs", action="store_true")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "vq_bench.json"))
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    embed = vo.synth_codebook(vo.DVAE_DIM, vo.DVAE_N_EMBED, 41)
    q = Quantize(vo.DVAE_DIM, vo.DVAE_N_EMBED)
    q.embed.copy_(torch.from_numpy(embed))
    q = q.eval().to(dev)
    plan = q.plan(dev)
    zs_host = [vo.synth_tokens(a.rows, embed, 500 + i, "near") for i in range(4)]
    zs = [torch.from_numpy(z).to(dev) for z in zs_host]
    e_dev = q.embed

    res = {"rows": a.rows, "dim": vo.DVAE_DIM, "n_embed": vo.DVAE_N_EMBED, "gpu": torch.cuda.get_device_name(0)}
    flop = 2.0 * a.rows * vo.DVAE_DIM * vo.DVAE_N_EMBED
    res["algorithmic_gflop"] = flop / 1e9

    ind = plan.nearest(zs[0])[0]
    plan.status()
    st = vo.check_indices(ind.cpu().numpy()[:1024], zs_host[0][:1024], embed)
    res["parity_sample"] = st
    assert st["bad"] == 0, st

    t = timed(lambda i: plan.nearest(zs[i % 4]), a.iters)
    # clocks under this load: a 1.2 s loop of the same call, sampled by nvidia-smi (bench.py's sampler)
    from bench import ClockSampler
    cs = ClockSampler(0)
    cs.start()
    for i in range(2500):
        plan.nearest(zs[i % 4])
    torch.cuda.synchronize()
    res["clocks"] = cs.stop()
    t["tflops_algorithmic"] = flop / (t["median_ms"] * 1e-3) / 1e12
    t["rows_per_s"] = a.rows / (t["median_ms"] * 1e-3)
    res["nearest_indices_only"] = t
    t = timed(lambda i: q(zs[i % 4]), a.iters)
    res["quantize_forward"] = t

    def torch_ref(i):                                        # the reference's lines on the library kernels
        flatten = zs[i % 4]
        dist = flatten.pow(2).sum(1, keepdim=True) - 2 * flatten @ e_dev + e_dev.pow(2).sum(0, keepdim=True)
        return (-dist).max(1)[1]
    t = timed(torch_ref, a.iters)
    t["tflops_algorithmic"] = flop / (t["median_ms"] * 1e-3) / 1e12
    res["reference_sequence_on_gpu"] = t
    ref_ind = torch_ref(0)
    res["agreement_with_reference_on_gpu"] = float((ref_ind == plan.nearest(zs[0])[0]).float().mean())

    if a.legs:
        legs = {}
        for name, flags in (("full", 0), ("no_epilogue", 4), ("no_converters", 1), ("no_codebook_loads", 2),
                            ("mma_only", 1 | 2 | 4), ("converters_only", 2 | 4 | 32), ("codebook_loads_only", 1 | 4 | 32),
                            ("no_mma", 32), ("one_product", 8), ("one_product_mma_only", 8 | 1 | 2 | 4),
                            ("full_again", 0)):
            os.environ["TTTS_VQ_EXP"] = str(flags)
            legs[name] = timed(lambda i: plan.nearest(zs[i % 4]), 8)["median_ms"]
        os.environ.pop("TTTS_VQ_EXP")
        plan.status()
        res["legs_ms"] = legs

    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
