"""Mint golden vectors for the codebook search by running the REFERENCE ITSELF (build container only).

    python oracle/gen_golden_vq.py            # writes tests/golden/vq_quantize.npz

Imports ``Quantize`` (/root/reference/ttts/vqvae/xtts_dvae.py:52-130) from the read-only reference checkout,
loads seeded codebooks into its ``embed`` buffer and records what its eval-mode ``forward`` and ``embed_code``
return.  Inputs are regenerated from the seeds by ``oracle.vq_oracle`` (numpy's legacy generator: version-stable),
so the full-size case (dim 512, 8192 codes: 16 MB of codebook) costs only its indices in the repository.

/root/reference does not exist on the GPU box; nothing at test/bench time reads it.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"
sys.path.insert(0, ROOT)

from oracle import vq_oracle as vo  # noqa: E402

# name: (dim, n_embed, codebook seed, duplicates, rows, token seed, kind, input shape)
CASES = {
    "small_near": (64, 512, 11, 0, 300, 12, "near", (3, 100, 64)),
    "small_randn": (64, 512, 11, 0, 257, 13, "randn", (257, 64)),
    "small_ties": (64, 512, 21, 8, 128, 22, "exact", (128, 64)),
    "odd_dims": (80, 1000, 31, 0, 131, 32, "near", (131, 80)),
    "dvae_near": (vo.DVAE_DIM, vo.DVAE_N_EMBED, 41, 0, 512, 42, "near", (2, 256, vo.DVAE_DIM)),
    "dvae_randn": (vo.DVAE_DIM, vo.DVAE_N_EMBED, 41, 0, 256, 43, "randn", (256, vo.DVAE_DIM)),
}


def main():
    sys.path.insert(0, REF)
    from ttts.vqvae.xtts_dvae import Quantize                  # the reference

    torch.set_num_threads(1)
    out = {"torch_version": np.array(torch.__version__)}
    for name, (dim, n_embed, cs, dup, rows, ts, kind, shape) in CASES.items():
        embed = vo.synth_codebook(dim, n_embed, cs, dup)
        z = vo.synth_tokens(rows, embed, ts, kind).reshape(shape)
        q = Quantize(dim, n_embed)
        q.embed.copy_(torch.from_numpy(embed))
        q.eval()
        with torch.no_grad():
            quantize, diff, ind = q(torch.from_numpy(z))
            code = q.embed_code(ind)
        assert ind.shape == tuple(shape[:-1]) and quantize.shape == tuple(shape)
        out[f"{name}.embed_ind"] = ind.numpy().astype(np.int32)
        out[f"{name}.diff"] = np.array(float(diff), dtype=np.float32)
        # the gathered rows are codebook columns and the straight-through value is input + (code - input): both
        # follow from the indices; keep a checksum so that a test can pin them without storing 1 MB per case
        out[f"{name}.quantize_sum"] = np.array(quantize.double().sum().item())
        out[f"{name}.code_sum"] = np.array(code.double().sum().item())
        if dim * rows <= 20000:
            out[f"{name}.quantize"] = quantize.numpy()
        st = vo.check_indices(ind.numpy(), z, embed)
        print(name, "reference vs exact:", st)
        assert st["bad"] == 0, "the acceptance rule must hold for the reference's own indices"
    path = os.path.join(ROOT, "tests", "golden", "vq_quantize.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
