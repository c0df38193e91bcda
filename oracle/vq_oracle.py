"""CPU oracle for the DVAE codebook search behind the log-mel path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s / ``tools/``' CPU legs may import this module; the
product package (``xtts_b200``) never does.

Parity status: PINNED by ``oracle/gen_golden_vq.py``, which imports the reference's own ``Quantize``
(/root/reference/ttts/vqvae/xtts_dvae.py:52-130) in the build container, runs its eval-mode forward and
``embed_code`` on the seeded inputs below and commits the outputs under ``tests/golden/vq_*.npz``.

Two layers, like ``mel_oracle``:

* ``callsite_quantize`` -- the reference's call sequence (xtts_dvae.py:87-93,119-120) with torch fp32 on the CPU.
* ``restated_distances`` -- |z - e_j|^2 in numpy fp64: the "truth" leg.  The fp32 distances of the reference are
  quantised to ulp(|z|^2 + |e|^2), so near-ties flip with the summation order of the sgemm behind ``@``; an index is
  therefore accepted when its exact distance is within ``tolerance()`` of the exact minimum, and must BE the exact
  arg-min (or an exact copy of that code) wherever the runner-up is further away than that (``check_indices``);
  ``differs_from_exact_argmin`` counts rows that do not return the LOWEST index of the exact minimum -- 0 for the CUDA
  path (bit-identical scores for identical codes), while the reference's own sgemm rounds two copies of a code
  differently now and then (MKL, 48 x 200: the higher copy for 4 rows in 300).  The reference's own indices are
  held to the same rule in tests/test_vq.py, so the rule is no looser than the reference's own rounding.
"""
from __future__ import annotations

from typing import Tuple

import numpy as np

DVAE_DIM, DVAE_N_EMBED = 512, 8192          # ttts/vqvae/config.json:19,22


def synth_codebook(dim: int, n_embed: int, seed: int, duplicates: int = 0) -> np.ndarray:
    """[dim, n_embed] float32 like the module's ``torch.randn(dim, n_embed)`` buffer (xtts_dvae.py:67); numpy's
    legacy generator, so the bits do not depend on the torch version.  ``duplicates`` copies that many codes to
    higher indices (exact ties: the lowest index must win)."""
    rs = np.random.RandomState(seed)
    e = rs.standard_normal((dim, n_embed)).astype(np.float32)
    for i in range(duplicates):
        src = (7 * i + 3) % (n_embed // 2)
        e[:, n_embed - 1 - i] = e[:, src]
    return e


def synth_tokens(rows: int, embed: np.ndarray, seed: int, kind: str = "near") -> np.ndarray:
    """[rows, dim] float32 encoder outputs.
    near: a code plus noise of half its norm (what a trained encoder emits: one clear winner, a crowd behind it);
    randn: independent of the codebook (all 8192 distances within a few percent: the hardest case for rounding);
    exact: rows equal to codes (distance 0; with duplicated codes an exact tie);
    tiny / huge: randn scaled by 1e-3 / 1e3."""
    rs = np.random.RandomState(seed)
    dim, n_embed = embed.shape
    if kind == "near":
        pick = rs.randint(0, n_embed, size=rows)
        return (embed[:, pick].T + 0.5 * rs.standard_normal((rows, dim))).astype(np.float32)
    if kind == "exact":
        pick = rs.randint(0, n_embed, size=rows)
        return np.ascontiguousarray(embed[:, pick].T)
    z = rs.standard_normal((rows, dim)).astype(np.float32)
    if kind == "tiny":
        z *= np.float32(1e-3)
    elif kind == "huge":
        z *= np.float32(1e3)
    elif kind != "randn":
        raise ValueError(kind)
    return z


def callsite_quantize(z, embed):
    """xtts_dvae.py:87-93,119-120 verbatim in torch fp32 on the CPU.  Returns (quantize, diff, embed_ind)."""
    import torch
    import torch.nn.functional as F
    input = torch.as_tensor(z)
    embed = torch.as_tensor(embed)
    dim = embed.shape[0]
    flatten = input.reshape(-1, dim)
    dist = flatten.pow(2).sum(1, keepdim=True) - 2 * flatten @ embed + embed.pow(2).sum(0, keepdim=True)
    soft_codes = -dist
    _, embed_ind = soft_codes.max(1)
    embed_ind = embed_ind.view(*input.shape[:-1])
    quantize = F.embedding(embed_ind, embed.transpose(0, 1))
    diff = (quantize.detach() - input).pow(2).mean()
    quantize = input + (quantize - input).detach()
    return quantize, diff, embed_ind


def restated_distances(z: np.ndarray, embed: np.ndarray) -> np.ndarray:
    """|z_i - e_j|^2 in fp64, [rows, n_embed] (blocked: the full-size table of 14 k rows would be 900 MB)."""
    z64 = z.reshape(-1, embed.shape[0]).astype(np.float64)
    e64 = embed.astype(np.float64)
    return (z64 * z64).sum(1, keepdims=True) - 2.0 * (z64 @ e64) + (e64 * e64).sum(0, keepdims=True)


def tolerance(z: np.ndarray, embed: np.ndarray, ulps: float = 8.0) -> np.ndarray:
    """Per-row slack on the exact distance: ``ulps`` fp32 ulps of |z|^2 + max|e|^2, the magnitude at which the
    reference's fp32 ``dist`` is rounded (three fp32 additions and an sgemm: a few ulps)."""
    z64 = z.reshape(-1, embed.shape[0]).astype(np.float64)
    mag = (z64 * z64).sum(1) + (embed.astype(np.float64) ** 2).sum(0).max()
    return ulps * mag * 2.0 ** -23


def check_indices(ind: np.ndarray, z: np.ndarray, embed: np.ndarray, ulps: float = 8.0, block: int = 2048) -> dict:
    """The acceptance rule above.  Returns counts; ``bad`` must be 0."""
    ind = np.asarray(ind).reshape(-1).astype(np.int64)
    zf = z.reshape(-1, embed.shape[0])
    n_embed = embed.shape[1]
    bad = flipped = decided = 0
    worst = 0.0
    for r0 in range(0, zf.shape[0], block):
        zb = zf[r0:r0 + block]
        d = restated_distances(zb, embed)
        tol = tolerance(zb, embed, ulps)
        ib = ind[r0:r0 + block]
        if ((ib < 0) | (ib >= n_embed)).any():
            bad += int(((ib < 0) | (ib >= n_embed)).sum())
            ib = np.clip(ib, 0, n_embed - 1)
        # the lowest index among the exact minima: duplicated codes tie exactly, but the fp64 table itself carries
        # ~1e-13 relative rounding that differs between columns, hence the 1e-4-of-tol band
        dmin0 = d.min(1)
        best = (d <= (dmin0 + 1e-4 * tol)[:, None]).argmax(1)
        dmin = d[np.arange(len(zb)), best]
        mine = d[np.arange(len(zb)), ib]
        d2 = np.where(d <= (dmin0 + 1e-4 * tol)[:, None], np.inf, d)
        gap = d2.min(1) - dmin                               # distance of the runner-up (exact copies of the winner aside)
        excess = mine - dmin
        worst = max(worst, float((excess / np.maximum(tol, 1e-300)).max()))
        bad += int((excess > tol).sum())
        clear = gap > tol
        decided += int(clear.sum())
        bad += int((clear & (mine > dmin0 + 1e-4 * tol)).sum())       # a clear winner (or an exact copy of it) is mandatory
        flipped += int((ib != best).sum())
    return {"rows": int(zf.shape[0]), "bad": bad, "differs_from_exact_argmin": flipped, "clear_rows": decided,
            "worst_excess_in_tol": worst}
