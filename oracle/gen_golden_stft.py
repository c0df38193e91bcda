"""Mint the STFT-family goldens by running the REFERENCE ITSELF (build container only):

    python oracle/gen_golden_stft.py        # writes tests/golden/stft_family.npz

``TorchSTFT``, ``STFTLoss`` and ``MultiScaleSTFTLoss`` are compiled out of /root/reference/ttts/hifigan/losses.py:9-218 with
``ast`` (the module's top-level ``import librosa`` fails here; none of the three classes touches librosa unless
``use_mel=True``); ``ISTFT`` is imported from /root/reference/ttts/vocoder/spectral_ops.py:7-60 as it is.
/root/reference does not exist on the GPU box; nothing at test/bench time reads it.
"""
from __future__ import annotations

import ast
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"
sys.path.insert(0, ROOT)

from oracle.mel_oracle import synth          # noqa: E402
from oracle.stft_oracle import MS_SHAPES, synth_spec   # noqa: E402


def _load_loss_classes():
    from torch import nn
    from torch.nn import functional as F
    ns = {"torch": torch, "nn": nn, "F": F, "Dict": dict, "Union": object}
    tree = ast.parse(open(os.path.join(REF, "ttts/hifigan/losses.py")).read())
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name in ("TorchSTFT", "STFTLoss", "MultiScaleSTFTLoss"):
            exec(compile(ast.Module([node], []), "losses.py", "exec"), ns)
    return ns["TorchSTFT"], ns["STFTLoss"], ns["MultiScaleSTFTLoss"]


def main():
    torch.manual_seed(0)
    torch.set_num_threads(1)
    TorchSTFT, STFTLoss, MultiScaleSTFTLoss = _load_loss_classes()
    sys.path.insert(0, REF)
    from ttts.vocoder.spectral_ops import ISTFT
    out = {"torch": torch.__version__}
    # forward magnitudes at the three MultiScaleSTFTLoss shapes, two lengths (one not a multiple of any hop)
    for L in (4000, 2499):
        x = synth("gauss", 2, L, 600 + L)
        out[f"x_{L}"] = x
        for n_fft, hop, win in MS_SHAPES:
            with torch.no_grad():
                out[f"mag_{L}_{n_fft}"] = TorchSTFT(n_fft, hop, win)(torch.from_numpy(x)).numpy()
    x = synth("uniform", 2, 3000, 77)
    out["x_opts"] = x
    with torch.no_grad():
        out["mag_power2"] = TorchSTFT(1024, 120, 600, power=2)(torch.from_numpy(x)).numpy()
        out["mag_normalized_db"] = TorchSTFT(512, 50, 240, normalized=True, do_amp_to_db=True, spec_gain=2.0)(torch.from_numpy(x)).numpy()
        out["mag_1024_256"] = TorchSTFT(1024, 256, 1024)(torch.from_numpy(x[:, None, :])).numpy()     # [B, 1, T] form
    # the losses
    y, y_hat = synth("gauss", 3, 6000, 5), synth("gauss", 3, 6000, 6)
    y_hat = (0.7 * y + 0.3 * y_hat).astype(np.float32)
    out["loss_y"], out["loss_y_hat"] = y, y_hat
    with torch.no_grad():
        lm, lsc = STFTLoss(1024, 120, 600)(torch.from_numpy(y_hat), torch.from_numpy(y))
        out["stft_loss"] = np.array([float(lm), float(lsc)])
        lm, lsc = MultiScaleSTFTLoss()(torch.from_numpy(y_hat), torch.from_numpy(y))
        out["ms_loss"] = np.array([float(lm), float(lsc)])
    # inverse: the Vocos head's shape (n_fft = win = 1024, hop 256), both paddings, and a small one
    for name, (n_fft, hop, T) in {"vocos": (1024, 256, 24), "small": (512, 128, 9)}.items():
        spec = synth_spec(2, n_fft, T, 900 + T)
        out[f"spec_{name}"] = spec
        for padding in ("same", "center"):
            with torch.no_grad():
                out[f"istft_{name}_{padding}"] = ISTFT(n_fft, hop, n_fft, padding=padding)(torch.from_numpy(spec)).numpy()
    path = os.path.join(ROOT, "tests", "golden", "stft_family.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(out), "arrays")


if __name__ == "__main__":
    main()
