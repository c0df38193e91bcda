"""Golden vectors for transform sizes other than 1024 / 256, minted by running the REFERENCE ITSELF (build container only):

    python oracle/gen_golden_general.py        # writes tests/golden/v1_general_shapes.npz

``TorchMelSpectrogram`` (/root/reference/ttts/utils/utils.py:486-522) is imported from the read-only reference checkout
and constructed with other filter_length / hop_length / win_length / n_mel_channels / sampling_rate values than its
defaults (its constructor takes any, :487-502).  Inputs are regenerated from their seed at test time (2 x 1.5 s clips).
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

from oracle.mel_oracle import default_mel_norms, synth  # noqa: E402

# (filter_length, hop_length, win_length, n_mel_channels, mel_fmin, mel_fmax, sampling_rate, normalize)
CASES = {
    "n2048_h512": (2048, 512, 2048, 80, 0, 8000, 22050, False),
    "n512_h128": (512, 128, 512, 80, 0, 8000, 22050, False),
    "n1024_h300_w800": (1024, 300, 800, 80, 0, 8000, 22050, False),
    "n400_h160_16k": (400, 160, 400, 64, 20, 7600, 16000, False),
    "n2048_h240_w1200_norm": (2048, 240, 1200, 100, 0, 12000, 24000, True),
}


def main():
    sys.path.insert(0, REF)
    from ttts.utils.utils import TorchMelSpectrogram   # the reference
    import torchaudio
    torch.manual_seed(0)
    torch.set_num_threads(1)
    out = dict(torch=torch.__version__, torchaudio=torchaudio.__version__,
               reference="boostpapa/xtts ttts/utils/utils.py:486-522 with non-default constructor arguments")
    for i, (name, (n_fft, hop, win, n_mels, fmin, fmax, sr, normalize)) in enumerate(CASES.items()):
        ref = TorchMelSpectrogram(filter_length=n_fft, hop_length=hop, win_length=win, n_mel_channels=n_mels,
                                  mel_fmin=fmin, mel_fmax=fmax, sampling_rate=sr, normalize=normalize, mel_norm_file=None)
        norms = default_mel_norms(n_mels) if i % 2 == 0 else None
        if norms is not None:
            ref.mel_norms = torch.from_numpy(norms)
        kind, L, seed = "uniform" if i % 2 == 0 else "gauss", int(1.5 * sr) + 7 * i, 4000 + i
        x = synth(kind, 2, L, seed)                     # regenerated from (kind, L, seed) at test time
        out[f"y_{name}"] = ref(torch.from_numpy(x)).contiguous().numpy()
        out[f"cfg_{name}"] = np.array([n_fft, hop, win, n_mels, fmin, fmax, sr, int(normalize), int(norms is not None),
                                       int(kind == "gauss"), L, seed])
    path = os.path.join(ROOT, "tests", "golden", "v1_general_shapes.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
