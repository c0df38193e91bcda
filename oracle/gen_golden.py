"""Mint golden vectors by running the REFERENCE ITSELF (build container only).

    python oracle/gen_golden.py            # writes tests/golden/*.npz

Imports ``TorchMelSpectrogram`` (/root/reference/ttts/utils/utils.py:486-522) and
``dvae_wav_to_mel`` (/root/reference/ttts/vqvae/xtts_dvae.py:28-49) from the read-only
reference checkout and records their outputs on seeded inputs.  The V2 module
(``MelSpectrogramFeatures``, ttts/vocoder/feature_extractors.py:73-99) cannot be imported
here (it needs ``encodec`` and ``librosa`` at module import), so its 10 lines are executed
from a source-extracted class body: see ``_load_v2_class``.

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

from oracle.mel_oracle import default_mel_norms, synth  # noqa: E402


def _load_v2_class():
    """Compile ONLY FeatureExtractor/MelSpectrogramFeatures/safe_log out of the reference's files
    (the module-level ``import encodec/librosa`` lines are what make a plain import fail)."""
    import torchaudio
    from torch import nn
    ns = {"torch": torch, "torchaudio": torchaudio, "nn": nn, "List": list}
    src_mod = open(os.path.join(REF, "ttts/vocoder/modules.py")).read()
    tree = ast.parse(src_mod)
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == "safe_log":
            exec(compile(ast.Module([node], []), "modules.py", "exec"), ns)
    src_fe = open(os.path.join(REF, "ttts/vocoder/feature_extractors.py")).read()
    tree = ast.parse(src_fe)
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name in ("FeatureExtractor", "MelSpectrogramFeatures"):
            exec(compile(ast.Module([node], []), "feature_extractors.py", "exec"), ns)
    return ns["MelSpectrogramFeatures"]


def main():
    sys.path.insert(0, REF)
    from ttts.utils.utils import TorchMelSpectrogram, make_pad_mask   # the reference
    from ttts.vqvae.xtts_dvae import dvae_wav_to_mel                  # its functional twin
    import torchaudio

    torch.manual_seed(0)
    torch.set_num_threads(1)
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    norms = default_mel_norms(80)
    meta = dict(torch=torch.__version__, torchaudio=torchaudio.__version__,
                reference="boostpapa/xtts ttts/utils/utils.py:486-522")

    ref = TorchMelSpectrogram(mel_norm_file=None)
    ref.mel_norms = torch.from_numpy(norms)
    ref_nonorm = TorchMelSpectrogram(mel_norm_file=None)

    # --- V1 headline config 1: one 10 s clip (input regenerated from the seed at test time)
    x = synth("uniform", 1, 220500, 1234)
    y = ref(torch.from_numpy(x)).contiguous().numpy()
    y0 = ref_nonorm(torch.from_numpy(x)).contiguous().numpy()
    np.savez_compressed(os.path.join(out_dir, "v1_c1_uniform_10s.npz"), kind="uniform", B=1, L=220500,
                        seed=1234, mel=y, mel_nonorm=y0, norms=norms, **meta)

    # --- V1 edge lengths (inputs stored: tiny)
    edge = {}
    for L in (513, 600, 767, 768, 769, 1023, 1024, 1025, 1279, 1280, 1281, 2047, 2048, 4099):
        x = synth("uniform", 2, L, 100 + L)
        edge[f"x_{L}"] = x
        edge[f"y_{L}"] = ref(torch.from_numpy(x)).contiguous().numpy()
    np.savez_compressed(os.path.join(out_dir, "v1_edge_lengths.npz"), norms=norms, **edge, **meta)

    # --- V1 other signal classes
    sig = {}
    for kind, B, L, seed in (("gauss", 3, 22050, 7), ("silence", 1, 2048, 0), ("tone", 2, 8192, 11)):
        x = synth(kind, B, L, seed)
        sig[f"x_{kind}"] = x
        sig[f"y_{kind}"] = ref(torch.from_numpy(x)).contiguous().numpy()
    # [B,1,L] input form (utils.py:510-511)
    x3 = synth("uniform", 2, 3000, 5)[:, None, :]
    sig["x_b1l"] = x3
    sig["y_b1l"] = ref(torch.from_numpy(x3)).contiguous().numpy()
    np.savez_compressed(os.path.join(out_dir, "v1_signals.npz"), norms=norms, **sig, **meta)

    # --- second importable oracle: dvae_wav_to_mel
    x = synth("uniform", 2, 30000, 21)
    y = dvae_wav_to_mel(torch.from_numpy(x), mel_norms=torch.from_numpy(norms)).contiguous().numpy()
    np.savez_compressed(os.path.join(out_dir, "v1_dvae_wav_to_mel.npz"), x=x, y=y, norms=norms, **meta)

    # --- make_pad_mask (utils.py:77-103)
    lens = torch.tensor([5, 3, 2, 9, 1], dtype=torch.int64)
    np.savez_compressed(os.path.join(out_dir, "pad_mask.npz"), lengths=lens.numpy(),
                        mask=make_pad_mask(lens).numpy(), mask12=make_pad_mask(lens, 12).numpy())

    # --- V2: MelSpectrogramFeatures (what save_mel_to_disk.py and the loaders call)
    V2 = _load_v2_class()
    v2 = {}
    m_center = V2()
    m_same = V2(padding="same")
    for L in (24000, 5000, 1024, 777):
        x = synth("uniform", 1, L, 300 + L)
        v2[f"x_{L}"] = x
        v2[f"center_{L}"] = m_center(torch.from_numpy(x)).contiguous().numpy()
        v2[f"same_{L}"] = m_same(torch.from_numpy(x)).contiguous().numpy()
    np.savez_compressed(os.path.join(out_dir, "v2_features.npz"), **v2, **meta)

    for f in sorted(os.listdir(out_dir)):
        print(f, os.path.getsize(os.path.join(out_dir, f)))


if __name__ == "__main__":
    main()
