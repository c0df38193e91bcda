"""Mint the ingestion golden vectors by running the REFERENCE's ``load_audio`` itself (build container only).

    python oracle/gen_golden_ingest.py        # writes tests/golden/ingest_load_audio.npz

Imports ``load_audio`` from /root/reference/ttts/utils/utils.py:32-47, writes seeded 16-bit PCM WAV files to a
temp dir and records what it returns.  ``torchaudio.load`` cannot run in this image (torchaudio 2.11 delegates
to TorchCodec, not installed), so that single call is served by ``ingest_oracle.torchaudio_load_pcm16`` (a
``wave``-module reader: int16 / 32768, float32 [C, L]); everything after it -- channel pick, the
``torchaudio.functional.resample`` call, ``clip_`` -- is the reference's code.  Inputs are regenerated from
their seeds at test time (``ingest_oracle.synth_pcm16``).

/root/reference does not exist on the GPU box; nothing at test/bench time reads it.
"""
from __future__ import annotations

import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"
sys.path.insert(0, ROOT)

from oracle import ingest_oracle as ing  # noqa: E402

# (name, source rate, target rate, frames, channels, seed, kind)
CASES = [
    ("44k1_stereo_to_22k05", 44100, 22050, 22050, 2, 11, "speechlike"),
    ("48k_mono_to_24k", 48000, 24000, 14401, 1, 12, "speechlike"),
    ("16k_mono_to_22k05", 16000, 22050, 4801, 1, 13, "speechlike"),
    ("24k_stereo_to_22k05_fullscale", 24000, 22050, 6000, 2, 14, "fullscale"),
    ("44k1_mono_to_24k", 44100, 24000, 8821, 1, 15, "uniform"),
    ("22k05_stereo_same_rate", 22050, 22050, 5000, 2, 16, "fullscale"),
    ("8k_three_channels_to_24k", 8000, 24000, 1603, 3, 17, "speechlike"),
    ("short_clip_5_frames", 44100, 22050, 5, 1, 18, "uniform"),
]


PROMPT_LENGTHS = (1000, 50000, 72000, 100000, 480000, 700000)
PROMPT_SEEDS = 5


def main():
    sys.path.insert(0, REF)
    import torchaudio
    torchaudio.load = ing.torchaudio_load_pcm16            # see the module docstring
    from ttts.utils.utils import get_prompt_slice, load_audio    # the reference
    import random
    torch.set_num_threads(1)
    out = {}
    names = []
    with tempfile.TemporaryDirectory() as tmp:
        for name, sr, target, frames, ch, seed, kind in CASES:
            pcm = ing.synth_pcm16(frames, ch, seed, kind)
            path = os.path.join(tmp, name + ".wav")
            ing.write_wav(path, pcm, sr)
            y = load_audio(path, target)
            assert y is not None and y.dim() == 2 and y.size(0) == 1
            out["y_" + name] = y.numpy()
            names.append(name)
            print(f"{name}: {sr} -> {target}, {frames} x {ch} -> {tuple(y.shape)}, max |y| {float(y.abs().max()):.4f}")
    # get_prompt_slice (utils.py:50-74): (L, is_eval, seed) -> (first sample, length) of the crop of arange(L)
    crops = []
    for L in PROMPT_LENGTHS:
        a = torch.arange(L, dtype=torch.float32)[None]
        for ev in (0, 1):
            for seed in range(PROMPT_SEEDS):
                random.seed(seed)
                c = get_prompt_slice(a, is_eval=bool(ev))
                crops.append([L, ev, seed, int(c[0, 0]) if c.shape[1] else 0, int(c.shape[1])])
    out["prompt_crops"] = np.array(crops, dtype=np.int64)
    dst = os.path.join(ROOT, "tests", "golden", "ingest_load_audio.npz")
    np.savez_compressed(dst, names=np.array(names), cases=np.array([list(map(str, c)) for c in CASES]),
                        torch=torch.__version__, torchaudio=torchaudio.__version__,
                        reference="boostpapa/xtts ttts/utils/utils.py:32-47", **out)
    print("wrote", dst, os.path.getsize(dst), "bytes")


if __name__ == "__main__":
    main()
