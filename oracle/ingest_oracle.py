"""CPU oracle for the ingestion step in front of the log-mel path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this module; the product
package (``xtts_b200``) never does.

Parity status: PINNED by ``oracle/gen_golden_ingest.py``, which imports the reference's own ``load_audio``
(/root/reference/ttts/utils/utils.py:32-47) in the build container, feeds it 16-bit PCM WAV files written from
seeded samples and commits its outputs under ``tests/golden/ingest_*.npz`` (``torchaudio.load`` needs TorchCodec
in torchaudio 2.11, absent from this image, so that one call is served by a ``wave``-module reader that returns
what torchaudio documents for 16-bit PCM with ``normalize=True``: ``int16 / 32768`` as float32 ``[C, L]``).

Where the arithmetic lives: an un-vendored, un-pinned third party -- ``torchaudio.functional.resample``
(here torchaudio 2.11.0+cu128, ``functional/functional.py`` ``_get_sinc_resample_kernel`` /
``_apply_sinc_resample_kernel``).  Two layers restate it, like ``mel_oracle``:

* ``callsite_load_audio`` / ``callsite_mel_extract`` -- the reference's own call sequence on the pinned third
  party (fp32; what parity is quoted against, and the timed CPU leg).
* ``restated_resample`` -- the published algorithm written out with numpy in fp64 (windowed-sinc taps,
  zero padding, strided correlation, truncation).  Independent of torchaudio: the "truth" leg.
"""
from __future__ import annotations

import math
import wave
from typing import Optional, Sequence

import numpy as np


def synth_pcm16(frames: int, channels: int, seed: int, kind: str = "speechlike") -> np.ndarray:
    """Seeded int16 PCM ``[frames, channels]`` (WAV sample order).  Channel 1.. differ from channel 0 so that a
    wrong channel pick shows."""
    rng = np.random.default_rng(seed)
    n = np.arange(frames, dtype=np.float64)
    out = np.empty((frames, channels), dtype=np.int16)
    for c in range(channels):
        if kind == "uniform":
            x = rng.uniform(-1.0, 1.0, frames)
        elif kind == "fullscale":          # hits -32768 / 32767: resampling overshoots, so clip_ matters
            x = np.sign(np.sin(2 * np.pi * n * (0.013 + 0.002 * c))) * 1.0
        else:                              # a few partials with a slow envelope + a little noise
            f0 = 0.011 + 0.003 * c
            x = sum(a * np.sin(2 * np.pi * n * f0 * h + rng.uniform(0, 6.28)) for h, a in
                    ((1, 0.5), (2, 0.25), (3.01, 0.12), (7.3, 0.06)))
            x = x * (0.6 + 0.4 * np.sin(2 * np.pi * n / 9000.0)) + rng.normal(0, 0.01, frames)
        out[:, c] = np.clip(np.round(x * 32767.0), -32768, 32767).astype(np.int16)
    return out


def write_wav(path: str, pcm: np.ndarray, sr: int) -> None:
    """16-bit PCM WAV from ``[frames, channels]`` int16."""
    pcm = np.ascontiguousarray(pcm.reshape(pcm.shape[0], -1), dtype="<i2")
    with wave.open(path, "wb") as w:
        w.setnchannels(pcm.shape[1])
        w.setsampwidth(2)
        w.setframerate(int(sr))
        w.writeframes(pcm.tobytes())


def read_wav(path: str):
    """(int16 ``[frames, channels]``, sample rate) of a 16-bit PCM WAV."""
    with wave.open(path, "rb") as w:
        assert w.getsampwidth() == 2, "16-bit PCM only"
        c, sr, n = w.getnchannels(), w.getframerate(), w.getnframes()
        pcm = np.frombuffer(w.readframes(n), dtype="<i2").reshape(n, c)
    return pcm, sr


def torchaudio_load_pcm16(path: str):
    """Stand-in for ``torchaudio.load(path)`` on a 16-bit PCM WAV (``normalize=True``, ``channels_first=True``):
    float32 ``[C, L]`` = int16 / 32768, and the sample rate."""
    import torch
    pcm, sr = read_wav(path)
    return torch.from_numpy(pcm.T.astype(np.float32) / np.float32(32768.0)), sr


def pcm16_to_float(pcm: np.ndarray) -> np.ndarray:
    """``[frames, channels]`` int16 -> channel 0 float32 ``[1, frames]`` (utils.py:33-37)."""
    pcm = pcm.reshape(pcm.shape[0], -1)
    return (pcm[:, 0].astype(np.float32) / np.float32(32768.0))[None, :]


def callsite_load_audio(pcm: np.ndarray, sr: int, sampling_rate: int) -> np.ndarray:
    """``load_audio`` (utils.py:32-47) from the decoded samples on: channel 0, functional.resample, clip_."""
    import torch
    import torchaudio
    audio = torch.from_numpy(pcm.reshape(pcm.shape[0], -1).T.astype(np.float32) / np.float32(32768.0))
    if audio.size(0) > 1:
        audio = audio[0].unsqueeze(0)
    if sr != sampling_rate:
        audio = torchaudio.functional.resample(audio, sr, sampling_rate)
    audio.clip_(-1, 1)
    return audio.numpy()


def callsite_mel_extract_wave(pcm: np.ndarray, sr: int, target: int = 24000) -> np.ndarray:
    """The ingestion half of ``process_mel`` (mel_extract.py:15-21): channel 0, transforms.Resample (taps
    evaluated in fp64), no clip."""
    import torch
    import torchaudio
    wav = torch.from_numpy(pcm.reshape(pcm.shape[0], -1).T.astype(np.float32) / np.float32(32768.0))
    if wav.size(0) > 1:
        wav = wav[0].unsqueeze(0)
    if sr != target:
        wav = torchaudio.transforms.Resample(sr, target)(wav)
    return wav.numpy()


def restated_taps(orig_freq: int, new_freq: int, lowpass_filter_width: int = 6, rolloff: float = 0.99):
    """fp64 windowed-sinc polyphase taps ``[new, 2*width + orig]``: tap k of phase p sits at
    ``tau = ((k - width) / orig - p / new) * cutoff`` with ``cutoff = min(orig, new) * rolloff``, clamped to
    +-lpw; value ``sinc(pi tau) * cos^2(pi tau / (2 lpw)) * cutoff / orig``."""
    g = math.gcd(int(orig_freq), int(new_freq))
    orig, new = int(orig_freq) // g, int(new_freq) // g
    cutoff = min(orig, new) * rolloff
    width = math.ceil(lowpass_filter_width * orig / cutoff)
    k = np.arange(-width, width + orig, dtype=np.float64)[None, :] / orig
    p = -np.arange(new, dtype=np.float64)[:, None] / new
    tau = np.clip((p + k) * cutoff, -lowpass_filter_width, lowpass_filter_width)
    hann = np.cos(tau * np.pi / lowpass_filter_width / 2.0) ** 2
    sinc = np.sinc(tau)                       # sin(pi tau) / (pi tau), 1 at 0
    return sinc * hann * (cutoff / orig), orig, new, width


def restated_resample(x: np.ndarray, orig_freq: int, new_freq: int, lowpass_filter_width: int = 6,
                      rolloff: float = 0.99) -> np.ndarray:
    """fp64: ``y[i*new + p] = sum_k xpad[i*orig + k] * taps[p, k]``, xpad = x with ``width`` zeros in front and
    ``width + orig`` behind, truncated to ``ceil(new * L / orig)``.  x ``[..., L]``."""
    taps, orig, new, width = restated_taps(orig_freq, new_freq, lowpass_filter_width, rolloff)
    x = np.asarray(x, dtype=np.float64)
    lead = x.shape[:-1]
    x2 = x.reshape(-1, x.shape[-1])
    L = x2.shape[1]
    xp = np.pad(x2, ((0, 0), (width, width + orig)))
    kw = taps.shape[1]
    rows = (xp.shape[1] - kw) // orig + 1
    win = np.lib.stride_tricks.sliding_window_view(xp, kw, axis=1)[:, ::orig][:, :rows]   # [n, rows, kw]
    y = np.einsum("nrk,pk->nrp", win, taps).reshape(x2.shape[0], rows * new)
    target = -((-new * L) // orig)
    return y[:, :target].reshape(*lead, target)


def out_len(length: int, orig_freq: int, new_freq: int) -> int:
    g = math.gcd(int(orig_freq), int(new_freq))
    orig, new = int(orig_freq) // g, int(new_freq) // g
    return -((-new * int(length)) // orig)


def callsite_load_audio_ragged(pcm_list: Sequence[np.ndarray], sr: int, sampling_rate: int):
    """One ``load_audio`` per clip, zero-padded into ``[B, max L']`` + lengths (the loaders' collate contract)."""
    outs = [callsite_load_audio(p, sr, sampling_rate)[0] for p in pcm_list]
    lens = np.array([o.shape[0] for o in outs], dtype=np.int64)
    y = np.zeros((len(outs), int(lens.max())), dtype=np.float32)
    for i, o in enumerate(outs):
        y[i, :o.shape[0]] = o
    return y, lens
