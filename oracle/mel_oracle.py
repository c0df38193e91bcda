"""CPU oracle for the batched log-mel front-end.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module; the product package
(``xtts_b200``) never does and fails loudly when its CUDA library is missing.

Parity status: PINNED.  The reference has no tests or golden vectors of its own
(SURVEY.md F3), so the pin is made here: ``oracle/gen_golden.py`` imports the
reference module itself (``/root/reference/ttts/utils/utils.py:486-522``
``TorchMelSpectrogram`` and ``ttts/vqvae/xtts_dvae.py:28-49`` ``dvae_wav_to_mel``),
runs it on seeded inputs in the build container and commits the outputs under
``tests/golden/``; ``tests/test_oracle.py`` checks both layers below against
those files.

Where the arithmetic lives: the reference's hot path is three calls into an
un-vendored, un-pinned third party -- ``torchaudio.transforms.MelSpectrogram``
(here torchaudio 2.11.0+cu128: ``transforms/_transforms.py:515-632``,
``functional/functional.py:54-145,425-591``) -> ``torch.stft`` (torch 2.11.0,
``functional.py:508-690``) -> ``matmul`` -- plus ``log(clamp(.,1e-5))`` and
``/ mel_norms`` in the reference itself.  Two layers restate it:

* ``callsite_logmel`` -- the reference's own call sequence on the pinned third
  party (what ``TorchMelSpectrogram.forward`` executes, line for line in
  meaning, not in text).  This is the fp32 oracle parity is quoted against and
  the timed CPU baseline (``kind: "port"``).
* ``restated_logmel`` -- the published algorithm written out with numpy in
  fp64 (explicit reflect pad, periodic Hann, real DFT, triangular filterbank,
  power, log, norm).  Independent of torchaudio; it is the "truth" leg of the
  dual-oracle rule for tonal inputs and gives stage-level
  oracles (frames, re/im, power, mel).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np


# --------------------------------------------------------------------------
# configuration rows
# --------------------------------------------------------------------------
@dataclass(frozen=True)
class MelConfig:
    """Parameters of one front-end variant.

    V1 = TorchMelSpectrogram defaults, ttts/utils/utils.py:487-502,518.
    V2 = MelSpectrogramFeatures defaults, ttts/vocoder/feature_extractors.py:74-91
         with safe_log clip 1e-7 from ttts/vocoder/modules.py:194-205.
    """
    sample_rate: int = 22050
    n_fft: int = 1024
    hop: int = 256
    win_length: int = 1024
    n_mels: int = 80
    f_min: float = 0.0
    f_max: Optional[float] = 8000.0
    power: float = 2.0
    norm: Optional[str] = "slaney"
    mel_scale: str = "htk"
    normalized: bool = False
    clamp: float = 1e-5
    center: bool = True
    pre_pad: int = 0            # "same" padding of V2: reflect-pad (win-hop)//2 first, center=False
    power_eps: float = 0.0      # V3: magnitude = sqrt(re^2 + im^2 + 1e-9), feature_extractors.py:46
    librosa_fb: bool = False    # V3: librosa.filters.mel (Slaney scale + norm) instead of torchaudio's fbanks

    @property
    def n_freqs(self) -> int:
        return self.n_fft // 2 + 1

    @property
    def pad(self) -> int:
        """Reflect padding applied on each side before framing."""
        return self.n_fft // 2 if self.center else self.pre_pad

    @property
    def fmax(self) -> float:
        return float(self.f_max) if self.f_max is not None else float(self.sample_rate // 2)


V1 = MelConfig()
V2 = MelConfig(sample_rate=24000, n_mels=100, f_max=None, power=1.0, norm=None, clamp=1e-7)
V2_SAME = MelConfig(sample_rate=24000, n_mels=100, f_max=None, power=1.0, norm=None, clamp=1e-7,
                    center=False, pre_pad=(1024 - 256) // 2)
# V3 = mel_spectrogram() / MelSpectrogramFeatures1 defaults, ttts/vocoder/feature_extractors.py:27-53,102-121.
# PARITY UNPINNED for this row: librosa (the source of its filterbank) is not installed here, so the filterbank is
# restated from librosa's published algorithm (melscale_fbanks64 with the Slaney scale) and no golden could be minted.
V3 = MelConfig(sample_rate=24000, n_mels=80, f_max=8000.0, power=1.0, norm="slaney", mel_scale="slaney", clamp=1e-5,
               center=False, pre_pad=(1024 - 256) // 2, power_eps=1e-9, librosa_fb=True)


def default_mel_norms(n_mels: int = 80) -> np.ndarray:
    """The fixed stand-in for ttts/data/mel_norms.pth (not shipped)."""
    return np.linspace(-6.0, -1.0, n_mels).astype(np.float32)


# --------------------------------------------------------------------------
# integer side: frame counts and padding masks (bit-exact contract)
# --------------------------------------------------------------------------
def frame_count(length: int, cfg: MelConfig = V1) -> int:
    """torch.stft frame count (torch/functional.py:648): 1 + (L + 2*pad - n_fft) // hop."""
    if length <= cfg.pad:
        raise ValueError(f"reflect padding {cfg.pad} needs length > {cfg.pad}, got {length}")
    return 1 + (length + 2 * cfg.pad - cfg.n_fft) // cfg.hop


def make_pad_mask(lengths: Sequence[int], max_len: int = 0) -> np.ndarray:
    """ttts/utils/utils.py:77-103: mask[b, t] = t >= lengths[b]  (True = padding)."""
    lengths = np.asarray(lengths, dtype=np.int64)
    max_len = int(max_len) if max_len > 0 else int(lengths.max())
    return np.arange(max_len, dtype=np.int64)[None, :] >= lengths[:, None]


# --------------------------------------------------------------------------
# restated algorithm, numpy fp64
# --------------------------------------------------------------------------
def hann_periodic(n: int) -> np.ndarray:
    """torch.hann_window(n, periodic=True): 0.5 - 0.5 cos(2 pi k / n)."""
    k = np.arange(n, dtype=np.float64)
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * k / n)


def full_window(cfg: MelConfig) -> np.ndarray:
    """Window as torch.stft sees it: win_length taps centred in n_fft."""
    w = hann_periodic(cfg.win_length)
    left = (cfg.n_fft - cfg.win_length) // 2
    out = np.zeros(cfg.n_fft, dtype=np.float64)
    out[left:left + cfg.win_length] = w
    return out


def reflect_pad(x: np.ndarray, pad: int) -> np.ndarray:
    """F.pad(x, (pad, pad), 'reflect') on the last axis (no edge repeat)."""
    if pad == 0:
        return x
    L = x.shape[-1]
    if L <= pad:
        raise ValueError("reflect pad needs length > pad")
    left = x[..., 1:pad + 1][..., ::-1]
    right = x[..., L - pad - 1:L - 1][..., ::-1]
    return np.concatenate([left, x, right], axis=-1)


def hz_to_mel(f: float, mel_scale: str) -> float:
    if mel_scale == "htk":
        return 2595.0 * math.log10(1.0 + f / 700.0)
    f_sp = 200.0 / 3
    if f >= 1000.0:
        return 1000.0 / f_sp + math.log(f / 1000.0) / (math.log(6.4) / 27.0)
    return f / f_sp


def mel_to_hz(m: np.ndarray, mel_scale: str) -> np.ndarray:
    m = np.asarray(m, dtype=np.float64)
    if mel_scale == "htk":
        return 700.0 * (10.0 ** (m / 2595.0) - 1.0)
    f_sp = 200.0 / 3
    f = f_sp * m
    min_log_mel = 1000.0 / f_sp
    logstep = math.log(6.4) / 27.0
    t = m >= min_log_mel
    f[t] = 1000.0 * np.exp(logstep * (m[t] - min_log_mel))
    return f


def melscale_fbanks64(cfg: MelConfig) -> np.ndarray:
    """Triangular filterbank [n_freqs, n_mels], fp64 (torchaudio functional.py:518-591)."""
    all_freqs = np.linspace(0.0, cfg.sample_rate // 2, cfg.n_freqs)
    m_pts = np.linspace(hz_to_mel(cfg.f_min, cfg.mel_scale), hz_to_mel(cfg.fmax, cfg.mel_scale),
                        cfg.n_mels + 2)
    f_pts = mel_to_hz(m_pts, cfg.mel_scale)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts[None, :] - all_freqs[:, None]
    down = -slopes[:, :-2] / f_diff[:-1]
    up = slopes[:, 2:] / f_diff[1:]
    fb = np.maximum(0.0, np.minimum(down, up))
    if cfg.norm == "slaney":
        fb = fb * (2.0 / (f_pts[2:cfg.n_mels + 2] - f_pts[:cfg.n_mels]))[None, :]
    return fb


@dataclass
class Stages:
    frames: np.ndarray = field(default=None)    # [T, n_fft] un-windowed
    re: np.ndarray = field(default=None)        # [T, n_freqs]
    im: np.ndarray = field(default=None)
    spec: np.ndarray = field(default=None)      # |X|^power  [T, n_freqs]
    mel: np.ndarray = field(default=None)       # [T, n_mels]
    logmel: np.ndarray = field(default=None)    # [T, n_mels] un-normalised


def restated_stages(x: np.ndarray, cfg: MelConfig = V1, fb: Optional[np.ndarray] = None) -> Stages:
    """All stages for ONE clip x[L], fp64."""
    x = np.asarray(x, dtype=np.float64)
    assert x.ndim == 1
    T = frame_count(x.shape[0], cfg)
    xp = reflect_pad(x, cfg.pad)
    idx = np.arange(T)[:, None] * cfg.hop + np.arange(cfg.n_fft)[None, :]
    frames = xp[idx]
    w = full_window(cfg)
    if cfg.normalized:
        w = w / math.sqrt(float((w ** 2).sum()))
    X = np.fft.rfft(frames * w[None, :], axis=-1)
    mag2 = X.real ** 2 + X.imag ** 2
    spec = mag2 if cfg.power == 2.0 else np.sqrt(mag2 + cfg.power_eps) ** cfg.power
    fb = melscale_fbanks64(cfg) if fb is None else np.asarray(fb, dtype=np.float64)
    mel = spec @ fb
    logmel = np.log(np.maximum(mel, cfg.clamp))
    return Stages(frames=frames, re=X.real, im=X.imag, spec=spec, mel=mel, logmel=logmel)


def restated_logmel(wav: np.ndarray, cfg: MelConfig = V1, mel_norms: Optional[np.ndarray] = None,
                    lengths: Optional[Sequence[int]] = None, fb: Optional[np.ndarray] = None):
    """fp64 log-mel of a batch wav[B, Lmax].

    Dense call (lengths None): returns mel[B, n_mels, T] like the reference.
    Ragged call: per-clip reflect at each clip's own length;
    returns (mel[B, n_mels, Tmax] with 0.0 in padded frames, n_frames[B] int64,
    pad_mask[B, Tmax] bool) -- the collate contract of ttts/gpt/dataset.py:95-130.
    """
    wav = np.asarray(wav)
    if wav.ndim == 3:
        wav = wav[:, 0, :]
    assert wav.ndim == 2
    B, Lmax = wav.shape
    lens = [Lmax] * B if lengths is None else [int(v) for v in lengths]
    n_frames = np.array([frame_count(v, cfg) for v in lens], dtype=np.int64)
    Tmax = frame_count(Lmax, cfg) if lengths is None else int(n_frames.max())
    out = np.zeros((B, cfg.n_mels, Tmax), dtype=np.float64)
    fb = melscale_fbanks64(cfg) if fb is None else fb
    for b in range(B):
        lm = restated_stages(wav[b, :lens[b]], cfg, fb).logmel      # [T, n_mels]
        if mel_norms is not None:
            lm = lm / np.asarray(mel_norms, dtype=np.float64)[None, :]
        out[b, :, :n_frames[b]] = lm.T
    if lengths is None:
        return out
    return out, n_frames, make_pad_mask(n_frames, Tmax)


# --------------------------------------------------------------------------
# the reference's call sequence on the pinned third party (fp32)
# --------------------------------------------------------------------------
_CALLSITE_CACHE = {}


def _callsite_transform(cfg: MelConfig):
    import torchaudio
    key = cfg
    if key not in _CALLSITE_CACHE:
        # ttts/utils/utils.py:498-502 (V1) and ttts/vocoder/feature_extractors.py:80-91 (V2)
        _CALLSITE_CACHE[key] = torchaudio.transforms.MelSpectrogram(
            n_fft=cfg.n_fft, hop_length=cfg.hop, win_length=cfg.win_length, power=cfg.power,
            normalized=cfg.normalized, sample_rate=cfg.sample_rate, f_min=cfg.f_min,
            f_max=cfg.f_max, n_mels=cfg.n_mels, norm=cfg.norm, mel_scale=cfg.mel_scale,
            center=cfg.center)
    return _CALLSITE_CACHE[key]


def callsite_logmel(wav, cfg: MelConfig = V1, mel_norms=None):
    """What TorchMelSpectrogram.forward (utils.py:509-522) / MelSpectrogramFeatures.forward
    (feature_extractors.py:93-99) compute, on torch CPU tensors, fp32.  Returns a torch tensor
    [B, n_mels, T]."""
    import torch
    x = torch.as_tensor(wav)
    if x.dim() == 3:
        x = x.squeeze(1)
    assert x.dim() == 2
    if cfg.librosa_fb:
        # mel_spectrogram(), feature_extractors.py:40-51
        basis = torch.from_numpy(melscale_fbanks64(cfg).T.astype(np.float32))          # librosa_mel_fn(...)
        y = torch.nn.functional.pad(x.unsqueeze(1), (cfg.pre_pad, cfg.pre_pad), mode="reflect").squeeze(1)
        spec = torch.stft(y, cfg.n_fft, hop_length=cfg.hop, win_length=cfg.win_length,
                          window=torch.hann_window(cfg.win_length), center=False, pad_mode="reflect",
                          normalized=False, onesided=True, return_complex=True)
        spec = torch.sqrt(torch.view_as_real(spec).pow(2).sum(-1) + cfg.power_eps)
        return torch.log(torch.clamp(torch.matmul(basis, spec), min=cfg.clamp))
    if cfg.pre_pad:
        x = torch.nn.functional.pad(x, (cfg.pre_pad, cfg.pre_pad), mode="reflect")
    mel = _callsite_transform(cfg)(x)
    mel = torch.log(torch.clamp(mel, min=cfg.clamp))
    if mel_norms is not None:
        mel = mel / torch.as_tensor(mel_norms).unsqueeze(0).unsqueeze(-1)
    return mel


def callsite_logmel_ragged(wav, lengths, cfg: MelConfig = V1, mel_norms=None):
    """The reference's preprocessing shape: one clip per call (ttts/prepare/mel_extract.py:15-27),
    then zero-pad the mels and carry integer lengths (ttts/gpt/dataset.py:95-130)."""
    import torch
    x = torch.as_tensor(wav)
    B = x.shape[0]
    n_frames = np.array([frame_count(int(v), cfg) for v in lengths], dtype=np.int64)
    Tmax = int(n_frames.max())
    out = torch.zeros(B, cfg.n_mels, Tmax, dtype=torch.float32)
    for b in range(B):
        out[b, :, :n_frames[b]] = callsite_logmel(x[b:b + 1, :int(lengths[b])], cfg, mel_norms)[0]
    return out, n_frames, make_pad_mask(n_frames, Tmax)


# --------------------------------------------------------------------------
# mel_norms statistics (SURVEY.md 8a row S4 -- no reference code; semantics defined there)
# --------------------------------------------------------------------------
def mel_stats(logmel_bct: np.ndarray, n_frames: Sequence[int]) -> np.ndarray:
    """acc[2*n_mels+1] fp64 = [sum_m ..., sumsq_m ..., count] over valid frames of UN-normalised
    log-mel laid out [B, n_mels, T]."""
    lm = np.asarray(logmel_bct, dtype=np.float64)
    B, C, _ = lm.shape
    acc = np.zeros(2 * C + 1, dtype=np.float64)
    for b in range(B):
        v = lm[b, :, :int(n_frames[b])]
        acc[:C] += v.sum(axis=1)
        acc[C:2 * C] += (v * v).sum(axis=1)
        acc[2 * C] += int(n_frames[b])
    return acc


def norms_from_stats(acc: np.ndarray) -> np.ndarray:
    """mel_norms[m] = mean over valid frames of log-mel[m] (tortoise convention)."""
    C = (acc.shape[0] - 1) // 2
    return (acc[:C] / acc[2 * C]).astype(np.float32)


# --------------------------------------------------------------------------
# seeded synthetic inputs shared by goldens, tests and bench
# --------------------------------------------------------------------------
def synth(kind: str, B: int, L: int, seed: int) -> np.ndarray:
    rng = np.random.default_rng(seed)
    if kind == "uniform":
        return (rng.random((B, L), dtype=np.float32) * 2.0 - 1.0).astype(np.float32)
    if kind == "gauss":
        return np.clip(rng.standard_normal((B, L), dtype=np.float32) * 0.1, -1.0, 1.0).astype(np.float32)
    if kind == "silence":
        return np.zeros((B, L), dtype=np.float32)
    if kind == "tone":
        t = np.arange(L, dtype=np.float64)[None, :]
        f = 440.0 * (1.0 + np.arange(B, dtype=np.float64)[:, None])
        x = 0.5 * np.sin(2 * np.pi * f * t / 22050.0) + 1e-3 * rng.standard_normal((B, L))
        return x.astype(np.float32)
    # clean high-dynamic-range inputs (SURVEY.md 8c ii: "pure tone", "1/f"), no added noise floor
    if kind == "puretone":
        t = np.arange(L, dtype=np.float64)[None, :]
        f = 440.0 * (1.0 + np.arange(B, dtype=np.float64)[:, None])
        return (0.5 * np.sin(2 * np.pi * f * t / 22050.0)).astype(np.float32)
    if kind == "tone_silence":      # half a second of tone, digital silence, tone again (speech-like on/off structure)
        x = synth("puretone", B, L, seed)
        seg = max(1, L // 3)
        x[:, seg:2 * seg] = 0.0
        return x
    if kind == "pink":              # 1/f noise: white spectrum shaped by f^-1/2 in amplitude, peak-normalised to 0.5
        w = rng.standard_normal((B, L))
        X = np.fft.rfft(w, axis=-1)
        f = np.arange(X.shape[-1], dtype=np.float64)
        f[0] = 1.0
        y = np.fft.irfft(X / np.sqrt(f)[None, :], n=L, axis=-1)
        return (0.5 * y / np.abs(y).max(axis=-1, keepdims=True)).astype(np.float32)
    raise ValueError(kind)
