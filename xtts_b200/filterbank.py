"""Host-side constants of the front-end: analysis window and mel filterbank.

The reference gets both from torchaudio at construction time
(``ttts/utils/utils.py:498-502`` -> torchaudio ``transforms/_transforms.py:86-87`` for the
window and ``functional/functional.py:518-591`` ``melscale_fbanks`` for ``fb``).  The product
path must not call torchaudio, so the same torch fp32 operations are issued here in the same
order; ``tests/test_filterbank.py`` asserts ``torch.equal`` against torchaudio's buffers.
"""
from __future__ import annotations

import math
from typing import Optional

import torch


def hz_to_mel(freq: float, mel_scale: str = "htk") -> float:
    if mel_scale not in ("slaney", "htk"):
        raise ValueError('mel_scale should be one of "htk" or "slaney".')
    if mel_scale == "htk":
        return 2595.0 * math.log10(1.0 + (freq / 700.0))
    f_sp = 200.0 / 3
    mels = freq / f_sp
    min_log_hz = 1000.0
    if freq >= min_log_hz:
        mels = min_log_hz / f_sp + math.log(freq / min_log_hz) / (math.log(6.4) / 27.0)
    return mels


def mel_to_hz(mels: torch.Tensor, mel_scale: str = "htk") -> torch.Tensor:
    if mel_scale == "htk":
        return 700.0 * (10.0 ** (mels / 2595.0) - 1.0)
    f_sp = 200.0 / 3
    freqs = f_sp * mels
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = math.log(6.4) / 27.0
    log_t = mels >= min_log_mel
    freqs[log_t] = min_log_hz * torch.exp(logstep * (mels[log_t] - min_log_mel))
    return freqs


def melscale_fbanks(n_freqs: int, f_min: float, f_max: float, n_mels: int, sample_rate: int,
                    norm: Optional[str] = None, mel_scale: str = "htk") -> torch.Tensor:
    """Triangular filterbank ``[n_freqs, n_mels]`` fp32, bit-identical to torchaudio's."""
    if norm is not None and norm != "slaney":
        raise ValueError('norm must be one of None or "slaney"')
    all_freqs = torch.linspace(0, sample_rate // 2, n_freqs)
    m_pts = torch.linspace(hz_to_mel(f_min, mel_scale), hz_to_mel(f_max, mel_scale), n_mels + 2)
    f_pts = mel_to_hz(m_pts, mel_scale)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts.unsqueeze(0) - all_freqs.unsqueeze(1)
    down_slopes = (-1.0 * slopes[:, :-2]) / f_diff[:-1]
    up_slopes = slopes[:, 2:] / f_diff[1:]
    fb = torch.max(torch.zeros(1), torch.min(down_slopes, up_slopes))
    if norm == "slaney":
        enorm = 2.0 / (f_pts[2:n_mels + 2] - f_pts[:n_mels])
        fb *= enorm.unsqueeze(0)
    return fb


def analysis_window(n_fft: int, win_length: Optional[int]) -> torch.Tensor:
    """The window torch.stft applies: periodic Hann of win_length taps centred in n_fft."""
    win_length = n_fft if win_length is None else win_length
    if win_length > n_fft:
        raise ValueError("win_length must be <= n_fft")
    w = torch.hann_window(win_length)
    if win_length == n_fft:
        return w
    left = (n_fft - win_length) // 2
    out = torch.zeros(n_fft)
    out[left:left + win_length] = w
    return out


def librosa_mel(sr: int, n_fft: int, n_mels: int, fmin: float = 0.0, fmax: Optional[float] = None) -> torch.Tensor:
    """``librosa.filters.mel(sr=, n_fft=, n_mels=, fmin=, fmax=)`` with its defaults (Slaney mel scale,
    ``norm='slaney'``, float32 result of an fp64 computation), restated from the published algorithm because
    librosa is not installed here.  The reference calls it in ``mel_spectrogram()``
    (ttts/vocoder/feature_extractors.py:36-37).  Returns ``[n_mels, 1 + n_fft//2]`` float32."""
    import numpy as np
    fmax = float(sr) / 2 if fmax is None else float(fmax)
    fftfreqs = np.linspace(0.0, float(sr) / 2, 1 + n_fft // 2)            # librosa.fft_frequencies

    def hz_to_mel_slaney(f):
        f = np.asanyarray(f, dtype=np.float64)
        f_sp = 200.0 / 3
        mels = f / f_sp
        min_log_hz, logstep = 1000.0, np.log(6.4) / 27.0
        min_log_mel = min_log_hz / f_sp
        return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-30) / min_log_hz) / logstep, mels)

    def mel_to_hz_slaney(m):
        m = np.asanyarray(m, dtype=np.float64)
        f_sp = 200.0 / 3
        min_log_hz, logstep = 1000.0, np.log(6.4) / 27.0
        min_log_mel = min_log_hz / f_sp
        return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)

    mel_f = mel_to_hz_slaney(np.linspace(hz_to_mel_slaney(fmin), hz_to_mel_slaney(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fftfreqs[None, :]
    weights = np.zeros((n_mels, 1 + n_fft // 2), dtype=np.float64)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        weights[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    weights *= enorm[:, None]
    return torch.from_numpy(weights.astype(np.float32))
