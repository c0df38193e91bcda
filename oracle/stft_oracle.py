"""CPU oracle for the STFT family.  TEST INFRASTRUCTURE ONLY -- only ``tests/`` and ``tools/`` import it;
the product package (``xtts_b200``) never does.

Parity status: PINNED.  ``oracle/gen_golden_stft.py`` runs the reference's own classes in the build container --
``TorchSTFT`` / ``STFTLoss`` / ``MultiScaleSTFTLoss`` compiled out of ``/root/reference/ttts/hifigan/losses.py:9-218`` with
``ast`` (the module imports librosa at the top, which is not installed) and ``ISTFT`` imported from
``/root/reference/ttts/vocoder/spectral_ops.py:7-60`` -- on seeded inputs and commits the outputs to
``tests/golden/stft_family.npz``; ``tests/test_stft.py`` checks both layers below against that file.

* ``callsite_*``  -- the reference's call sequences on the pinned third party (torch 2.11 ``torch.stft`` / ``irfft`` /
  ``fold``), fp32: what parity is quoted against.
* ``restated_*``  -- the published algorithm in numpy fp64 (explicit reflect pad, centred window, DFT, overlap-add).
"""
from __future__ import annotations

import numpy as np

MS_SHAPES = ((1024, 120, 600), (2048, 240, 1200), (512, 50, 240))      # losses.py:201


# ---------------------------------------------------------------- forward: TorchSTFT.__call__, losses.py:107-148
def callsite_stft_mag(x, n_fft, hop, win, power=None, normalized=False):
    import torch
    x = torch.as_tensor(x)
    if x.ndim == 2:
        x = x.unsqueeze(1)
    o = torch.stft(x.squeeze(1), n_fft, hop, win, torch.hann_window(win), center=True, pad_mode="reflect",
                   normalized=normalized, onesided=True, return_complex=True)
    o = torch.view_as_real(o)
    S = torch.sqrt(torch.clamp(o[..., 0] ** 2 + o[..., 1] ** 2, min=1e-8))
    if power is not None:
        S = S ** power
    return S


def restated_stft_mag(x, n_fft, hop, win, power=None, normalized=False):
    x = np.asarray(x, dtype=np.float64)
    pad = n_fft // 2
    L = x.shape[-1]
    xp = np.concatenate([x[..., 1:pad + 1][..., ::-1], x, x[..., L - pad - 1:L - 1][..., ::-1]], axis=-1)
    T = 1 + L // hop
    w = np.zeros(n_fft)
    k = np.arange(win, dtype=np.float64)
    w[(n_fft - win) // 2:(n_fft - win) // 2 + win] = 0.5 - 0.5 * np.cos(2.0 * np.pi * k / win)
    if normalized:
        w = w / np.sqrt(n_fft)
    idx = np.arange(T)[:, None] * hop + np.arange(n_fft)[None, :]
    X = np.fft.rfft(xp[..., idx] * w, axis=-1)                      # [B, T, bins]
    S = np.sqrt(np.maximum(X.real ** 2 + X.imag ** 2, 1e-8))
    if power is not None:
        S = S ** power
    return np.swapaxes(S, -1, -2)                                   # [B, bins, T]


def callsite_stft_loss(y_hat, y, n_fft, hop, win):
    """STFTLoss.forward, losses.py:188-194."""
    import torch
    a, b = callsite_stft_mag(y_hat, n_fft, hop, win), callsite_stft_mag(y, n_fft, hop, win)
    return (torch.nn.functional.l1_loss(torch.log(b), torch.log(a)),
            torch.norm(b - a, p="fro") / torch.norm(b, p="fro"))


def callsite_multiscale_loss(y_hat, y, shapes=MS_SHAPES):
    """MultiScaleSTFTLoss.forward, losses.py:207-218."""
    lm = lsc = 0
    for n_fft, hop, win in shapes:
        a, b = callsite_stft_loss(y_hat, y, n_fft, hop, win)
        lm, lsc = lm + a, lsc + b
    return lm / len(shapes), lsc / len(shapes)


# ---------------------------------------------------------------- inverse: ISTFT.forward, spectral_ops.py:32-60
def callsite_istft(spec, n_fft, hop, win, padding="same"):
    import torch
    spec = torch.as_tensor(spec)
    window = torch.hann_window(win)
    if padding == "center":
        return torch.istft(spec, n_fft, hop, win, window, center=True)
    pad = (win - hop) // 2
    B, N, T = spec.shape
    ifft = torch.fft.irfft(spec, n_fft, dim=1, norm="backward") * window[None, :, None]
    output_size = (T - 1) * hop + win
    y = torch.nn.functional.fold(ifft, output_size=(1, output_size), kernel_size=(1, win), stride=(1, hop))[:, 0, 0, pad:-pad]
    window_sq = window.square().expand(1, T, -1).transpose(1, 2)
    env = torch.nn.functional.fold(window_sq, output_size=(1, output_size), kernel_size=(1, win),
                                   stride=(1, hop)).squeeze()[pad:-pad]
    assert (env > 1e-11).all()
    return y / env


def restated_istft(spec, n_fft, hop, win, padding="same"):
    spec = np.asarray(spec, dtype=np.complex128)
    B, N, T = spec.shape
    k = np.arange(win, dtype=np.float64)
    w = 0.5 - 0.5 * np.cos(2.0 * np.pi * k / win)
    frames = np.fft.irfft(spec, n_fft, axis=1) * w[None, :, None]          # [B, n_fft, T]
    size = (T - 1) * hop + win
    y, env = np.zeros((B, size)), np.zeros(size)
    for t in range(T):
        y[:, t * hop:t * hop + win] += frames[:, :, t]
        env[t * hop:t * hop + win] += w * w
    pad = (win - hop) // 2 if padding == "same" else n_fft // 2
    return y[:, pad:size - pad] / env[pad:size - pad]


def synth_spec(B, n_fft, T, seed):
    """A complex spectrogram as a Vocos head emits it: magnitude exp(N(0,1)) clipped at 1e2, uniform phase."""
    rng = np.random.default_rng(seed)
    N = n_fft // 2 + 1
    mag = np.minimum(np.exp(rng.standard_normal((B, N, T))), 1e2)
    ph = rng.uniform(-np.pi, np.pi, (B, N, T))
    return (mag * np.exp(1j * ph)).astype(np.complex64)
