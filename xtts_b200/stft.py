"""STFT family behind the log-mel path (first slice -- FORWARD ONLY, no autograd).

Host-side mirrors of (file:line in /root/reference):

* ``TorchSTFT``            ttts/hifigan/losses.py:9-172   (``__call__`` :107-148; ``use_mel`` needs librosa and is not built)
* ``STFTLoss``             ttts/hifigan/losses.py:180-194
* ``MultiScaleSTFTLoss``   ttts/hifigan/losses.py:196-218 ((n_fft, hop, win) = (1024, 120, 600), (2048, 240, 1200), (512, 50, 240))
* ``ISTFT``                ttts/vocoder/spectral_ops.py:7-60 (Vocos head, ``heads.py:40``: n_fft = win_length)

None of these shapes has hop = n_fft / 4, which the fused log-mel kernel's slot sharing assumes, so they run on the
library's general dense contraction instead (``ttts_resample_forward``: [strides x taps] x [taps x phases] on tcgen05,
3xTF32): stride = hop, taps = win_length, phases = (re, im) of every bin -- and for the inverse stride = taps = 2 * bins,
phases = n_fft -- with the small kernels of ``include/ttts_stft.h`` around it.  The losses need a backward pass to be
useful in training; this slice gives the forward values only (the reference's two scalars) and says so.
"""
from __future__ import annotations

import ctypes
import math
from typing import Optional

import numpy as np
import torch
from torch import nn

from . import _cabi


class _Contraction:
    """y[b, s * n_out + p] = sum_k x[b, s * stride + k - width] * taps[p, k]  (one ``ttts_resample_plan``)."""

    def __init__(self, taps: np.ndarray, stride: int, width: int, device: torch.device):
        self.n_out, self.kw = int(taps.shape[0]), int(taps.shape[1])
        assert self.kw == 2 * width + stride
        self.stride, self.width, self.device = int(stride), int(width), device
        self._lib = _cabi.lib()
        self._h = None
        host = torch.from_numpy(np.ascontiguousarray(taps, dtype=np.float32))
        h = ctypes.c_void_p()
        with torch.cuda.device(device):
            _cabi.check(self._lib.ttts_resample_plan_create(host.data_ptr(), self.stride, self.n_out, self.width,
                                                            ctypes.byref(h)))
        self._h = h

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.ttts_resample_plan_destroy(h)
            except Exception:
                pass

    def __call__(self, x: torch.Tensor, length: int) -> torch.Tensor:
        """x [B, >= length] float32 (any row stride); returns [B, rows * n_out], rows = ceil(length / stride)."""
        B = int(x.shape[0])
        rows = (int(length) + self.stride - 1) // self.stride
        cols = rows * self.n_out
        pitch = (cols + 3) // 4 * 4
        with torch.cuda.device(self.device):
            out = torch.empty((B, pitch), dtype=torch.float32, device=self.device)
            stream = torch.cuda.current_stream(self.device).cuda_stream
            _cabi.check(self._lib.ttts_resample_forward(self._h, x.data_ptr(), int(x.stride(0)) if B > 1 else int(x.shape[1]),
                                                        None, B, int(length), out.data_ptr(), pitch, 0, stream))
        return out


def windowed_dft_taps(window: np.ndarray, n_fft: int) -> np.ndarray:
    """[2 * bins, win] float64: row 2k = w[n] cos(2 pi k m / N), row 2k + 1 = -w[n] sin(.), m = n + (N - win) / 2 (the
    window sits centred in the n_fft frame, as torch.stft pads it)."""
    N, W = int(n_fft), int(window.shape[0])
    nb = N // 2 + 1
    n = np.arange(W, dtype=np.int64) + (N - W) // 2
    k = np.arange(nb, dtype=np.int64)
    ang = 2.0 * np.pi * ((k[:, None] * n[None, :]) % N) / N
    taps = np.empty((2 * nb, W), dtype=np.float64)
    taps[0::2] = np.cos(ang) * window[None, :]
    taps[1::2] = -np.sin(ang) * window[None, :]
    return taps


def framed_contraction(lib, con: "_Contraction", x: torch.Tensor, lengths_dev: Optional[torch.Tensor], N: int, H: int,
                       W: int, pad: int, device: torch.device):
    """Reflect-pad every clip by ``pad`` at its OWN length (torch.stft(center=True): pad = N / 2) and run the windowed-DFT
    contraction over it.  Returns (y [B, rows * 2 * bins], d, T): frame t of clip b is row d + t of y[b], T = the frame
    count of the full width.  The contraction's zero-padding convention puts d = ceil(width / H) rows before frame 0, so
    the padded clip is offset (and given a zero margin when the window fills the frame) to make row d + t frame t."""
    B, L = int(x.shape[0]), int(x.shape[1])
    if L <= pad or L + 2 * pad < N:
        raise RuntimeError(f"Argument #4: Padding size should be less than the corresponding input dimension, but got: "
                           f"padding ({pad}, {pad}) at dimension 2 of input of length {L}")
    if pad % 4:
        raise NotImplementedError("the reflect padding must be a multiple of 4 samples here")
    T = 1 + (L + 2 * pad - N) // H
    width = con.width
    d = -(-width // H)                                  # leading rows of the contraction that precede frame 0
    off0 = (N - W) // 2 - (d * H - width)               # where row 0 of the contraction starts in the padded clip
    margin = (max(0, -off0) + 3) // 4 * 4               # rows before frame 0 may start left of it: zeros there
    length = (T + d) * H + width                        # samples the contraction may read per clip
    rows = -(-length // H)
    Lp = (max(off0, 0) + rows * H + W + 8 + margin + 3) // 4 * 4
    Lp = max(Lp, (L + 2 * pad + margin + 3) // 4 * 4)
    stream = torch.cuda.current_stream(device).cuda_stream
    flat = torch.empty(B * Lp + margin, dtype=torch.float32, device=device)
    if margin:
        flat[:margin].zero_()
    _cabi.check(lib.ttts_reflect_pad(x.data_ptr(), int(x.stride(0)) if B > 1 else L,
                                     lengths_dev.data_ptr() if lengths_dev is not None else None, B, L, pad,
                                     flat.data_ptr() + 4 * margin, Lp, stream))
    xv = flat.as_strided((B, Lp - max(off0, 0)), (Lp, 1), margin + off0)
    return con(xv, length), d, T


def _device_of(x: torch.Tensor) -> torch.device:
    if not x.is_cuda:
        raise RuntimeError("xtts_b200.stft runs on CUDA sm_100a only; there is no CPU fallback")
    return x.device if x.device.index is not None else torch.device("cuda", torch.cuda.current_device())


class TorchSTFT(nn.Module):
    """ttts/hifigan/losses.py:9-172 (magnitude spectrogram by torch.stft) on the B200 library.  Forward only."""

    def __init__(self, n_fft, hop_length, win_length, pad_wav=False, window="hann_window", sample_rate=None,
                 mel_fmin=0, mel_fmax=None, n_mels=80, use_mel=False, do_amp_to_db=False, spec_gain=1.0, power=None,
                 use_htk=False, mel_norm="slaney", normalized=False):
        super().__init__()
        self.n_fft, self.hop_length, self.win_length = n_fft, hop_length, win_length
        self.pad_wav, self.sample_rate = pad_wav, sample_rate
        self.mel_fmin, self.mel_fmax, self.n_mels, self.use_mel = mel_fmin, mel_fmax, n_mels, use_mel
        self.do_amp_to_db, self.spec_gain, self.power = do_amp_to_db, spec_gain, power
        self.use_htk, self.mel_norm, self.normalized = use_htk, mel_norm, normalized
        self.window = nn.Parameter(getattr(torch, window)(win_length), requires_grad=False)
        self.mel_basis = None
        if use_mel:
            raise NotImplementedError("TorchSTFT(use_mel=True) builds its basis with librosa.filters.mel (losses.py:150-160); "
                                      "the STFT losses of the reference do not use it and it is not built")
        if (n_fft - win_length) % 2 or win_length > n_fft or n_fft % 8:
            raise ValueError("need win_length <= n_fft, (n_fft - win_length) even and n_fft % 8 == 0")
        if (win_length - hop_length) % 2 or hop_length > win_length:
            raise ValueError("need hop_length <= win_length and (win_length - hop_length) even")

    # ---- plan: windowed DFT basis as the taps of the contraction, one per device
    def _plan(self, device: torch.device) -> _Contraction:
        plans = self.__dict__.setdefault("_plans", {})
        key = (device.index, self.window._version)
        if key not in plans:
            N, W = self.n_fft, self.win_length
            w = self.window.detach().double().cpu().numpy()
            if self.normalized:
                w = w / math.sqrt(N)                          # torch.stft(normalized=True): frame_length ** -0.5
            taps = windowed_dft_taps(w, N)
            plans[key] = _Contraction(taps, self.hop_length, (W - self.hop_length) // 2, device)
        return plans[key]

    def __getstate__(self):
        state = self.__dict__.copy()
        state.pop("_plans", None)
        return state

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        """x [B, T] or [B, 1, T] -> magnitude spectrogram [B, n_fft // 2 + 1, 1 + T' // hop] float32."""
        if x.ndim == 2:
            x = x.unsqueeze(1)
        device = _device_of(x)
        x = x.squeeze(1)
        if x.dtype != torch.float32:
            x = x.float()
        if x.stride(1) != 1:
            x = x.contiguous()
        lib = _cabi.lib()
        N, H, W = self.n_fft, self.hop_length, self.win_length
        with torch.cuda.device(device):
            stream = torch.cuda.current_stream(device).cuda_stream
            if self.pad_wav:                                    # losses.py:124-126: reflect (n_fft - hop) / 2 first
                p0 = int((N - H) / 2)
                if p0 % 4:
                    raise NotImplementedError("pad_wav needs (n_fft - hop_length) / 2 to be a multiple of 4 here")
                B0, L0 = int(x.shape[0]), int(x.shape[1])
                Lp0 = (L0 + 2 * p0 + 3) // 4 * 4
                x0 = torch.empty((B0, Lp0), dtype=torch.float32, device=device)
                _cabi.check(lib.ttts_reflect_pad(x.data_ptr(), int(x.stride(0)) if B0 > 1 else L0, None, B0, L0, p0,
                                                 x0.data_ptr(), Lp0, stream))
                x = x0[:, :L0 + 2 * p0]
            B, L = int(x.shape[0]), int(x.shape[1])
            con = self._plan(device)
            y, d, T = framed_contraction(lib, con, x, None, N, H, W, N // 2, device)   # row d + t = frame t, (re, im) pairs
            nb = N // 2 + 1
            out = torch.empty((B, nb, T), dtype=torch.float32, device=device)
            power = float(self.power) if self.power is not None else 0.0
            _cabi.check(lib.ttts_stft_magnitude(y.data_ptr(), int(y.stride(0)), d, 2 * nb, B, T, nb, 1e-8, power,
                                                1 if self.do_amp_to_db else 0, float(self.spec_gain), out.data_ptr(), stream))
        return out

    forward = __call__

    @staticmethod
    def _amp_to_db(x, spec_gain=1.0):
        return torch.log(torch.clamp(x, min=1e-5) * spec_gain)

    @staticmethod
    def _db_to_amp(x, spec_gain=1.0):
        return torch.exp(x) / spec_gain


class STFTLoss(nn.Module):
    """ttts/hifigan/losses.py:180-194: L1 of the log magnitudes and the spectral-convergence ratio.  FORWARD VALUES ONLY
    (no autograd through the library): the magnitudes come from the tcgen05 contraction, the two reductions over them
    are plain device reductions (as in the reference)."""

    def __init__(self, n_fft, hop_length, win_length):
        super().__init__()
        self.n_fft, self.hop_length, self.win_length = n_fft, hop_length, win_length
        self.stft = TorchSTFT(n_fft, hop_length, win_length)

    def forward(self, y_hat, y):
        y_hat_M = self.stft(y_hat)
        y_M = self.stft(y)
        loss_mag = torch.nn.functional.l1_loss(torch.log(y_M), torch.log(y_hat_M))
        loss_sc = torch.norm(y_M - y_hat_M, p="fro") / torch.norm(y_M, p="fro")
        return loss_mag, loss_sc


class MultiScaleSTFTLoss(nn.Module):
    """ttts/hifigan/losses.py:196-218."""

    def __init__(self, n_ffts=(1024, 2048, 512), hop_lengths=(120, 240, 50), win_lengths=(600, 1200, 240)):
        super().__init__()
        self.loss_funcs = nn.ModuleList()
        for n_fft, hop_length, win_length in zip(n_ffts, hop_lengths, win_lengths):
            self.loss_funcs.append(STFTLoss(n_fft, hop_length, win_length))

    def forward(self, y_hat, y):
        N = len(self.loss_funcs)
        loss_sc = 0
        loss_mag = 0
        for f in self.loss_funcs:
            lm, lsc = f(y_hat, y)
            loss_mag += lm
            loss_sc += lsc
        loss_sc /= N
        loss_mag /= N
        return loss_mag, loss_sc


class ISTFT(nn.Module):
    """ttts/vocoder/spectral_ops.py:7-60: inverse STFT with "same" (CNN-like) or "center" padding.
    spec (B, N, T) complex64 -> (B, L) float32.  The inverse real DFT, the window, the overlap-add, the window-envelope
    normalisation and the trim of the reference (irfft, two ``fold`` calls, a divide) are one contraction + one kernel."""

    def __init__(self, n_fft: int, hop_length: int, win_length: int, padding: str = "same"):
        super().__init__()
        if padding not in ["center", "same"]:
            raise ValueError("Padding must be 'center' or 'same'.")
        self.padding = padding
        self.n_fft, self.hop_length, self.win_length = n_fft, hop_length, win_length
        self.register_buffer("window", torch.hann_window(win_length))
        if win_length != n_fft:
            # the reference multiplies the n_fft-long irfft by the window directly (spectral_ops.py:53): equal sizes only
            raise ValueError("ISTFT needs win_length == n_fft (as the reference's own broadcast does)")

    def _plan(self, device: torch.device) -> _Contraction:
        plans = self.__dict__.setdefault("_plans", {})
        key = device.index
        if key not in plans:
            N = self.n_fft
            nb = N // 2 + 1
            w = self.window.detach().double().cpu().numpy()
            n = np.arange(N, dtype=np.int64)
            k = np.arange(nb, dtype=np.int64)
            ang = 2.0 * np.pi * ((n[:, None] * k[None, :]) % N) / N
            coef = np.full(nb, 2.0)
            coef[0] = 1.0
            if N % 2 == 0:
                coef[-1] = 1.0
            taps = np.empty((N, 2 * nb), dtype=np.float64)      # irfft, norm="backward": 1/N; Im of bins 0 and N/2 is ignored
            taps[:, :nb] = np.cos(ang) * coef[None, :]
            simag = -np.sin(ang) * coef[None, :]
            simag[:, 0] = 0.0
            if N % 2 == 0:
                simag[:, -1] = 0.0
            taps[:, nb:] = simag
            taps *= (w / N)[:, None]
            plans[key] = _Contraction(taps, 2 * nb, 0, device)
        return plans[key]

    def __getstate__(self):
        state = self.__dict__.copy()
        state.pop("_plans", None)
        return state

    def forward(self, spec: torch.Tensor) -> torch.Tensor:
        assert spec.dim() == 3, "Expected a 3D tensor as input"
        device = _device_of(spec)
        if not spec.is_complex():
            raise RuntimeError("ISTFT expects a complex spectrogram")
        spec = spec.to(torch.complex64).contiguous()
        B, N, T = (int(v) for v in spec.shape)
        if N != self.n_fft // 2 + 1:
            raise RuntimeError(f"expected {self.n_fft // 2 + 1} frequency bins, got {N}")
        pad = (self.win_length - self.hop_length) // 2 if self.padding == "same" else self.n_fft // 2
        out_len = (T - 1) * self.hop_length + self.n_fft - 2 * pad
        if out_len <= 0:
            raise RuntimeError("too few frames for this padding")
        lib = _cabi.lib()
        with torch.cuda.device(device):
            stream = torch.cuda.current_stream(device).cuda_stream
            x = torch.empty((B, T * 2 * N), dtype=torch.float32, device=device)
            _cabi.check(lib.ttts_spec_pack(torch.view_as_real(spec).data_ptr(), B, N, T, x.data_ptr(), stream))
            v = self._plan(device)(x, T * 2 * N)                 # [B, T * n_fft] windowed inverse-DFT frames
            if int(v.stride(0)) != T * self.n_fft and B > 1:
                v = v.contiguous()
            pitch = (out_len + 3) // 4 * 4
            out = torch.empty((B, pitch), dtype=torch.float32, device=device)[:, :out_len]
            status = torch.empty(1, dtype=torch.int32, device=device)
            win = self.window.to(device=device, dtype=torch.float32).contiguous()
            _cabi.check(lib.ttts_istft_overlap_add(v.data_ptr(), B, T, self.n_fft, self.hop_length, pad, win.data_ptr(),
                                                   out.data_ptr(), pitch, status.data_ptr(), stream))
            self._last_status = status
        return out

    def check_envelope(self) -> None:
        """The reference asserts ``(window_envelope > 1e-11).all()`` inside forward (a sync); here it is on request."""
        st = getattr(self, "_last_status", None)
        if st is not None and int(st.item()) != 0:
            raise AssertionError("window envelope <= 1e-11 (NOLA violated)")
