"""Host-side mirror of the reference's log-mel modules, running on the sm_100a library.

Drop-in for (same class names, constructor arguments, attributes, call semantics and error
behaviour; file:line in /root/reference):

* ``TorchMelSpectrogram``      ttts/utils/utils.py:486-522           (V1, 22.05 kHz / 80 mel / power 2)
* ``MelSpectrogramFeatures``   ttts/vocoder/feature_extractors.py:73-99 (V2, 24 kHz / 100 mel / power 1)
* ``MelSpectrogramFeatures1`` / ``mel_spectrogram``  ttts/vocoder/feature_extractors.py:27-53,102-121 (V3, librosa mel)
* ``dvae_wav_to_mel``          ttts/vqvae/xtts_dvae.py:28-49
* ``make_pad_mask``            ttts/utils/utils.py:77-103

plus one additive keyword, ``forward(wav, lengths=...)``, for ragged batches: every clip is
reflect-padded at its own end (the reference processes one clip per call,
ttts/prepare/mel_extract.py:15-27) and the result follows the loaders' collate contract
(zero-padded mels + integer lengths + ``make_pad_mask``, ttts/gpt/dataset.py:95-130).

PyTorch is used for device memory and streams only; the arithmetic is in ``libttts_mel.so``.
CUDA tensors only -- there is no CPU path here.
"""
from __future__ import annotations

import ctypes
import math
import os
from typing import Optional, Sequence, Tuple, Union

import torch
from torch import nn

from . import _cabi
from .filterbank import analysis_window, librosa_mel, melscale_fbanks

DEFAULT_MEL_NORM_FILE = os.path.join(os.path.dirname(os.path.realpath(__file__)), "data", "mel_norms.pth")


def make_pad_mask(lengths: torch.Tensor, max_len: int = 0) -> torch.Tensor:
    """mask[b, t] = t >= lengths[b]  (True = padded part), as ttts/utils/utils.py:77-103."""
    batch_size = lengths.size(0)
    max_len = max_len if max_len > 0 else int(lengths.max().item())
    seq_range = torch.arange(0, max_len, dtype=torch.int64, device=lengths.device)
    return seq_range.unsqueeze(0).expand(batch_size, max_len) >= lengths.unsqueeze(-1)


class MelPlan:
    """One immutable device plan (folded basis, mel table) + the forward call.  One per device."""

    def __init__(self, *, sample_rate: int, n_fft: int, hop_length: int, win_length: Optional[int],
                 n_mels: int, f_min: float, f_max: Optional[float], power: float, norm: Optional[str],
                 mel_scale: str = "htk", normalized: bool = False, clamp: float = 1e-5,
                 center: bool = True, pre_pad: int = 0, power_eps: float = 0.0,
                 fb: Optional[torch.Tensor] = None, device: Union[str, torch.device, int] = "cuda",
                 lab: bool = False):
        self._lib = _cabi.lab() if lab else _cabi.lib()      # lab: tests/tools only (include/ttts_mel_lab.h)
        self.lab = bool(lab)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("MelPlan needs a CUDA device (sm_100a); there is no CPU fallback")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.n_fft, self.hop, self.n_mels = int(n_fft), int(hop_length), int(n_mels)
        self.pad = self.n_fft // 2 if center else int(pre_pad)
        self.power, self.clamp = float(power), float(clamp)
        f_max = float(f_max) if f_max is not None else float(sample_rate // 2)
        if fb is None:
            fb = melscale_fbanks(self.n_fft // 2 + 1, float(f_min), f_max, self.n_mels, int(sample_rate), norm, mel_scale)
        self.fb = fb.to(torch.float32).contiguous()          # [n_freqs, n_mels]
        self.window = analysis_window(self.n_fft, win_length).contiguous()
        spec_scale = 1.0
        if normalized:      # torchaudio functional.py:139-140: spec_f /= window.pow(2).sum().sqrt()
            spec_scale = 1.0 / math.sqrt(float(self.window.pow(2.0).sum()))
        cfg = _cabi.MelConfig(self.n_fft, self.hop, self.n_mels, self.pad, self.power, self.clamp, spec_scale,
                              float(power_eps))
        handle = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _cabi.check(self._lib.ttts_mel_plan_create(ctypes.byref(cfg), self.fb.data_ptr(),
                                                       self.window.data_ptr(), ctypes.byref(handle)))
        self._h = handle
        self.first_bin = int(self._lib.ttts_mel_first_bin(self._h))
        self.num_bins = int(self._lib.ttts_mel_num_bins(self._h))
        self.live_bins = int(self._lib.ttts_mel_live_bins(self._h))
        self.last_launches = 0

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                self._lib.ttts_mel_plan_destroy(h)
            except Exception:
                pass
            self._h = None

    def set_norms(self, norms: Optional[torch.Tensor]) -> None:
        with torch.cuda.device(self.device):
            if norms is None:
                _cabi.check(self._lib.ttts_mel_plan_set_norms(self._h, None))
                return
            host = norms.detach().to("cpu", torch.float32).contiguous()
            if host.numel() != self.n_mels:
                raise ValueError(f"mel_norms has {host.numel()} entries, expected {self.n_mels}")
            # the plan is read by kernels already in flight on this device's streams
            torch.cuda.synchronize(self.device)
            _cabi.check(self._lib.ttts_mel_plan_set_norms(self._h, host.data_ptr()))

    def frames(self, length: int) -> int:
        """torch.stft frame count; raises like F.pad(mode='reflect') when the clip is too short."""
        if length <= self.pad or length + 2 * self.pad < self.n_fft:
            raise RuntimeError(
                f"Argument #4: Padding size should be less than the corresponding input dimension, but got: "
                f"padding ({self.pad}, {self.pad}) at dimension 1 of input of length {length}")
        return 1 + (length + 2 * self.pad - self.n_fft) // self.hop

    def forward(self, wav: torch.Tensor, lengths: Optional[torch.Tensor] = None, *, want_mask: bool = False,
                impl: str = "tcgen05", debug_reim: bool = False, t_max: Optional[int] = None,
                stats_acc: Optional[torch.Tensor] = None):
        """wav [B, L] fp32 CUDA (last dim contiguous).  Returns (mel_btc [B, T, n_mels] fp32,
        n_frames [B] int32, mask [B, T] uint8 | None[, reim]).  ``stats_acc`` (fp64 [2*n_mels+1], device): the
        per-channel statistics of this batch's un-normalised log-mel are added to it inside the same kernel
        (``ttts_mel_forward_stats``)."""
        if wav.device != self.device:
            raise RuntimeError(f"input on {wav.device}, plan on {self.device}")
        if wav.dtype != torch.float32 or wav.dim() != 2:
            raise RuntimeError("MelPlan.forward expects a 2-D float32 tensor")
        if wav.stride(1) != 1 or (wav.shape[0] > 1 and wav.stride(0) < wav.shape[1]):
            wav = wav.contiguous()
        B, Lmax = int(wav.shape[0]), int(wav.shape[1])
        if B == 0:
            raise RuntimeError("empty batch")
        T = self.frames(Lmax) if t_max is None else int(t_max)
        lib = self._lib
        len_dev = None
        if lengths is not None:
            len_dev = lengths.to(device=self.device, dtype=torch.int64).contiguous()
            if len_dev.numel() != B:
                raise ValueError("lengths must have one entry per clip")
        with torch.cuda.device(self.device):
            row_stride = int(wav.stride(0)) if B > 1 else Lmax      # a size-1 batch dimension may carry any stride
            if impl == "tcgen05":       # in-place inputs do not need the padded copy (half the transient memory at 32 x 120 s)
                ws_bytes = int(lib.ttts_mel_workspace_bytes_for(self._h, B, Lmax, wav.data_ptr(), row_stride))
            else:
                ws_bytes = int(lib.ttts_mel_workspace_bytes(self._h, B, Lmax))
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=self.device)
            out = torch.empty((B, T, self.n_mels), dtype=torch.float32, device=self.device)
            n_frames = torch.empty((B,), dtype=torch.int32, device=self.device)
            mask = torch.empty((B, T), dtype=torch.uint8, device=self.device) if want_mask else None
            stream = torch.cuda.current_stream(self.device).cuda_stream
            args = [self._h, wav.data_ptr(), row_stride, len_dev.data_ptr() if len_dev is not None else None,
                    B, Lmax, T, out.data_ptr(), n_frames.data_ptr(), mask.data_ptr() if mask is not None else None,
                    ws.data_ptr(), ws_bytes, stream]
            reim = None
            if debug_reim:
                total = B * self.frames(Lmax)
                reim = torch.zeros((total, 2 * self.num_bins), dtype=torch.float32, device=self.device)
                if not self.lab:
                    raise RuntimeError("the DFT-stage dump exists in the lab build only: MelPlan(..., lab=True)")
                _cabi.check(lib.ttts_mel_forward_debug(*args, reim.data_ptr()))
            elif stats_acc is not None:
                if impl != "tcgen05":
                    raise ValueError("fused statistics run on the tcgen05 kernel only")
                assert stats_acc.dtype == torch.float64 and stats_acc.numel() == 2 * self.n_mels + 1 \
                    and stats_acc.is_contiguous() and stats_acc.device == self.device
                _cabi.check(lib.ttts_mel_forward_stats(*args, stats_acc.data_ptr()))
            elif impl == "tcgen05":
                _cabi.check(lib.ttts_mel_forward(*args))
            elif impl == "simt":
                if not self.lab:
                    raise RuntimeError("the CUDA-core cross-check kernel exists in the lab build only: MelPlan(..., lab=True)")
                _cabi.check(lib.ttts_mel_forward_simt(*args))
            else:
                raise ValueError(f"unknown impl {impl!r}")
            self.last_launches = int(lib.ttts_mel_last_launch_count())
            self._last_ws = ws
        if debug_reim:
            return out, n_frames, mask, reim
        return out, n_frames, mask

    def status(self) -> None:
        """Synchronise and raise if the last forward reported a device-side error."""
        ws = getattr(self, "_last_ws", None)
        if ws is None:
            return
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream(self.device).cuda_stream
            _cabi.check(self._lib.ttts_mel_status(ws.data_ptr(), stream))

    def profile_enable(self, max_calls: int) -> None:
        """Measurement hook (include/ttts_mel_lab.h): time the fused kernel of the next ``max_calls`` forwards of this plan."""
        _cabi.check(self._lib.ttts_mel_profile_enable(self._h, int(max_calls)))

    def profile_collect(self, capacity: int):
        buf = (ctypes.c_float * int(capacity))()
        n = int(self._lib.ttts_mel_profile_collect(self._h, buf, int(capacity)))
        if n < 0:
            _cabi.check(n)
        return [float(buf[i]) for i in range(n)]

    def stats_accumulate(self, logmel_btc: torch.Tensor, n_frames: torch.Tensor, acc: torch.Tensor) -> None:
        """acc[2*n_mels+1] (fp64, device) += [sum | sumsq | count] of un-normalised log-mel."""
        assert logmel_btc.is_contiguous() and logmel_btc.dtype == torch.float32
        assert acc.dtype == torch.float64 and acc.numel() == 2 * self.n_mels + 1 and acc.is_contiguous()
        nf = n_frames.to(device=self.device, dtype=torch.int32).contiguous()
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream(self.device).cuda_stream
            _cabi.check(self._lib.ttts_mel_stats_accumulate(
                self._h, logmel_btc.data_ptr(), nf.data_ptr(), int(logmel_btc.shape[0]), int(logmel_btc.shape[1]),
                acc.data_ptr(), stream))


class GeneralMelPlan:
    """The same call as ``MelPlan.forward`` for transform sizes the fused kernel is not built for (the reference's
    constructor takes any filter_length / hop_length / win_length, utils.py:487-507; every shipped config uses 1024 / 256 and
    runs on ``MelPlan``).  Three launches instead of one fused kernel: per-clip reflect pad (``ttts_reflect_pad``), the
    windowed DFT as the library's dense tcgen05 contraction (``ttts_resample_forward``, 3xTF32, as ``stft.TorchSTFT``),
    and ``ttts_stft_logmel`` (power -> filterbank -> log(clamp) -> / mel_norms).  The spectrum makes one HBM round trip."""

    def __init__(self, *, sample_rate: int, n_fft: int, hop_length: int, win_length: Optional[int], n_mels: int,
                 f_min: float, f_max: Optional[float], power: float, norm: Optional[str], mel_scale: str = "htk",
                 normalized: bool = False, clamp: float = 1e-5, center: bool = True, pre_pad: int = 0,
                 power_eps: float = 0.0, fb: Optional[torch.Tensor] = None,
                 device: Union[str, torch.device, int] = "cuda", lab: bool = False):
        from .stft import _Contraction, windowed_dft_taps
        self._lib = _cabi.lib()
        self.lab = False
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("GeneralMelPlan needs a CUDA device (sm_100a); there is no CPU fallback")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.n_fft, self.hop, self.n_mels = int(n_fft), int(hop_length), int(n_mels)
        self.win_length = int(win_length) if win_length is not None else self.n_fft
        N, H, W = self.n_fft, self.hop, self.win_length
        if N % 8 or W > N or (N - W) % 2 or H > W or (W - H) % 2 or H < 1:
            raise NotImplementedError(f"general log-mel path: need n_fft % 8 == 0, hop <= win_length <= n_fft and "
                                      f"(n_fft - win_length), (win_length - hop) even; got {N} / {H} / {W}")
        self.pad = N // 2 if center else int(pre_pad)
        self.power, self.clamp, self.power_eps = float(power), float(clamp), float(power_eps)
        if self.power not in (1.0, 2.0):
            raise ValueError("power must be 1 or 2")
        f_max = float(f_max) if f_max is not None else float(sample_rate // 2)
        if fb is None:
            fb = melscale_fbanks(N // 2 + 1, float(f_min), f_max, self.n_mels, int(sample_rate), norm, mel_scale)
        self.fb = fb.to(torch.float32).contiguous()          # [n_freqs, n_mels]
        window = torch.hann_window(W, periodic=True, dtype=torch.float32)   # torchaudio's default window_fn
        self.window = window
        w = window.double().numpy()
        if normalized:      # torchaudio functional.py:139-140: spec_f /= window.pow(2).sum().sqrt()
            w = w / math.sqrt(float(window.pow(2.0).sum()))
        # bin pruning (as the fused kernel): only the DFT rows some filter weighs are computed
        live = torch.nonzero((self.fb != 0).any(dim=1)).flatten()
        if live.numel() == 0:
            raise ValueError("filterbank is all zero")
        self.first_bin, kmax = int(live[0]), int(live[-1])
        self.live_bins = self.num_bins = kmax - self.first_bin + 1
        fbl = self.fb[self.first_bin:kmax + 1]               # [live bins, n_mels]
        with torch.cuda.device(self.device):
            taps = windowed_dft_taps(w, N)[2 * self.first_bin:2 * (kmax + 1)]
            self._con = _Contraction(taps, H, (W - H) // 2, self.device)
            nz = fbl != 0
            any_nz = nz.any(dim=0)
            k_idx = torch.arange(fbl.shape[0])
            lo = torch.where(nz, k_idx[:, None], fbl.shape[0]).min(dim=0).values
            hi = torch.where(nz, k_idx[:, None], -1).max(dim=0).values
            lo = torch.where(any_nz, lo, torch.ones_like(lo))
            hi = torch.where(any_nz, hi, torch.zeros_like(hi))
            self._krange = torch.stack([lo, hi], dim=1).to(torch.int32).contiguous().to(self.device)
            self._fbT = fbl.t().contiguous().to(self.device)
        self._norms: Optional[torch.Tensor] = None
        self.last_launches = 0

    def set_norms(self, norms: Optional[torch.Tensor]) -> None:
        if norms is None:
            self._norms = None
            return
        if norms.numel() != self.n_mels:
            raise ValueError(f"mel_norms has {norms.numel()} entries, expected {self.n_mels}")
        self._norms = norms.detach().to(device=self.device, dtype=torch.float32).contiguous().clone()

    def frames(self, length: int) -> int:
        if length <= self.pad or length + 2 * self.pad < self.n_fft:
            raise RuntimeError(
                f"Argument #4: Padding size should be less than the corresponding input dimension, but got: "
                f"padding ({self.pad}, {self.pad}) at dimension 1 of input of length {length}")
        return 1 + (length + 2 * self.pad - self.n_fft) // self.hop

    def forward(self, wav: torch.Tensor, lengths: Optional[torch.Tensor] = None, *, want_mask: bool = False,
                impl: str = "tcgen05", debug_reim: bool = False, t_max: Optional[int] = None,
                stats_acc: Optional[torch.Tensor] = None):
        from .stft import framed_contraction
        if impl != "tcgen05" or debug_reim or stats_acc is not None:
            raise NotImplementedError("the general-shape path has the tcgen05 contraction only (no cross-check kernel, "
                                      "no DFT-stage dump, no fused statistics)")
        if wav.device != self.device:
            raise RuntimeError(f"input on {wav.device}, plan on {self.device}")
        if wav.dtype != torch.float32 or wav.dim() != 2:
            raise RuntimeError("GeneralMelPlan.forward expects a 2-D float32 tensor")
        if wav.stride(1) != 1 or (wav.shape[0] > 1 and wav.stride(0) < wav.shape[1]):
            wav = wav.contiguous()
        B, Lmax = int(wav.shape[0]), int(wav.shape[1])
        if B == 0:
            raise RuntimeError("empty batch")
        T_full = self.frames(Lmax)
        T = T_full if t_max is None else int(t_max)
        lib = self._lib
        N, H, W = self.n_fft, self.hop, self.win_length
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream(self.device).cuda_stream
            len_dev = None
            if lengths is not None:
                len_dev = lengths.to(device=self.device, dtype=torch.int64).contiguous()
                if len_dev.numel() != B:
                    raise ValueError("lengths must have one entry per clip")
                nf64 = 1 + torch.div(len_dev + (2 * self.pad - N), H, rounding_mode="floor")
                ok = (len_dev > self.pad) & (len_dev + 2 * self.pad >= N) & (len_dev <= Lmax)
                n_frames = torch.where(ok, nf64, torch.zeros_like(nf64)).clamp_(max=T).to(torch.int32)
            else:
                n_frames = torch.full((B,), min(T_full, T), dtype=torch.int32, device=self.device)
            y, d, Trows = framed_contraction(lib, self._con, wav, len_dev, N, H, W, self.pad, self.device)
            out = torch.empty((B, T, self.n_mels), dtype=torch.float32, device=self.device)
            nb = self.live_bins
            _cabi.check(lib.ttts_stft_logmel(y.data_ptr(), int(y.stride(0)), d, 2 * nb, B, min(Trows, T),
                                             n_frames.data_ptr(), T, nb, self._fbT.data_ptr(), self._krange.data_ptr(),
                                             self.n_mels, self.power, self.power_eps, self.clamp,
                                             self._norms.data_ptr() if self._norms is not None else None,
                                             out.data_ptr(), stream))
            mask = None
            if want_mask:
                mask = (torch.arange(T, device=self.device)[None, :] >= n_frames[:, None]).to(torch.uint8)
            self.last_launches = 4
        return out, n_frames, mask

    def status(self) -> None:
        """Nothing device-side to report: every length is checked where the frame counts are computed."""
        torch.cuda.synchronize(self.device)


def make_plan(**kwargs):
    """``MelPlan`` (one fused kernel) for the transform size it is built for, ``GeneralMelPlan`` for every other one."""
    if int(kwargs["n_fft"]) == 1024 and int(kwargs["hop_length"]) == 256:
        return MelPlan(**kwargs)
    return GeneralMelPlan(**kwargs)


class _MelModuleBase(nn.Module):
    """Shared forward logic of the two reference modules."""

    def _plan_kwargs(self) -> dict:
        raise NotImplementedError

    def _norms(self) -> Optional[torch.Tensor]:
        return None

    def _plan_for(self, device: torch.device):
        plans = self.__dict__.setdefault("_plans", {})
        key = (device.type, device.index if device.index is not None else torch.cuda.current_device())
        plan = plans.get(key)
        if plan is None:
            plan = make_plan(device=torch.device(*key), **self._plan_kwargs())
            plans[key] = plan
            plan._norms_key = None
        norms = self._norms()
        nkey = None if norms is None else (id(norms), norms._version, norms.data_ptr())
        if nkey != plan._norms_key:
            plan.set_norms(norms)
            plan._norms_key = nkey
        return plan

    # The cached plans hold library handles (ctypes pointers): a copy / pickle of the module (EMA copies, torch.save(model),
    # spawn workers -- the reference module supports all of them) drops them and rebuilds lazily.
    def __getstate__(self):
        state = self.__dict__.copy()
        state.pop("_plans", None)
        return state

    def __deepcopy__(self, memo):
        import copy
        new = self.__class__.__new__(self.__class__)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            if k != "_plans":
                new.__dict__[k] = copy.deepcopy(v, memo)
        return new

    def status(self, device=None) -> None:
        """Synchronise and raise if the last forward on ``device`` reported a device-side error.  With ``lengths`` given
        as a CUDA tensor nothing is validated on the host (no sync in ``forward``): a clip that is too short or longer
        than the batch width yields all-zero rows and is reported here (the reference raises inside its call)."""
        for key, plan in self.__dict__.get("_plans", {}).items():
            if device is None or torch.device(*key) == torch.device(device):
                plan.status()

    def _run(self, inp: torch.Tensor, lengths=None, impl: str = "tcgen05"):
        if len(inp.shape) == 3:   # squeeze the channel dimension (mono), utils.py:510-511
            inp = inp.squeeze(1)
        assert len(inp.shape) == 2
        if not inp.is_cuda:
            raise RuntimeError(f"{type(self).__name__} (xtts_b200) runs on CUDA sm_100a only; got a {inp.device} tensor "
                               "and there is no CPU fallback")
        if inp.dtype == torch.float64:
            raise RuntimeError("expected scalar type Float but found Double")   # torch.stft's complaint
        if inp.dtype != torch.float32:
            inp = inp.float()       # fp16/bf16 promote to an fp32 result in the reference too
        plan = self._plan_for(inp.device)
        if lengths is None:
            mel, _, _ = plan.forward(inp, impl=impl)
            return mel.transpose(1, 2)          # [B, n_mels, T] view with the reference's strides
        lens = torch.as_tensor(lengths)
        if lens.dim() != 1 or lens.numel() != inp.shape[0]:
            raise ValueError("lengths must be 1-D with one entry per clip")
        if not lens.is_cuda:
            lmin, lmax = int(lens.min()), int(lens.max())
            if lmax > inp.shape[1]:
                raise ValueError(f"a length ({lmax}) exceeds the padded width {inp.shape[1]}")
            plan.frames(lmin)                    # raises for clips too short to reflect-pad
            t_max = plan.frames(lmax)
        else:
            t_max = None
        mel, n_frames, mask = plan.forward(inp, lens, want_mask=True, impl=impl, t_max=t_max)
        return mel.transpose(1, 2), n_frames.to(torch.int64), mask.bool()


class TorchMelSpectrogram(_MelModuleBase):
    """ttts/utils/utils.py:486-522 on the B200 library."""

    def __init__(self, filter_length=1024, hop_length=256, win_length=1024, n_mel_channels=80, mel_fmin=0,
                 mel_fmax=8000, sampling_rate=22050, normalize=False, mel_norm_file=DEFAULT_MEL_NORM_FILE):
        super().__init__()
        self.filter_length = filter_length
        self.hop_length = hop_length
        self.win_length = win_length
        self.n_mel_channels = n_mel_channels
        self.mel_fmin = mel_fmin
        self.mel_fmax = mel_fmax
        self.sampling_rate = sampling_rate
        self.normalize = normalize
        self.mel_stft = None            # the reference keeps a torchaudio transform here; nothing to keep
        self.mel_norm_file = mel_norm_file
        if self.mel_norm_file is not None:
            self.mel_norms = torch.load(self.mel_norm_file)
        else:
            self.mel_norms = None

    def _plan_kwargs(self) -> dict:
        return dict(sample_rate=self.sampling_rate, n_fft=self.filter_length, hop_length=self.hop_length,
                    win_length=self.win_length, n_mels=self.n_mel_channels, f_min=self.mel_fmin,
                    f_max=self.mel_fmax, power=2.0, norm="slaney", mel_scale="htk",
                    normalized=bool(self.normalize), clamp=1e-5, center=True)

    def _norms(self):
        return self.mel_norms

    def forward(self, inp, lengths=None):
        return self._run(inp, lengths)


class MelSpectrogramFeatures(_MelModuleBase):
    """ttts/vocoder/feature_extractors.py:73-99 (+ safe_log, ttts/vocoder/modules.py:194-205)."""

    def __init__(self, sample_rate=24000, n_fft=1024, hop_length=256, win_length=None, n_mels=100, mel_fmin=0,
                 mel_fmax=None, normalize=False, padding="center"):
        super().__init__()
        if padding not in ["center", "same"]:
            raise ValueError("Padding must be 'center' or 'same'.")
        self.padding = padding
        self.sample_rate, self.n_fft, self.hop_length = sample_rate, n_fft, hop_length
        self.win_length = win_length if win_length is not None else n_fft
        self.n_mels, self.mel_fmin, self.mel_fmax, self.normalize = n_mels, mel_fmin, mel_fmax, normalize

    def _plan_kwargs(self) -> dict:
        same = self.padding == "same"
        return dict(sample_rate=self.sample_rate, n_fft=self.n_fft, hop_length=self.hop_length,
                    win_length=self.win_length, n_mels=self.n_mels, f_min=self.mel_fmin, f_max=self.mel_fmax,
                    power=1.0, norm=None, mel_scale="htk", normalized=bool(self.normalize), clamp=1e-7,
                    center=not same, pre_pad=(self.win_length - self.hop_length) // 2 if same else 0)

    def forward(self, audio, lengths=None, **kwargs):
        return self._run(audio, lengths)


_V3_PLANS = {}


def mel_spectrogram(y, n_fft, num_mels, sampling_rate, hop_size, win_size, fmin, fmax, center=False):
    """ttts/vocoder/feature_extractors.py:27-53: reflect-pad (n_fft-hop)/2, STFT with center=False,
    sqrt(re^2+im^2+1e-9), librosa (Slaney) mel basis, log(clamp(., 1e-5)).  Returns [B, num_mels, L // hop]."""
    if center:
        raise NotImplementedError("mel_spectrogram(center=True) is not used by the reference and is not built")
    if not y.is_cuda:
        raise RuntimeError("mel_spectrogram (xtts_b200) runs on CUDA sm_100a only; there is no CPU fallback")
    dev = y.device if y.device.index is not None else torch.device("cuda", torch.cuda.current_device())
    key = (n_fft, num_mels, sampling_rate, hop_size, win_size, fmin, fmax, dev.index)
    plan = _V3_PLANS.get(key)
    if plan is None:
        fb = librosa_mel(sampling_rate, n_fft, num_mels, fmin, fmax).t().contiguous()
        plan = _V3_PLANS[key] = make_plan(sample_rate=sampling_rate, n_fft=n_fft, hop_length=hop_size, win_length=win_size,
                                       n_mels=num_mels, f_min=fmin, f_max=fmax, power=1.0, norm=None, clamp=1e-5,
                                       center=False, pre_pad=int((n_fft - hop_size) / 2), power_eps=1e-9, fb=fb, device=dev)
    if y.dim() == 3:
        y = y.squeeze(1)
    mel, _, _ = plan.forward(y.float())
    return mel.transpose(1, 2)


class MelSpectrogramFeatures1(nn.Module):
    """ttts/vocoder/feature_extractors.py:102-121 (selected by mel_type == "librosa", ttts/vqvae/dataset.py:27-28)."""

    def __init__(self, sample_rate=24000, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, mel_fmin=0,
                 mel_fmax=8000, normalize=False, padding="center"):
        super().__init__()
        if padding not in ["center", "same"]:
            raise ValueError("Padding must be 'center' or 'same'.")
        self.padding = padding
        self.sampling_rate, self.n_fft, self.num_mels = sample_rate, n_fft, n_mels
        self.hop_size, self.win_size, self.fmin, self.fmax = hop_length, win_length, mel_fmin, mel_fmax

    def forward(self, audio, **kwargs):
        return mel_spectrogram(audio, self.n_fft, self.num_mels, self.sampling_rate, self.hop_size, self.win_size,
                               self.fmin, self.fmax, center=False)


_DVAE_MODULES = {}


def dvae_wav_to_mel(wav, mel_norms_file="../experiments/clips_mel_norms.pth", mel_norms=None,
                    device=torch.device("cuda")):
    """ttts/vqvae/xtts_dvae.py:28-49 (the reference rebuilds its transform every call; the plan is cached)."""
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("dvae_wav_to_mel (xtts_b200) runs on CUDA only; there is no CPU fallback")
    mod = _DVAE_MODULES.get("v1")
    if mod is None:
        mod = _DVAE_MODULES["v1"] = TorchMelSpectrogram(mel_norm_file=None)
    if mel_norms is None:
        mel_norms = torch.load(mel_norms_file, map_location=device)
    mod.mel_norms = mel_norms
    return mod(wav.to(device))
