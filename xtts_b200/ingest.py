"""Host-side mirror of the reference's audio ingestion in front of the log-mel path.

Reference (file:line in /root/reference):

* ``load_audio``            ttts/utils/utils.py:32-47   torchaudio.load -> channel 0 -> functional.resample -> clip_(-1, 1)
* ``process_mel`` loop body ttts/prepare/mel_extract.py:15-21   torchaudio.load -> channel 0 -> transforms.Resample

The file decode (``torchaudio.load``: container parsing, host I/O) stays with the caller; what arrives here
is what a 16-bit PCM WAV holds -- interleaved int16 samples -- already in HBM.  ``decode_pcm16`` is the
``int16 / 32768`` + channel-0 step, ``Resampler`` the polyphase windowed-sinc FIR torchaudio applies
(``sinc_interp_hann``, lowpass_filter_width 6, rolloff 0.99: ``functional.resample`` defaults), both as sm_100a
kernels behind ``include/ttts_ingest.h``.  CUDA tensors only -- there is no CPU path here.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple, Union

import torch

from . import _cabi


def sinc_resample_taps(orig_freq: int, new_freq: int, lowpass_filter_width: int = 6, rolloff: float = 0.99,
                       dtype: Optional[torch.dtype] = torch.float32) -> Tuple[torch.Tensor, int, int, int]:
    """The polyphase FIR of torchaudio's ``_get_sinc_resample_kernel`` (``sinc_interp_hann``), built with the
    same torch ops on the CPU so that the taps are bit-identical to the reference's call.

    ``dtype=torch.float32`` is what ``torchaudio.functional.resample`` passes for a float32 waveform
    (utils.py:41); ``dtype=None`` is ``torchaudio.transforms.Resample`` (mel_extract.py:19): taps evaluated in
    float64, then cast.  Returns (taps [new, 2*width + orig] float32, orig, new, width) with both rates
    divided by their gcd.
    """
    if int(orig_freq) != orig_freq or int(new_freq) != new_freq:
        raise Exception("Frequencies must be of integer type to ensure quality resampling computation.")
    if orig_freq <= 0.0 or new_freq <= 0.0:
        raise ValueError("Original frequency and desired frequecy should be positive")
    if lowpass_filter_width <= 0:
        raise ValueError("Low pass filter width should be positive.")
    g = math.gcd(int(orig_freq), int(new_freq))
    orig, new = int(orig_freq) // g, int(new_freq) // g
    cutoff = min(orig, new) * rolloff
    width = math.ceil(lowpass_filter_width * orig / cutoff)
    pos_dtype = dtype if dtype is not None else torch.float64
    # tap k of phase p sits at time (k - width) / orig - p / new (in input-sample periods / orig)
    pos = torch.arange(-width, width + orig, dtype=pos_dtype)[None, None] / orig
    t = torch.arange(0, -new, -1, dtype=dtype)[:, None, None] / new + pos
    t *= cutoff
    t = t.clamp_(-lowpass_filter_width, lowpass_filter_width)
    hann = torch.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t *= math.pi
    taps = torch.where(t == 0, torch.tensor(1.0).to(t), t.sin() / t)
    taps *= hann * (cutoff / orig)
    if dtype is None:
        taps = taps.to(dtype=torch.float32)
    return taps.reshape(new, 2 * width + orig).contiguous(), orig, new, width


def _resolve(device: Union[str, torch.device, int]) -> torch.device:
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("the ingestion kernels need a CUDA device (sm_100a); there is no CPU fallback")
    if device.index is None:
        device = torch.device("cuda", torch.cuda.current_device())
    return device


def _check_pcm(pcm: torch.Tensor, frames: Optional[torch.Tensor], want_device: Optional[torch.device] = None):
    if pcm.dtype != torch.int16:
        raise TypeError(f"expected int16 PCM, got {pcm.dtype}")
    if pcm.dim() == 1:
        pcm = pcm[None, :, None]
    elif pcm.dim() == 2:
        pcm = pcm[None]
    elif pcm.dim() != 3:
        raise RuntimeError("pcm must be [F], [F, C] or [B, F, C]")
    device = _resolve(pcm.device)
    if want_device is not None and device != want_device:
        raise RuntimeError(f"input on {device}, plan on {want_device}")
    pcm = pcm.contiguous()
    B, F, C = (int(s) for s in pcm.shape)
    if B == 0 or F == 0 or C == 0:
        raise RuntimeError("empty PCM batch")
    fr = None
    if frames is not None:
        fr = frames.to(device=device, dtype=torch.int64).contiguous()
        if fr.numel() != B:
            raise ValueError("frames must have one entry per clip")
    return pcm, B, F, C, fr


def decode_pcm16(pcm: torch.Tensor, frames: Optional[torch.Tensor] = None) -> torch.Tensor:
    """int16 PCM -> float32 channel 0, as ``torchaudio.load`` + ``audio[0]`` (utils.py:33-37).

    pcm: CUDA int16, ``[F]`` / ``[F, C]`` (one clip, WAV sample order) or ``[B, F, C]``; ``frames`` [B] holds the
    valid frames of each clip of a ragged batch (the rest of the row is zero-filled).  Returns ``[B, F]``
    (``[1, F]`` for one clip, the reference's shape).
    """
    pcm, B, F, C, fr = _check_pcm(pcm, frames)
    device = pcm.device
    lib = _cabi.lib()
    with torch.cuda.device(device):
        # rows on a 16-byte pitch (a view of [B, pitch]): the log-mel kernel then reads the batch in place instead of
        # copying it into its padded workspace first
        pitch = (F + 3) // 4 * 4
        out = torch.empty((B, pitch), dtype=torch.float32, device=device)[:, :F]
        stream = torch.cuda.current_stream(device).cuda_stream
        _cabi.check(lib.ttts_pcm16_to_f32(pcm.data_ptr(), F * C, C, fr.data_ptr() if fr is not None else None,
                                          B, F, out.data_ptr(), pitch, stream))
    return out


class Resampler:
    """``torchaudio.functional.resample(x, orig_freq, new_freq)`` (``kernel_dtype=torch.float32``) or
    ``torchaudio.transforms.Resample(orig_freq, new_freq)`` (``kernel_dtype=None``) for a batch of clips.
    One immutable device plan per (rates, device)."""

    def __init__(self, orig_freq: int, new_freq: int, *, lowpass_filter_width: int = 6, rolloff: float = 0.99,
                 kernel_dtype: Optional[torch.dtype] = torch.float32, device: Union[str, torch.device, int] = "cuda"):
        self.device = _resolve(device)
        g = math.gcd(int(orig_freq), int(new_freq))
        o, n = int(orig_freq) // g, int(new_freq) // g
        w = math.ceil(lowpass_filter_width * o / (min(o, n) * rolloff))
        if n * (2 * w + o) > (64 << 20):
            raise ValueError(f"resample {orig_freq} -> {new_freq}: {n} phases x {2 * w + o} taps do not fit a table "
                             "(rates that share no large common divisor are not supported)")
        taps, self.orig, self.new, self.width = sinc_resample_taps(orig_freq, new_freq, lowpass_filter_width,
                                                                   rolloff, kernel_dtype)
        self.taps = taps
        self._lib = _cabi.lib()          # the library this plan lives in (tests may select the lab build)
        self._h = None
        import ctypes
        h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _cabi.check(self._lib.ttts_resample_plan_create(taps.data_ptr(), self.orig, self.new, self.width,
                                                              ctypes.byref(h)))
        self._h = h

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.ttts_resample_plan_destroy(h)
            except Exception:
                pass

    def out_len(self, length: int) -> int:
        """``ceil(new * length / orig)``: torchaudio's ``target_length``."""
        return (int(length) * self.new + self.orig - 1) // self.orig

    def __call__(self, x: torch.Tensor, lengths: Optional[torch.Tensor] = None, *, clip: bool = False):
        """x ``[B, L]`` (or ``[L]``) float32 CUDA.  Returns y ``[B, L']`` -- and the per-clip output lengths when
        ``lengths`` is given (each clip is resampled as if alone: zero padding starts at its own end)."""
        if not x.is_floating_point():
            raise TypeError(f"Expected floating point type for waveform tensor, but received {x.dtype}.")
        if x.dtype != torch.float32:
            raise TypeError("the resample kernel is float32 (the reference's waveform dtype)")
        if x.device != self.device:
            raise RuntimeError(f"input on {x.device}, plan on {self.device}")
        squeeze = x.dim() == 1
        if squeeze:
            x = x[None]
        if x.dim() != 2:
            raise RuntimeError("expected [B, L] or [L]")
        if x.stride(1) != 1 or (x.shape[0] > 1 and x.stride(0) < x.shape[1]):
            x = x.contiguous()
        B, L = int(x.shape[0]), int(x.shape[1])
        if B == 0 or L == 0:
            raise RuntimeError("empty batch")
        len_dev = None
        if lengths is not None:
            len_dev = lengths.to(device=self.device, dtype=torch.int64).contiguous()
            if len_dev.numel() != B:
                raise ValueError("lengths must have one entry per clip")
        out = self._alloc(B, self.out_len(L))
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream(self.device).cuda_stream
            row_stride = int(x.stride(0)) if B > 1 else L
            _cabi.check(self._lib.ttts_resample_forward(
                self._h, x.data_ptr(), row_stride, len_dev.data_ptr() if len_dev is not None else None, B, L,
                out.data_ptr(), self._row_stride(out), 1 if clip else 0, stream))
        if squeeze:
            out = out[0]
        if len_dev is not None:
            return out, (len_dev * self.new + (self.orig - 1)) // self.orig
        return out

    def _alloc(self, B: int, Lo: int) -> torch.Tensor:
        """[B, Lo] float32 whose rows start 16-byte aligned (row pitch padded to a multiple of 4 floats)."""
        pitch = (Lo + 3) // 4 * 4 if B > 1 else Lo
        with torch.cuda.device(self.device):
            return torch.empty((B, pitch), dtype=torch.float32, device=self.device)[:, :Lo]

    @staticmethod
    def _row_stride(out: torch.Tensor) -> int:
        return int(out.stride(0)) if out.shape[0] > 1 else int(out.shape[1])

    def from_pcm16(self, pcm: torch.Tensor, frames: Optional[torch.Tensor] = None, *, clip: bool = False):
        """Decode and resample in one kernel: ``pcm`` int16 CUDA ``[B, F, C]`` (channel 0 is read, x / 32768 while
        staging), same result as ``self(decode_pcm16(pcm, frames), frames)`` bit for bit."""
        pcm, B, F, C, fr = _check_pcm(pcm, frames, self.device)
        out = self._alloc(B, self.out_len(F))
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream(self.device).cuda_stream
            _cabi.check(self._lib.ttts_resample_pcm16_forward(
                self._h, pcm.data_ptr(), F * C, C, fr.data_ptr() if fr is not None else None, B, F,
                out.data_ptr(), self._row_stride(out), 1 if clip else 0, stream))
        if fr is not None:
            return out, (fr * self.new + (self.orig - 1)) // self.orig
        return out


_resamplers: dict = {}


def _resampler_for(sr: int, sampling_rate: int, device: torch.device, kernel_dtype) -> Resampler:
    key = (int(sr), int(sampling_rate), device, kernel_dtype)
    r = _resamplers.get(key)
    if r is None:
        r = _resamplers[key] = Resampler(sr, sampling_rate, kernel_dtype=kernel_dtype, device=device)
    return r


def load_audio_pcm16(pcm: torch.Tensor, sr: int, sampling_rate: int, frames: Optional[torch.Tensor] = None,
                     *, kernel_dtype: Optional[torch.dtype] = torch.float32):
    """``load_audio`` (utils.py:32-47) from the decoded PCM onwards: channel 0, resample to ``sampling_rate``
    when ``sr`` differs, clip to [-1, 1].  One clip (``[F]`` / ``[F, C]``) returns ``[1, L']`` like the
    reference; a batch ``[B, F, C]`` returns ``[B, L']`` -- plus the output lengths when ``frames`` is given."""
    out_len = None
    if int(sr) != int(sampling_rate):
        rs = _resampler_for(sr, sampling_rate, _resolve(pcm.device), kernel_dtype)
        if frames is not None:
            wav, out_len = rs.from_pcm16(pcm, frames, clip=True)
        else:
            wav = rs.from_pcm16(pcm, clip=True)
    else:
        wav = decode_pcm16(pcm, frames)
        wav.clip_(-1, 1)      # int16 / 32768 is already inside [-1, 1): kept for the contract
        if frames is not None:
            out_len = frames.to(device=wav.device, dtype=torch.int64)
    if frames is not None:
        return wav, out_len
    return wav


def get_prompt_slice(audio: torch.Tensor, max_audio_length: float = 20, min_audio_length: float = 3,
                     sample_rate: int = 24000, is_eval: bool = False) -> torch.Tensor:
    """Conditioning-prompt crop of ``[C, L]`` audio (utils.py:50-74), a view -- works on device tensors as is.

    Training: a length drawn uniformly from [min, max] seconds (half the clip when the clip is shorter than the
    draw), at a uniformly drawn start; eval: the mid length (the whole clip when shorter) from sample 0.  Draws
    come from Python's ``random`` in the reference's order (length, then start), so a seeded run crops identically.
    """
    import random
    longest, shortest = max_audio_length * sample_rate, min_audio_length * sample_rate
    total = audio.shape[-1]
    want = int((shortest + longest) / 2) if is_eval else random.randint(shortest, longest)
    if want > total:
        want = total if is_eval else total // 2
    start = 0 if is_eval else random.randint(0, total - want)
    return audio[:, start:start + want]
