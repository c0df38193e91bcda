"""Ingestion step: int16 PCM -> float32 channel 0 -> sinc resample -> clip.

CPU part: the oracle against the golden vectors minted from the reference's ``load_audio``
(oracle/gen_golden_ingest.py), the fp64 restatement against both, the host mirror's taps.
GPU part (``-m gpu``): the CUDA kernels through the C ABI against the oracle and the goldens.
Tolerance: the conversion is exact (x * 2^-15); the FIR sums <= 475 fp32 products of |x| <= 1 samples in a
different order than conv1d does, so 2e-6 absolute on O(1) samples (measured: ~3e-7).
"""
import os

import numpy as np
import pytest
import torch

from oracle import ingest_oracle as ing
from oracle.gen_golden_ingest import CASES

RESAMPLE_TOL = 2e-6      # CUDA kernel vs the reference's call (same fp32 taps, different summation order)
TRUTH_TOL = 5e-5         # either of them vs the fp64 restatement: torchaudio evaluates the tap positions in fp32
                         # (argument up to 6 pi x cutoff at fp32 spacing), measured 1.6e-5 for 320 -> 441


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "ingest_load_audio.npz"))


def test_golden_file_matches_case_table(gold):
    assert [str(n) for n in gold["names"]] == [c[0] for c in CASES]
    assert "utils.py:32-47" in str(gold["reference"])


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_callsite_oracle_reproduces_reference_load_audio(gold, case):
    name, sr, target, frames, ch, seed, kind = case
    pcm = ing.synth_pcm16(frames, ch, seed, kind)
    y = ing.callsite_load_audio(pcm, sr, target)
    g = gold["y_" + name]
    assert y.shape == g.shape == (1, ing.out_len(frames, sr, target))
    assert np.array_equal(y, g)
    assert np.abs(y).max() <= 1.0


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_fp64_restatement_agrees_with_reference(gold, case):
    name, sr, target, frames, ch, seed, kind = case
    pcm = ing.synth_pcm16(frames, ch, seed, kind)
    x = ing.pcm16_to_float(pcm)
    y = x.astype(np.float64) if sr == target else ing.restated_resample(x, sr, target)
    y = np.clip(y, -1.0, 1.0)
    assert np.abs(y - gold["y_" + name]).max() <= TRUTH_TOL


def test_wav_round_trip_and_load_stand_in(tmp_path):
    pcm = ing.synth_pcm16(1000, 2, 5)
    p = str(tmp_path / "a.wav")
    ing.write_wav(p, pcm, 44100)
    back, sr = ing.read_wav(p)
    assert sr == 44100 and np.array_equal(back, pcm)
    t, sr2 = ing.torchaudio_load_pcm16(p)
    assert t.shape == (2, 1000) and t.dtype == torch.float32 and sr2 == 44100
    assert np.array_equal(t[0].numpy(), ing.pcm16_to_float(pcm)[0])
    assert t.min() >= -1.0 and t.max() < 1.0


@pytest.mark.parametrize("rates", [(44100, 22050), (48000, 24000), (16000, 22050), (22050, 24000), (44100, 24000),
                                   (8000, 24000), (24000, 22050)])
def test_host_mirror_taps_are_torchaudios(rates):
    import math
    import torchaudio.functional.functional as F
    from xtts_b200 import sinc_resample_taps
    o, n = rates
    g = math.gcd(o, n)
    for dtype in (torch.float32, None):
        k, w = F._get_sinc_resample_kernel(o, n, g, dtype=dtype)
        taps, orig, new, width = sinc_resample_taps(o, n, dtype=dtype)
        assert (orig, new, width) == (o // g, n // g, w)
        assert taps.dtype == torch.float32 and taps.shape == (new, 2 * width + orig)
        assert torch.equal(taps, k.reshape(taps.shape))
    ref, *_ = ing.restated_taps(o, n)
    # even the "fp64" taps of transforms.Resample carry an fp32-rounded phase offset (int64 arange / new_freq is a
    # float32 division in torch): up to 1.3e-5 from the exact windowed sinc at 320 -> 441
    assert np.abs(taps.numpy() - ref).max() <= 2e-5
    # each phase is a unit-DC-gain low-pass to within the window's ripple
    assert np.abs(ref.sum(axis=1) - 1.0).max() < 5e-3


def test_host_mirror_argument_errors():
    from xtts_b200 import decode_pcm16, sinc_resample_taps, Resampler
    with pytest.raises(Exception, match="integer"):
        sinc_resample_taps(44100, 5512.5)
    with pytest.raises(ValueError):
        sinc_resample_taps(44100, 22050, lowpass_filter_width=0)
    with pytest.raises(TypeError):
        decode_pcm16(torch.zeros(10, dtype=torch.float32))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        decode_pcm16(torch.zeros(10, dtype=torch.int16))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Resampler(44100, 22050, device="cpu")


def test_get_prompt_slice_crops_like_the_reference(gold):
    import random
    from xtts_b200.ingest import get_prompt_slice
    crops = gold["prompt_crops"]
    assert len(crops) == 60
    for L, ev, seed, first, length in crops.tolist():
        random.seed(seed)
        c = get_prompt_slice(torch.arange(L, dtype=torch.float32)[None], is_eval=bool(ev))
        assert c.shape == (1, length) and (length == 0 or int(c[0, 0]) == first)
    a = torch.zeros(2, 100000)
    assert get_prompt_slice(a, is_eval=True).shape == (2, 100000)           # shorter than the mid length: whole clip
    assert get_prompt_slice(torch.zeros(1, 700000), is_eval=True).shape == (1, 276000)


def test_out_len_is_ceil():
    for o, n in [(44100, 22050), (16000, 22050), (44100, 24000)]:
        for L in (1, 2, 3, 319, 320, 321, 44099, 44100, 44101):
            x = np.zeros((1, L), dtype=np.float32)
            assert ing.restated_resample(x, o, n).shape[1] == ing.out_len(L, o, n)
            y = ing.callsite_load_audio(np.zeros((L, 1), dtype=np.int16), o, n)
            assert y.shape[1] == ing.out_len(L, o, n)


# ------------------------------------------------------------------------------------------------ GPU parity

gpu = pytest.mark.gpu


@gpu
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_gpu_load_audio_matches_reference_golden(gold, case):
    from xtts_b200 import load_audio_pcm16
    name, sr, target, frames, ch, seed, kind = case
    pcm = ing.synth_pcm16(frames, ch, seed, kind)
    y = load_audio_pcm16(torch.from_numpy(pcm).cuda(), sr, target)
    g = gold["y_" + name]
    assert tuple(y.shape) == g.shape and y.dtype == torch.float32
    err = np.abs(y.cpu().numpy() - g).max()
    if sr == target:
        assert err == 0.0
    else:
        assert err <= RESAMPLE_TOL, err


@gpu
@pytest.mark.parametrize("channels", [1, 2, 3, 6])
@pytest.mark.parametrize("frames", [1, 3, 4, 5, 1023, 4096, 100003])
def test_gpu_decode_pcm16_is_exact(channels, frames):
    from xtts_b200 import decode_pcm16
    pcm = ing.synth_pcm16(frames, channels, 100 + frames + channels, "uniform")
    pcm[0, 0] = -32768
    pcm[-1, 0] = 32767
    y = decode_pcm16(torch.from_numpy(pcm).cuda())
    assert np.array_equal(y.cpu().numpy(), ing.pcm16_to_float(pcm))
    if channels == 1:
        y1 = decode_pcm16(torch.from_numpy(pcm[:, 0].copy()).cuda())
        assert torch.equal(y1, y)


@gpu
def test_gpu_decode_pcm16_ragged_batch_zero_fills():
    from xtts_b200 import decode_pcm16
    B, F, C = 5, 3001, 2
    frames = [3001, 1, 0, 2999, 1500]
    pcm = np.stack([ing.synth_pcm16(F, C, 200 + b, "uniform") for b in range(B)])
    y = decode_pcm16(torch.from_numpy(pcm).cuda(), torch.tensor(frames)).cpu().numpy()
    for b, n in enumerate(frames):
        assert np.array_equal(y[b, :n], ing.pcm16_to_float(pcm[b])[0, :n])
        assert not y[b, n:].any()


@gpu
@pytest.mark.parametrize("rates", [(44100, 22050), (48000, 24000), (16000, 22050), (22050, 24000), (44100, 24000),
                                   (8000, 24000), (24000, 22050), (44100, 16000), (48000, 16000), (32000, 8000),
                                   (48000, 22050), (16000, 24000), (25, 16), (16, 25), (50, 33), (7, 5), (5, 7),
                                   (1000, 999)])
@pytest.mark.parametrize("kernel_dtype", [torch.float32, None], ids=["functional", "transform"])
def test_gpu_resampler_matches_torchaudio_call(rates, kernel_dtype):
    """Dense batch against the reference's two call forms: functional.resample (utils.py:41) and
    transforms.Resample (mel_extract.py:19)."""
    import torchaudio
    from xtts_b200 import Resampler
    o, n = rates
    rng = np.random.default_rng(o + n)
    x = rng.uniform(-1, 1, (3, 20011)).astype(np.float32)
    rs = Resampler(o, n, kernel_dtype=kernel_dtype)
    y = rs(torch.from_numpy(x).cuda()).cpu().numpy()
    if kernel_dtype is None:
        ref = torchaudio.transforms.Resample(o, n)(torch.from_numpy(x)).numpy()
    else:
        ref = torchaudio.functional.resample(torch.from_numpy(x), o, n).numpy()
    assert y.shape == ref.shape == (3, ing.out_len(20011, o, n))
    assert np.abs(y - ref).max() <= RESAMPLE_TOL
    truth = ing.restated_resample(x, o, n)
    # torchaudio's fp32 tap positions get coarser as the reduced rates grow (argument ~ 6 pi x cutoff): at 1000:999 its
    # own output is 2e-4 from the exact windowed sinc, and this kernel follows torchaudio, not the truth
    assert np.abs(y - truth).max() <= (TRUTH_TOL if max(rs.orig, rs.new) <= 441 else 10 * TRUTH_TOL)


@gpu
@pytest.mark.parametrize("rates", [(44100, 22050), (16000, 22050), (44100, 24000)])
def test_gpu_ragged_batch_equals_clip_by_clip(rates):
    from xtts_b200 import load_audio_pcm16
    o, n = rates
    frames = [30011, 1, 17, 12000, 29999, 640]
    F = max(frames)
    pcms = [ing.synth_pcm16(f, 2, 300 + i, "fullscale" if i % 2 else "speechlike") for i, f in enumerate(frames)]
    batch = np.zeros((len(frames), F, 2), dtype=np.int16)
    for i, p in enumerate(pcms):
        batch[i, :p.shape[0]] = p
        batch[i, p.shape[0]:] = 12345           # garbage behind the clip must not leak in
    y, lens = load_audio_pcm16(torch.from_numpy(batch).cuda(), o, n, torch.tensor(frames))
    ref, ref_lens = ing.callsite_load_audio_ragged(pcms, o, n)
    assert lens.cpu().tolist() == ref_lens.tolist()
    y = y.cpu().numpy()
    assert y.shape[1] == ing.out_len(F, o, n)
    assert np.abs(y[:, :ref.shape[1]] - ref).max() <= RESAMPLE_TOL
    for b, l in enumerate(ref_lens):
        assert not y[b, l:].any()
    assert np.abs(y).max() <= 1.0


@gpu
@pytest.mark.parametrize("rates", [(44100, 22050), (48000, 16000), (32000, 8000), (44100, 24000), (16000, 22050)])
@pytest.mark.parametrize("channels", [1, 2, 3])
def test_gpu_fused_decode_resample_is_bit_identical_to_two_steps(rates, channels):
    from xtts_b200 import Resampler, decode_pcm16
    o, n = rates
    frames = [25001, 9, 24000, 13333]
    F = max(frames)
    pcm = np.stack([ing.synth_pcm16(F, channels, 400 + b, "uniform") for b in range(len(frames))])
    t = torch.from_numpy(pcm).cuda()
    fr = torch.tensor(frames)
    rs = Resampler(o, n)
    y2, l2 = rs(decode_pcm16(t, fr), fr, clip=True)
    y1, l1 = rs.from_pcm16(t, fr, clip=True)
    assert torch.equal(l1, l2) and torch.equal(y1, y2)
    assert torch.equal(rs.from_pcm16(t), rs(decode_pcm16(t)))


@gpu
def test_gpu_decimator_agrees_with_general_kernel(monkeypatch):
    """k_decimate (2:1 / 3:1 / 4:1 fast path) against k_resample on the same plan inputs."""
    from xtts_b200 import Resampler, _cabi
    x = torch.rand((5, 50021), device="cuda") * 2 - 1
    lens = torch.tensor([50021, 1, 2, 33333, 4099])
    for o, n in ((44100, 22050), (48000, 16000), (32000, 8000)):
        fast = Resampler(o, n)                                # product library: its default kernel choice
        _cabi.select_lab(True)                                # kernel-selection knobs exist in the lab build only
        monkeypatch.setenv("TTTS_RESAMPLE_GENERIC", "1")
        try:
            slow = Resampler(o, n)
        finally:
            monkeypatch.delenv("TTTS_RESAMPLE_GENERIC")
            _cabi.select_lab(False)
        ya, la = fast(x, lens)
        yb, lb = slow(x, lens)
        assert torch.equal(la, lb)
        assert float((ya - yb).abs().max()) <= RESAMPLE_TOL


@gpu
@pytest.mark.parametrize("rates", [(44100, 24000), (16000, 22050), (24000, 22050), (44100, 16000), (48000, 22050)])
def test_gpu_tensor_core_resampler_agrees_with_fma_kernel(monkeypatch, rates):
    """k_resample_tc (tcgen05, 3xTF32) against k_resample (fp32 FMA) on the same plan inputs, ragged."""
    from xtts_b200 import Resampler, _cabi
    o, n = rates
    x = torch.rand((6, 40013), device="cuda") * 2 - 1
    lens = torch.tensor([40013, 1, 2, 33333, 4099, 20000])
    fast = Resampler(o, n)                                    # product library: its default kernel choice
    _cabi.select_lab(True)                                    # kernel-selection knobs exist in the lab build only
    monkeypatch.setenv("TTTS_RESAMPLE_SIMT", "1")
    try:
        slow = Resampler(o, n)
    finally:
        monkeypatch.delenv("TTTS_RESAMPLE_SIMT")
        _cabi.select_lab(False)
    ya, la = fast(x, lens, clip=True)
    yb, lb = slow(x, lens, clip=True)
    assert torch.equal(la, lb)
    assert float((ya - yb).abs().max()) <= RESAMPLE_TOL
    for b, l in enumerate(la.cpu().tolist()):
        assert not ya[b, l:].any()


@gpu
def test_gpu_full_size_linearity_and_dc_gain():
    """BASELINE-sized batch (64 clips x 10 s at 44.1 kHz): size-independent properties -- linearity, and a
    constant stays constant away from the edges (unit DC gain of every phase)."""
    from xtts_b200 import Resampler
    rs = Resampler(44100, 22050)
    g = torch.Generator(device="cuda").manual_seed(3)
    a = torch.rand((64, 441000), device="cuda", generator=g) * 2 - 1
    b = torch.rand((64, 441000), device="cuda", generator=g) * 2 - 1
    ya, yb, yab = rs(a), rs(b), rs(a + 0.5 * b)
    assert ya.shape == (64, 220500)
    assert float((yab - (ya + 0.5 * yb)).abs().max()) <= 4e-6
    c = rs(torch.full((2, 44100), 0.25, device="cuda"))
    assert float((c[:, 50:-50] - 0.25).abs().max()) < 2e-3


@gpu
def test_gpu_ingest_feeds_the_mel_path():
    """PCM -> load_audio_pcm16 -> TorchMelSpectrogram equals the reference chain on the CPU oracle."""
    from oracle import mel_oracle as orc
    from xtts_b200 import TorchMelSpectrogram, load_audio_pcm16
    pcm = ing.synth_pcm16(88200, 2, 77, "uniform")      # broadband: the tonal limit of the fp32 mel path is not the topic
    wav = load_audio_pcm16(torch.from_numpy(pcm).cuda(), 44100, 22050)
    mod = TorchMelSpectrogram(mel_norm_file=None)
    mel = mod(wav).cpu().numpy()
    ref_wav = ing.callsite_load_audio(pcm, 44100, 22050)
    assert np.abs(wav.cpu().numpy() - ref_wav).max() <= RESAMPLE_TOL
    # the log amplifies the (tolerated) 3e-7 waveform difference in near-silent bins, so the mel stage is checked on
    # the waveform it was actually given, and the whole chain at a looser bound
    ref = orc.callsite_logmel(wav.cpu().numpy(), orc.V1, None).numpy()
    assert mel.shape == ref.shape
    assert np.abs(mel - ref).max() <= 1e-4
    chain = orc.callsite_logmel(ref_wav, orc.V1, None).numpy()
    assert np.abs(mel - chain).max() <= 2e-3


@gpu
def test_gpu_c_abi_argument_errors():
    from xtts_b200 import _cabi, Resampler
    lib = _cabi.lib()
    x = torch.zeros(8, device="cuda")
    assert lib.ttts_pcm16_to_f32(None, 8, 1, None, 1, 8, x.data_ptr(), 8, None) == _cabi.EINVAL
    p = torch.zeros(8, dtype=torch.int16, device="cuda")
    assert lib.ttts_pcm16_to_f32(p.data_ptr(), 4, 1, None, 1, 8, x.data_ptr(), 8, None) == _cabi.EINVAL
    assert b"in_row_stride" in lib.ttts_mel_last_error()
    rs = Resampler(2, 1)
    assert lib.ttts_resample_forward(rs._h, x.data_ptr(), 4, None, 1, 8, x.data_ptr(), 8, 0, None) == _cabi.EINVAL
    assert lib.ttts_resample_forward(rs._h, x.data_ptr(), 8, None, 1, 8, x.data_ptr(), 3, 0, None) == _cabi.EINVAL
    assert lib.ttts_resample_out_len(rs._h, 9) == 5


@gpu
@pytest.mark.parametrize("rates", [(44100, 22050), (48000, 16000), (44100, 24000), (16000, 22050), (16000, 24000)])
def test_gpu_kernels_stay_inside_their_rows(rates):
    """Guard bands (compute-sanitizer is not available on the pool): every output row sits in a sentinel-filled
    buffer with an odd pitch; nothing outside [0, out_len) of a row may change, for fp32 and int16 input."""
    from xtts_b200 import _cabi, Resampler
    lib = _cabi.lib()
    o, n = rates
    rs = Resampler(o, n)
    B, L, C = 3, 10007, 2
    Lo = rs.out_len(L)
    pitch, lead = Lo + 13, 5
    stream = torch.cuda.current_stream().cuda_stream
    x = torch.rand((B, L), device="cuda") * 2 - 1
    pcm = torch.randint(-32768, 32767, (B, L, C), dtype=torch.int16, device="cuda")
    lens = torch.tensor([L, 1, 5000], dtype=torch.int64, device="cuda")
    for src in ("f32", "pcm16"):
        buf = torch.full((B * pitch + lead + 7,), 777.0, device="cuda")
        out_ptr = buf.data_ptr() + 4 * lead
        if src == "f32":
            _cabi.check(lib.ttts_resample_forward(rs._h, x.data_ptr(), L, lens.data_ptr(), B, L, out_ptr, pitch, 0, stream))
        else:
            _cabi.check(lib.ttts_resample_pcm16_forward(rs._h, pcm.data_ptr(), L * C, C, lens.data_ptr(), B, L, out_ptr,
                                                        pitch, 0, stream))
        torch.cuda.synchronize()
        body = buf[lead:lead + B * pitch].view(B, pitch)
        assert bool((buf[:lead] == 777.0).all()) and bool((buf[lead + B * pitch:] == 777.0).all())
        assert bool((body[:, Lo:] == 777.0).all())
        assert bool((body[:, :Lo] != 777.0).all())           # every sample of every row was written (zeros past a clip)
        for b, l in enumerate(lens.cpu().tolist()):
            assert not body[b, rs.out_len(l):Lo].any()
    # decode: same idea
    buf = torch.full((B * (L + 9) + 3,), 777.0, device="cuda")
    _cabi.check(lib.ttts_pcm16_to_f32(pcm.data_ptr(), L * C, C, lens.data_ptr(), B, L, buf.data_ptr() + 12, L + 9, stream))
    torch.cuda.synchronize()
    body = buf[3:].view(B, L + 9)
    assert bool((buf[:3] == 777.0).all()) and bool((body[:, L:] == 777.0).all()) and bool((body[:, :L] != 777.0).all())


@gpu
def test_gpu_unreduced_or_huge_ratios_fail_loudly():
    """Coprime rates one apart need a 44101 x 44114 tap table: the plan refuses instead of allocating gigabytes."""
    import ctypes
    from xtts_b200 import _cabi
    lib = _cabi.lib()
    h = ctypes.c_void_p()
    taps = torch.zeros(16)
    assert lib.ttts_resample_plan_create(taps.data_ptr(), 44100, 44101, 7, ctypes.byref(h)) == _cabi.EINVAL
    assert b"gcd" in lib.ttts_mel_last_error()
    assert lib.ttts_resample_plan_create(taps.data_ptr(), 0, 3, 7, ctypes.byref(h)) == _cabi.EINVAL


@gpu
def test_gpu_plans_of_different_sizes_coexist():
    """A plan with few phases created after one with many must not shrink the first one's launch limits
    (the dynamic shared-memory attribute is per kernel, not per plan)."""
    from xtts_b200 import Resampler
    x = torch.rand((2, 30000), device="cuda") * 2 - 1
    big = Resampler(16000, 22050)
    y0 = big(x)
    small = Resampler(25, 16)
    small(x)
    tiny = Resampler(16000, 24000)
    tiny(x)
    assert torch.equal(big(x), y0)

