"""ttts_mel_forward_static (include/ttts_mel.h): the log-mel forward with its workspace in a file-scope device array and
no allocation on the call path.  The entry point has no counterpart in the reference; it is checked against the existing
entry point (ttts_mel_forward) on the same inputs, through the C ABI, on two calls in a row that share the array.

Collected last on purpose."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-4                      # the parity tolerance of tests/test_gpu_parity.py
STATIC_WS_BYTES = 8 << 20       # TTTS_MEL_STATIC_WS_BYTES


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import xtts_b200._cabi as cabi
    cabi.lib()                  # fail loudly if the extension is missing
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def plan(dev):
    from oracle import mel_oracle as orc
    from xtts_b200 import MelPlan
    p = MelPlan(device=dev, sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0,
                f_max=8000, power=2.0, norm="slaney")
    p.set_norms(torch.from_numpy(orc.default_mel_norms(80)))
    return p


def _static(plan, wav, lengths, T, want_mask):
    """One call of the new entry point, straight through ctypes: the caller passes inputs and outputs only."""
    from xtts_b200 import _cabi
    lib, dev = plan._lib, wav.device
    B, Lmax = int(wav.shape[0]), int(wav.shape[1])
    out = torch.full((B, T, plan.n_mels), float("nan"), dtype=torch.float32, device=dev)
    n_frames = torch.full((B,), -1, dtype=torch.int32, device=dev)
    mask = torch.full((B, T), 7, dtype=torch.uint8, device=dev) if want_mask else None
    ws = ctypes.c_void_p()
    stream = torch.cuda.current_stream(dev).cuda_stream
    rc = lib.ttts_mel_forward_static(plan._h, wav.data_ptr(), int(wav.stride(0)) if B > 1 else Lmax,
                                     lengths.data_ptr() if lengths is not None else None, B, Lmax, T, out.data_ptr(),
                                     n_frames.data_ptr(), mask.data_ptr() if mask is not None else None, stream,
                                     ctypes.byref(ws))
    _cabi.check(rc)
    assert ws.value, "the entry point reports the device address of its static workspace"
    _cabi.check(lib.ttts_mel_status(ws, stream))          # synchronises; device-side status word of this call
    return out, n_frames, mask, ws.value


def _same(a, b):
    a, b = a.detach().cpu().numpy(), b.detach().cpu().numpy()
    assert np.isfinite(a).all() and np.isfinite(b).all()
    return float(np.abs(a - b).max())


def test_static_workspace_forward_equals_the_caller_workspace_forward_twice_in_a_row(dev, plan):
    g = torch.Generator(device=dev).manual_seed(2026)

    # call 1: dense batch, 16-byte addressable rows (the kernel reads the waveform in place)
    x1 = torch.rand(6, 40000, device=dev, generator=g) * 2 - 1
    ref1, nf1, _ = plan.forward(x1)
    plan.status()
    torch.cuda.synchronize(dev)
    y1, n1, _, ws1 = _static(plan, x1, None, plan.frames(40000), False)
    assert _same(y1, ref1) <= TOL
    assert torch.equal(n1, nf1)

    # call 2, right behind it on the same array: ragged batch with a mask, rows NOT 16-byte addressable (odd row stride,
    # so the padded copy of the batch lives in the static workspace too), different batch size and clip lengths
    base = torch.rand(5, 30001, device=dev, generator=g) * 2 - 1
    x2 = base[:, :30000]
    assert x2.stride(0) % 4 != 0
    lengths = torch.tensor([30000, 513, 12345, 29999, 7000], dtype=torch.int64, device=dev)
    ref2, nf2, m2 = plan.forward(x2, lengths, want_mask=True)
    plan.status()
    torch.cuda.synchronize(dev)
    free_before = torch.cuda.mem_get_info(dev)[0]
    y2, n2, k2, ws2 = _static(plan, x2, lengths, plan.frames(30000), True)
    free_after = torch.cuda.mem_get_info(dev)[0]
    assert ws2 == ws1, "one array per device"
    assert free_after >= free_before, "the call allocated device memory"
    assert _same(y2, ref2) <= TOL
    assert torch.equal(n2, nf2) and torch.equal(k2, m2)
    assert [int(v) for v in n2.cpu()] == [plan.frames(int(l)) for l in lengths.cpu()]

    # and the first batch again: nothing of call 2 is left behind in what call 3 reads
    y3, n3, _, _ = _static(plan, x1, None, plan.frames(40000), False)
    assert torch.equal(y3, y1) and torch.equal(n3, n1)


def test_static_workspace_rejects_batches_beyond_its_size(dev, plan):
    from xtts_b200 import _cabi
    lib = plan._lib
    # 16 x 10 s, rows not addressable in place: the padded copy alone is 14 MB > 8 MiB
    x = torch.zeros(16, 220501, device=dev)[:, :220500]
    need = int(lib.ttts_mel_workspace_bytes_for(plan._h, 16, 220500, x.data_ptr(), int(x.stride(0))))
    assert need > STATIC_WS_BYTES
    T = plan.frames(220500)
    out = torch.zeros((16, T, plan.n_mels), device=dev)
    ws = ctypes.c_void_p(1)
    rc = lib.ttts_mel_forward_static(plan._h, x.data_ptr(), int(x.stride(0)), None, 16, 220500, T, out.data_ptr(), None,
                                     None, torch.cuda.current_stream(dev).cuda_stream, ctypes.byref(ws))
    assert rc == _cabi.ENOMEM
    assert not ws.value
    msg = lib.ttts_mel_last_error().decode()
    assert str(need) in msg and str(STATIC_WS_BYTES) in msg, msg
    with pytest.raises(_cabi.MelError):
        _cabi.check(rc)
    torch.cuda.synchronize(dev)
    assert float(out.abs().max()) == 0.0                    # nothing was launched
    # the same batch in place (aligned rows) fits: only offsets, counts and edges live in the workspace
    xa = torch.zeros(16, 220500, device=dev)
    assert int(lib.ttts_mel_workspace_bytes_for(plan._h, 16, 220500, xa.data_ptr(), 220500)) <= STATIC_WS_BYTES
