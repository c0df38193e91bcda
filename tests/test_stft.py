"""STFT family (first slice): the oracle against goldens minted from the reference's own classes (CPU), and
the CUDA path through the C ABI against both (``-m gpu``)."""
import numpy as np
import pytest
import torch

from oracle import mel_oracle as orc
from oracle import stft_oracle as so

gpu = pytest.mark.gpu


@pytest.fixture(scope="module")
def g(golden_dir):
    return np.load(f"{golden_dir}/stft_family.npz")


# ------------------------------------------------------------------ CPU: the oracle is pinned on the reference
def test_oracle_forward_matches_reference_goldens(g):
    for L in (4000, 2499):
        x = g[f"x_{L}"]
        for n_fft, hop, win in so.MS_SHAPES:
            ref = g[f"mag_{L}_{n_fft}"]
            assert ref.shape == (2, n_fft // 2 + 1, 1 + L // hop)
            assert np.abs(so.callsite_stft_mag(x, n_fft, hop, win).numpy() - ref).max() <= 1e-6 * ref.max()
            assert np.abs(so.restated_stft_mag(x, n_fft, hop, win) - ref).max() <= 2e-5 * ref.max()
    x = g["x_opts"]
    assert np.abs(so.restated_stft_mag(x, 1024, 120, 600, power=2) - g["mag_power2"]).max() <= 2e-5 * g["mag_power2"].max()
    db = np.log(np.maximum(so.restated_stft_mag(x, 512, 50, 240, normalized=True), 1e-5) * 2.0)
    assert np.abs(db - g["mag_normalized_db"]).max() <= 2e-4
    assert np.abs(so.restated_stft_mag(x, 1024, 256, 1024) - g["mag_1024_256"]).max() <= 2e-5 * g["mag_1024_256"].max()


def test_oracle_losses_match_reference_goldens(g):
    y, y_hat = torch.from_numpy(g["loss_y"]), torch.from_numpy(g["loss_y_hat"])
    lm, lsc = so.callsite_stft_loss(y_hat, y, 1024, 120, 600)
    assert np.allclose([float(lm), float(lsc)], g["stft_loss"], rtol=1e-6)
    lm, lsc = so.callsite_multiscale_loss(y_hat, y)
    assert np.allclose([float(lm), float(lsc)], g["ms_loss"], rtol=1e-6)


def test_oracle_inverse_matches_reference_goldens(g):
    for name, (n_fft, hop) in {"vocos": (1024, 256), "small": (512, 128)}.items():
        spec = g[f"spec_{name}"]
        T = spec.shape[2]
        for padding in ("same", "center"):
            ref = g[f"istft_{name}_{padding}"]
            assert ref.shape == (2, T * hop if padding == "same" else (T - 1) * hop)
            scale = np.abs(ref).max()
            assert np.abs(so.callsite_istft(spec, n_fft, hop, n_fft, padding).numpy() - ref).max() <= 1e-6 * scale
            assert np.abs(so.restated_istft(spec, n_fft, hop, n_fft, padding) - ref).max() <= 1e-5 * scale


def test_no_cpu_fallback_and_constructor_surface():
    import inspect
    from xtts_b200 import ISTFT, MultiScaleSTFTLoss, STFTLoss, TorchSTFT
    assert list(inspect.signature(TorchSTFT.__init__).parameters)[1:] == [
        "n_fft", "hop_length", "win_length", "pad_wav", "window", "sample_rate", "mel_fmin", "mel_fmax", "n_mels", "use_mel",
        "do_amp_to_db", "spec_gain", "power", "use_htk", "mel_norm", "normalized"]          # losses.py:67-85
    assert list(inspect.signature(ISTFT.__init__).parameters)[1:] == ["n_fft", "hop_length", "win_length", "padding"]
    assert list(inspect.signature(MultiScaleSTFTLoss.__init__).parameters)[1:] == ["n_ffts", "hop_lengths", "win_lengths"]
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        TorchSTFT(1024, 120, 600)(torch.zeros(1, 4000))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ISTFT(1024, 256, 1024)(torch.zeros(1, 513, 4, dtype=torch.complex64))
    with pytest.raises(ValueError):
        ISTFT(1024, 256, 1024, padding="valid")
    with pytest.raises(NotImplementedError):
        TorchSTFT(1024, 256, 1024, use_mel=True, sample_rate=22050)
    assert len(MultiScaleSTFTLoss().loss_funcs) == 3 and isinstance(MultiScaleSTFTLoss().loss_funcs[0], STFTLoss)


# ------------------------------------------------------------------ GPU: the CUDA path through the C ABI
@gpu
def test_gpu_forward_against_reference_goldens(g):
    from xtts_b200 import TorchSTFT
    dev = torch.device("cuda:0")
    for L in (4000, 2499):
        x = torch.from_numpy(g[f"x_{L}"]).to(dev)
        for n_fft, hop, win in so.MS_SHAPES:
            ref = g[f"mag_{L}_{n_fft}"]
            y = TorchSTFT(n_fft, hop, win)(x)
            assert tuple(y.shape) == ref.shape and y.dtype == torch.float32
            assert np.abs(y.cpu().numpy() - ref).max() <= 2e-5 * ref.max(), (L, n_fft)
    x = torch.from_numpy(g["x_opts"]).to(dev)
    assert np.abs(TorchSTFT(1024, 120, 600, power=2)(x).cpu().numpy() - g["mag_power2"]).max() <= 2e-5 * g["mag_power2"].max()
    db = TorchSTFT(512, 50, 240, normalized=True, do_amp_to_db=True, spec_gain=2.0)(x).cpu().numpy()
    assert np.abs(db - g["mag_normalized_db"]).max() <= 2e-4
    y = TorchSTFT(1024, 256, 1024)(x[:, None, :])
    assert np.abs(y.cpu().numpy() - g["mag_1024_256"]).max() <= 2e-5 * g["mag_1024_256"].max()


@gpu
def test_gpu_losses_against_reference_goldens(g):
    from xtts_b200 import MultiScaleSTFTLoss, STFTLoss
    dev = torch.device("cuda:0")
    y, y_hat = torch.from_numpy(g["loss_y"]).to(dev), torch.from_numpy(g["loss_y_hat"]).to(dev)
    lm, lsc = STFTLoss(1024, 120, 600)(y_hat, y)
    assert np.allclose([float(lm), float(lsc)], g["stft_loss"], rtol=2e-5)
    lm, lsc = MultiScaleSTFTLoss()(y_hat, y)
    assert np.allclose([float(lm), float(lsc)], g["ms_loss"], rtol=2e-5)


@gpu
def test_gpu_inverse_against_reference_goldens(g):
    from xtts_b200 import ISTFT
    dev = torch.device("cuda:0")
    for name, (n_fft, hop) in {"vocos": (1024, 256), "small": (512, 128)}.items():
        spec = torch.from_numpy(g[f"spec_{name}"]).to(dev)
        for padding in ("same", "center"):
            ref = g[f"istft_{name}_{padding}"]
            m = ISTFT(n_fft, hop, n_fft, padding=padding)
            y = m(spec)
            m.check_envelope()
            assert tuple(y.shape) == ref.shape
            assert np.abs(y.cpu().numpy() - ref).max() <= 2e-5 * np.abs(ref).max(), (name, padding)


@gpu
def test_gpu_full_size_against_the_oracle_and_round_trip():
    """64 x 10 s at 22.05 kHz (the headline batch) through the (1024, 120, 600) loss shape against the fp64 restatement on a
    4-clip sample, and STFT -> ISTFT with the Vocos shape returns the waveform (window / envelope identity)."""
    from xtts_b200 import ISTFT, TorchSTFT
    dev = torch.device("cuda:0")
    x = orc.synth("gauss", 64, 220500, 31)
    y = TorchSTFT(1024, 120, 600)(torch.from_numpy(x).to(dev))
    assert tuple(y.shape) == (64, 513, 1 + 220500 // 120)
    for b in (0, 21, 42, 63):
        ref = so.restated_stft_mag(x[b:b + 1], 1024, 120, 600)
        assert np.abs(y[b:b + 1].cpu().numpy() - ref).max() <= 2e-5 * ref.max(), b
    # round trip: torch.stft (library, fp32) -> our ISTFT "center" == the waveform
    xs = torch.from_numpy(x[:4, :65536]).to(dev)
    spec = torch.stft(xs, 1024, 256, 1024, torch.hann_window(1024, device=dev), center=True, return_complex=True)
    back = ISTFT(1024, 256, 1024, padding="center")(spec)
    assert float((back - xs[:, :back.shape[1]]).abs().max()) <= 2e-5
