"""The oracle against the golden vectors minted from the reference itself (oracle/gen_golden.py)."""
import numpy as np
import pytest
import torch

from oracle import mel_oracle as orc


def test_callsite_layer_reproduces_reference_c1(golden_dir):
    g = np.load(f"{golden_dir}/v1_c1_uniform_10s.npz")
    assert str(g["torchaudio"]).startswith("2.")        # the pinned third party the vectors came from
    x = orc.synth("uniform", 1, 220500, 1234)
    y = orc.callsite_logmel(x, orc.V1, g["norms"])
    assert tuple(y.shape) == (1, 80, 862) and y.stride() == (80 * 862, 1, 80)
    # same library, same machine class: bit-equal here; allow the batch-shape jitter of SURVEY F12 elsewhere
    assert np.abs(y.numpy() - g["mel"]).max() <= 1e-6
    y0 = orc.callsite_logmel(x, orc.V1, None)
    assert np.abs(y0.numpy() - g["mel_nonorm"]).max() <= 1e-6


def test_restated_fp64_layer_c1(golden_dir):
    g = np.load(f"{golden_dir}/v1_c1_uniform_10s.npz")
    x = orc.synth("uniform", 1, 220500, 1234)
    r = orc.restated_logmel(x, orc.V1, g["norms"])
    assert r.shape == g["mel"].shape
    assert np.abs(r - g["mel"]).max() <= 3e-5           # the fp32 reference's own rounding noise


@pytest.mark.parametrize("L", [513, 600, 767, 768, 769, 1023, 1024, 1025, 1279, 1280, 1281, 2047, 2048, 4099])
def test_edge_lengths(golden_dir, L):
    g = np.load(f"{golden_dir}/v1_edge_lengths.npz")
    x, y = g[f"x_{L}"], g[f"y_{L}"]
    assert y.shape == (2, 80, orc.frame_count(L)) == (2, 80, 1 + L // 256)
    assert np.abs(orc.callsite_logmel(x, orc.V1, g["norms"]).numpy() - y).max() <= 1e-6
    assert np.abs(orc.restated_logmel(x, orc.V1, g["norms"]) - y).max() <= 5e-5


def test_too_short_raises():
    with pytest.raises(ValueError):
        orc.frame_count(512)
    with pytest.raises(RuntimeError):
        orc.callsite_logmel(np.zeros((1, 512), np.float32))


def test_signal_classes(golden_dir):
    g = np.load(f"{golden_dir}/v1_signals.npz")
    for kind in ("gauss", "silence", "b1l"):
        r = orc.restated_logmel(g[f"x_{kind}"], orc.V1, g["norms"])
        assert np.abs(r - g[f"y_{kind}"]).max() <= 5e-5, kind
    # silence is exactly log(1e-5) / norm
    expect = (np.float32(np.log(np.float32(1e-5))) / g["norms"])[None, :, None]
    assert np.array_equal(g["y_silence"], np.broadcast_to(expect, g["y_silence"].shape).astype(np.float32))
    # tonal input: the fp32 reference is itself ~1e-4..1e-3 from the truth
    r = orc.restated_logmel(g["x_tone"], orc.V1, g["norms"])
    assert np.abs(r - g["y_tone"]).max() <= 5e-3


def test_second_reference_entry_point(golden_dir):
    g = np.load(f"{golden_dir}/v1_dvae_wav_to_mel.npz")
    assert np.abs(orc.callsite_logmel(g["x"], orc.V1, g["norms"]).numpy() - g["y"]).max() <= 1e-6


def test_v2_features(golden_dir):
    g = np.load(f"{golden_dir}/v2_features.npz")
    for L in (24000, 5000, 1024, 777):
        x = g[f"x_{L}"]
        assert g[f"center_{L}"].shape == (1, 100, orc.frame_count(L, orc.V2))
        assert g[f"same_{L}"].shape == (1, 100, orc.frame_count(L, orc.V2_SAME)) == (1, 100, L // 256)
        assert np.abs(orc.callsite_logmel(x, orc.V2).numpy() - g[f"center_{L}"]).max() <= 2e-6
        assert np.abs(orc.callsite_logmel(x, orc.V2_SAME).numpy() - g[f"same_{L}"]).max() <= 2e-6
        assert np.abs(orc.restated_logmel(x, orc.V2) - g[f"center_{L}"]).max() <= 5e-5
        assert np.abs(orc.restated_logmel(x, orc.V2_SAME) - g[f"same_{L}"]).max() <= 5e-5


def test_pad_mask(golden_dir):
    g = np.load(f"{golden_dir}/pad_mask.npz")
    assert np.array_equal(orc.make_pad_mask(g["lengths"]), g["mask"])
    assert np.array_equal(orc.make_pad_mask(g["lengths"], 12), g["mask12"])
    from xtts_b200.mel import make_pad_mask
    assert np.array_equal(make_pad_mask(torch.from_numpy(g["lengths"])).numpy(), g["mask"])
    assert np.array_equal(make_pad_mask(torch.from_numpy(g["lengths"]), 12).numpy(), g["mask12"])


def test_ragged_contract():
    lens = [513, 3000, 1024]
    x = orc.synth("uniform", 3, 3000, 2)
    y, nf, mask = orc.restated_logmel(x, orc.V1, None, lengths=lens)
    yc, nfc, maskc = orc.callsite_logmel_ragged(x, lens, orc.V1, None)
    assert nf.tolist() == nfc.tolist() == [3, 12, 5]
    assert np.array_equal(mask, maskc) and mask.shape == (3, 12)
    assert np.abs(y - yc.numpy()).max() <= 5e-5
    assert (y[0, :, 3:] == 0).all()
    # zero-padding the WAVEFORM instead (what a naive batched call does) changes the last frames
    xz = x.copy(); xz[0, 513:] = 0
    naive = orc.callsite_logmel(xz, orc.V1, None).numpy()
    assert np.abs(naive[0, :, :3] - yc.numpy()[0, :, :3]).max() > 1e-2


def test_stage_oracles_consistent():
    x = orc.synth("gauss", 1, 5000, 9)[0]
    st = orc.restated_stages(x, orc.V1)
    T = orc.frame_count(5000)
    assert st.frames.shape == (T, 1024) and st.re.shape == (T, 513) and st.mel.shape == (T, 80)
    # even/odd fold identity used by the kernel
    w = orc.full_window(orc.V1)
    f = st.frames
    n = np.arange(1, 513)
    E = f[:, n] + f[:, 1024 - n]; O = f[:, n] - f[:, 1024 - n]
    k = np.arange(513)
    wn = w[n] * np.where(n == 512, 0.5, 1.0)
    C = wn[:, None] * np.cos(2 * np.pi * np.outer(n, k) / 1024)
    S = wn[:, None] * np.sin(2 * np.pi * np.outer(n, k) / 1024)
    assert np.abs(E @ C - st.re).max() <= 1e-10 and np.abs(O @ S + st.im).max() <= 1e-10
    # live bins of the V1 filterbank
    fb = orc.melscale_fbanks64(orc.V1)
    live = np.nonzero(fb.any(axis=1))[0]
    assert live.min() == 1 and live.max() == 371


def test_mel_stats():
    rng = np.random.default_rng(0)
    lm = rng.standard_normal((3, 80, 20))
    acc = orc.mel_stats(lm, [20, 7, 1])
    assert acc[160] == 28
    assert np.allclose(acc[:80], lm[0].sum(1) + lm[1, :, :7].sum(1) + lm[2, :, :1].sum(1))
    assert np.allclose(orc.norms_from_stats(acc), acc[:80] / 28, rtol=1e-6)


def test_v3_layers_agree_and_filterbank_restatement():
    """V3 (mel_spectrogram, feature_extractors.py:27-53): parity UNPINNED (no librosa here) -- the two oracle layers
    and the product's librosa_mel restatement must at least agree with each other and with torchaudio's Slaney bank."""
    import torchaudio
    from xtts_b200.filterbank import librosa_mel
    fb = librosa_mel(24000, 1024, 80, 0, 8000)
    assert fb.shape == (80, 513) and fb.dtype == torch.float32
    assert np.abs(fb.numpy() - orc.melscale_fbanks64(orc.V3).T).max() < 1e-7
    ta = torchaudio.functional.melscale_fbanks(513, 0.0, 8000.0, 80, 24000, "slaney", "slaney").t()
    assert float((fb - ta).abs().max()) < 2e-7
    x = orc.synth("uniform", 2, 6000, 12)
    a = orc.callsite_logmel(x, orc.V3).numpy()
    assert a.shape == (2, 80, 6000 // 256)
    assert np.abs(orc.restated_logmel(x, orc.V3) - a).max() <= 5e-5


def test_general_shape_goldens_match_both_oracle_layers(golden_dir):
    """tests/golden/v1_general_shapes.npz: the reference's TorchMelSpectrogram built with other filter_length / hop_length /
    win_length / n_mel_channels / sampling_rate values (oracle/gen_golden_general.py)."""
    g = np.load(f"{golden_dir}/v1_general_shapes.npz")
    names = [k[4:] for k in g.files if k.startswith("cfg_")]
    assert len(names) == 5
    for name in names:
        n_fft, hop, win, n_mels, fmin, fmax, sr, normalize, has_norms, gauss, L, seed = (int(v) for v in g[f"cfg_{name}"])
        cfg = orc.MelConfig(sample_rate=sr, n_fft=n_fft, hop=hop, win_length=win, n_mels=n_mels, f_min=float(fmin),
                            f_max=float(fmax), normalized=bool(normalize))
        norms = orc.default_mel_norms(n_mels) if has_norms else None
        x = orc.synth("gauss" if gauss else "uniform", 2, L, seed)
        ref = g[f"y_{name}"]
        y = orc.callsite_logmel(x, cfg, norms).numpy()
        assert y.shape == ref.shape and np.array_equal(y, ref), name              # same call sequence: bit for bit
        assert ref.shape[2] == orc.frame_count(L, cfg)
        r = orc.restated_logmel(x, cfg, norms)
        assert np.abs(r - ref).max() <= 2e-4, (name, float(np.abs(r - ref).max()))  # fp64 restatement vs the fp32 reference
