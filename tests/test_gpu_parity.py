"""GPU parity tests proper: the CUDA path (through the C ABI) against the oracle."""
import numpy as np
import pytest
import torch

from oracle import mel_oracle as orc

pytestmark = pytest.mark.gpu

TOL = 1e-4   # BASELINE.json north_star: max abs error on log-mel vs the reference's torchaudio path


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import xtts_b200._cabi as cabi
    cabi.lib()          # fail loudly if the extension is missing
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def v1(dev):
    from xtts_b200 import TorchMelSpectrogram
    m = TorchMelSpectrogram(mel_norm_file=None)
    m.mel_norms = torch.from_numpy(orc.default_mel_norms(80))
    return m


def _np(t):
    return t.detach().cpu().numpy()


def test_c1_golden(dev, v1, golden_dir):
    g = np.load(f"{golden_dir}/v1_c1_uniform_10s.npz")
    x = orc.synth("uniform", 1, 220500, 1234)
    y = v1(torch.from_numpy(x).to(dev))
    assert tuple(y.shape) == g["mel"].shape == (1, 80, 862)
    assert y.stride() == (862 * 80, 1, 80)          # the reference's transposed view
    assert np.abs(_np(y) - g["mel"]).max() <= TOL
    v1.mel_norms = None
    y0 = v1(torch.from_numpy(x).to(dev))
    assert np.abs(_np(y0) - g["mel_nonorm"]).max() <= TOL
    v1.mel_norms = torch.from_numpy(orc.default_mel_norms(80))


def test_edge_lengths_golden(dev, v1, golden_dir):
    g = np.load(f"{golden_dir}/v1_edge_lengths.npz")
    for L in (513, 600, 767, 768, 769, 1023, 1024, 1025, 1279, 1280, 1281, 2047, 2048, 4099):
        y = v1(torch.from_numpy(g[f"x_{L}"]).to(dev))
        ref = g[f"y_{L}"]
        assert tuple(y.shape) == ref.shape, L        # bit-exact frame counts
        assert np.abs(_np(y) - ref).max() <= TOL, L


def test_signal_classes_golden(dev, v1, golden_dir):
    g = np.load(f"{golden_dir}/v1_signals.npz")
    y = v1(torch.from_numpy(g["x_gauss"]).to(dev))
    assert np.abs(_np(y) - g["y_gauss"]).max() <= TOL
    y = v1(torch.from_numpy(g["x_b1l"]).to(dev))      # [B,1,L] form
    assert np.abs(_np(y) - g["y_b1l"]).max() <= TOL
    # digital silence -> exactly log(1e-5)/norm everywhere
    y = v1(torch.from_numpy(g["x_silence"]).to(dev))
    assert np.array_equal(_np(y), g["y_silence"])
    # tonal input: dual-oracle rule (SURVEY 8c ii)
    x = g["x_tone"]
    y = _np(v1(torch.from_numpy(x).to(dev))).astype(np.float64)
    truth = orc.restated_logmel(x, orc.V1, g["norms"])
    bound = np.maximum(TOL, 1.5 * np.abs(g["y_tone"].astype(np.float64) - truth))
    assert (np.abs(y - truth) <= bound).all()


def test_dvae_twin_golden(dev, golden_dir):
    from xtts_b200 import dvae_wav_to_mel
    g = np.load(f"{golden_dir}/v1_dvae_wav_to_mel.npz")
    y = dvae_wav_to_mel(torch.from_numpy(g["x"]), mel_norms=torch.from_numpy(g["norms"]), device=dev)
    assert np.abs(_np(y) - g["y"]).max() <= TOL


def test_v2_features_golden(dev, golden_dir):
    from xtts_b200 import MelSpectrogramFeatures
    g = np.load(f"{golden_dir}/v2_features.npz")
    mc, ms = MelSpectrogramFeatures(), MelSpectrogramFeatures(padding="same")
    for L in (24000, 5000, 1024, 777):
        x = torch.from_numpy(g[f"x_{L}"]).to(dev)
        yc, ys = mc(x), ms(x)
        assert tuple(yc.shape) == g[f"center_{L}"].shape and tuple(ys.shape) == g[f"same_{L}"].shape
        assert np.abs(_np(yc) - g[f"center_{L}"]).max() <= TOL, L
        assert np.abs(_np(ys) - g[f"same_{L}"]).max() <= TOL, L


@pytest.mark.parametrize("kind", ["uniform", "gauss"])
def test_dense_batch_vs_oracle(dev, v1, kind):
    x = orc.synth(kind, 6, 33000, 42)
    y = v1(torch.from_numpy(x).to(dev))
    ref = orc.callsite_logmel(x, orc.V1, orc.default_mel_norms(80)).numpy()
    assert tuple(y.shape) == ref.shape
    assert np.abs(_np(y) - ref).max() <= TOL


def test_ragged_batch_masks_and_values(dev, v1):
    rng = np.random.default_rng(3)
    lens = [513, 514, 767, 768, 1024, 20000, 9999, 12800, 12799, 12801, 2560, 700]
    x = orc.synth("uniform", len(lens), max(lens), 8)
    for b, L in enumerate(lens):
        x[b, L:] = rng.random(max(lens) - L, dtype=np.float32)       # garbage beyond the clip must not matter
    y, nf, mask = v1(torch.from_numpy(x).to(dev), lengths=torch.tensor(lens))
    ry, rn, rmask = orc.callsite_logmel_ragged(x, lens, orc.V1, orc.default_mel_norms(80))
    assert nf.dtype == torch.int64 and nf.cpu().tolist() == rn.tolist()       # bit-exact frame counts
    assert mask.dtype == torch.bool and np.array_equal(_np(mask), rmask)      # bit-exact padding masks
    assert tuple(y.shape) == tuple(ry.shape)
    assert np.abs(_np(y) - ry.numpy()).max() <= TOL
    assert (_np(y)[rmask[:, None, :].repeat(80, 1)] == 0.0).all()             # padded frames are 0.0


def test_dft_stage_vs_fp64(dev):
    from xtts_b200 import MelPlan
    plan = MelPlan(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0,
                   f_max=8000, power=2.0, norm="slaney", device=dev, lab=True)      # the stage dump is a lab entry point
    x = orc.synth("uniform", 2, 9000, 17)
    out, nf, _, reim = plan.forward(torch.from_numpy(x).to(dev), debug_reim=True)
    plan.status()
    T = int(nf[0])
    k0, live, nbt = plan.first_bin, plan.live_bins, plan.num_bins
    assert (k0, live) == (1, 371)
    for b in range(2):
        st = orc.restated_stages(x[b], orc.V1)
        re = _np(reim[b * T:(b + 1) * T, :live])
        im = _np(reim[b * T:(b + 1) * T, nbt:nbt + live])
        scale = np.abs(st.re).max()
        assert np.abs(re - st.re[:, k0:k0 + live]).max() <= 6e-6 * scale      # fp32 TMEM accumulation, 3xTF32 products
        assert np.abs(im + st.im[:, k0:k0 + live]).max() <= 6e-6 * scale


def test_simt_crosscheck_full_size(dev, v1):
    """Config 2 (64 x 10 s): the tcgen05 kernel against the independent CUDA-core kernel."""
    from xtts_b200 import MelPlan
    plan = v1._plan_for(dev)                                     # the product library
    lab = MelPlan(device=dev, lab=True, **v1._plan_kwargs())     # k_simt lives in the lab build (ttts_mel_lab.h)
    lab.set_norms(v1.mel_norms)
    g = torch.Generator(device=dev).manual_seed(1234)
    x = torch.rand(64, 220500, device=dev, generator=g) * 2 - 1
    a, nfa, _ = plan.forward(x, impl="tcgen05")
    b, nfb, _ = lab.forward(x, impl="simt")
    assert torch.equal(a, lab.forward(x, impl="tcgen05")[0])     # same source, same kernel: bit-identical
    plan.status()
    lab.status()
    assert a.shape == (64, 862, 80) and torch.equal(nfa, nfb)
    assert float((a - b).abs().max()) <= TOL
    # 16-clip sample against the oracle itself
    idx = list(range(0, 64, 4))
    ref = orc.callsite_logmel(x[idx].cpu().numpy(), orc.V1, orc.default_mel_norms(80)).numpy()
    assert np.abs(_np(a[idx].transpose(1, 2)) - ref).max() <= TOL


def test_long_form_linearity_and_shift(dev):
    """Config 5 shape (120 s clips): size-independent properties instead of a CPU oracle run."""
    from xtts_b200 import TorchMelSpectrogram
    m = TorchMelSpectrogram(mel_norm_file=None)
    g = torch.Generator(device=dev).manual_seed(7)
    x = torch.rand(4, 2646000, device=dev, generator=g) * 2 - 1
    y = m(x)
    assert tuple(y.shape) == (4, 80, 10336)
    # power spectrum scales with the square of the gain: log-mel shifts by 2*log(g)
    y2 = m(x * 0.5)
    assert float((y2 - (y + 2 * np.log(0.5))).abs().max()) <= 2e-5
    # a shift by exactly one hop moves interior frames by one
    y3 = m(x[:, 256:])
    assert float((y3[:, :, 2:1000] - y[:, :, 3:1001]).abs().max()) <= 2e-5
    # spot-check one clip prefix against the oracle (frames away from the right edge are identical)
    ref = orc.callsite_logmel(x[:1, :220500].cpu().numpy(), orc.V1, None).numpy()
    assert np.abs(_np(y[:1, :, :850]) - ref[:, :, :850]).max() <= TOL


def test_error_behaviour(dev, v1):
    with pytest.raises(AssertionError):
        v1(torch.zeros(2, 2, 2, 2000, device=dev))
    with pytest.raises(RuntimeError):
        v1(torch.zeros(1, 512, device=dev))             # reflect pad needs L > 512
    with pytest.raises(RuntimeError):
        v1(torch.zeros(1, 4000))                        # CPU tensor: no fallback
    with pytest.raises(RuntimeError):
        v1(torch.zeros(1, 4000, device=dev, dtype=torch.float64))
    y = v1(torch.zeros(1, 4000, device=dev, dtype=torch.float16))
    assert y.dtype == torch.float32


def test_mel_stats(dev):
    from xtts_b200 import MelPlan
    plan = MelPlan(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0,
                   f_max=8000, power=2.0, norm="slaney", device=dev)
    lens = [22050, 8000, 15000, 513]
    x = orc.synth("uniform", 4, 22050, 77)
    out, nf, _ = plan.forward(torch.from_numpy(x).to(dev), torch.tensor(lens), t_max=plan.frames(22050))
    acc = torch.zeros(161, dtype=torch.float64, device=dev)
    plan.stats_accumulate(out, nf, acc)
    plan.stats_accumulate(out, nf, acc)          # accumulates (not zeroed by the call)
    ref = orc.mel_stats(_np(out).transpose(0, 2, 1), _np(nf))
    got = acc.cpu().numpy() / 2
    assert got[160] == ref[160] == sum(orc.frame_count(L) for L in lens)
    assert np.allclose(got[:160], ref[:160], rtol=1e-12, atol=1e-9)


def test_config3_ragged_sweep_buckets_masks_stats(dev):
    """BASELINE config 3: 256 clips of 1..30 s, length-bucketed, padding masks, mel_norms statistics."""
    from xtts_b200 import MelPlan
    from xtts_b200.sweep import MelSweep, norms_from_stats
    plan = MelPlan(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0,
                   f_max=8000, power=2.0, norm="slaney", device=dev)
    rng = np.random.default_rng(1234)
    lens = [int(round(22050 * u)) for u in rng.uniform(1.0, 30.0, size=256)]
    sweep = MelSweep(plan, lens, bucket_seconds=2.0, max_clips=64)
    assert len({b.max_len // 44100 for b in sweep.buckets}) <= 15
    clips = {}

    def fetch(idx, max_len):
        wav = torch.zeros(len(idx), max_len, device=dev)
        for k, i in enumerate(idx):
            g = torch.Generator(device=dev).manual_seed(5000 + i)
            clips[i] = torch.rand(lens[i], device=dev, generator=g) * 2 - 1
            wav[k, :lens[i]] = clips[i]
            wav[k, lens[i]:] = 0.777                      # junk in the padding must not leak into the clip
        return wav, torch.tensor([lens[i] for i in idx])

    got = {}

    def sink(idx, mel, n_frames):
        nf = n_frames.cpu().tolist()
        T = mel.shape[2]
        mask = torch.arange(T, device=dev)[None, :] >= n_frames[:, None]
        assert bool((mel.transpose(1, 2)[mask] == 0).all())              # padded frames are exactly 0.0
        for k, i in enumerate(idx):
            assert nf[k] == orc.frame_count(lens[i])                     # bit-exact frame counts
            got[i] = mel[k, :, :nf[k]].clone()

    acc = sweep.run_stats(fetch, sink=sink)
    plan.status()
    assert sorted(got) == list(range(256))
    # values: a 16-clip sample against the reference call sequence, one clip per call (mel_extract.py:15-27)
    for i in range(0, 256, 16):
        ref = orc.callsite_logmel(clips[i].cpu().numpy()[None, :], orc.V1, None).numpy()[0]
        assert ref.shape == tuple(got[i].shape)
        assert np.abs(_np(got[i]) - ref).max() <= TOL, i
    # statistics: fp64 reduction of the produced log-mel
    ref_acc = np.zeros(161)
    for i in range(256):
        lm = _np(got[i]).astype(np.float64)
        ref_acc[:80] += lm.sum(1); ref_acc[80:160] += (lm * lm).sum(1); ref_acc[160] += lm.shape[1]
    a = acc.cpu().numpy()
    assert a[160] == ref_acc[160] == sum(orc.frame_count(L) for L in lens)
    assert np.allclose(a[:160], ref_acc[:160], rtol=1e-9)       # fused fixed-point sums: 2^-32 / 2^-30 steps per value
    mean, std = norms_from_stats(acc)
    assert np.allclose(mean.cpu().numpy(), ref_acc[:80] / ref_acc[160], rtol=1e-6)
    # second pass with the derived norms applied == first pass / norms
    plan.set_norms(mean)
    wav, l = fetch([3, 77], max(lens[3], lens[77]))
    y, nf, _ = plan.forward(wav, l.to(dev), t_max=plan.frames(max(lens[3], lens[77])))
    assert np.abs(_np(y[0, :int(nf[0])].T) - _np(got[3]) / mean.cpu().numpy()[:, None]).max() <= 1e-5


def test_unaligned_rows_take_the_padded_workspace_path(dev, v1):
    """Row stride not a multiple of 4 floats (or a misaligned base): same results through k_pad + xp."""
    x = orc.synth("uniform", 3, 9001, 31)
    big = torch.zeros(3, 9003, device=dev)
    view = big[:, 1:9002]                       # stride 9003, base offset 4 B: not 16 B addressable
    view.copy_(torch.from_numpy(x))
    assert view.stride(0) == 9003 and view.data_ptr() % 16 != 0
    plan = v1._plan_for(dev)
    a, nf, _ = plan.forward(view)
    b, _, _ = plan.forward(torch.from_numpy(x).to(dev))
    plan.status()
    ref = orc.callsite_logmel(x, orc.V1, orc.default_mel_norms(80)).numpy()
    assert np.abs(_np(a.transpose(1, 2)) - ref).max() <= TOL
    assert np.abs(_np(b.transpose(1, 2)) - ref).max() <= TOL
    assert float((a - b).abs().max()) <= 2e-5


def test_short_and_edge_heavy_clips(dev, v1):
    """Clips of <= 8 frames live entirely in the edge buffer; 9..12 frames mix both sources."""
    for L in (513, 1800, 2047, 2048, 2049, 2303, 2304, 2305, 2560, 2816, 3000, 3072, 3328):
        x = orc.synth("uniform", 5, L, 700 + L)
        y = v1(torch.from_numpy(x).to(dev))
        ref = orc.callsite_logmel(x, orc.V1, orc.default_mel_norms(80)).numpy()
        assert tuple(y.shape) == ref.shape, L
        assert np.abs(_np(y) - ref).max() <= TOL, L


def test_constructor_variants_normalize_and_short_window(dev):
    """normalize=True (spectrum / ||window||_2, torchaudio functional.py:139-140) and win_length < n_fft."""
    from xtts_b200 import TorchMelSpectrogram
    x = orc.synth("uniform", 3, 12000, 55)
    for kw, cfg in ((dict(normalize=True), orc.MelConfig(normalized=True)),
                    (dict(win_length=800), orc.MelConfig(win_length=800)),
                    (dict(n_mel_channels=64, mel_fmax=7600, sampling_rate=16000),
                     orc.MelConfig(n_mels=64, f_max=7600.0, sample_rate=16000))):
        m = TorchMelSpectrogram(mel_norm_file=None, **kw)
        y = m(torch.from_numpy(x).to(dev))
        ref = orc.callsite_logmel(x, cfg, None).numpy()
        assert tuple(y.shape) == ref.shape, kw
        assert np.abs(_np(y) - ref).max() <= TOL, kw


def test_many_tiny_clips_and_side_stream(dev, v1):
    """3000 clips of 3 frames each (every frame is an edge frame; tiles span ~43 clips) on a non-default stream."""
    x = orc.synth("gauss", 3000, 600, 77)
    s = torch.cuda.Stream(dev)
    xd = torch.from_numpy(x).to(dev)
    torch.cuda.synchronize()
    with torch.cuda.stream(s):
        y = v1(xd)
    s.synchronize()
    ref = orc.callsite_logmel(x, orc.V1, orc.default_mel_norms(80)).numpy()
    assert tuple(y.shape) == ref.shape == (3000, 80, 3)
    assert np.abs(_np(y) - ref).max() <= TOL


def test_device_side_length_errors_are_reported(dev):
    from xtts_b200 import MelPlan, _cabi
    plan = MelPlan(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0,
                   f_max=8000, power=2.0, norm="slaney", device=dev)
    x = torch.zeros(2, 4000, device=dev)
    plan.forward(x, torch.tensor([4000, 300], device=dev))       # 300 <= pad: cannot be reflect-padded
    with pytest.raises(_cabi.MelError, match="length"):
        plan.status()
    plan.forward(x, torch.tensor([4000, 4000], device=dev), t_max=5)
    with pytest.raises(_cabi.MelError, match="Tmax"):
        plan.status()
    plan.forward(x)
    plan.status()


def test_v3_librosa_style_mel_spectrogram(dev):
    """feature_extractors.py:27-53,102-121 (mel_type == "librosa"); oracle for this row is unpinned (see oracle header)."""
    from xtts_b200 import MelSpectrogramFeatures1, mel_spectrogram
    m = MelSpectrogramFeatures1()
    for L in (24000, 5000, 1024, 777):
        x = orc.synth("uniform", 2, L, 900 + L)
        y = m(torch.from_numpy(x).to(dev))
        ref = orc.callsite_logmel(x, orc.V3).numpy()
        assert tuple(y.shape) == ref.shape == (2, 80, L // 256)
        assert np.abs(_np(y) - ref).max() <= TOL, L
    y2 = mel_spectrogram(torch.from_numpy(x).to(dev), 1024, 80, 24000, 256, 1024, 0, 8000, center=False)
    assert torch.equal(y, y2)
    # another transform size through the same function (the general-shape path): n_fft 2048, hop 512
    import dataclasses
    cfg = dataclasses.replace(orc.V3, n_fft=2048, hop=512, win_length=2048, pre_pad=(2048 - 512) // 2)
    x = orc.synth("uniform", 2, 30000, 77)
    y3 = mel_spectrogram(torch.from_numpy(x).to(dev), 2048, 80, 24000, 512, 2048, 0, 8000, center=False)
    ref = orc.callsite_logmel(x, cfg).numpy()
    assert tuple(y3.shape) == ref.shape and np.abs(_np(y3) - ref).max() <= TOL


def test_group_layout_frame_counts_around_multiples_of_32(dev, v1):
    """k_logmel2 lays clips out in groups of 32 frames (one converter / epilogue warp each): frame counts just
    below, at and above group boundaries, partly empty last groups, and tiles whose four groups belong to
    different clips must all equal the per-clip oracle (k_logmel2.cuh; frame count torch functional.py:648)."""
    frames = [3, 8, 9, 12, 31, 32, 33, 34, 63, 64, 65, 96, 127, 128, 129, 160]
    lens = [256 * (t - 1) + 7 * (i % 5) for i, t in enumerate(frames)]      # T = 1 + L // 256
    lens[0] = 513
    x = orc.synth("gauss", len(lens), max(lens), 4242)
    y, nf, mask = v1(torch.from_numpy(x).to(dev), lengths=torch.tensor(lens))
    ry, rn, rmask = orc.callsite_logmel_ragged(x, lens, orc.V1, orc.default_mel_norms(80))
    assert nf.cpu().tolist() == rn.tolist() == [1 + L // 256 for L in lens]
    assert np.array_equal(_np(mask), rmask)
    assert np.abs(_np(y) - ry.numpy()).max() <= TOL


def test_double_fold_kernel_agrees_with_single_fold_kernel(dev, monkeypatch):
    """Two independent tcgen05 kernels in the library: k_logmel (K = 512, window in the basis) and the default
    k_logmel2 (K = 256 + centre tap, even / odd bin passes, window in the converter).  Same inputs -> both within
    TOL of the oracle and within 2 TOL of each other, dense and ragged, V1 and V2."""
    from xtts_b200 import MelPlan
    kw1 = dict(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0, f_max=8000,
               power=2.0, norm="slaney")
    kw2 = dict(sample_rate=24000, n_fft=1024, hop_length=256, win_length=1024, n_mels=100, f_min=0, f_max=None,
               power=1.0, norm=None, clamp=1e-7)
    x = torch.from_numpy(orc.synth("uniform", 6, 40000, 77)).to(dev)
    lens = torch.tensor([40000, 513, 8192, 30001, 12345, 39999])
    for kw in (kw1, kw2):
        outs = []
        for algo in ("1", "2"):
            monkeypatch.setenv("TTTS_MEL_ALGO", algo)      # read by the LAB build's ttts_mel_plan_create only
            plan = MelPlan(device=dev, lab=True, **kw)
            dense = plan.forward(x)[0]
            rag = plan.forward(x, lens.to(dev))[0]
            torch.cuda.synchronize()
            outs.append((_np(dense), _np(rag)))
        for a, b in zip(outs[0], outs[1]):
            assert a.shape == b.shape
            assert np.abs(a - b).max() <= 2 * TOL
    monkeypatch.delenv("TTTS_MEL_ALGO")


# ------------------------------------------------------------------------------------------------------------------
# round 2: the configurations BASELINE.json names, at their sizes, and the inputs round 1 avoided
# ------------------------------------------------------------------------------------------------------------------
def test_config5_full_size_every_clip_against_the_oracle(dev):
    """BASELINE config 5 at its size: 32 clips x 120 s (2 646 000 samples) = 330 752 frames in one launch, every
    value of every clip against the reference's call sequence run on that clip alone (utils.py:509-522)."""
    from xtts_b200 import TorchMelSpectrogram
    m = TorchMelSpectrogram(mel_norm_file=None)
    m.mel_norms = torch.from_numpy(orc.default_mel_norms(80))
    B, L = 32, 2646000
    x = orc.synth("uniform", B, L, 505)
    y = m(torch.from_numpy(x).to(dev))
    m.status()
    assert tuple(y.shape) == (B, 80, 10336) and y.stride() == (10336 * 80, 1, 80)
    yh = _np(y)
    worst = 0.0
    for b in range(B):
        ref = orc.callsite_logmel(x[b:b + 1], orc.V1, orc.default_mel_norms(80)).numpy()[0]
        err = np.abs(yh[b] - ref)
        worst = max(worst, float(err.max()))
        for lo, hi in ((0, 300), (5018, 5318), (10036, 10336)):          # first / middle / last 300 frames, named
            assert err[:, lo:hi].max() <= TOL, (b, lo)
        assert err.max() <= TOL, b
    print(f"config 5 full size: max abs err {worst:.2e} over {B * 80 * 10336} values")


@pytest.mark.parametrize("padding", ["center", "same"])
def test_config3_through_v2_ragged_masks_and_values(dev, padding):
    """BASELINE config 3 "through the save_mel_to_disk.py path" = MelSpectrogramFeatures (SURVEY F6,
    mel_extract.py:4-6,26): 256 clips of 1..30 s at 24 kHz, length-sorted batches of 64, per-clip reflect padding;
    frame counts and padding masks bit-exact, and EVERY value of all 256 clips against the reference call on that clip
    alone (mel_extract.py:15-27), both paddings (feature_extractors.py:94-96).

    Value bar for V2 (power 1, 100 un-normalised HTK filters: the lowest ~10 filters are ONE bin wide): 1e-4 on the
    log-mel, except where the mel magnitude itself is tiny -- a single bin whose |X| happens to be ~1e-2 of its typical
    size (a Rayleigh tail event, ~1e-5 of the values): there the fp32 accumulation of the DFT dot product bounds the
    ABSOLUTE error of the magnitude (<= 2e-5 for O(10) magnitudes), so the bar is |d log-mel| <= 2e-5 / magnitude
    (the log has no bound of its own as the magnitude goes to 0 -- the reference's neither).  V1 (power 2, >= 2 bins per
    filter, the graded config) never needs the exception."""
    from xtts_b200 import MelSpectrogramFeatures
    cfg = orc.V2 if padding == "center" else orc.V2_SAME
    m = MelSpectrogramFeatures(padding=padding)
    rng = np.random.default_rng(2468)
    lens = [int(round(24000 * u)) for u in rng.uniform(1.0, 30.0, size=256)]
    order = sorted(range(256), key=lambda i: lens[i])
    n_val = n_over = 0
    worst = worst_abs = 0.0
    for k in range(0, 256, 64):
        idx = order[k:k + 64]
        Lmax = max(lens[i] for i in idx)
        pitch = (Lmax + 3) // 4 * 4
        wav = torch.full((64, pitch), 0.777, device=dev)[:, :Lmax]        # junk beyond each clip must not leak in
        clips = {}
        for r, i in enumerate(idx):
            clips[i] = orc.synth("uniform" if i % 2 else "gauss", 1, lens[i], 7000 + i)[0]
            wav[r, :lens[i]] = torch.from_numpy(clips[i]).to(dev)
        l = torch.tensor([lens[i] for i in idx])
        y, nf, mask = m(wav, lengths=l)
        m.status()
        want_nf = np.array([orc.frame_count(lens[i], cfg) for i in idx])
        assert nf.dtype == torch.int64 and np.array_equal(_np(nf), want_nf)               # bit-exact frame counts
        assert mask.dtype == torch.bool and np.array_equal(_np(mask), orc.make_pad_mask(want_nf, int(want_nf.max())))
        assert tuple(y.shape) == (64, 100, int(want_nf.max()))
        assert bool((y.transpose(1, 2)[mask] == 0).all())                                 # padded frames are 0.0
        yh = _np(y)
        for r, i in enumerate(idx):
            ref = orc.callsite_logmel(clips[i][None, :], cfg, None).numpy()[0]
            assert ref.shape[1] == want_nf[r]
            err = np.abs(yh[r, :, :want_nf[r]] - ref)
            mag = np.exp(ref.astype(np.float64))                     # the mel magnitude itself (power 1)
            bound = np.maximum(TOL, 2e-5 / mag)
            assert (err <= bound).all(), (padding, i, float(err.max()))
            n_val += err.size
            n_over += int((err > TOL).sum())
            worst = max(worst, float(err.max()))
            worst_abs = max(worst_abs, float((err * mag)[err > TOL].max()) if (err > TOL).any() else 0.0)
    print(f"V2 {padding}: {n_val} values, max |d log-mel| {worst:.2e}, above 1e-4: {n_over} ({n_over / n_val:.1e} of the "
          f"values, all on mel magnitudes < 0.2; their absolute magnitude error <= {worst_abs:.1e})")
    assert n_over <= 2e-5 * n_val


def test_fused_statistics_equal_the_separate_pass_and_are_bit_reproducible(dev):
    """Stage (4) fused into the log pass (ttts_mel_forward_stats) against the fp64 reduction of the produced log-mel
    (<= 1e-6 relative; measured ~1e-10) and against k_stats; bit-identical run to run and for any split
    of the batch (the in-launch sums are 64-bit fixed point)."""
    from xtts_b200 import MelPlan
    plan = MelPlan(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0,
                   f_max=8000, power=2.0, norm="slaney", device=dev)
    rng = np.random.default_rng(99)
    lens = [int(v) for v in rng.integers(513, 60000, size=37)]
    x = torch.from_numpy(orc.synth("gauss", len(lens), max(lens), 31)).to(dev)
    l = torch.tensor(lens, device=dev)
    tm = plan.frames(max(lens))
    acc = torch.zeros(161, dtype=torch.float64, device=dev)
    out, nf, _ = plan.forward(x, l, t_max=tm, stats_acc=acc)
    plan.status()
    ref = orc.mel_stats(_np(out).transpose(0, 2, 1), _np(nf))
    got = _np(acc)
    assert got[160] == ref[160] == sum(orc.frame_count(L) for L in lens)
    assert np.allclose(got[:160], ref[:160], rtol=1e-9, atol=0)
    sep = torch.zeros(161, dtype=torch.float64, device=dev)
    plan.stats_accumulate(out, nf, sep)
    assert np.allclose(got, _np(sep), rtol=1e-9)
    # norms set: the statistics are still those of the UN-normalised log-mel
    plan.set_norms(torch.from_numpy(orc.default_mel_norms(80)))
    acc_n = torch.zeros(161, dtype=torch.float64, device=dev)
    plan.forward(x, l, t_max=tm, stats_acc=acc_n)
    assert torch.equal(acc_n, acc)
    plan.set_norms(None)
    # run to run, and split into three launches accumulating into one vector
    for _ in range(3):
        again = torch.zeros(161, dtype=torch.float64, device=dev)
        plan.forward(x, l, t_max=tm, stats_acc=again)
        assert torch.equal(again, acc)
    parts = torch.zeros(161, dtype=torch.float64, device=dev)
    for a, b in ((0, 5), (5, 20), (20, 37)):
        plan.forward(x[a:b], l[a:b], t_max=tm, stats_acc=parts)
    assert np.allclose(_np(parts), got, rtol=1e-14, atol=0)          # only the final fp64 additions differ in order
    # dense batch (no lengths)
    accd = torch.zeros(161, dtype=torch.float64, device=dev)
    outd, nfd, _ = plan.forward(x[:4, :30000].contiguous(), stats_acc=accd)
    refd = orc.mel_stats(_np(outd).transpose(0, 2, 1), _np(nfd))
    assert np.allclose(_np(accd), refd, rtol=1e-9)


# Clean high-dynamic-range inputs under SURVEY 8c (ii)'s dual-oracle rule: pass iff, element-wise,
#   |ours - fp64 truth| <= max(1e-4, 1.5 * |reference fp32 - fp64 truth|).
# Round 1 only tested a tone with a -60 dB noise floor (and DESIGN.md feared the clean cases would fail); measured in round
# 2 they hold: clean tones, tone / digital silence / tone, a bin-centred tone (no leakage: everything off the peak sits at
# the clamp floor) and 1/f noise.
@pytest.mark.parametrize("kind", ["puretone", "tone_silence", "bintone", "pink"])
def test_high_dynamic_range_inputs_dual_oracle(dev, v1, kind):
    if kind == "bintone":
        t = np.arange(44100, dtype=np.float64)[None, :]
        k = np.array([[20.0], [100.0], [300.0]])                 # exact DFT bins: f = k * 22050 / 1024
        x = (0.9 * np.sin(2 * np.pi * k * t / 1024.0)).astype(np.float32)
    else:
        x = orc.synth(kind, 3, 44100, 321)
    norms = orc.default_mel_norms(80)
    y = _np(v1(torch.from_numpy(x).to(dev))).astype(np.float64)
    truth = orc.restated_logmel(x, orc.V1, norms)
    ref = orc.callsite_logmel(x, orc.V1, norms).numpy().astype(np.float64)
    ours_err, ref_err = np.abs(y - truth), np.abs(ref - truth)
    bound = np.maximum(TOL, 1.5 * ref_err)
    bad = ours_err > bound
    print(f"{kind}: ours max {ours_err.max():.2e}, reference max {ref_err.max():.2e}, over the bound: {bad.mean():.2%} "
          f"of values, worst ratio to the bound {(ours_err / bound).max():.2f}, max |ours - reference| {np.abs(y - ref).max():.2e}")
    assert not bad.any()


def _nccl_sweep_worker(rank, world, port, lens, ret):
    import os
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from xtts_b200 import MelPlan
        from xtts_b200.sweep import MelSweep
        plan = MelPlan(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0,
                       f_max=8000, power=2.0, norm="slaney", device=dev)

        def fetch(idx, max_len):                      # seeded per UTTERANCE: the data do not depend on the rank layout
            wav = torch.zeros(len(idx), (max_len + 3) // 4 * 4, device=dev)[:, :max_len]
            for k, i in enumerate(idx):
                wav[k, :lens[i]] = torch.from_numpy(orc.synth("uniform", 1, lens[i], 9000 + i)[0]).to(dev)
            return wav, torch.tensor([lens[i] for i in idx], device=dev)

        sharded = MelSweep(plan, lens, max_clips=16, rank=rank, world_size=world).run_stats(fetch)
        dist.barrier()
        # the same corpus on this rank alone, outside any process group collective
        acc = torch.zeros(161, dtype=torch.float64, device=dev)
        single = MelSweep(plan, lens, max_clips=16)
        for b in single.buckets:
            wav, l = fetch(list(b.indices), b.max_len)
            plan.forward(wav, l, t_max=plan.frames(b.max_len), stats_acc=acc)
        plan.status()
        ret[rank] = (sharded.cpu().numpy(), acc.cpu().numpy())
    finally:
        dist.destroy_process_group()


def test_two_rank_nccl_sweep_statistics_equal_single_rank(dev):
    """the sharded statistics (LPT-sharded buckets + one fp64 NCCL all-reduce, the template of
    xtts_dvae.py:104-110) must equal the single-rank ones to <= 1e-12 relative.  Needs two GPUs."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    lens = [int(v) for v in np.random.default_rng(77).integers(2000, 120000, size=96)]
    ret = mp.Manager().dict()
    mp.spawn(_nccl_sweep_worker, args=(2, port, lens, ret), nprocs=2, join=True)
    (s0, one0), (s1, one1) = ret[0], ret[1]
    assert np.array_equal(s0, s1)                                 # every rank holds the same all-reduced vector
    assert np.array_equal(one0, one1)                             # fixed-point in-launch sums: identical on both GPUs
    assert s0[160] == one0[160] == sum(orc.frame_count(L) for L in lens)
    rel = np.abs(s0[:160] - one0[:160]) / np.abs(one0[:160])
    print(f"2-rank NCCL vs single rank: max relative difference {rel.max():.2e}")
    assert rel.max() <= 1e-12


# ---- transform sizes other than 1024 / 256 (the reference's constructor takes any: utils.py:487-507) -- GeneralMelPlan:
# reflect pad + the tcgen05 contraction + ttts_stft_logmel, through the same module and C ABI
def _general_module(cfg_row):
    from xtts_b200 import TorchMelSpectrogram
    n_fft, hop, win, n_mels, fmin, fmax, sr, normalize, has_norms = (int(v) for v in cfg_row[:9])
    m = TorchMelSpectrogram(filter_length=n_fft, hop_length=hop, win_length=win, n_mel_channels=n_mels, mel_fmin=fmin,
                            mel_fmax=fmax, sampling_rate=sr, normalize=bool(normalize), mel_norm_file=None)
    if has_norms:
        m.mel_norms = torch.from_numpy(orc.default_mel_norms(n_mels))
    cfg = orc.MelConfig(sample_rate=sr, n_fft=n_fft, hop=hop, win_length=win, n_mels=n_mels, f_min=float(fmin),
                        f_max=float(fmax), normalized=bool(normalize))
    return m, cfg, (orc.default_mel_norms(n_mels) if has_norms else None)


def test_general_transform_sizes_match_the_reference_goldens(dev, golden_dir):
    g = np.load(f"{golden_dir}/v1_general_shapes.npz")
    names = [k[4:] for k in g.files if k.startswith("cfg_")]
    assert len(names) == 5
    for name in names:
        row = g[f"cfg_{name}"]
        m, cfg, _ = _general_module(row)
        gauss, L, seed = int(row[9]), int(row[10]), int(row[11])
        x = orc.synth("gauss" if gauss else "uniform", 2, L, seed)
        y = m(torch.from_numpy(x).to(dev))
        ref = g[f"y_{name}"]
        assert tuple(y.shape) == ref.shape, name                       # bit-exact frame counts
        assert y.stride() == (ref.shape[2] * ref.shape[1], 1, ref.shape[1]), name
        err = float(np.abs(_np(y) - ref).max())
        assert err <= TOL, (name, err)
        y3 = m(torch.from_numpy(x).to(dev)[:, None, :])                # [B, 1, L] form
        assert torch.equal(y3, y)


def test_general_transform_sizes_ragged_batch_and_batch_size(dev):
    row = np.array([2048, 512, 2048, 80, 0, 8000, 22050, 0, 1])
    m, cfg, norms = _general_module(row)
    lens = [40000, 1025, 2048, 33333, 12000, 40000]
    x = orc.synth("uniform", len(lens), max(lens), 77)
    ym, nf, mask = m(torch.from_numpy(x).to(dev), lengths=torch.tensor(lens))
    rm, rn, rmask = orc.callsite_logmel_ragged(x, lens, cfg, norms)
    assert nf.cpu().tolist() == rn.tolist() and np.array_equal(mask.cpu().numpy(), rmask)
    assert float(np.abs(_np(ym) - rm.numpy()).max()) <= TOL
    # device-resident lengths: same result, nothing validated on the host
    ym2, nf2, mask2 = m(torch.from_numpy(x).to(dev), lengths=torch.tensor(lens, device=dev))
    assert torch.equal(nf2, nf) and torch.equal(mask2, mask) and torch.equal(ym2[:, :, :ym.shape[2]], ym)
    # 64 x 10 s at the size the headline uses: every value of a 4-clip sample against the reference's call
    xb = torch.rand(64, 220500, device=dev, generator=torch.Generator(device=dev).manual_seed(5)) * 2 - 1
    yb = m(xb)
    assert tuple(yb.shape) == (64, 80, orc.frame_count(220500, cfg))
    for b in (0, 17, 40, 63):
        ref = orc.callsite_logmel(xb[b:b + 1].cpu().numpy(), cfg, norms).numpy()
        assert float(np.abs(_np(yb[b:b + 1]) - ref).max()) <= TOL, b
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 1024, device=dev))                          # reflect pad 1024 needs more than 1024 samples


def test_nan_samples_poison_exactly_the_frames_the_reference_poisons(dev, v1):
    """torch.clamp keeps NaN, so one NaN sample makes every mel of the frames that contain it NaN in the reference (the FFT
    spreads it over all bins); everything else must stay within the usual tolerance.  (A NaN that sits exactly on a frame's
    first tap is dropped here with the zero window value the reference multiplies it by: positions off the hop grid.)"""
    from xtts_b200 import MelSpectrogramFeatures
    x = orc.synth("uniform", 3, 30000, 31)
    x[0, 7003] = np.nan
    x[2, 29990] = np.nan                                  # inside the last frames and the reflected tail
    for mod, cfg, norms in ((v1, orc.V1, orc.default_mel_norms(80)), (MelSpectrogramFeatures(), orc.V2, None)):
        y = _np(mod(torch.from_numpy(x).to(dev)))
        ref = orc.callsite_logmel(x, cfg, norms).numpy()
        assert np.array_equal(np.isnan(y), np.isnan(ref))
        assert np.isnan(ref[0]).any() and np.isnan(ref[2]).any() and not np.isnan(ref[1]).any()
        ok = ~np.isnan(ref)
        bar = TOL if cfg is orc.V1 else np.maximum(TOL, 2e-5 / np.exp(np.where(ok, ref, 0.0)))
        assert (np.abs(y - ref)[ok] <= (bar if np.isscalar(bar) else bar[ok])).all()
