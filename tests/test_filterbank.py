"""Host constants of the product path are bit-identical to what the reference's torchaudio builds."""
import torch
import torchaudio

from oracle import mel_oracle as orc
from xtts_b200.filterbank import analysis_window, melscale_fbanks


def test_v1_filterbank_bit_identical():
    ta = torchaudio.transforms.MelSpectrogram(n_fft=1024, hop_length=256, win_length=1024, power=2, normalized=False,
                                              sample_rate=22050, f_min=0, f_max=8000, n_mels=80, norm="slaney")
    fb = melscale_fbanks(513, 0.0, 8000.0, 80, 22050, "slaney", "htk")
    assert torch.equal(fb, ta.mel_scale.fb)
    assert torch.equal(analysis_window(1024, 1024), ta.spectrogram.window)
    nz = fb.ne(0)
    assert int(nz.sum()) == 729 and int(nz.any(1).nonzero().min()) == 1 and int(nz.any(1).nonzero().max()) == 371


def test_v2_filterbank_bit_identical():
    ta = torchaudio.transforms.MelSpectrogram(sample_rate=24000, n_fft=1024, hop_length=256, win_length=None, power=1,
                                              f_min=0, f_max=None, n_mels=100)
    fb = melscale_fbanks(513, 0.0, 12000.0, 100, 24000, None, "htk")
    assert torch.equal(fb, ta.mel_scale.fb)


def test_slaney_scale_variant_bit_identical():
    ta = torchaudio.functional.melscale_fbanks(513, 0.0, 8000.0, 80, 22050, "slaney", "slaney")
    assert torch.equal(melscale_fbanks(513, 0.0, 8000.0, 80, 22050, "slaney", "slaney"), ta)


def test_fp64_restatement_close():
    fb = melscale_fbanks(513, 0.0, 8000.0, 80, 22050, "slaney", "htk").double().numpy()
    assert abs(fb - orc.melscale_fbanks64(orc.V1)).max() < 2e-7     # fp32 rounding of weights up to 0.043


def test_two_adjacent_filters_per_bin():
    """The kernel's sparse mel table relies on this structure of triangular filterbanks."""
    for fb in (melscale_fbanks(513, 0.0, 8000.0, 80, 22050, "slaney", "htk"),
               melscale_fbanks(513, 0.0, 12000.0, 100, 24000, None, "htk")):
        prev = 0
        for row in fb:
            idx = row.nonzero().flatten().tolist()
            assert len(idx) <= 2
            if idx:
                assert idx[-1] - idx[0] <= 1 and idx[0] >= prev
                prev = idx[0]


def test_short_window_is_centred():
    w = analysis_window(1024, 800)
    assert w.shape == (1024,) and float(w[:112].abs().max()) == 0.0 and float(w[912:].abs().max()) == 0.0
