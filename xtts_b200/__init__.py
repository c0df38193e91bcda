"""xtts_b200 -- B200-native (sm_100a) batched log-mel front-end of TTTS / xtts.

Only the hot path: ``TorchMelSpectrogram`` / ``MelSpectrogramFeatures`` / ``dvae_wav_to_mel`` /
``make_pad_mask`` (host mirror of the reference interface, ``mel.py``) on top of
``libttts_mel.so`` (``csrc/ttts_mel.cu`` behind the C ABI of ``include/ttts_mel.h``).
"""
from .mel import (DEFAULT_MEL_NORM_FILE, MelPlan, MelSpectrogramFeatures, MelSpectrogramFeatures1,
                  TorchMelSpectrogram, dvae_wav_to_mel, make_pad_mask, mel_spectrogram)

__all__ = ["DEFAULT_MEL_NORM_FILE", "MelPlan", "MelSpectrogramFeatures", "MelSpectrogramFeatures1",
           "TorchMelSpectrogram", "dvae_wav_to_mel", "make_pad_mask", "mel_spectrogram"]
