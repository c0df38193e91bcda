"""xtts_b200 -- B200-native (sm_100a) batched log-mel front-end of TTTS / xtts.

Only the hot path: ``TorchMelSpectrogram`` / ``MelSpectrogramFeatures`` / ``dvae_wav_to_mel`` /
``make_pad_mask`` (host mirror of the reference interface, ``mel.py``) on top of
``libttts_mel.so`` (``csrc/ttts_mel.cu`` behind the C ABI of ``include/ttts_mel.h``), and the ingestion step in
front of it (``ingest.py``: int16 PCM -> float32 channel 0 -> sinc resample -> clip, ``include/ttts_ingest.h``)
and the DVAE codebook search behind it (``vq.py``: ``Quantize`` in eval mode, ``include/ttts_vq.h``), and the STFT family
next to it (``stft.py``: ``TorchSTFT`` / ``STFTLoss`` / ``MultiScaleSTFTLoss`` forward values, Vocos ``ISTFT``; ``include/ttts_stft.h``).
"""
from .ingest import Resampler, decode_pcm16, get_prompt_slice, load_audio_pcm16, sinc_resample_taps
from .stft import ISTFT, MultiScaleSTFTLoss, STFTLoss, TorchSTFT
from .vq import Quantize, VQPlan, get_codebook_indices
from .mel import (DEFAULT_MEL_NORM_FILE, MelPlan, MelSpectrogramFeatures, MelSpectrogramFeatures1,
                  TorchMelSpectrogram, dvae_wav_to_mel, make_pad_mask, mel_spectrogram)

__all__ = ["DEFAULT_MEL_NORM_FILE", "MelPlan", "MelSpectrogramFeatures", "MelSpectrogramFeatures1",
           "TorchMelSpectrogram", "dvae_wav_to_mel", "make_pad_mask", "mel_spectrogram",
           "Resampler", "decode_pcm16", "get_prompt_slice", "load_audio_pcm16", "sinc_resample_taps",
           "Quantize", "VQPlan", "get_codebook_indices",
           "TorchSTFT", "STFTLoss", "MultiScaleSTFTLoss", "ISTFT"]
