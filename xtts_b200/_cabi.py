"""ctypes binding of the C ABI in ``include/ttts_mel.h``, ``include/ttts_ingest.h`` and ``include/ttts_vq.h`` (``libttts_mel.so``, built in-tree).

There is no CPU implementation behind this module: if the shared library is missing or a call
fails, an exception is raised.

``lib()`` is the product library.  ``lab()`` is the same source compiled with ``-DTTTS_MEL_LAB``
(``libttts_mel_lab.so``, ``include/ttts_mel_lab.h``): it adds the cross-check kernels, the DFT-stage dump,
the leg-removal switch and the environment knobs; only ``tests/`` and ``tools/`` load it.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import (POINTER, Structure, c_char_p, c_double, c_float, c_int32, c_int64, c_size_t,
                    c_uint8, c_void_p)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libttts_mel.so")
LAB_PATH = os.path.join(_HERE, "libttts_mel_lab.so")

OK, EINVAL, ECUDA, ENOMEM, EDEVICE = 0, -1, -2, -3, -4


class MelConfig(Structure):
    _fields_ = [("n_fft", c_int32), ("hop", c_int32), ("n_mels", c_int32), ("pad", c_int32),
                ("power", c_float), ("clamp", c_float), ("spec_scale", c_float), ("power_eps", c_float)]


class MelError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"ttts_mel error {code}: {msg}")
        self.code = code


_FWD_ARGS = [c_void_p, c_void_p, c_int64, c_void_p, c_int32, c_int64, c_int32, c_void_p, c_void_p,
             c_void_p, c_void_p, c_size_t, c_void_p]

SYMBOLS = {
    "ttts_mel_plan_create": (ctypes.c_int, [POINTER(MelConfig), c_void_p, c_void_p, POINTER(c_void_p)]),
    "ttts_mel_plan_set_norms": (ctypes.c_int, [c_void_p, c_void_p]),
    "ttts_mel_plan_destroy": (None, [c_void_p]),
    "ttts_mel_frames": (c_int64, [c_void_p, c_int64]),
    "ttts_mel_workspace_bytes": (c_size_t, [c_void_p, c_int32, c_int64]),
    "ttts_mel_workspace_bytes_for": (c_size_t, [c_void_p, c_int32, c_int64, c_void_p, c_int64]),
    "ttts_mel_forward": (ctypes.c_int, _FWD_ARGS),
    "ttts_mel_forward_stats": (ctypes.c_int, _FWD_ARGS + [c_void_p]),
    "ttts_mel_forward_static": (ctypes.c_int, _FWD_ARGS[:10] + [c_void_p, POINTER(c_void_p)]),
    "ttts_mel_config_size": (c_size_t, []),
    "ttts_mel_first_bin": (c_int32, [c_void_p]),
    "ttts_mel_num_bins": (c_int32, [c_void_p]),
    "ttts_mel_live_bins": (c_int32, [c_void_p]),
    "ttts_mel_status": (ctypes.c_int, [c_void_p, c_void_p]),
    "ttts_mel_stats_accumulate": (ctypes.c_int, [c_void_p, c_void_p, c_void_p, c_int32, c_int32, c_void_p, c_void_p]),
    "ttts_mel_last_launch_count": (c_int32, []),
    # include/ttts_mel_lab.h, measurement hooks (both libraries)
    "ttts_mel_profile_enable": (ctypes.c_int, [c_void_p, c_int32]),
    "ttts_mel_profile_collect": (ctypes.c_int, [c_void_p, c_void_p, c_int32]),
    # include/ttts_ingest.h
    "ttts_pcm16_to_f32": (ctypes.c_int, [c_void_p, c_int64, c_int32, c_void_p, c_int32, c_int64, c_void_p, c_int64, c_void_p]),
    "ttts_resample_plan_create": (ctypes.c_int, [c_void_p, c_int32, c_int32, c_int32, POINTER(c_void_p)]),
    "ttts_resample_plan_destroy": (None, [c_void_p]),
    "ttts_resample_out_len": (c_int64, [c_void_p, c_int64]),
    "ttts_resample_forward": (ctypes.c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int32, c_int64, c_void_p,
                                             c_int64, c_int32, c_void_p]),
    "ttts_resample_pcm16_forward": (ctypes.c_int, [c_void_p, c_void_p, c_int64, c_int32, c_void_p, c_int32, c_int64,
                                                   c_void_p, c_int64, c_int32, c_void_p]),
    # include/ttts_stft.h
    "ttts_reflect_pad": (ctypes.c_int, [c_void_p, c_int64, c_void_p, c_int32, c_int64, c_int32, c_void_p, c_int64, c_void_p]),
    "ttts_stft_magnitude": (ctypes.c_int, [c_void_p, c_int64, c_int32, c_int32, c_int32, c_int32, c_int32, c_float, c_float,
                                           c_int32, c_float, c_void_p, c_void_p]),
    "ttts_stft_logmel": (ctypes.c_int, [c_void_p, c_int64, c_int32, c_int32, c_int32, c_int32, c_void_p, c_int32, c_int32,
                                        c_void_p, c_void_p, c_int32, c_float, c_float, c_float, c_void_p, c_void_p, c_void_p]),
    "ttts_spec_pack": (ctypes.c_int, [c_void_p, c_int32, c_int32, c_int32, c_void_p, c_void_p]),
    "ttts_istft_overlap_add": (ctypes.c_int, [c_void_p, c_int32, c_int32, c_int32, c_int32, c_int32, c_void_p, c_void_p,
                                              c_int64, c_void_p, c_void_p]),
    # include/ttts_vq.h
    "ttts_vq_plan_create": (ctypes.c_int, [c_void_p, c_int32, c_int32, POINTER(c_void_p)]),
    "ttts_vq_plan_destroy": (None, [c_void_p]),
    "ttts_vq_workspace_bytes": (c_size_t, [c_void_p, c_int64]),
    "ttts_vq_nearest": (ctypes.c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p,
                                       c_void_p, c_size_t, c_void_p]),
    "ttts_vq_nearest_bct": (ctypes.c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_void_p,
                                           c_void_p, c_size_t, c_void_p]),
    "ttts_vq_embed_code": (ctypes.c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p]),
    "ttts_vq_status": (ctypes.c_int, [c_void_p, c_void_p]),
    "ttts_mel_last_error": (c_char_p, []),
    "ttts_mel_version": (c_char_p, []),
}

# include/ttts_mel_lab.h, lab build only
LAB_SYMBOLS = {
    "ttts_mel_forward_simt": (ctypes.c_int, _FWD_ARGS),
    "ttts_mel_forward_debug": (ctypes.c_int, _FWD_ARGS + [c_void_p]),
    "ttts_mel_set_experiment": (None, [c_int32]),
}

_lib = None
_lab = None


def _load(path: str, symbols: dict) -> ctypes.CDLL:
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  There is no CPU fallback for this path.")
    handle = ctypes.CDLL(path)
    for name, (res, args) in symbols.items():
        fn = getattr(handle, name)      # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    return handle


_use_lab = False


def select_lab(on: bool) -> None:
    """tests/ and tools/ only: make ``lib()`` return the lab build, so that the ingestion / codebook plans honour the
    kernel-selection environment knobs (objects created before the switch keep their library)."""
    global _use_lab
    _use_lab = bool(on)


def lib() -> ctypes.CDLL:
    """Load the product library once; raise loudly when it has not been built."""
    global _lib
    if _use_lab:
        return lab()
    if _lib is None:
        _lib = _load(LIB_PATH, SYMBOLS)
    return _lib


def lab() -> ctypes.CDLL:
    """The lab build (tests/ and tools/ only): every product symbol plus include/ttts_mel_lab.h's."""
    global _lab
    if _lab is None:
        _lab = _load(LAB_PATH, {**SYMBOLS, **LAB_SYMBOLS})
    return _lab


def check(rc: int) -> None:
    if rc != OK:
        raise MelError(int(rc), lib().ttts_mel_last_error().decode("utf-8", "replace"))
