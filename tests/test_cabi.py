"""The C-ABI library builds for sm_100a, loads, and exports every symbol include/*.h declares.
No compute calls here (no GPU in this container)."""
import ctypes
import os
import re
import subprocess

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    names = set()
    for h in ("ttts_mel.h", "ttts_ingest.h", "ttts_vq.h", "ttts_stft.h"):
        names |= _symbols_of(h)
    names |= _symbols_of("ttts_mel_lab.h", lab=False)        # the measurement hooks both libraries export
    return sorted(names)


def _symbols_of(header, lab=None):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    if lab is not None:                                      # split ttts_mel_lab.h at its #ifdef TTTS_MEL_LAB block
        m = re.search(r"#ifdef TTTS_MEL_LAB\n(.*?)#endif", src, flags=re.S)
        src = m.group(1) if lab else src.replace(m.group(0), "")
    return set(re.findall(r"\b(ttts_[a-z0-9_]+)\s*\(", src))


def test_library_exports_every_declared_symbol(built_lib):
    names = _header_symbols()
    assert len(names) >= 23
    for n in names:
        assert getattr(built_lib, n) is not None, n
    from xtts_b200 import _cabi
    assert sorted(_cabi.SYMBOLS) == names       # the ctypes table and the header agree


def test_lab_library_adds_only_the_lab_header(built_lib):
    """libttts_mel_lab.so = the product symbols + include/ttts_mel_lab.h's lab block; the product library exports none of
    the lab block (no leg-removal switch, no environment knob can change what it computes)."""
    from xtts_b200 import _cabi
    lab_only = sorted(_symbols_of("ttts_mel_lab.h", lab=True))
    assert lab_only == sorted(_cabi.LAB_SYMBOLS) and "ttts_mel_set_experiment" in lab_only
    lab = _cabi.lab()
    for n in lab_only:
        assert getattr(lab, n) is not None
        assert not hasattr(built_lib, n), n
    strings = subprocess.run(["strings", _cabi.LIB_PATH], capture_output=True, text=True).stdout
    assert "TTTS_MEL_ALGO" not in strings and "TTTS_VQ_EXP" not in strings and "TTTS_RESAMPLE_EXP" not in strings


def test_integration_stub_matches_the_header(built_lib):
    """the ctypes block a maintainer would paste is executed against the built library
    (round 1 shipped a 7-field struct there while the header had 8)."""
    from xtts_b200 import _cabi
    md = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    block = next(b for b in re.findall(r"```python\n(.*?)```", md, flags=re.S) if "class _Cfg(Structure)" in b)
    ns = {}
    exec(block.replace('ctypes.CDLL("libttts_mel.so")', f'ctypes.CDLL({_cabi.LIB_PATH!r})'), ns)
    assert ctypes.sizeof(ns["_Cfg"]) == built_lib.ttts_mel_config_size() == ctypes.sizeof(_cabi.MelConfig) == 32
    assert [f[0] for f in ns["_Cfg"]._fields_] == [f[0] for f in _cabi.MelConfig._fields_]
    hdr = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "ttts_mel.h")).read(), flags=re.S)
    body = re.search(r"typedef struct ttts_mel_config \{(.*?)\} ttts_mel_config;", hdr, flags=re.S).group(1)
    assert re.findall(r"\b(\w+);", body) == [f[0] for f in _cabi.MelConfig._fields_]


def test_library_is_sm100a_with_tcgen05_and_tma(built_lib):
    from xtts_b200 import _cabi
    sass = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-sass", _cabi.LIB_PATH], capture_output=True, text=True)
    if sass.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    out = sass.stdout
    assert "sm_100a" in out or "SM100a" in out or "sm_100" in out
    assert "UTCHMMA" in out           # tcgen05.mma
    assert "LDTM" in out and "STTM" in out    # tcgen05.ld / tcgen05.st
    assert "UBLKCP" in out            # cp.async.bulk (TMA)
    assert "HMMA." not in out.replace("UTCHMMA", "")   # no legacy mma.sync path


def test_host_only_entry_points(built_lib):
    assert built_lib.ttts_mel_version().decode().startswith("ttts_mel")
    assert built_lib.ttts_mel_workspace_bytes(None, 4, 1000) == 0
    assert built_lib.ttts_mel_frames(None, 1000) < 0
    assert b"null plan" in built_lib.ttts_mel_last_error()


def test_product_path_has_no_cpu_fallback():
    from xtts_b200 import MelSpectrogramFeatures, TorchMelSpectrogram, dvae_wav_to_mel
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        TorchMelSpectrogram(mel_norm_file=None)(torch.zeros(1, 4000))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        MelSpectrogramFeatures()(torch.zeros(1, 4000))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        dvae_wav_to_mel(torch.zeros(1, 4000), mel_norms=torch.ones(80), device=torch.device("cpu"))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "xtts_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("dual-oracle", ""), f
                assert "torchaudio" not in "\n".join(l for l in text.splitlines() if l.strip().startswith(("import", "from"))), f


def test_reference_constructor_surface():
    import inspect
    from xtts_b200 import MelSpectrogramFeatures, TorchMelSpectrogram
    sig = inspect.signature(TorchMelSpectrogram.__init__)
    assert list(sig.parameters)[1:] == ["filter_length", "hop_length", "win_length", "n_mel_channels", "mel_fmin",
                                        "mel_fmax", "sampling_rate", "normalize", "mel_norm_file"]
    d = {k: v.default for k, v in sig.parameters.items() if k != "self"}
    assert (d["filter_length"], d["hop_length"], d["win_length"], d["n_mel_channels"], d["mel_fmin"], d["mel_fmax"],
            d["sampling_rate"], d["normalize"]) == (1024, 256, 1024, 80, 0, 8000, 22050, False)
    with pytest.raises(FileNotFoundError):
        TorchMelSpectrogram()        # the default mel_norms.pth is not shipped -- same as the reference
    m = TorchMelSpectrogram(mel_norm_file=None)
    for a in ("filter_length", "hop_length", "win_length", "n_mel_channels", "mel_fmin", "mel_fmax", "sampling_rate",
              "mel_norm_file", "mel_norms"):
        assert hasattr(m, a)
    sig2 = inspect.signature(MelSpectrogramFeatures.__init__)
    assert list(sig2.parameters)[1:] == ["sample_rate", "n_fft", "hop_length", "win_length", "n_mels", "mel_fmin",
                                         "mel_fmax", "normalize", "padding"]
    with pytest.raises(ValueError):
        MelSpectrogramFeatures(padding="valid")
