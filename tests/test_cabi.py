"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol that
include/pac_b200.h declares (no compute calls: there is no GPU here)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    lib = ctypes.CDLL(os.path.join(ROOT, "audio_codec_b200", "libpac_b200.so"))
    hdr = open(os.path.join(ROOT, "include", "pac_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(pac_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 12
    for n in sorted(names):
        assert hasattr(lib, n), f"{n} declared in pac_b200.h but not exported"
    from audio_codec_b200 import _lib
    assert set(_lib.EXPORTS) == names


def test_no_cpu_fallback_and_no_oracle_in_product():
    """The product package never imports the oracle, and refuses to run without CUDA."""
    import pytest
    import torch
    pkg = os.path.join(ROOT, "audio_codec_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in src.replace("oracle/", ""), fn
    if not torch.cuda.is_available():
        from audio_codec_b200 import _lib
        with pytest.raises(_lib.PacError):
            _lib.Context(48000)


def test_error_strings():
    from audio_codec_b200 import _lib
    assert _lib.lib.pac_strerror(0) == b"ok"
    assert b"argument" in _lib.lib.pac_strerror(-1)
    assert _lib.lib.pac_create(0, None, None) == -1


def test_create_refuses_configurations_the_quantisers_cannot_hold():
    """pac_create argument checks run before any CUDA call: n_scale_bits = 6 needs 63 + 16 magnitude bits
    (the reference uses Python bigints, coder/quantize.py:99-125; the device quantisers are 64-bit), so it is
    refused as unsupported instead of silently coding garbage; so does a .pac header carrying it."""
    import ctypes as C
    from audio_codec_b200 import _lib
    h = C.c_void_p()
    for nS, want in ((6, -3), (8, -3), (9, -1), (0, -1)):
        cfg = _lib.PacConfig(48000, 2, 1024, nS, 12, 0, 1, 128 / 48.0, 1)
        assert _lib.lib.pac_create(0, C.byref(cfg), C.byref(h)) == want, nS
    cfg = _lib.PacConfig(48000, 2, 1024, 4, 12, 0, 1, 128 / 48.0, 7)          # bad precision: refused before allocating
    assert _lib.lib.pac_create(0, C.byref(cfg), C.byref(h)) == -1
    assert _lib.lib.pac_set_option(None, 0, 1.0) == -1
