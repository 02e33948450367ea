"""ctypes binding of libpac_b200.so (include/pac_b200.h).

There is NO CPU fallback: importing this module without the built library, or
calling an op without a CUDA device, raises.
"""
import ctypes as C
import os
from pathlib import Path

import numpy as np

_PKG = Path(__file__).resolve().parent
_LIB_PATH = _PKG / os.environ.get("PAC_LIB_NAME", "libpac_b200.so")   # import-time only: kernel experiments (tools/)

F32, F64 = 0, 1
W_SINE, W_START, W_STOP, W_STARTSTOP, W_KAISER, W_NONE, W_HANN = range(7)
PSY_SPL, PSY_INTENSITY, PSY_BARK, PSY_THRESH = range(4)


class PacConfig(C.Structure):
    _fields_ = [("sample_rate", C.c_int), ("n_channels", C.c_int), ("n_mdct_lines", C.c_int),
                ("n_scale_bits", C.c_int), ("n_mant_size_bits", C.c_int), ("use_sbr", C.c_int),
                ("use_vq", C.c_int), ("target_bits_per_sample", C.c_double), ("precision", C.c_int)]


class PacError(RuntimeError):
    pass


def _load():
    if not _LIB_PATH.exists():
        raise ImportError(
            f"{_LIB_PATH} is missing: build it with `python -m audio_codec_b200.build` "
            "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(str(_LIB_PATH))
    vp, ll, i, d = C.c_void_p, C.c_longlong, C.c_int, C.c_double
    pi = C.POINTER(C.c_int)
    sig = {
        "pac_create": (i, [i, C.POINTER(PacConfig), C.POINTER(vp)]),
        "pac_destroy": (None, [vp]),
        "pac_strerror": (C.c_char_p, [i]),
        "pac_set_option": (i, [vp, i, d]),
        "pac_get_bands": (i, [vp, i, C.POINTER(i), C.POINTER(i)]),
        "pac_mdct_forward": (i, [vp, i, vp, ll, i, i, i, i, vp, vp]),
        "pac_imdct_ola": (i, [vp, i, vp, i, i, i, i, vp, ll, vp]),
        "pac_mdct_direct": (i, [vp, i, vp, ll, i, i, i, vp, vp]),
        "pac_header_bytes": (i, [vp]),
        "pac_encode_bound": (ll, [vp, C.POINTER(i), i]),
        "pac_encode_streams": (i, [vp, vp, C.POINTER(i), C.POINTER(i), i, vp, ll, C.POINTER(ll), C.POINTER(i), vp]),
        "pac_encode_host": (i, [vp, vp, C.POINTER(i), C.POINTER(i), i, vp, ll, C.POINTER(ll), C.POINTER(i), vp]),
        "pac_decode_host": (i, [vp, vp, C.POINTER(ll), i, vp, ll, C.POINTER(ll), C.POINTER(i), vp]),
        "pac_last_kernel_ms": (i, [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
        "pac_kernel_stats": (i, [vp, i, C.POINTER(C.c_float), C.POINTER(i)]),
        "pac_kernel_launches": (ll, [vp]),
        "pac_decode_plan": (i, [vp, vp, C.POINTER(ll), i, C.POINTER(ll), vp]),
        "pac_decode_streams": (i, [vp, vp, C.POINTER(ll), i, vp, ll, C.POINTER(ll), C.POINTER(i), vp]),
        # per-function ops
        "pac_window": (i, [vp, i, vp, ll, i, i, i, d, vp, vp]),
        "pac_transient_detect": (i, [vp, vp, ll, i, i, d, vp, vp]),
        "pac_calc_smr": (i, [vp, i, vp, vp, vp, ll, i, i, pi, pi, vp, vp, vp]),
        "pac_psy_map": (i, [vp, i, vp, ll, vp, vp]),
        "pac_pcm_to_float": (i, [vp, vp, ll, vp, vp]),
        "pac_selftest_division": (ll, [vp, ll, C.c_ulonglong, vp]),
        "pac_bit_alloc": (i, [vp, vp, i, i, pi, vp, ll, vp, vp]),
        "pac_quantize_uniform": (i, [vp, vp, ll, i, vp, vp]),
        "pac_dequantize_uniform": (i, [vp, vp, ll, i, vp, vp]),
        "pac_scale_factor": (i, [vp, vp, ll, i, i, vp, vp]),
        "pac_mantissa": (i, [vp, vp, ll, i, i, i, vp, vp]),
        "pac_dequantize": (i, [vp, i, vp, ll, i, i, vp, vp]),
        "pac_gain_shape_quantize": (i, [vp, vp, i, ll, vp, i, vp, vp, vp, i, pi, vp]),
        "pac_gain_shape_dequantize": (i, [vp, vp, ll, vp, vp, i, ll, vp, vp, pi, vp]),
        "pac_gain_shape_dequantize_static": (i, [vp, vp, ll, vp, vp, i, ll, vp, vp, pi, vp]),
        "pac_gain_shape_dequantize_static_max_vectors": (ll, [i]),
        "pac_pack_bits": (i, [vp, vp, vp, vp, i, ll, i, vp, pi, vp]),
        "pac_unpack_bits": (i, [vp, vp, i, vp, vp, vp, i, ll, vp, vp]),
        "pac_record_stride": (i, []),
        "pac_encode_blocks": (i, [vp, vp, vp, ll, vp, vp, vp, pi, vp]),
        "pac_encode_frames": (i, [vp, vp, i, vp, i, ll, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, pi, vp]),
        "pac_decode_blocks": (i, [vp, vp, i, vp, ll, vp, pi, vp]),
        "pac_decode_frames": (i, [vp, vp, i, vp, i, vp, ll, vp, vp, pi, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()
EXPORTS = ["pac_create", "pac_destroy", "pac_strerror", "pac_set_option", "pac_get_bands", "pac_mdct_forward", "pac_imdct_ola", "pac_mdct_direct",
           "pac_header_bytes", "pac_encode_bound", "pac_encode_streams", "pac_decode_plan", "pac_decode_streams", "pac_last_kernel_ms",
           "pac_kernel_stats", "pac_kernel_launches",
           "pac_encode_host", "pac_decode_host",
           "pac_window", "pac_transient_detect", "pac_calc_smr", "pac_psy_map", "pac_pcm_to_float", "pac_selftest_division", "pac_bit_alloc", "pac_quantize_uniform",
           "pac_dequantize_uniform", "pac_scale_factor", "pac_mantissa", "pac_dequantize", "pac_gain_shape_quantize",
           "pac_gain_shape_dequantize", "pac_gain_shape_dequantize_static", "pac_gain_shape_dequantize_static_max_vectors",
           "pac_pack_bits", "pac_unpack_bits", "pac_record_stride", "pac_encode_blocks",
           "pac_encode_frames", "pac_decode_blocks", "pac_decode_frames"]


def check(rc):
    if rc < 0:
        raise PacError(f"libpac_b200: {lib.pac_strerror(rc).decode()} ({rc})")
    return rc


def check_soft(rc, err):
    """For calls that report per-item soft errors: PAC_ERR_SOFT raises with the flags."""
    if rc == -4:
        raise PacError(f"libpac_b200: soft error flags {err.value:#x} (input whose behaviour is undefined in the "
                       "reference: degenerate PVQ vector, field wider than 64 bits, or malformed bitstream)")
    return check(rc)


class Context:
    """One per (GPU, configuration).  Mirrors the reference's CodingParams fields
    (coder/pacfile.py:157-174, :699-707)."""

    def __init__(self, sample_rate=48000, n_channels=2, rate_kbps=128.0, use_vq=True, use_sbr=None,
                 n_scale_bits=4, n_mant_size_bits=12, device=None, target_bits_per_sample=None, precision="f64"):
        import torch
        if not torch.cuda.is_available():
            raise PacError("no CUDA device: audio_codec_b200 has no CPU path")
        self.device = torch.cuda.current_device() if device is None else int(device)
        if use_sbr is None:
            use_sbr = rate_kbps < 128
        if target_bits_per_sample is None:
            target_bits_per_sample = float(rate_kbps) / (sample_rate / 1000)     # coder/pacfile.py:702
        self.cfg = PacConfig(int(sample_rate), int(n_channels), 1024, int(n_scale_bits), int(n_mant_size_bits),
                             int(bool(use_sbr)), int(bool(use_vq)), float(target_bits_per_sample),
                             {"f32": 0, "f64": 1}[precision])
        self.precision = precision
        self.rate_kbps = rate_kbps
        h = C.c_void_p()
        check(lib.pac_create(self.device, C.byref(self.cfg), C.byref(h)))
        self.h = h

    OPTIONS = {"batch_cb": 0, "smem_pad": 1, "host_chunk_mb": 2, "variant": 3}

    def set_option(self, name, value):
        """Tuning knobs / test hooks (pac_set_option); 0 restores the default."""
        check(lib.pac_set_option(self.h, self.OPTIONS[name], float(value)))

    def bands(self, short=False):
        n = (C.c_int * 32)()
        o = (C.c_int * 32)()
        nb = check(lib.pac_get_bands(self.h, int(short), n, o))
        return np.array(n[:nb]), np.array(o[:nb])

    def last_kernel_ms(self):
        e, d = C.c_float(), C.c_float()
        check(lib.pac_last_kernel_ms(self.h, C.byref(e), C.byref(d)))
        return e.value, d.value

    def kernel_stats(self, kind):
        """(device ms, launches) of kernel `kind` (0 k_encode, 1 k_pvq, 2 k_dequant, 3 k_decode) in the last call."""
        ms, n = C.c_float(), C.c_int()
        check(lib.pac_kernel_stats(self.h, kind, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def kernel_launches(self):
        return int(lib.pac_kernel_launches(self.h))

    def close(self):
        if getattr(self, "h", None):
            lib.pac_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default = {}


def default_context(sample_rate=48000, **kw):
    """Cached context for the stateless drop-in functions (mdct.MDCT ...)."""
    import torch
    key = (torch.cuda.current_device() if torch.cuda.is_available() else -1, sample_rate,
           tuple(sorted(kw.items())))
    if key not in _default:
        _default[key] = Context(sample_rate=sample_rate, **kw)
    return _default[key]


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
