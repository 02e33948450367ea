"""Drop-in for the reference's `quantize` module (coder/quantize.py:14-274), on the GPU:
midtread uniform quantiser and the block-floating-point scale factor / mantissa codes."""
import numpy as np
import torch

from . import _lib
from ._util import from_device, ptr, to_device

__all__ = ["QuantizeUniform", "DequantizeUniform", "vQuantizeUniform", "vDequantizeUniform", "ScaleFactor", "Mantissa",
           "Dequantize", "vMantissa", "vDequantize", "MantissaVector"]


def _ctx():
    return _lib.default_context()


def vQuantizeUniform(aNumVec, nBits):
    """Sign-magnitude midtread codes of signed fractions, vectorised (coder/quantize.py:61-78)."""
    x, was_np = to_device(aNumVec, torch.float64)
    out = torch.zeros(x.shape, dtype=torch.int64, device=x.device)
    if nBits > 0:
        _lib.check(_lib.lib.pac_quantize_uniform(_ctx().h, ptr(x), x.numel(), int(nBits), ptr(out), _lib.stream_ptr()))
    return from_device(out, was_np)


def vDequantizeUniform(aQuantizedNumVec, nBits):
    """Signed fractions of sign-magnitude midtread codes (coder/quantize.py:82-95)."""
    c, was_np = to_device(aQuantizedNumVec, torch.int64)
    out = torch.zeros(c.shape, dtype=torch.float64, device=c.device)
    if nBits > 0:
        _lib.check(_lib.lib.pac_dequantize_uniform(_ctx().h, ptr(c), c.numel(), int(nBits), ptr(out), _lib.stream_ptr()))
    return from_device(out, was_np)


def QuantizeUniform(aNum, nBits):
    """Scalar form (coder/quantize.py:14-36)."""
    return int(vQuantizeUniform(np.array([aNum], dtype=np.float64), nBits)[0])


def DequantizeUniform(aQuantizedNum, nBits):
    """Scalar form (coder/quantize.py:40-57)."""
    return float(vDequantizeUniform(np.array([aQuantizedNum], dtype=np.int64), nBits)[0])


def ScaleFactor(aNum, nScaleBits=3, nMantBits=5):
    """Block-floating-point scale: leading zeros of the (2^nScaleBits - 1 + nMantBits)-bit code of |aNum|,
    capped at 2^nScaleBits - 1 (coder/quantize.py:99-125).  Arrays give one scale per element."""
    scalar = np.isscalar(aNum) or (not isinstance(aNum, torch.Tensor) and np.ndim(aNum) == 0)
    x, was_np = to_device(np.atleast_1d(np.asarray(aNum, dtype=np.float64)) if not isinstance(aNum, torch.Tensor) else aNum,
                          torch.float64)
    out = torch.empty(x.shape, dtype=torch.int64, device=x.device)
    _lib.check(_lib.lib.pac_scale_factor(_ctx().h, ptr(x), x.numel(), int(nScaleBits), int(nMantBits), ptr(out),
                                         _lib.stream_ptr()))
    r = from_device(out, was_np)
    return int(r[0]) if scalar else r


def vMantissa(aNumVec, scale, nScaleBits=3, nMantBits=5):
    """Mantissa codes of a vector sharing one scale factor (coder/quantize.py:229-250)."""
    x, was_np = to_device(aNumVec, torch.float64)
    out = torch.empty(x.shape, dtype=torch.int64, device=x.device)
    _lib.check(_lib.lib.pac_mantissa(_ctx().h, ptr(x), x.numel(), int(scale), int(nScaleBits), int(nMantBits), ptr(out),
                                     _lib.stream_ptr()))
    return from_device(out, was_np)


def vDequantize(scale, mantissaVec, nScaleBits=3, nMantBits=5):
    """Signed fractions of block-floating-point codes (coder/quantize.py:254-274)."""
    m, was_np = to_device(mantissaVec, torch.int64)
    out = torch.empty(m.shape, dtype=torch.float64, device=m.device)
    _lib.check(_lib.lib.pac_dequantize(_ctx().h, int(scale), ptr(m), m.numel(), int(nScaleBits), int(nMantBits), ptr(out),
                                       _lib.stream_ptr()))
    return from_device(out, was_np)


def Mantissa(aNum, scale, nScaleBits=3, nMantBits=5):
    """Scalar form (coder/quantize.py:178-198)."""
    return int(vMantissa(np.array([aNum], dtype=np.float64), scale, nScaleBits, nMantBits)[0])


def Dequantize(scale, mantissa, nScaleBits=3, nMantBits=5):
    """Scalar form (coder/quantize.py:201-225)."""
    return float(vDequantize(scale, np.array([mantissa], dtype=np.int64), nScaleBits, nMantBits)[0])


MantissaVector = vMantissa      # name used by BASELINE.json's north star
