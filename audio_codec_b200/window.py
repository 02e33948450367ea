"""Drop-in for the reference's `window` module (coder/window.py:14-92), on the GPU.

Each function returns a windowed copy of its argument, as the reference does.
Arguments may be NumPy arrays or CUDA tensors of any leading batch shape; the
window runs along the last axis.  float32 input selects the fp32 table.
"""
import torch

from . import _lib
from ._util import from_device, ptr, to_device

__all__ = ["SineWindow", "HanningWindow", "KBDWindow", "StartWindow", "StopWindow", "StartStopWindow"]


def _apply(x, kind, n_short=0, alpha=0.0):
    t, was_np = to_device(x)
    if t.dtype == torch.float32:
        code = _lib.F32
    else:
        code = _lib.F64
        t = t.to(torch.float64)
    if t.dim() == 0:
        raise TypeError("len() of unsized object")
    N = t.shape[-1]
    out = torch.empty_like(t)
    if t.numel():
        ctx = _lib.default_context()
        _lib.check(_lib.lib.pac_window(ctx.h, code, ptr(t), t.numel() // N, N, kind, int(n_short), float(alpha),
                                       ptr(out), _lib.stream_ptr()))
    return from_device(out, was_np)


def SineWindow(dataSampleArray):
    """sin(pi (n + 1/2) / N) (coder/window.py:14-25)."""
    return _apply(dataSampleArray, _lib.W_SINE)


def HanningWindow(dataSampleArray):
    """(1 - cos(2 pi (n + 1/2) / N)) / 2 (coder/window.py:29-40)."""
    return _apply(dataSampleArray, _lib.W_HANN)


def KBDWindow(dataSampleArray, alpha=4.):
    """The reference's "KBD" window is a plain Kaiser window I0(pi alpha sqrt(1 - ((2n+1)/N - 1)^2)) / I0(pi alpha)
    (coder/window.py:45-58); reproduced as written."""
    return _apply(dataSampleArray, _lib.W_KAISER, alpha=alpha)


def _edler(x, kind, N_long, N_short):
    n = x.shape[-1] if hasattr(x, "shape") else len(x)
    if n != N_long:
        raise ValueError(f"operands could not be broadcast together with shapes ({N_long},) ({n},)")
    return _apply(x, kind, n_short=N_short)


def StartWindow(dataSampleArray, N_long, N_short):
    """Long -> short transition window of an Edler block switch (coder/window.py:61-71)."""
    return _edler(dataSampleArray, _lib.W_START, N_long, N_short)


def StopWindow(dataSampleArray, N_long, N_short):
    """Short -> long transition window (coder/window.py:73-80)."""
    return _edler(dataSampleArray, _lib.W_STOP, N_long, N_short)


def StartStopWindow(dataSampleArray, N_long, N_short):
    """Window of a long block between two short ones (coder/window.py:82-92)."""
    return _edler(dataSampleArray, _lib.W_STARTSTOP, N_long, N_short)
