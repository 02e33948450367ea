"""Drop-in for the reference's `mdct` module (coder/mdct.py), on the GPU.

MDCT(data, a, b, isInverse=False) / IMDCT(data, a, b) keep the reference's
signature and scaling (2/N on the forward, coder/mdct.py:63-69 / :56-62) but run
the fused sm_100a kernel (N/4-point complex FFT).  Inputs may be NumPy arrays or
CUDA tensors, with optional leading batch dimensions; float32 input selects the
fp32 fast path, anything else the fp64 path.  The codec's block sizes (a == b,
N in {2048, 256}) run the FFT kernels; every other split (any a, b with a + b
even, e.g. the reference's self-test sizes N = 8 and N = 20, coder/mdct.py:86-107)
runs a direct-sum kernel.  There is no CPU fallback.
"""
import torch

from . import _lib
from ._util import from_device, ptr, to_device

__all__ = ["MDCT", "IMDCT", "MDCTslow", "mdct_frames", "imdct_ola"]


def _dt(t):
    return (_lib.F32, torch.float32) if t.dtype == torch.float32 else (_lib.F64, torch.float64)


def MDCTslow(data, a, b, isInverse=False):
    """coder/mdct.py:11-39: the defining sums, for any a, b with a + b even (direct-sum kernel)."""
    N = int(a) + int(b)
    if N < 2 or N % 2:
        raise ValueError("MDCT needs an even block length a + b")
    t, was_np = to_device(data)
    code, dt = _dt(t)
    t = t.to(dt)
    n_in, n_out = (N // 2, N) if isInverse else (N, N // 2)
    if t.shape[-1] != n_in:
        raise ValueError(f"expected {n_in} values, got {t.shape[-1]}")
    rows = t.reshape(-1, n_in).contiguous()
    out = torch.empty((rows.shape[0], n_out), dtype=dt, device=t.device)
    ctx = _lib.default_context()
    _lib.check(_lib.lib.pac_mdct_direct(ctx.h, code, ptr(rows), rows.shape[0], int(a), int(b), int(bool(isInverse)),
                                        ptr(out), _lib.stream_ptr()))
    return from_device(out.reshape(*t.shape[:-1], n_out), was_np)


def MDCT(data, a, b, isInverse=False):
    N = a + b
    if a != b or N not in (2048, 256):
        return MDCTslow(data, a, b, isInverse)
    t, was_np = to_device(data)
    code, dt = _dt(t)
    t = t.to(dt)
    ctx = _lib.default_context()
    if not isInverse:
        if t.shape[-1] != N:
            raise ValueError(f"expected {N} samples, got {t.shape[-1]}")
        rows = t.reshape(-1, N)
        out = torch.empty((rows.shape[0], N // 2), dtype=dt, device=t.device)
        _lib.check(_lib.lib.pac_mdct_forward(ctx.h, code, ptr(rows), N, rows.shape[0], 1, N, _lib.W_NONE,
                                             ptr(out), _lib.stream_ptr()))
        return from_device(out.reshape(*t.shape[:-1], N // 2), was_np)
    if t.shape[-1] != N // 2:
        raise ValueError(f"expected {N // 2} coefficients, got {t.shape[-1]}")
    rows = t.reshape(-1, N // 2)
    out = torch.empty((rows.shape[0], N), dtype=dt, device=t.device)
    _lib.check(_lib.lib.pac_imdct_ola(ctx.h, code, ptr(rows), rows.shape[0], 1, N, _lib.W_NONE, ptr(out), N,
                                      _lib.stream_ptr()))
    return from_device(out.reshape(*t.shape[:-1], N), was_np)


def IMDCT(data, a, b):
    return MDCT(data, a, b, isInverse=True)


def mdct_frames(x, N=2048, window=_lib.W_SINE, ctx=None):
    """Batched analysis of whole streams: x [rows, (F+1)*N/2] -> [rows, F, N/2];
    frame f is the hop-N/2 view x[:, f*N/2 : f*N/2+N] (never materialised)."""
    t, was_np = to_device(x)
    code, dt = _dt(t)
    t = t.to(dt)
    M = N // 2
    F = t.shape[-1] // M - 1
    rows = t.reshape(-1, t.shape[-1])
    out = torch.empty((rows.shape[0], F, M), dtype=dt, device=t.device)
    ctx = ctx or _lib.default_context()
    _lib.check(_lib.lib.pac_mdct_forward(ctx.h, code, ptr(rows), rows.stride(0), rows.shape[0], F, N, window,
                                         ptr(out), _lib.stream_ptr()))
    return from_device(out, was_np)


def imdct_ola(X, N=2048, window=_lib.W_SINE, ctx=None):
    """Batched synthesis: X [rows, F, N/2] -> y [rows, (F+1)*N/2] (windowed IMDCT,
    overlap-added; coder/pacfile.py:289-295)."""
    t, was_np = to_device(X)
    code, dt = _dt(t)
    t = t.to(dt)
    M = N // 2
    rows, F = t.shape[0], t.shape[1]
    out = torch.empty((rows, (F + 1) * M), dtype=dt, device=t.device)
    ctx = ctx or _lib.default_context()
    _lib.check(_lib.lib.pac_imdct_ola(ctx.h, code, ptr(t), rows, F, N, window, ptr(out), out.stride(0),
                                      _lib.stream_ptr()))
    return from_device(out, was_np)
