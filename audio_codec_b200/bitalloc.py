"""Drop-in for `bitalloc.BitAlloc` / `BitAlloc_SBR` (coder/bitalloc.py:62-145), on the GPU:
the iterative water-filling loop runs one warp per frame."""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._util import from_device, ptr, to_device

__all__ = ["BitAlloc", "BitAlloc_SBR", "DBTOBITS"]

DBTOBITS = 6.2


def BitAlloc(bitBudget, maxMantBits, nBands, nLines, SMR):
    """Mantissa bits per band that spread the quantisation noise evenly below the masked
    threshold within `bitBudget` (coder/bitalloc.py:62-121).  SMR may carry leading batch
    dimensions (bitBudget then per frame or scalar); returns an int array [.., nBands]."""
    smr, was_np = to_device(SMR, torch.float64)
    if smr.shape[-1] != nBands:
        raise ValueError("SMR must have nBands entries per frame")
    s2 = smr.reshape(-1, nBands)
    n = s2.shape[0]
    if isinstance(bitBudget, torch.Tensor):
        b = bitBudget.to(device=smr.device, dtype=torch.float64).reshape(-1)
    else:
        b = torch.as_tensor(np.asarray(bitBudget, dtype=np.float64).reshape(-1), device=smr.device)
    if b.numel() == 1 and n > 1:
        b = b.expand(n)
    b = b.contiguous()
    nl = np.ascontiguousarray(np.asarray(nLines, dtype=np.int32))
    if nl.size != nBands or nBands > 32:
        raise ValueError("nLines must have nBands (<= 32) entries")
    bits = torch.empty((n, nBands), dtype=torch.int32, device=smr.device)
    ctx = _lib.default_context()
    _lib.check(_lib.lib.pac_bit_alloc(ctx.h, ptr(b), int(maxMantBits), int(nBands), nl.ctypes.data_as(C.POINTER(C.c_int)),
                                      ptr(s2), n, ptr(bits), _lib.stream_ptr()))
    out = bits.reshape(*smr.shape[:-1], nBands)
    return from_device(out, was_np).astype(np.int64) if was_np else out


def BitAlloc_SBR(bitBudget, maxMantBits, nBands, nLines, SMR, omittedBands):
    """BitAlloc with every SBR-omitted band counted as ONE line.  As in the reference
    (coder/bitalloc.py:142-143) the caller's nLines array is modified in place."""
    for b in omittedBands:
        nLines[b] = 1
    return BitAlloc(bitBudget, maxMantBits, nBands, nLines, SMR)
