"""Drop-in for `gain_shape_quantize.quantize_gain_shape` / `dequantize_gain_shape`
(coder/gain_shape_quantize.py:476-538), on the GPU: mu-law gain + pyramid-VQ shape with
the recursive mid/side split, one warp per vector."""
import numpy as np
import torch

from . import _lib
from ._util import from_device, ptr, to_device

__all__ = ["quantize_gain_shape", "dequantize_gain_shape", "quantize_gain_shape_batch", "SPLIT_BITS"]

SPLIT_BITS = 32
_MAX_L = 1024


def _fine(k_fine):
    if k_fine != 0:
        raise NotImplementedError("k_fine != 0 is never used by the codec (coder/codec.py:26 K_FINE = 0)")


def quantize_gain_shape_batch(x, bit_alloc):
    """x [n, L] (CUDA tensor or array), bit_alloc [n] -> (indices int64 [n, F] holding uint64 bit
    patterns, bits int32 [n, F], nfields int32 [n]) as CUDA tensors."""
    t, _ = to_device(x, torch.float64)
    n, L = t.shape
    if not 1 <= L <= _MAX_L:
        raise ValueError(f"band length {L} outside 1..{_MAX_L}")
    ba = torch.as_tensor(np.asarray(bit_alloc, dtype=np.int32) if not isinstance(bit_alloc, torch.Tensor) else bit_alloc,
                         device=t.device).to(torch.int32).reshape(-1).contiguous()
    mx = int(ba.max()) if n else 0
    maxf = max(8, mx // 2 + 8)           # every field but empty leaves spends >= 1 bit; generous
    idx = torch.zeros((n, maxf), dtype=torch.int64, device=t.device)
    bits = torch.zeros((n, maxf), dtype=torch.int32, device=t.device)
    nf = torch.zeros(n, dtype=torch.int32, device=t.device)
    err = _lib.C.c_int(0)
    rc = _lib.lib.pac_gain_shape_quantize(_lib.default_context().h, ptr(t), L, n, ptr(ba), max(mx, 0), ptr(idx), ptr(bits),
                                          ptr(nf), maxf, _lib.C.byref(err), _lib.stream_ptr())
    _lib.check_soft(rc, err)
    return idx, bits, nf


def quantize_gain_shape(x, bit_alloc, k_fine=0):
    """One band: returns (indices, bits), the fields in stream order [theta, mid.., side..]*, gain
    (coder/gain_shape_quantize.py:476-512).  An all-zero band gives ([0], [0])."""
    _fine(k_fine)
    v = np.asarray(x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else x, dtype=np.float64).reshape(1, -1)
    idx, bits, nf = quantize_gain_shape_batch(v, [int(bit_alloc)])
    k = int(nf[0])
    iv = idx[0, :k].cpu().numpy()
    return [int(a) & 0xFFFFFFFFFFFFFFFF for a in iv], [int(b) for b in bits[0, :k].cpu().numpy()]


def dequantize_gain_shape(pb, bit_alloc, L, k_fine=0):
    """Parses one band from the PackedBits cursor and returns gain * shape
    (coder/gain_shape_quantize.py:515-538); the cursor is left after the band's fields."""
    _fine(k_fine)
    if not 1 <= L <= _MAX_L:
        raise ValueError(f"band length {L} outside 1..{_MAX_L}")
    d = pb.device_data()
    pos = torch.tensor([pb.iByte * 8 + pb.iBit], dtype=torch.int64, device=d.device)
    ba = torch.tensor([int(bit_alloc)], dtype=torch.int32, device=d.device)
    out = torch.empty((1, L), dtype=torch.float64, device=d.device)
    pos_out = torch.zeros(1, dtype=torch.int64, device=d.device)
    err = _lib.C.c_int(0)
    rc = _lib.lib.pac_gain_shape_dequantize(_lib.default_context().h, ptr(d), d.numel(), ptr(pos), ptr(ba), int(L), 1,
                                            ptr(out), ptr(pos_out), _lib.C.byref(err), _lib.stream_ptr())
    _lib.check_soft(rc, err)
    pb.iByte, pb.iBit = divmod(int(pos_out[0]), 8)
    return from_device(out[0], True)
