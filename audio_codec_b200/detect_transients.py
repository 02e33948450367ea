"""Drop-in for `detect_transients.parTransientDetect` (coder/detect_transients.py:5-23)."""
import torch

from . import _lib
from ._util import ptr, to_device

__all__ = ["parTransientDetect"]


def parTransientDetect(block, thresh=4.5, axis=1):
    """Peak-to-average-ratio transient verdict of a [channels, samples] block.

    A 2-D block returns what the reference returns (0 when the block is silent,
    else True / False); a 3-D batch [blocks, channels, samples] returns a bool
    CUDA tensor of verdicts.
    """
    t, _ = to_device(block, torch.float64)
    if t.dim() == 2:
        if axis not in (1, -1):
            t = t.transpose(0, 1).contiguous()
        single = True
        t = t.unsqueeze(0)
    elif t.dim() == 3:
        if axis not in (1, -1, 2):
            raise ValueError("batched parTransientDetect works along the last axis")
        single = False
    else:
        raise ValueError("expected [channels, samples] or [blocks, channels, samples]")
    nb, C, n = t.shape
    flags = torch.zeros(nb, dtype=torch.uint8, device=t.device)
    ctx = _lib.default_context()
    _lib.check(_lib.lib.pac_transient_detect(ctx.h, ptr(t), nb, C, n, float(thresh), ptr(flags), _lib.stream_ptr()))
    if not single:
        return flags.bool()
    if not bool((t != 0).any()):
        return 0                      # `if np.any(avg == 0): return 0`
    return bool(flags[0])
