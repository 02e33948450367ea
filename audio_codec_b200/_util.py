"""Host-side plumbing shared by the drop-in modules: NumPy <-> CUDA tensor."""
import ctypes as C

import numpy as np


def to_device(x, dtype=None):
    """NumPy array / sequence / torch tensor -> contiguous CUDA tensor.
    Returns (tensor, was_numpy)."""
    import torch
    if isinstance(x, torch.Tensor):
        t = x
        was_np = False
    else:
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(x)))
        was_np = True
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    if not t.is_cuda:
        t = t.cuda()
    return t.contiguous(), was_np


def from_device(t, was_np):
    return t.cpu().numpy() if was_np else t


def ptr(t):
    return C.c_void_p(t.data_ptr())
