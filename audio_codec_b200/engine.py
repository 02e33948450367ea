"""Batched stream encoder / decoder over libpac_b200 (the L4 driver of SURVEY.md 1).

encode_streams() takes whole PCM streams and returns their .pac images; every
stage (transient detection, window + MDCT, psychoacoustic model, bit
allocation, quantisation, bit packing, byte layout) runs on the GPU.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib


def pack_pcm(pcm_list, n_channels):
    """List of int16 [C, n_s] arrays -> (planar block-padded int16 tensor on the
    host, n_blocks[S]).  Stream s channel c lives at ((blkOff[s]*C + c*nb[s])*1024."""
    nbs = [max(1, -(-p.shape[1] // 1024)) for p in pcm_list]
    tot = sum(nbs) * n_channels * 1024
    buf = np.zeros(tot, dtype=np.int16)
    o = 0
    for p, nb in zip(pcm_list, nbs):
        assert p.shape[0] == n_channels
        for c in range(n_channels):
            buf[o:o + p.shape[1]] = p[c]
            o += nb * 1024
    return buf, np.array(nbs, dtype=np.int32)


def encode_device(ctx, d_pcm, n_blocks, hdr_samples, d_out=None):
    """d_pcm: int16 CUDA tensor in pack_pcm layout.  Returns (d_out, stream_off[S+1], err)."""
    S = len(n_blocks)
    nb = (C.c_int * S)(*[int(v) for v in n_blocks])
    hs = (C.c_int * S)(*[int(v) for v in hdr_samples])
    if d_out is None:
        cap = _lib.lib.pac_encode_bound(ctx.h, nb, S)
        d_out = torch.empty(int(cap), dtype=torch.uint8, device=d_pcm.device)
    off = (C.c_longlong * (S + 1))()
    err = C.c_int(0)
    rc = _lib.lib.pac_encode_streams(ctx.h, C.c_void_p(d_pcm.data_ptr()), nb, hs, S,
                                     C.c_void_p(d_out.data_ptr()), d_out.numel(), off, C.byref(err),
                                     _lib.stream_ptr())
    if rc != 0 and rc != -4:
        _lib.check(rc)
    return d_out, np.array(off[:]), err.value


def encode_streams(ctx, pcm_list, hdr_samples=None):
    """Host convenience: list of int16 [C, n] arrays -> list of .pac bytes."""
    Cn = ctx.cfg.n_channels
    buf, nbs = pack_pcm(pcm_list, Cn)
    if hdr_samples is None:
        hdr_samples = [p.shape[1] for p in pcm_list]
    d_pcm = torch.from_numpy(buf).cuda()
    d_out, off, err = encode_device(ctx, d_pcm, nbs, hdr_samples)
    if err:
        raise _lib.PacError(f"soft error flags {err:#x} (degenerate PVQ input / field wider than 64 bits: "
                            "undefined in the reference)")
    host = d_out[:int(off[-1])].cpu().numpy().tobytes()
    return [host[int(off[s]):int(off[s + 1])] for s in range(len(pcm_list))]
