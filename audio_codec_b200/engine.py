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


def parse_header(buf):
    """coder/pacfile.py:136-151 -> dict of the header fields.  Besides master's layout (8 fields: ..., useSBR,
    useVQ) the two older generations whose files the reference ships are recognised (README.md:4-5 names the
    branches): 7 fields (test_decoded/: no useVQ field - block floating point with block switching, 18 merged
    bands) and 6 fields (test_decoded96/, test_decoded_vq/: neither flag, the 25 unmerged Zwicker bands).  "layout"
    is the number of fields, "header_bytes" the size of the header."""
    import struct
    if buf[:4] != b"PAC ":
        raise RuntimeError("Tried to read a non-PAC file into a PACFile object")
    fs, C_, ns, nl, nsb, nmsb = struct.unpack_from("<LHLLHH", buf, 4)
    # a layout is recognised by a band table whose line counts add up to nMDCTLines; a file in which no layout
    # does is read as master's (its band table is then refused by context_for_header)
    for layout, extra, strict in ((8, 2, True), (7, 1, True), (6, 0, True), (8, 2, False)):
        o = 22 + 2 * extra
        if len(buf) < o + 4:
            continue
        nb = struct.unpack_from("<L", buf, o)[0]
        if not 1 <= nb <= 64 or len(buf) < o + 4 + 2 * nb:
            continue
        lines = struct.unpack_from("<%dH" % nb, buf, o + 4)
        if strict and sum(lines) != nl:
            continue
        fl = struct.unpack_from("<%dH" % extra, buf, 22) if extra else ()
        return dict(sample_rate=fs, n_channels=C_, num_samples=ns, n_mdct_lines=nl, n_scale_bits=nsb,
                    n_mant_size_bits=nmsb, use_sbr=bool(fl[0]) if extra >= 1 else False,
                    use_vq=bool(fl[1]) if extra >= 2 else False, n_lines=lines, layout=layout, header_bytes=o + 4 + 2 * nb)
    raise RuntimeError("unrecognised .pac header layout")


def upgrade_legacy_stream(buf):
    """A 7-field stream (test_decoded/ generation) in master's layout: the same records behind a header with the
    useVQ field added (= 0: block floating point).  The 6-field generation is refused: its 25-band payload was
    written by code that is not in the checkout and does not parse as either of master's two quantisers
    (DESIGN.md 6.1 #4)."""
    h = parse_header(buf)
    if h["layout"] == 8:
        return buf
    if h["layout"] == 6:
        raise _lib.PacError("6-field .pac generation (25 unmerged bands, test_decoded96/ and test_decoded_vq/): unsupported, "
                            "its writer is not in the reference checkout")
    return bytes(buf[:24]) + b"\0\0" + bytes(buf[24:])


def context_for_header(h, precision="f64"):
    ctx = _lib.default_context(h["sample_rate"], n_channels=h["n_channels"], use_vq=h["use_vq"],
                               use_sbr=h["use_sbr"], n_scale_bits=h["n_scale_bits"],
                               n_mant_size_bits=h["n_mant_size_bits"], precision=precision)
    nl, _ = ctx.bands()
    if tuple(int(v) for v in nl) != tuple(h["n_lines"]) or h["n_mdct_lines"] != 1024:
        raise _lib.PacError("unsupported band layout in .pac header")
    return ctx


def decode_device(ctx, d_pac, stream_off, d_pcm=None):
    """d_pac: uint8 CUDA tensor holding the streams back to back.  Returns
    (d_pcm int16, pcm_off[S+1] in samples, err)."""
    S = len(stream_off) - 1
    so = (C.c_longlong * (S + 1))(*[int(v) for v in stream_off])
    po = (C.c_longlong * (S + 1))()
    if d_pcm is None:
        _lib.check(_lib.lib.pac_decode_plan(ctx.h, C.c_void_p(d_pac.data_ptr()), so, S, po, _lib.stream_ptr()))
        d_pcm = torch.empty(int(po[S]), dtype=torch.int16, device=d_pac.device)
    err = C.c_int(0)
    rc = _lib.lib.pac_decode_streams(ctx.h, C.c_void_p(d_pac.data_ptr()), so, S, C.c_void_p(d_pcm.data_ptr()),
                                     d_pcm.numel(), po, C.byref(err), _lib.stream_ptr())
    if rc != 0 and rc != -4:
        _lib.check(rc)
    return d_pcm, np.array(po[:]), err.value


def block_records(buf):
    """Split one .pac stream into its per-(block, channel) records (the bytes after each <L
    length prefix, coder/pacfile.py:239-256).  Host helper for comparing two encodings."""
    import struct
    h = parse_header(buf)
    pos = h["header_bytes"]
    out = []
    while pos + 4 <= len(buf):
        n = struct.unpack_from("<L", buf, pos)[0]
        out.append(bytes(buf[pos + 4:pos + 4 + n]))
        pos += 4 + n
    return out


def record_mismatch_rate(a, b):
    """Fraction of (block, channel) records that differ between two encodings of the same PCM."""
    ra, rb = block_records(a), block_records(b)
    n = max(len(ra), len(rb))
    same = sum(1 for x, y in zip(ra, rb) if x == y)
    return (n - same) / max(n, 1)


def check_same_configuration(pac_list):
    """One decode call shares one context: every stream's header must describe the same codec
    configuration as the first (all fields but numSamples).  Returns the first header."""
    h = parse_header(pac_list[0])
    for i, b in enumerate(pac_list[1:], 1):
        g = parse_header(b)
        bad = [k for k in h if k not in ("num_samples", "layout", "header_bytes") and h[k] != g[k]]
        if bad:
            raise _lib.PacError(f"stream {i} differs from stream 0 in {', '.join(bad)}: decode streams of "
                                "different configurations in separate calls")
    return h


def decode_streams(pac_list, precision="f64", ctx=None):
    """Host convenience: list of .pac bytes (same configuration) -> list of int16 [C, n]."""
    pac_list = [upgrade_legacy_stream(b) for b in pac_list]          # older header generations -> master's layout
    h = check_same_configuration(pac_list)
    if ctx is None:
        ctx = context_for_header(h, precision)
    off = np.concatenate([[0], np.cumsum([len(b) for b in pac_list])])
    d_pac = torch.from_numpy(np.frombuffer(b"".join(pac_list), dtype=np.uint8).copy()).cuda()
    d_pcm, po, err = decode_device(ctx, d_pac, off)
    if err:
        raise _lib.PacError(f"soft error flags {err:#x} while decoding")
    pcm = d_pcm.cpu().numpy()
    Cn = h["n_channels"]
    return [pcm[int(po[s]):int(po[s + 1])].reshape(Cn, -1) for s in range(len(pac_list))]


def encode_host(ctx, h_pcm, n_blocks, hdr_samples, h_out):
    """Host buffers in and out (torch CPU tensors, ideally pinned): int16 PCM in the
    pack_pcm layout -> .pac bytes in h_out.  Returns (stream_off[S+1], err)."""
    S = len(n_blocks)
    nb = (C.c_int * S)(*[int(v) for v in n_blocks])
    hs = (C.c_int * S)(*[int(v) for v in hdr_samples])
    off = (C.c_longlong * (S + 1))()
    err = C.c_int(0)
    rc = _lib.lib.pac_encode_host(ctx.h, C.c_void_p(h_pcm.data_ptr()), nb, hs, S, C.c_void_p(h_out.data_ptr()),
                                  h_out.numel(), off, C.byref(err), _lib.stream_ptr())
    if rc != 0 and rc != -4:
        _lib.check(rc)
    return np.array(off[:]), err.value


def decode_host(ctx, h_pac, stream_off, h_pcm):
    """Host buffers in and out: .pac bytes -> int16 PCM.  Returns (pcm_off[S+1], err)."""
    S = len(stream_off) - 1
    so = (C.c_longlong * (S + 1))(*[int(v) for v in stream_off])
    po = (C.c_longlong * (S + 1))()
    err = C.c_int(0)
    rc = _lib.lib.pac_decode_host(ctx.h, C.c_void_p(h_pac.data_ptr()), so, S,
                                  C.c_void_p(h_pcm.data_ptr()) if h_pcm is not None else None,
                                  h_pcm.numel() if h_pcm is not None else 0, po, C.byref(err), _lib.stream_ptr())
    if rc != 0 and rc != -4:
        _lib.check(rc)
    return np.array(po[:]), err.value


# ---------------------------------------------------------------------- multi-GPU sharding
def shard_streams(durations, world_size):
    """Longest-processing-time-first assignment of independent streams to ranks:
    returns a list of index lists, one per rank.  Streams share nothing (no bit reservoir,
    coder/codec.py:288-299), so this is the whole of the N > 1 data path: no collective."""
    order = sorted(range(len(durations)), key=lambda i: (-durations[i], i))
    load = [0.0] * world_size
    out = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += durations[i]
    return [sorted(v) for v in out]


def gather_manifest(local_rows, group=None):
    """All-gather the per-stream (stream_id, n_bytes, error_flag) rows of every rank into one
    table sorted by stream id: the only cross-rank exchange of a sharded run (a few bytes per
    stream; works on any torch.distributed backend)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    rows = [None] * world
    dist.all_gather_object(rows, [tuple(int(x) for x in r) for r in local_rows], group=group)
    table = sorted(r for part in rows for r in part)
    return table
