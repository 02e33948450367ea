"""Drop-in for the reference's `codec` module (coder/codec.py): per-channel frame
encode / decode, with and without SBR.  One call = one fused kernel launch (window,
MDCT, psychoacoustic model, bit allocation, quantisation; or parse, dequantise, SBR,
IMDCT, window), one CTA per frame."""
import numpy as np
import torch

from . import _lib, bitpack
from ._util import ptr

__all__ = ["Encode", "EncodeSingleChannel", "Encode_SBR", "EncodeSingleChannel_SBR", "Decode", "Decode_SBR",
           "DecodeSingleChannel", "getCorrectWindow", "K_FINE", "SHORT", "encode_frames"]

K_FINE = 0
SHORT = 256
_LOG_CAP = 320


def getCorrectWindow(lastTrans, curTrans, nextTrans, Nlong=2048):
    """Window function for a block given the transient flags (coder/codec.py:30-44)."""
    from . import window as w
    if curTrans:
        return w.SineWindow
    if lastTrans and nextTrans:
        return lambda x: w.StartStopWindow(x, Nlong, SHORT)
    if lastTrans:
        return lambda x: w.StopWindow(x, Nlong, SHORT)
    if nextTrans:
        return lambda x: w.StartWindow(x, Nlong, SHORT)
    return w.SineWindow


def context_for(codingParams, sbr):
    """The device context matching a CodingParams bag (coder/pacfile.py:699-707)."""
    return _lib.default_context(int(codingParams.sampleRate), n_channels=int(getattr(codingParams, "nChannels", 1)),
                                use_vq=bool(codingParams.useVQ), use_sbr=bool(sbr),
                                n_scale_bits=int(codingParams.nScaleBits),
                                n_mant_size_bits=int(codingParams.nMantSizeBits),
                                target_bits_per_sample=float(getattr(codingParams, "targetBitsPerSample", 0.0)))


def _check_bands(ctx, codingParams, short):
    nl, _ = ctx.bands(short)
    sf = codingParams.sfBandsShort if short else codingParams.sfBands
    have = np.asarray(sf.nLines)
    # BitAlloc_SBR leaves nLines[omitted] = 1 on the caller's object (coder/bitalloc.py:142-143)
    if have.shape != nl.shape or not np.all((have == nl) | (have == 1)):
        raise _lib.PacError("codingParams scale factor bands differ from ScaleFactorBands("
                            "AssignMDCTLinesFromFreqLimits(nMDCTLines, sampleRate))")
    return nl


def _flag(last, cur, nxt):
    return int(bool(last)) | (int(bool(cur)) << 1) | (int(bool(nxt)) << 2)


def encode_frames(data, codingParams, flags, sbr=False):
    """Batched frame encoder: data [n, N] (N = 2 nMDCTLines), flags [n] (last | cur<<1 | next<<2).
    Returns a dict of CUDA tensors: scale [n], bitAlloc [n, 32], scaleFactor [n, 32], bandBits
    [n, 32], rec uint8 [n, stride] (the frame's packed fields), recBits [n], and for VQ the
    field log logVal / logWidth [n, 32, cap], logCount [n, 32]."""
    dev = torch.device("cuda", torch.cuda.current_device())
    x = torch.as_tensor(np.asarray(data, dtype=np.float64) if not isinstance(data, torch.Tensor) else data,
                        device=dev).to(torch.float64).contiguous()
    n, N = x.shape
    if N != 2 * codingParams.nMDCTLines or N not in (2048, 256):
        raise ValueError("expected frames of 2 * nMDCTLines samples with nMDCTLines in {1024, 128}")
    ctx = context_for(codingParams, sbr)
    _check_bands(ctx, codingParams, N == 256)
    fl = torch.as_tensor(np.asarray(flags, dtype=np.uint8), device=dev).contiguous()
    stride = _lib.lib.pac_record_stride()
    out = dict(scale=torch.zeros(n, dtype=torch.int32, device=dev),
               bitAlloc=torch.zeros((n, 32), dtype=torch.int32, device=dev),
               scaleFactor=torch.zeros((n, 32), dtype=torch.int32, device=dev),
               bandBits=torch.zeros((n, 32), dtype=torch.int32, device=dev),
               rec=torch.zeros((n, stride), dtype=torch.uint8, device=dev),
               recBits=torch.zeros(n, dtype=torch.int32, device=dev))
    lv = lw = lc = None
    if codingParams.useVQ:
        out["logVal"] = torch.zeros((n, 32, _LOG_CAP), dtype=torch.int64, device=dev)
        out["logWidth"] = torch.zeros((n, 32, _LOG_CAP), dtype=torch.int32, device=dev)
        out["logCount"] = torch.zeros((n, 32), dtype=torch.int32, device=dev)
        lv, lw, lc = ptr(out["logVal"]), ptr(out["logWidth"]), ptr(out["logCount"])
    err = _lib.C.c_int(0)
    rc = _lib.lib.pac_encode_frames(ctx.h, ptr(x), N, ptr(fl), int(bool(sbr)), n, ptr(out["scale"]),
                                    ptr(out["bitAlloc"]), ptr(out["scaleFactor"]), ptr(out["bandBits"]), ptr(out["rec"]),
                                    ptr(out["recBits"]), lv, lw, lc, _LOG_CAP, _lib.C.byref(err), _lib.stream_ptr())
    _lib.check_soft(rc, err)
    out["ctx"] = ctx
    return out


def _single(data, codingParams, lastTrans, curTrans, nextTrans, sbr):
    halfN = codingParams.nMDCTLines
    short = halfN == 128
    if bool(curTrans) != short:
        raise ValueError("curTrans frames are the 256-sample sub-frames (nMDCTLines = 128), as PACFile.WriteDataBlock "
                         "calls them (coder/pacfile.py:489-547)")
    r = encode_frames(np.asarray(data, dtype=np.float64).reshape(1, -1), codingParams, [_flag(lastTrans, curTrans, nextTrans)],
                      sbr)
    ctx = r["ctx"]
    nl, om = ctx.bands(short)
    nB = len(nl)
    ba = r["bitAlloc"][0, :nB].cpu().numpy().astype(np.int64)
    scale = int(r["scale"][0])
    if codingParams.useVQ:
        cnt = r["logCount"][0, :nB].cpu().numpy()
        lv = r["logVal"][0, :nB].cpu().numpy()
        lw = r["logWidth"][0, :nB].cpu().numpy()
        all_idx, all_bits = [], []
        for b in range(nB):
            if ba[b]:
                k = int(cnt[b])
                all_idx.append([int(v) & 0xFFFFFFFFFFFFFFFF for v in lv[b, :k]])
                all_bits.append([int(w) for w in lw[b, :k]])
        if sbr:                                   # side effect of BitAlloc_SBR (coder/bitalloc.py:142-143)
            for b in codingParams.omittedBands:
                codingParams.sfBands.nLines[b] = 1
        return (ba, all_idx, all_bits, scale)
    sf = r["scaleFactor"][0, :nB].cpu().numpy().astype(np.int32)
    # mantissas: read them back from the frame's packed fields (scale | per band: ba, sf, mantissas)
    nS, nM = codingParams.nScaleBits, codingParams.nMantSizeBits
    widths, keep = [nS], []
    for b in range(nB):
        widths += [nM, nS]
        if ba[b]:
            keep += list(range(len(widths), len(widths) + int(nl[b])))
            widths += [int(ba[b])] * int(nl[b])
    vals = bitpack.unpack_rows(r["rec"][:1], [0], [widths], [len(widths)])[0].cpu().numpy()
    mant = vals[keep].astype(np.int32) if keep else np.empty(0, dtype=np.int32)
    return (sf, ba, mant, scale)


def EncodeSingleChannel(data, codingParams, lastTrans=False, curTrans=False, nextTrans=False):
    """One channel, one frame (coder/codec.py:266-380).  VQ: (bitAlloc, all_indices, all_idx_bits,
    overallScale); block floating point: (scaleFactor, bitAlloc, mantissa, overallScale)."""
    return _single(data, codingParams, lastTrans, curTrans, nextTrans, False)


def EncodeSingleChannel_SBR(data, codingParams, lastTrans=False, curTrans=False, nextTrans=False):
    """SBR variant (coder/codec.py:426-555): omitted bands carry one envelope value."""
    if not codingParams.useVQ:
        raise _lib.PacError("block floating point + SBR is undefined in the reference (it crashes, coder/codec.py:545)")
    return _single(data, codingParams, lastTrans, curTrans, nextTrans, True)


def _multi(fn, data, codingParams, lastTrans, curTrans, nextTrans):
    cols = [[], [], [], []]
    for iCh in range(codingParams.nChannels):
        for c, v in zip(cols, fn(data[iCh], codingParams, lastTrans, curTrans, nextTrans)):
            c.append(v)
    return tuple(cols)


def Encode(data, codingParams, lastTrans=False, curTrans=False, nextTrans=False):
    """All channels of a block (coder/codec.py:225-263)."""
    return _multi(EncodeSingleChannel, data, codingParams, lastTrans, curTrans, nextTrans)


def Encode_SBR(data, codingParams, lastTrans=False, curTrans=False, nextTrans=False):
    """All channels of a block, SBR (coder/codec.py:383-423)."""
    return _multi(EncodeSingleChannel_SBR, data, codingParams, lastTrans, curTrans, nextTrans)


def _decode(scaleFactor, bitAlloc, mantissa, overallScaleFactor, pb, codingParams, lastTrans, curTrans, nextTrans, sbr):
    halfN = codingParams.nMDCTLines
    N = 2 * halfN
    short = halfN == 128
    ctx = context_for(codingParams, sbr)
    nl = _check_bands(ctx, codingParams, short)
    nB = len(nl)
    nS, nM = int(codingParams.nScaleBits), int(codingParams.nMantSizeBits)
    ba = [int(v) for v in np.asarray(bitAlloc).reshape(-1)[:nB]]
    fields = [(int(overallScaleFactor), nS)]
    payload_start = None
    if codingParams.useVQ:
        fields += [((b - 1) if b else 0, nM) for b in ba]
        # the band payloads follow in pb from its cursor: re-read them as 64-bit chunks
        start = pb.iByte * 8 + pb.iBit
        avail = max(0, pb.nBytes * 8 - start)
        n64 = min((avail + 63) // 64, 500)      # one channel-block record is < 4 KB
        chunks = bitpack.unpack_rows(pb.device_data().reshape(1, -1), [start], [[64] * max(n64, 1)], [n64])[0]
        payload_start = sum(w for _, w in fields)
        vals = torch.cat([torch.tensor([v for v, _ in fields], dtype=torch.int64, device=chunks.device), chunks[:n64]])
        wid = [w for _, w in fields] + [64] * n64
    else:
        m = [int(v) for v in np.asarray(mantissa).reshape(-1)]
        sfs = [int(v) for v in np.asarray(scaleFactor).reshape(-1)[:nB]]
        k = 0
        for b in range(nB):
            fields += [((ba[b] - 1) if ba[b] else 0, nM), (sfs[b], nS)]
            if ba[b]:
                fields += [(m[k + i], ba[b]) for i in range(int(nl[b]))]
                k += int(nl[b])
        vals = torch.tensor([v for v, _ in fields], dtype=torch.int64, device="cuda")
        wid = [w for _, w in fields]
    nbytes = (sum(wid) + 7) // 8 + 8
    rec = bitpack.pack_rows(vals.reshape(1, -1), [wid], [len(wid)], nbytes)
    dev = rec.device
    rb = torch.tensor([rec.shape[1]], dtype=torch.int32, device=dev)
    fl = torch.tensor([_flag(lastTrans, curTrans, nextTrans)], dtype=torch.uint8, device=dev)
    out = torch.empty((1, N), dtype=torch.float64, device=dev)
    used = torch.zeros(1, dtype=torch.int32, device=dev)
    err = _lib.C.c_int(0)
    rc = _lib.lib.pac_decode_frames(ctx.h, ptr(rec), rec.shape[1], ptr(rb), N, ptr(fl), 1, ptr(out), ptr(used),
                                    _lib.C.byref(err), _lib.stream_ptr())
    _lib.check_soft(rc, err)
    if payload_start is not None:                       # advance the caller's cursor past the band payloads
        adv = int(used[0]) - payload_start
        pos = pb.iByte * 8 + pb.iBit + adv
        pb.iByte, pb.iBit = divmod(pos, 8)
    return out[0].cpu().numpy()


def Decode(scaleFactor, bitAlloc, mantissa, overallScaleFactor, pb, codingParams, lastTrans=False, curTrans=False,
           nextTrans=False):
    """One channel, one frame: dequantise, undo the overall scale, IMDCT, synthesis window
    (coder/codec.py:47-92).  VQ frames are parsed from `pb`, whose cursor is advanced."""
    return _decode(scaleFactor, bitAlloc, mantissa, overallScaleFactor, pb, codingParams, lastTrans, curTrans, nextTrans,
                   False)


def Decode_SBR(scaleFactor, bitAlloc, mantissa, overallScaleFactor, pb, codingParams, lastTrans=False, curTrans=False,
               nextTrans=False):
    """SBR variant (coder/codec.py:95-222): the omitted bands are rebuilt from the low band and
    the transmitted envelope."""
    return _decode(scaleFactor, bitAlloc, mantissa, overallScaleFactor, pb, codingParams, lastTrans, curTrans, nextTrans,
                   True)


DecodeSingleChannel = Decode      # name used by BASELINE.json's north star
