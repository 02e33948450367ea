"""Oracle: whole-stream .pac encode / decode (bytes in, bytes out).
TEST INFRASTRUCTURE ONLY.

Follows coder/pacfile.py: header :300-340 / :136-175, block framing
:449-610 / :231-298, the driver loop :711-752 and Close :612-625, including the
quirks listed in SURVEY.md A.8 (last block coded twice, zero flush block,
numSamples padding rule, dropped transient blocks, +4 flag bits).
"""
import struct

import numpy as np

from . import codec, dsp
from .bits import BitReader, BitWriter

SHORT = 128


def _frame_bits(p, ba, cur):
    """coder/pacfile.py:342-361 (getNumBytesNeeded; the name lies, it is bits)."""
    bands = p.sfBandsShort if cur else p.sfBands
    n = p.nScaleBits
    for ib in range(bands.nBands):
        n += p.nMantSizeBits + p.nScaleBits
        if ba[ib]:
            if ib in p.omittedBands and not cur:
                n += int(ba[ib])
            else:
                n += int(ba[ib]) * int(bands.nLines[ib])
    return n


def _write_frame(pb, p, enc, cur):
    """coder/pacfile.py:363-402 (VQ) and :404-447 (block floating point)."""
    bands = p.sfBandsShort if cur else p.sfBands
    ba = enc["bitAlloc"]
    pb.WriteBits(enc["scale"], p.nScaleBits)
    if p.useVQ:
        for ib in range(bands.nBands):
            pb.WriteBits(int(ba[ib]) - 1 if ba[ib] else 0, p.nMantSizeBits)
        ctr = 0
        for ib in range(bands.nBands):
            if ba[ib]:
                f = enc["fields"][ctr]
                ctr += 1
                if ib in p.omittedBands and not cur:
                    f = f[-1:]
                for v, n in f:
                    pb.WriteBits(v, n)
        return
    im = 0
    for ib in range(bands.nBands):
        pb.WriteBits(int(ba[ib]) - 1 if ba[ib] else 0, p.nMantSizeBits)
        pb.WriteBits(int(enc["sf"][ib]), p.nScaleBits)
        if ba[ib]:
            for m in enc["mant"][im]:
                pb.WriteBits(int(m), int(ba[ib]))
            im += 1


def encode_block(prior, data, p, last, cur, nxt):
    """coder/pacfile.py:449-610: one block (all channels) -> bytes, or b'' when
    the reference silently drops the block (a zero short sub-frame)."""
    full = [np.concatenate((prior[c], data[c])) for c in range(p.nChannels)]
    if not cur:
        sbr = p.useSBR
        frames = [[codec.encode_channel(full[c], p, last, cur, nxt, sbr=sbr)
                   for c in range(p.nChannels)]]
    else:
        frames = []
        for n in range(448, 2048 - SHORT - 448, SHORT):
            for c in range(p.nChannels):
                if np.all(full[c][n:n + 2 * SHORT] == 0):
                    return b""
            frames.append([codec.encode_channel(full[c][n:n + 2 * SHORT], p, last, cur, nxt)
                           for c in range(p.nChannels)])
    out = b""
    for c in range(p.nChannels):
        nbits = sum(_frame_bits(p, fr[c]["bitAlloc"], cur) for fr in frames) + 4
        nbytes = (nbits + 7) // 8
        pb = BitWriter(nbytes)
        pb.WriteBits(int(bool(last)), 1)
        pb.WriteBits(int(bool(cur)), 1)
        pb.WriteBits(int(bool(nxt)), 1)
        for fr in frames:
            _write_frame(pb, p, fr[c], cur)
        out += struct.pack("<L", nbytes) + pb.bytes()
    return out


def header(p, num_samples):
    """coder/pacfile.py:306-333 incl. the numSamples rule (:309-315)."""
    n = num_samples
    if not n % p.nMDCTLines:
        n += p.nMDCTLines - n % p.nMDCTLines
    n += p.nMDCTLines
    h = b"PAC " + struct.pack("<LHLLHHHH", p.sampleRate, p.nChannels, n, p.nMDCTLines,
                              p.nScaleBits, p.nMantSizeBits, int(p.useSBR), int(p.useVQ))
    h += struct.pack("<L", p.sfBands.nBands)
    h += struct.pack("<%dH" % p.sfBands.nBands, *[int(v) for v in p.sfBands.nLines])
    return h


def block_flags(x, n_blocks):
    """Transient verdict of every PCM block: the detector always sees
    [block || zeros] (coder/pacfile.py:717-731)."""
    C = x.shape[0]
    z = np.zeros((C, 1024))
    return [dsp.transient_detect(np.concatenate([x[:, k * 1024:(k + 1) * 1024], z], axis=1))
            for k in range(n_blocks)]


def read_wav_like_reference(path):
    """16-bit WAV -> (fs, pcm[C, n_read], n_header) the way the reference reads it.

    coder/pcmfile.py:32-64 takes numSamples from the data-chunk size, but
    PACFile.WriteFileHeader (coder/pacfile.py:309-315) then ADDS up to 2048 to
    that same codingParams.numSamples before the read loop runs, so
    PCMFile.ReadDataBlock (:66-77) keeps reading past the data chunk: any
    trailing RIFF chunk bytes are coded as if they were PCM (harpsichord.wav
    has 168 such bytes).  n_header is the data-chunk count the header is
    computed from."""
    raw = open(path, "rb").read()
    assert raw[:4] == b"RIFF" and raw[8:12] == b"WAVE"
    pos = 12
    while raw[pos:pos + 4] != b"fmt ":
        pos += 4
    _, tag, C, fs, _, _, bps = struct.unpack_from("<LHHLLHH", raw, pos + 4)
    assert tag == 1 and bps == 16
    pos += 24
    while raw[pos:pos + 4] != b"data":
        pos += 4
    n = struct.unpack_from("<L", raw, pos + 4)[0] // (C * 2)
    pos += 8
    n_inflated = n + 1024 + (1024 if n % 1024 == 0 else 0)
    body = raw[pos:pos + n_inflated * C * 2]
    body = body + b"\0" * (-len(body) % (2 * C))          # short read is zero padded (:78-80)
    pcm = np.frombuffer(body, dtype="<i2").reshape(-1, C).T.copy()
    return fs, pcm, n


def encode_pcm(pcm, fs, rate_kbps, use_vq=True, use_sbr=None, header_samples=None):
    """int16 PCM [C, n] -> .pac bytes, as coder/pacfile.py:674-756 ("Encode").
    header_samples: sample count the header is derived from when it differs
    from pcm.shape[1] (see read_wav_like_reference)."""
    pcm = np.asarray(pcm)
    C, n = pcm.shape
    p = codec.Params(fs, rate_kbps, C, use_vq, use_sbr)
    nb = -(-n // 1024)
    x = np.zeros((C, nb * 1024))
    x[:, :n] = dsp.pcm_to_float(pcm)
    det = block_flags(x, nb)
    out = [header(p, n if header_samples is None else header_samples)]
    prior = [np.zeros(1024) for _ in range(C)]
    last = cur = False
    # blocks 0..nb-1, then the last block AGAIN (:725-743), then Close's zeros
    order = list(range(nb)) + [nb - 1]
    for j, k in enumerate(order):
        nxt = det[k] if j < nb else False
        data = [x[c, k * 1024:(k + 1) * 1024] for c in range(C)]
        out.append(encode_block(prior, data, p, last, cur, nxt))
        prior = data
        last, cur = cur, nxt
    zero = [np.zeros(1024) for _ in range(C)]
    out.append(encode_block(prior, zero, p, False, False, False))     # Close(): flags default
    return b"".join(out)


def parse_header(buf):
    """coder/pacfile.py:136-175"""
    assert buf[:4] == b"PAC ", "not a PAC file"
    fs, C, ns, nl, nsb, nmsb, sbr, vq = struct.unpack_from("<LHLLHHHH", buf, 4)
    nb = struct.unpack_from("<L", buf, 26)[0]
    lines = struct.unpack_from("<%dH" % nb, buf, 30)
    p = codec.Params(fs, 128, C, bool(vq), bool(sbr), nl, nsb, nmsb)
    from . import psycho
    p.sfBands = psycho.Bands(lines)
    om = psycho.omitted_bands(p.sfBands)
    p.omittedBands = [int(b) for b in om] if p.useSBR else []
    p.numSamples = ns
    return p, 30 + 2 * nb


def _decode_frame(pb, p, half, last, cur, nxt):
    """coder/pacfile.py:177-229"""
    bands = p.sfBandsShort if cur else p.sfBands
    scale = pb.ReadBits(p.nScaleBits)
    ba, sf = [], []
    mant = np.zeros(half, np.int64)
    for ib in range(bands.nBands):
        b = pb.ReadBits(p.nMantSizeBits)
        if b:
            b += 1
        ba.append(b)
        if not p.useVQ:
            sf.append(pb.ReadBits(p.nScaleBits))
            if b:
                lo = int(bands.lowerLine[ib])
                for j in range(int(bands.nLines[ib])):
                    mant[lo + j] = pb.ReadBits(b)
    if not p.useVQ:
        return codec.decode_channel_bfp(ba, sf, mant, scale, p, half, last, cur, nxt)
    if p.useSBR and not cur and any(ba[b] != 0 for b in p.omittedBands):   # pacfile.py:660-664
        return codec.decode_channel_sbr(pb, ba, scale, p, half, last, cur, nxt)
    return codec.decode_channel_vq(pb, ba, scale, p, half, last, cur, nxt)


def decode_pac(buf):
    """.pac bytes -> (fs, int16 PCM [C, n]) as coder/pacfile.py:231-298 plus the
    PCM writer's quantiser (coder/pcmfile.py:118-147)."""
    p, pos = parse_header(buf)
    C = p.nChannels
    ola = [np.zeros(1024) for _ in range(C)]
    chunks = []
    while pos < len(buf):
        blk = []
        for c in range(C):
            nbytes = struct.unpack_from("<L", buf, pos)[0]
            pb = BitReader(buf[pos + 4:pos + 4 + nbytes])
            pos += 4 + nbytes
            last, cur, nxt = pb.ReadBits(1), pb.ReadBits(1), pb.ReadBits(1)
            if not cur:
                y = _decode_frame(pb, p, 1024, last, cur, nxt)
            else:
                y = np.zeros(2048)
                for n in range(448, 2048 - SHORT - 448, SHORT):
                    y[n:n + 2 * SHORT] += _decode_frame(pb, p, SHORT, last, cur, nxt)
            blk.append(ola[c] + y[:1024])
            ola[c] = y[1024:]
        chunks.append(np.stack(blk))
    chunks.append(np.stack(ola))                       # final overlap flush (:247-251)
    x = np.concatenate(chunks, axis=1)
    return p.sampleRate, dsp.float_to_pcm(x)
