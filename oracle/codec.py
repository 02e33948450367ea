"""Oracle: one channel x one MDCT frame encode / decode.  TEST INFRASTRUCTURE ONLY.

Follows coder/codec.py (EncodeSingleChannel :266-380, EncodeSingleChannel_SBR
:426-555, Decode :47-92, Decode_SBR :95-222).
"""
from math import floor

import numpy as np
from scipy import interpolate
from scipy.ndimage import gaussian_filter1d

from . import dsp, psycho, pvq


class Params:
    """The fields of the reference's CodingParams bag that the frame codec reads
    (coder/pacfile.py:157-174, :699-707, :326-339)."""

    def __init__(self, fs, rate_kbps, n_channels=2, use_vq=True, use_sbr=None,
                 n_lines=1024, n_scale_bits=4, n_mant_size_bits=12):
        self.sampleRate = fs
        self.nChannels = n_channels
        self.nMDCTLines = n_lines
        self.nScaleBits = n_scale_bits
        self.nMantSizeBits = n_mant_size_bits
        self.targetBitsPerSample = rate_kbps / (fs / 1000)
        self.useVQ = use_vq
        self.useSBR = (rate_kbps < 128) if use_sbr is None else use_sbr
        self.sfBands = psycho.Bands(psycho.assign_lines(n_lines, fs))
        self.sfBandsShort = psycho.Bands(psycho.assign_lines(128, fs))
        om = psycho.omitted_bands(self.sfBands)
        self.omittedBands = [int(b) for b in om] if self.useSBR else []


def _budget(p, bands, half, last, cur, nxt, sbr):
    """coder/codec.py:288-299 (and :446-453 for SBR: no 0.85/1.45 factors)."""
    if sbr:
        nb = half
    else:
        nb = int(1.45 * half) if cur else half
        if last or nxt:
            nb = int(0.85 * nb)
    b = p.targetBitsPerSample * nb
    if p.useVQ:
        b -= p.nScaleBits
    else:
        b -= p.nScaleBits * (bands.nBands + 1)
    b -= p.nMantSizeBits * bands.nBands
    return b


def encode_channel(data, p, last=False, cur=False, nxt=False, sbr=False):
    """Returns dict(scale, bitAlloc, fields=[(value,width)...] per coded band)
    for VQ, or dict(scale, bitAlloc, sf, mant) for block floating point."""
    half = len(data) // 2
    N = 2 * half
    bands = p.sfBandsShort if cur else p.sfBands
    omitted = p.omittedBands if sbr else []
    max_mant = min(1 << p.nMantSizeBits, 16)
    budget = _budget(p, bands, half, last, cur, nxt, sbr)
    lines = dsp.mdct(dsp.select_window(last, cur, nxt, N) * data, half, half)[:half]
    max_line = np.max(np.abs(lines))
    if sbr:
        fft_lines = np.abs(np.fft.rfft(dsp.hann_window(N) * data)) / half   # codec.py:460-463
        max_line = max(max_line, np.max(fft_lines))
    scale = dsp.scale_factor(max_line, p.nScaleBits)
    lines = lines * (1 << scale)
    if sbr:
        fft_lines = fft_lines * (1 << scale)
    smr = psycho.calc_smrs(data, lines, scale, p.sampleRate, bands)
    n_lines = bands.nLines.copy()
    for b in omitted:
        n_lines[b] = 1                                   # bitalloc.py:142-143
    ba = psycho.bit_alloc(budget, max_mant, bands.nBands, n_lines, smr)
    if not p.useVQ:
        if sbr:
            raise RuntimeError("BFP + SBR crashes in the reference (codec.py:545); undefined")
        sf = np.zeros(bands.nBands, dtype=np.int64)
        mant = []
        for ib in range(bands.nBands):
            lo, hi = bands.lowerLine[ib], bands.upperLine[ib] + 1
            sf[ib] = dsp.scale_factor(np.max(np.abs(lines[lo:hi])), p.nScaleBits, int(ba[ib]))
            if ba[ib]:
                mant.append(dsp.vmantissa(lines[lo:hi], sf[ib], p.nScaleBits, int(ba[ib])))
        return dict(scale=scale, bitAlloc=ba, sf=sf, mant=mant)
    fields = []
    for ib in range(bands.nBands):
        if not ba[ib]:
            continue
        lo, hi = bands.lowerLine[ib], bands.upperLine[ib] + 1
        if ib in omitted:
            x = np.array([np.mean(np.abs(fft_lines[lo:hi]))])            # codec.py:508-510
        else:
            x = lines[lo:hi]
        idx, bits = pvq.quantize_gain_shape(x, int(ba[ib] * n_lines[ib]), 0)
        tot = sum(bits)
        if tot == 0:
            ba[ib] = 0
        else:
            fields.append(list(zip(idx, bits)))
        if sbr and tot < ba[ib] and ib < bands.nBands - 2:                # codec.py:522-524
            ba[ib + 1] += ba[ib] - tot
    return dict(scale=scale, bitAlloc=ba, fields=fields)


def _synth(lines, scale, half, last, cur, nxt):
    lines = lines / (1. * (1 << scale))
    return dsp.select_window(last, cur, nxt, 2 * half) * dsp.imdct(lines, half, half)


def decode_channel_vq(pb, ba, scale, p, half, last, cur, nxt):
    """coder/codec.py:47-92 with useVQ: parses the band fields from `pb`."""
    bands = p.sfBandsShort if cur else p.sfBands
    lines = np.zeros(half)
    for ib in range(bands.nBands):
        n = int(bands.nLines[ib])
        lo = int(bands.lowerLine[ib])
        if ba[ib]:
            lines[lo:lo + n] = pvq.dequantize_gain_shape(pb, int(ba[ib] * n), n, 0)
    return _synth(lines, scale, half, last, cur, nxt)


def decode_channel_bfp(ba, sf, mant, scale, p, half, last, cur, nxt):
    """coder/codec.py:47-92 with useVQ False; `mant` is the full-length array."""
    bands = p.sfBandsShort if cur else p.sfBands
    lines = np.zeros(half)
    for ib in range(bands.nBands):
        n = int(bands.nLines[ib])
        lo = int(bands.lowerLine[ib])
        if ba[ib]:
            lines[lo:lo + n] = dsp.vdequantize(int(sf[ib]), mant[lo:lo + n], p.nScaleBits, int(ba[ib]))
    return _synth(lines, scale, half, last, cur, nxt)


def decode_channel_sbr(pb, ba, scale, p, half, last, cur, nxt):
    """coder/codec.py:95-222 (VQ): omitted bands carry one envelope value each,
    placed at consecutive lines from the cut-off; high band = 2:1 transposed low
    band shaped by the Gaussian-smoothed envelope."""
    bands = p.sfBands
    lines = np.zeros(half)
    pos = 0
    for ib in range(bands.nBands):
        n = 1 if ib in p.omittedBands else int(bands.nLines[ib])
        if ba[ib]:
            lines[pos:pos + n] = pvq.dequantize_gain_shape(pb, int(ba[ib] * n), n, 0)
        pos += n
    cut = int(bands.lowerLine[p.omittedBands[0]])
    n_om = half - cut
    env = lines[cut:]
    env = np.append(env, np.zeros_like(env))
    smooth = gaussian_filter1d(env, sigma=200)
    up = floor(half / n_om)
    mf = (np.arange(half) + 1 / 2) * (p.sampleRate / (2 * half))
    ii = np.arange(cut // up - 1, half // up + 1)
    fn = interpolate.interp1d(mf[ii], lines[ii], kind='slinear')
    lines[cut:] = fn(mf[cut:] / up)
    for ib in p.omittedBands:
        lo, hi = int(bands.lowerLine[ib]), int(bands.upperLine[ib]) + 1
        if np.max(np.abs(lines[lo:hi])) > 0:
            lines[lo:hi] *= smooth[lo - cut:hi - cut] / np.mean(np.abs(lines[lo:hi]))
    return _synth(lines, scale, half, last, cur, nxt)
