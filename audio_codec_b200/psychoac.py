"""Drop-in for the reference's `psychoac` module (coder/psychoac.py), on the GPU.

CalcSMRs / getMaskedThreshold run the fused Hann-FFT + masker + spreading kernel
(one CTA per frame); SPL / Intensity / Bark / Thresh are elementwise kernels.  The
scale-factor-band tables (AssignMDCTLinesFromFreqLimits, ScaleFactorBands) are
config-time integers and are built on the host, as in the reference.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._util import from_device, ptr, to_device

__all__ = ["SPL", "Intensity", "Thresh", "Bark", "cbFreqLimits", "AssignMDCTLinesFromFreqLimits", "ScaleFactorBands",
           "getMaskedThreshold", "CalcSMRs"]


def _map(op, x):
    scalar = np.isscalar(x)
    t, was_np = to_device(np.atleast_1d(np.asarray(x, dtype=np.float64)) if not isinstance(x, torch.Tensor) else x,
                          torch.float64)
    out = torch.empty_like(t)
    ctx = _lib.default_context()
    _lib.check(_lib.lib.pac_psy_map(ctx.h, op, ptr(t), t.numel(), ptr(out), _lib.stream_ptr()))
    r = from_device(out, was_np)
    return float(r[0]) if scalar else r


def SPL(intensity):
    """dB SPL of an intensity, clamped at -30 dB; zeros count as 1e-8 (coder/psychoac.py:10-25, array form)."""
    return _map(_lib.PSY_SPL, intensity)


def Intensity(spl):
    """10^((spl - 96) / 10) (coder/psychoac.py:28-32)."""
    return _map(_lib.PSY_INTENSITY, spl)


def Thresh(f):
    """Threshold in quiet in dB SPL, frequencies below 10 Hz clamped (coder/psychoac.py:35-40)."""
    return _map(_lib.PSY_THRESH, f)


def Bark(f):
    """Bark scale 13 atan(0.76 f / 1000) + 3.5 atan((f / 7500)^2) (coder/psychoac.py:43-48)."""
    return _map(_lib.PSY_BARK, f)


# the 25 Zwicker critical band upper limits in Hz (coder/psychoac.py:100-103)
cbFreqLimits = np.array([100, 200, 300, 400, 510, 630, 770, 920, 1080, 1270, 1480, 1720, 2000, 2320, 2700, 3150, 3700,
                         4400, 5300, 6400, 7700, 9500, 12000, 15500, 24000])


def AssignMDCTLinesFromFreqLimits(nMDCTLines, sampleRate, flimit=cbFreqLimits):
    """Number of MDCT lines per band for band edges `flimit` (coder/psychoac.py:106-124):
    the last line of a band is floor(f / binwidth - 1/2); the first band whose edge lies above
    Nyquist takes whatever is left."""
    flimit = np.asarray(flimit)
    width = sampleRate / (2 * nMDCTLines)
    last = np.floor(flimit / width - 0.5)
    counts = np.diff(np.concatenate(([-1.0], last)))
    above = np.nonzero(flimit > sampleRate / 2)[0]
    if above.size:
        k = int(above[0])
        counts[k] = nMDCTLines - counts[:k].sum()
        counts[k + 1:] = 0
    return counts


class ScaleFactorBands:
    """Scale factor bands from a list of line counts; a band of 12 lines or fewer is merged
    into its upper neighbour (coder/psychoac.py:127-160).  Attributes: nBands, nLines,
    lowerLine, upperLine."""

    def __init__(self, nLines):
        n = [int(v) for v in nLines]
        i = 1
        while i < len(n):
            if n[i - 1] <= 12:
                n[i] += n[i - 1]
                del n[i - 1]
            else:
                i += 1
        self.nLines = np.array(n, dtype=int)
        self.nBands = len(n)
        self.lowerLine = np.concatenate(([0], np.cumsum(self.nLines)[:-1])).astype(int)
        self.upperLine = (self.lowerLine + self.nLines - 1).astype(int)


def _int_array(v):
    a = np.ascontiguousarray(np.asarray(v, dtype=np.int32))
    return a, a.ctypes.data_as(C.POINTER(C.c_int))


def _frames(data, MDCTdata, MDCTscale):
    d, was_np = to_device(data, torch.float64)
    N = d.shape[-1]
    if N not in (2048, 256):
        raise ValueError("audio_codec_b200.psychoac: only the codec's block sizes N = 2048 and N = 256 run on the GPU")
    d2 = d.reshape(-1, N)
    n = d2.shape[0]
    m = s = None
    if MDCTdata is not None:
        m, _ = to_device(MDCTdata, torch.float64)
        if m.shape[-1] != N // 2 or m.numel() != n * (N // 2):
            raise ValueError("MDCTdata must hold N/2 lines per frame of data")
        m = m.reshape(-1, N // 2)
        if isinstance(MDCTscale, torch.Tensor):
            s = MDCTscale.to(device=d.device, dtype=torch.int32).reshape(-1).contiguous()
        else:
            s = torch.as_tensor(np.asarray(MDCTscale, dtype=np.int32).reshape(-1), device=d.device)
        if s.numel() == 1 and n > 1:
            s = s.expand(n).contiguous()
    return d, d2, m, s, n, N, was_np


def getMaskedThreshold(data, MDCTdata, MDCTscale, sampleRate, sfBands):
    """Masked threshold in dB SPL at the N/2 MDCT line frequencies (coder/psychoac.py:163-217):
    Hann-windowed FFT intensity, every strict local maximum a tonal masker, max over the
    spreading curves and the threshold in quiet."""
    d, d2, _, _, n, N, was_np = _frames(data, None, None)
    ctx = _lib.default_context(int(sampleRate))
    mask = torch.empty((n, N // 2), dtype=torch.float64, device=d.device)
    _lib.check(_lib.lib.pac_calc_smr(ctx.h, ctx.cfg.precision, ptr(d2), None, None, n, N, 0, None, None, None, ptr(mask),
                                     _lib.stream_ptr()))
    return from_device(mask.reshape(*d.shape[:-1], N // 2), was_np)


def CalcSMRs(data, MDCTdata, MDCTscale, sampleRate, sfBands):
    """Per-band signal-to-mask ratio: max over the band's lines of SPL(4 (X / 2^scale)^2) minus the masked
    threshold (coder/psychoac.py:220-291).  Leading batch dimensions are allowed (MDCTscale per frame)."""
    d, d2, m, s, n, N, was_np = _frames(data, MDCTdata, MDCTscale)
    ctx = _lib.default_context(int(sampleRate))
    lo, lo_p = _int_array(sfBands.lowerLine)
    nl, nl_p = _int_array(sfBands.nLines)
    smr = torch.empty((n, sfBands.nBands), dtype=torch.float64, device=d.device)
    _lib.check(_lib.lib.pac_calc_smr(ctx.h, ctx.cfg.precision, ptr(d2), ptr(m), ptr(s), n, N, int(sfBands.nBands), lo_p, nl_p, ptr(smr),
                                     None, _lib.stream_ptr()))
    return from_device(smr.reshape(*d.shape[:-1], sfBands.nBands), was_np)
