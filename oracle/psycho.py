"""Oracle: scale-factor bands, psychoacoustic model (SMRs), bit allocation.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
"""
import numpy as np

from .dsp import hann_window

CB_FREQ_LIMITS = np.array([  # coder/psychoac.py:100-103 (Zwicker critical bands)
    100, 200, 300, 400, 510, 630, 770, 920, 1080, 1270, 1480, 1720, 2000, 2320,
    2700, 3150, 3700, 4400, 5300, 6400, 7700, 9500, 12000, 15500, 24000])
EPS = np.finfo(float).eps
DBTOBITS = 6.2  # coder/bitalloc.py:61


def assign_lines(n_lines, fs, flimit=CB_FREQ_LIMITS):
    """coder/psychoac.py:106-124"""
    binw = fs / (2 * n_lines)
    c = np.floor(flimit / binw - 0.5)
    a = c - np.concatenate([[-1], c[:-1]])
    for i in range(len(a)):
        if flimit[i] > fs / 2:
            a[i] = n_lines - np.sum(a[0:i])
            a[i + 1:] = 0
            break
    return a


class Bands:
    """coder/psychoac.py:127-160 (merge of bands with <= 12 lines included)."""

    def __init__(self, n_lines):
        nl = [int(v) for v in np.asarray(n_lines)]
        i = 1
        while i < len(nl):
            if nl[i - 1] <= 12:
                nl[i] += nl[i - 1]
                del nl[i - 1]
            else:
                i += 1
        self.nLines = np.array(nl, dtype=np.int64)
        self.nBands = len(nl)
        self.lowerLine = np.concatenate([[0], np.cumsum(self.nLines)[:-1]]).astype(np.int64)
        self.upperLine = self.lowerLine + self.nLines - 1


def omitted_bands(bands, factor=2):
    """coder/sbr.py:6-9"""
    return np.where(bands.lowerLine >= bands.upperLine[-1] // factor)[0]


def spl(intensity):
    """coder/psychoac.py:10-25, array form (zeros -> 1e-8, clamp at -30 dB)."""
    i = np.where(intensity == 0, 1e-8, intensity)
    s = 96 + 10 * np.log10(np.abs(i) + EPS)
    return np.where(s < -30, -30.0, s)


def spl_scalar(i):
    """coder/psychoac.py:10-25, scalar form (exact zero -> -30)."""
    if i == 0:
        return -30
    s = 96 + 10 * np.log10(abs(i) + EPS)
    return -30 if s < -30 else s


def bark(f):
    """coder/psychoac.py:43-48"""
    return 13.0 * np.arctan(0.76 * f / 1000.0) + 3.5 * np.arctan((f / 7500.0) ** 2)


def thresh(f):
    """coder/psychoac.py:35-40 (f < 10 Hz clamped to 10)."""
    f = np.where(f < 10, 10.0, f)
    return (3.64 * (f / 1000) ** (-0.8) - 6.5 * np.exp(-0.6 * (f / 1000 - 3.3) ** 2)
            + 10 ** (-3) * (f / 1000) ** 4)


def fft_intensity(data):
    """coder/psychoac.py:170-179: Hann-FFT intensity spectrum, N/2+1 bins."""
    N = len(data)
    norm = 4 / (N ** 2 * np.mean(np.hanning(N) ** 2))
    X = np.fft.rfft(hann_window(N) * data)
    return norm * abs(X) ** 2


def find_peaks(I, freqs):
    """coder/psychoac.py:308-329: strict local maxima (last bin one-sided),
    each merged with the bin BELOW only; returns (freq, SPL) lists."""
    n = len(I)
    up = I[1:] > I[:-1]                       # I[k] > I[k-1], k = 1..n-1
    dn = np.concatenate([I[1:-1] > I[2:], [True]])
    pk = np.nonzero(up & dn)[0] + 1
    fr, sp = [], []
    for k in pk:
        ip = I[k - 1] + I[k]
        sp.append(spl_scalar(ip))
        fr.append((freqs[k - 1] * I[k - 1] + freqs[k] * I[k]) / ip)
    return fr, sp


def masked_threshold(data, n_mdct, fs):
    """coder/psychoac.py:163-217.  Tonal maskers only: the noise-masker loop
    at :197-211 never reaches `complete_mask` and is omitted."""
    N = len(data)
    I = fft_intensity(data)
    freqs = np.fft.rfftfreq(N, d=1 / fs)
    pf, ps = find_peaks(I, freqs)
    mf = (fs / (2 * n_mdct)) * (np.arange(n_mdct) + 0.5)
    z = bark(mf)
    th = thresh(mf)
    if not pf:
        return th
    pf = np.array(pf, dtype=np.float64)
    ps = np.array(ps, dtype=np.float64)
    zb = bark(pf)
    dz = z[None, :] - zb[:, None]
    adz = np.abs(dz) - 0.5
    up = (-27 + 0.367 * np.maximum(ps - 40, 0))[:, None] * adz
    gain = np.where(dz < -0.5, -27 * adz, np.where(dz > 0.5, up, 0.0))
    s = ps[:, None] + gain - 16                      # Masker.drop = 16 (tonal)
    curves = spl(10 ** ((s - 96) / 10))              # SPL(Intensity(.)) round trip
    return np.maximum(np.max(curves, axis=0), th)


def calc_smrs(data, mdct_scaled, scale, fs, bands):
    """coder/psychoac.py:220-291"""
    mask = masked_threshold(data, len(mdct_scaled), fs)
    m = mdct_scaled / 2 ** scale
    ms = spl(m ** 2 * 4.0)
    d = ms - mask
    return np.array([np.amax(d[lo:hi + 1]) for lo, hi in zip(bands.lowerLine, bands.upperLine)])


def bit_alloc(budget, max_mant, n_bands, n_lines, smr):
    """coder/bitalloc.py:62-121 (water-filling with the Nflip needle, np.round
    half-to-even, 201-pass cap)."""
    max_mant = min(max_mant, 16)
    n_lines = np.asarray(n_lines)
    bits = np.zeros(n_bands, dtype=np.int64)
    flagged = np.zeros(n_bands, dtype=bool)
    nflip = 0
    count = 0
    while True:
        valid = ~flagged
        N = np.sum(n_lines[valid])
        if N == 0:
            N = N + 1e-12
        ropt = budget / N + (1.0 / DBTOBITS) * (smr[valid] - np.sum(n_lines[valid] * smr[valid]) / N)
        fr = ropt - np.floor(ropt) - 0.5
        dopt = np.sort(fr[fr > 0])
        if nflip > len(dopt):
            nflip -= 1
        else:
            level = 0.0 if nflip == 0 else dopt[nflip - 1]
            bits[valid] = np.round(ropt - level)
        bits[bits > max_mant] = max_mant
        flagged = bits < 2
        bits[flagged] = 0
        stable = np.logical_xor(flagged, valid).all()
        tot = np.sum(bits * n_lines)
        if stable and tot <= budget:
            break
        if stable and tot > budget:
            nflip += 1
        count += 1
        if count > 200:
            break
    return bits
