"""Oracle: windows, MDCT/IMDCT, transient detector, uniform / block-FP quantisers.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  All arithmetic is fp64 with
the same NumPy calls the reference makes, so results are bit-identical to it
on the same machine.
"""
import numpy as np

N_SHORT = 256  # coder/codec.py:27


# ---------------------------------------------------------------- windows
def sine_window(N):
    """coder/window.py:21-24"""
    return np.sin(np.pi * (np.arange(N) + 0.5) / N)


def hann_window(N):
    """coder/window.py:36-39 (half-sample offset, not np.hanning)"""
    return 0.5 * (1 - np.cos(2 * np.pi * (np.arange(N) + 0.5) / N))


def kaiser_window(N, alpha=4.0):
    """coder/window.py:52-57 -- named KBDWindow there, but a plain Kaiser."""
    n = np.arange(N)
    return np.i0(np.pi * alpha * np.sqrt(1 - ((2 * n + 1) / N - 1) ** 2)) / np.i0(np.pi * alpha)


def start_window(N, Ns=N_SHORT):
    """coder/window.py:66-69"""
    pad = N // 4 - Ns // 4
    return np.concatenate((sine_window(N)[:N // 2], np.ones(pad),
                           sine_window(Ns)[Ns // 2:], np.zeros(pad)))


def stop_window(N, Ns=N_SHORT):
    """coder/window.py:79"""
    return np.flip(start_window(N, Ns))


def startstop_window(N, Ns=N_SHORT):
    """coder/window.py:87-90"""
    pad = N // 4 - Ns // 4
    return np.concatenate((np.zeros(pad), sine_window(Ns)[:Ns // 2], np.ones(2 * pad),
                           sine_window(Ns)[Ns // 2:], np.zeros(pad)))


def select_window(last, cur, nxt, N):
    """coder/codec.py:30-44: window shape from the three transient flags."""
    if cur:
        return sine_window(N)
    if last and nxt:
        return startstop_window(N)
    if last:
        return stop_window(N)
    if nxt:
        return start_window(N)
    return sine_window(N)


# ---------------------------------------------------------------- MDCT
def mdct(x, a, b):
    """coder/mdct.py:63-69 (forward, N-point complex FFT + twiddles)."""
    N = a + b
    n0 = (b + 1) / 2
    pre = x * np.exp(-1j * 2 * np.pi * np.arange(N) / (2 * N))
    f = np.fft.fft(pre, N)[:N // 2]
    out = (f * np.exp(-1j * 2 * np.pi * n0 * (np.arange(N // 2) + 0.5) / N)).real
    out *= 2 / N
    return out


def imdct(X, a, b):
    """coder/mdct.py:56-62 (inverse; ifft zero-pads N/2 -> N and carries 1/N)."""
    N = a + b
    n0 = (b + 1) / 2
    pre = X * np.exp(1j * 2 * np.pi * np.arange(N // 2) * n0 / N)
    t = np.fft.ifft(pre, N)
    out = (t * np.exp(1j * 2 * np.pi * (np.arange(N) + n0) / (2 * N))).real
    out *= 2 * N
    return out


def mdct_slow(x, a, b, inverse=False):
    """coder/mdct.py:22-38 (O(N^2) definition, used as a cross-check)."""
    N = a + b
    n0 = (b + 1) / 2.0
    if inverse:
        n = np.arange(N)[:, None]
        k = np.arange(N // 2)[None, :]
        return 2 * np.sum(x[None, :] * np.cos(2 * np.pi / N * (n + n0) * (k + 0.5)), axis=1)
    k = np.arange(N // 2)[:, None]
    n = np.arange(N)[None, :]
    return 2 / N * np.sum(x[None, :] * np.cos(2 * np.pi / N * (n + n0) * (k + 0.5)), axis=1)


# ---------------------------------------------------------------- transient detector
def transient_detect(block, thresh=4.5):
    """coder/detect_transients.py:5-23 on a (C, 2N) look-ahead buffer."""
    a = np.abs(block)
    mymax = np.max(a, axis=1)
    argmax = np.argmax(a, axis=1) + 500
    n = min(max(argmax), block.shape[1])
    avg = np.mean(np.abs(np.take(block, np.arange(n), axis=1))) if n else np.mean(a)
    if avg == 0:
        return False
    return bool(np.any(mymax / avg > thresh))


# ---------------------------------------------------------------- uniform quantiser
def quantize_uniform(x, nbits):
    """coder/quantize.py:14-36 (scalar midtread, sign-magnitude)."""
    if nbits <= 0:
        return 0
    s = 1 if x < 0 else 0
    if abs(x) >= 1:
        code = 2 ** (nbits - 1) - 1
    else:
        code = int(((2 ** nbits - 1) * abs(x) + 1) // 2)
    return int((s << (nbits - 1)) + code)


def dequantize_uniform(code, nbits):
    """coder/quantize.py:40-57"""
    if nbits <= 0:
        return 0
    s = code & (1 << (nbits - 1))
    mag = code & (2 ** (nbits - 1) - 1)
    sign = 1 if s == 0 else -1
    return sign * 2 * mag / (2 ** nbits - 1)


def vquantize_uniform(x, nbits):
    """coder/quantize.py:61-78"""
    x = np.asarray(x, dtype=np.float64)
    code = (((2 ** nbits - 1) * np.abs(x) + 1) // 2).astype(np.int64)
    code[np.abs(x) >= 1] = 2 ** (nbits - 1) - 1
    code[x < 0] += 1 << (nbits - 1)
    return code


def vdequantize_uniform(code, nbits):
    """coder/quantize.py:82-95"""
    code = np.asarray(code, dtype=np.int64)
    s = code & (1 << (nbits - 1))
    mag = code & (2 ** (nbits - 1) - 1)
    sign = np.where(s != 0, -1, 1)
    return sign * 2 * mag / (2 ** nbits - 1)


def scale_factor(x, nscale=3, nmant=5):
    """coder/quantize.py:99-125: leading zeros of the R-bit magnitude code."""
    R = 2 ** nscale - 1 + nmant
    code = quantize_uniform(x, R) & (2 ** (R - 1) - 1)
    zeros = (R - 1) - code.bit_length()
    return int(min(zeros, 2 ** nscale - 1))


def vmantissa(x, scale, nscale=3, nmant=5):
    """coder/quantize.py:229-250"""
    R = 2 ** nscale - 1 + nmant
    sc = vquantize_uniform(x, R)
    s = np.where((sc & (1 << (R - 1))) > 0, 1 << (nmant - 1), 0)
    code = sc & (2 ** (R - 1) - 1)
    if scale == 2 ** nscale - 1:
        return s + (code & (2 ** (nmant - 1) - 1))
    return s + ((code >> (R - scale - nmant)) & (2 ** (nmant - 1) - 1))


def vdequantize(scale, mant, nscale=3, nmant=5):
    """coder/quantize.py:254-274"""
    R = 2 ** nscale - 1 + nmant
    mant = np.asarray(mant, dtype=np.int64)
    s = np.where((mant & (1 << (nmant - 1))) > 0, 1 << (R - 1), 0)
    code = mant & (2 ** (nmant - 1) - 1)
    v = s + (code << max(R - scale - nmant, 0))
    if scale < 2 ** nscale - 1:
        v = v + np.where(code > 0, 1 << (R - scale - nmant - 1), 0)
    return vdequantize_uniform(v, R)


# ---------------------------------------------------------------- PCM mapping
def pcm_to_float(pcm):
    """coder/pcmfile.py:89-98: int16 -> signed fraction (and -32768 -> 0.0)."""
    c = np.asarray(pcm).astype(np.int64)
    neg = c < 0
    mag = np.where(neg, -c, c)
    out = vdequantize_uniform(mag, 16)
    out = np.where(neg, out * -1.0, out)
    return out + 0.0  # -0.0 -> 0.0 is value-equal; keep fp64


def float_to_pcm(x):
    """coder/pcmfile.py:127-134: signed fraction -> int16."""
    x = np.asarray(x, dtype=np.float64)
    neg = np.signbit(x)
    code = vquantize_uniform(np.abs(x), 16).astype(np.int16)
    return np.where(neg, -code, code).astype(np.int16)
