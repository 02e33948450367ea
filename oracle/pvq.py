"""Oracle: gain-shape vector quantiser (mu-law gain + pyramid VQ shape with
recursive mid/side band splitting).  TEST INFRASTRUCTURE ONLY.

Follows coder/gain_shape_quantize.py.  Field readers take any object with a
ReadBits(n) method (oracle.bits.BitReader).
"""
import numpy as np

from .dsp import dequantize_uniform, quantize_uniform

SPLIT_BITS = 32  # coder/gain_shape_quantize.py:27
EPS = np.finfo(float).eps

_rows = {}  # L -> python list N(L, 0..k)


def _row(L, kmax):
    """N(L, k) for k <= kmax by the Fischer recurrence
    N(L,K) = N(L-1,K) + N(L-1,K-1) + N(L,K-1); N(L,0)=1; N(0,K>=1)=0
    (coder/gain_shape_quantize.py:87-99)."""
    r = _rows.get(L)
    if r is None:
        r = _rows[L] = [1]
    if len(r) > kmax:
        return r
    if L == 0:
        r.extend([0] * (kmax + 1 - len(r)))
        return r
    up = _row(L - 1, kmax)
    for k in range(len(r), kmax + 1):
        r.append(up[k] + up[k - 1] + r[k - 1])
    return r


def codebook_size(L, K):
    return _row(L, K)[K]


def k_for_bits(L, max_bits):
    """coder/gain_shape_quantize.py:275-291: largest K with N(L,K) <= 2^bits and
    the field width ceil(log2(N + eps)) (which is 1, not 0, when N == 1)."""
    lim = 2 ** max_bits
    if L <= 1 and max_bits >= 1:
        # N(1,k) = 2 for every k: the reference loops forever here
        raise ValueError("degenerate PVQ leaf (L<=1 with bits): undefined in the reference")
    k = 1
    while codebook_size(L, k) <= lim:
        k += 1
    n = codebook_size(L, k - 1)
    return k - 1, int(np.ceil(np.log2(n + EPS)))


def pvq_search(x, K):
    """coder/gain_shape_quantize.py:41-54: floor then greedy pulse placement."""
    l1 = np.sum(np.abs(x))
    if l1 == 0:
        # reference divides by zero -> NaN -> garbage int cast (undefined)
        raise ValueError("degenerate PVQ input (all-zero shape): undefined in the reference")
    xh = np.abs(K * x / l1)
    p = np.floor(xh)
    rem = K - np.sum(np.abs(p))
    while rem > 0:
        i = np.argmax(np.abs(xh) - p)
        p[i] += 1
        rem -= 1
    p *= np.sign(x)
    return p.astype(np.int64)


def pvq_index(y, K):
    """coder/gain_shape_quantize.py:105-124: Fischer enumeration index."""
    b = 0
    k = K
    l = len(y)
    for v in y:
        v = int(v)
        a = abs(v)
        if a == 1:
            b += codebook_size(l - 1, k) + (codebook_size(l - 1, k - 1) if v < 0 else 0)
        elif a > 1:
            b += codebook_size(l - 1, k) + 2 * sum(codebook_size(l - 1, k - j) for j in range(1, a))
            if v < 0:
                b += codebook_size(l - 1, k - a)
        k -= a
        l -= 1
        if k == 0:
            break
    return int(b)


def pvq_vector(b, L, K):
    """coder/gain_shape_quantize.py:127-186 (`decoder.step_1..5`), as a loop."""
    x = np.zeros(L, dtype=np.int64)
    xb = 0
    k, l, i = K, L, 0
    while True:
        if b == xb:                                   # step 1
            x[i] = 0
            if k > 0:                                 # step 5
                x[L - 1] = k - abs(x[i])
            return x
        n0 = codebook_size(l - 1, k)                  # step 2
        if b < xb + n0:
            x[i] = 0
        else:
            xb += n0
            j = 1
            while True:                               # step 3
                j = min(j, k)
                nj = codebook_size(l - 1, k - j)
                if b < xb + 2 * codebook_size(l - 1, max(k - j, 0)):
                    if b >= xb + nj:
                        x[i] = -j
                    elif xb <= b:
                        x[i] = j
                    break
                xb += 2 * nj
                if j < k:
                    j += 1
                    continue
                # j == k and still not found: reference jumps back to step 1
                j = None
                break
            if j is None:
                continue
        k -= abs(x[i])                                # step 4
        l -= 1
        i += 1
        if k <= 0:
            return x


def quantize_pvq(x, nbits):
    """coder/gain_shape_quantize.py:249-260"""
    nrm = np.linalg.norm(x)
    assert np.allclose(nrm, 1.0) or np.allclose(nrm, 0.0), f"not a unit vector: {nrm}"
    K, used = k_for_bits(len(x), nbits)
    return pvq_index(pvq_search(x, K), K), used


def dequantize_pvq(idx, L, K):
    """coder/gain_shape_quantize.py:263-272"""
    x = pvq_vector(idx, L, K).astype(float)
    n = np.linalg.norm(x)
    if n != 0:
        x = x / np.linalg.norm(x)
    return x


def gain_shape_alloc(R, L, k_fine=0):
    """coder/gain_shape_quantize.py:57-62"""
    rg = np.floor(R / L + 0.5 * np.log2(L) - k_fine)
    return int(rg), int(max(R - rg, 0))


def mu_law(x, mu=255):
    """coder/gain_shape_quantize.py:294-295"""
    return np.sign(x) * np.log(1 + mu * np.abs(x)) / np.log(1 + mu)


def inv_mu_law(y, mu=255):
    """coder/gain_shape_quantize.py:298-299"""
    return np.sign(y) / mu * ((1 + mu) ** np.abs(y) - 1)


def ms_alloc(a_ms, theta, length):
    """coder/gain_shape_quantize.py:302-312"""
    if theta == 0:
        a_mid = 0
    else:
        a_mid = int(np.floor((a_ms - (length - 1) * np.log2(np.tan(abs(theta)) + EPS)) / 2))
    a_mid = min(max(a_mid, 0), a_ms)
    return a_mid, max(a_ms - a_mid, 0)


def split_encode(x, nbits, k_fine=0):
    """coder/gain_shape_quantize.py:315-403: fields in pre-order
    [theta, mid..., side...]; leaves are PVQ indices of <= 32 bits."""
    if nbits <= SPLIT_BITS:
        idx, used = quantize_pvq(x, min(32, nbits))
        return [idx], [used]
    n = len(x)
    mid = n // 2
    half = int(np.ceil(n / 2))
    left = np.concatenate([x[:mid], [0]]) if half > mid else x[:mid]
    right = x[mid:]
    M = (left + right) / 2
    S = (left - right) / 2
    nm, ns = np.linalg.norm(M), np.linalg.norm(S)
    m = M / nm if nm != 0 else M
    s = S / ns if ns != 0 else S
    theta = 0 if nm == 0 else np.arctan(ns / nm)
    a_theta, a_ms = gain_shape_alloc(nbits, half, k_fine)
    t_idx = quantize_uniform(theta / (np.pi / 2), a_theta)
    t_q = dequantize_uniform(t_idx, a_theta) * (np.pi / 2)
    a_mid, a_side = ms_alloc(a_ms, t_q, half)
    idx, bits = [t_idx], [a_theta]
    for vec, a in ((m, a_mid), (s, a_side)):
        if a > SPLIT_BITS:
            i2, b2 = split_encode(vec, a, k_fine)
            assert sum(b2) <= a
            idx += i2
            bits += b2
        elif a > 0:
            i1, b1 = quantize_pvq(vec, a)
            idx.append(i1)
            bits.append(b1)
    return idx, bits


def split_decode(pb, nbits, n, k_fine=0):
    """coder/gain_shape_quantize.py:406-473"""
    if nbits <= SPLIT_BITS:
        K, used = k_for_bits(n, nbits)
        x = dequantize_pvq(pb.ReadBits(used), n, K)
        nrm = np.linalg.norm(x)
        if nrm != 0:
            x /= nrm
        return x, used
    half = int(np.ceil(n / 2))
    a_theta, a_ms = gain_shape_alloc(nbits, half, k_fine)
    th = dequantize_uniform(pb.ReadBits(a_theta), a_theta) * (np.pi / 2)
    a_mid, a_side = ms_alloc(a_ms, th, half)
    total = a_theta
    parts = []
    for a in (a_mid, a_side):
        if a > SPLIT_BITS:
            v, used = split_decode(pb, a, half, k_fine)
            total += used
        elif a > 0:
            K, used = k_for_bits(half, a)
            v = dequantize_pvq(pb.ReadBits(used), half, K)
            total += used
        else:
            v = np.zeros((half,))
        parts.append(v)
    mh, sh = parts
    left = mh * np.cos(th) + sh * np.sin(th)
    right = mh * np.cos(th) - sh * np.sin(th)
    left /= np.sqrt(2)
    right /= np.sqrt(2)
    if half > n // 2:
        left = left[:len(left) - 1]
    x = np.concatenate([left, right])
    nrm = np.linalg.norm(x)
    if nrm != 0:
        x /= nrm
    return x, total


def quantize_gain_shape(x, nbits, k_fine=0):
    """coder/gain_shape_quantize.py:476-512 -> (indices, widths)."""
    L = len(x)
    bg, bs = gain_shape_alloc(nbits, L, k_fine)
    gain = np.linalg.norm(x)
    if gain == 0:
        return [0], [0]
    if bs != 0:
        idx, bits = split_encode(x / gain, bs, k_fine)
        used = sum(bits)
        assert used <= bs, "Used more bits than allocated in quantize"
        bg += bs - used
    else:
        idx, bits = [], []
    g = mu_law(gain / L)
    if bg < 0:
        bg = 0
    return idx + [quantize_uniform(g, bg)], bits + [bg]


def dequantize_gain_shape(pb, nbits, L, k_fine=0):
    """coder/gain_shape_quantize.py:515-538"""
    bg, bs = gain_shape_alloc(nbits, L, k_fine)
    if bs != 0:
        shape, used = split_decode(pb, bs, L, k_fine)
        assert used <= bs, "Used more bits than allocated in dequantize"
        bg += bs - used
    else:
        shape = np.ones((L,))
    g = dequantize_uniform(pb.ReadBits(bg), bg)
    return inv_mu_law(g) * L * shape
