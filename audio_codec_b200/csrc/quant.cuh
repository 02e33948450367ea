// Uniform midtread quantiser, block-floating-point scale factor / mantissa,
// and MSB-first bit writers / readers over 32-bit words.
// Functional spec: coder/quantize.py:14-125, :229-274; coder/bitpack.py:37-175.
#pragma once
#include "pac_common.cuh"

// (2^n - 1) as the double Python's int*float multiplication sees it
// (int -> float conversion rounds to nearest even; exact for n <= 53).
__device__ __forceinline__ double pow2m1(int n) {
    if (n <= 53) return (double)((1ull << n) - 1ull);
    return ldexp(1.0, n);            // 2^n - 1 rounds up to 2^n
}

// Magnitude code of QuantizeUniform(|a|, nBits): coder/quantize.py:30-33.
// Valid for 1 <= nBits <= 64 (fields wider than 64 bits are flagged by callers).
__device__ __forceinline__ unsigned long long quant_mag(double a, int nBits) {
    a = fabs(a);
    if (a >= 1.0) return (nBits >= 64) ? 0x7fffffffffffffffull : ((1ull << (nBits - 1)) - 1ull);
    double t = floor((pow2m1(nBits) * a + 1.0) * 0.5);
    if (t >= 18446744073709551615.0) return ~0ull;
    return (unsigned long long)t;
}

__device__ __forceinline__ unsigned long long quantize_uniform(double a, int nBits) {
    if (nBits <= 0) return 0ull;
    unsigned long long code = quant_mag(a, nBits);
    if (a < 0.0) code += 1ull << (nBits - 1);
    return code;
}

// DequantizeUniform: sign * 2 * code / (2^n - 1), coder/quantize.py:40-57.  Python
// divides two exact ints (correctly rounded); for n <= 53 both are exact doubles so
// the IEEE division is the same value.  For wider fields 2^n - 1 is not
// representable and the quotient is taken against 2^n (error < 1 ulp).
__device__ __forceinline__ double dequantize_uniform(unsigned long long q, int nBits) {
    if (nBits <= 0) return 0.0;
    unsigned long long s = q & (1ull << (nBits - 1));
    unsigned long long code = q & ((1ull << (nBits - 1)) - 1ull);
    double v;
    if (nBits <= 53) v = (double)(2ull * code) / (double)((1ull << nBits) - 1ull);
    else v = ldexp((double)code, 1 - nBits);
    return s ? -v : v;
}

// ScaleFactor(aNum, nScaleBits, nMantBits): coder/quantize.py:99-125.
__device__ __forceinline__ int scale_factor(double a, int nScale, int nMant) {
    int R = (1 << nScale) - 1 + nMant;
    unsigned long long code = quant_mag(a, R) & ((1ull << (R - 1)) - 1ull);
    int blen = code ? 64 - __clzll((long long)code) : 0;
    int zeros = (R - 1) - blen;
    int cap = (1 << nScale) - 1;
    return zeros < cap ? zeros : cap;
}

// vMantissa for one value: coder/quantize.py:229-250.
__device__ __forceinline__ unsigned mantissa_bfp(double a, int scale, int nScale, int nMant) {
    int R = (1 << nScale) - 1 + nMant;
    unsigned long long code = quant_mag(a, R);
    unsigned s = a < 0.0 ? (1u << (nMant - 1)) : 0u;
    unsigned m = (1u << (nMant - 1)) - 1u;
    if (scale == (1 << nScale) - 1) return s + ((unsigned)code & m);
    return s + ((unsigned)(code >> (R - scale - nMant)) & m);
}

// vDequantize for one value: coder/quantize.py:254-274.
__device__ __forceinline__ double dequantize_bfp(int scale, unsigned mant, int nScale, int nMant) {
    int R = (1 << nScale) - 1 + nMant;
    unsigned long long code = mant & ((1u << (nMant - 1)) - 1u);
    int sh = R - scale - nMant;
    unsigned long long v = code << (sh > 0 ? sh : 0);
    if (scale < (1 << nScale) - 1 && code > 0) v += 1ull << (sh - 1);
    if (mant & (1u << (nMant - 1))) v |= 1ull << (R - 1);
    return dequantize_uniform(v, R);
}

// ---- MSB-first bit I/O on big-endian-in-word u32 buffers (bit 0 of the stream is
// bit 31 of word 0).  Converting words with __byte_perm gives the byte order of
// coder/bitpack.py.
struct BitW {
    unsigned* w;
    int pos;
};
// single writer (one lane); buffer must be zero-initialised
__device__ __forceinline__ void bw_put(BitW& b, unsigned long long v, int n) {
    while (n > 0) {
        int off = b.pos & 31, room = 32 - off;
        int take = n < room ? n : room;
        unsigned chunk = (unsigned)((v >> (n - take)) & ((take == 32) ? 0xffffffffull : ((1ull << take) - 1ull)));
        b.w[b.pos >> 5] |= chunk << (room - take);
        b.pos += take;
        n -= take;
    }
}
// concurrent writers (disjoint bit ranges): atomicOr at an absolute position
__device__ __forceinline__ void bw_put_at(unsigned* w, int pos, unsigned long long v, int n) {
    while (n > 0) {
        int off = pos & 31, room = 32 - off;
        int take = n < room ? n : room;
        unsigned chunk = (unsigned)((v >> (n - take)) & ((take == 32) ? 0xffffffffull : ((1ull << take) - 1ull)));
        if (chunk) atomicOr(&w[pos >> 5], chunk << (room - take));
        pos += take;
        n -= take;
    }
}

struct BitR {
    const unsigned char* p;   // byte stream (coder/bitpack.py layout)
    long long pos;            // bit cursor
    long long end;            // total bits available
    int cap;                  // 0, or: p is 8-byte aligned and cap (a multiple of 8) bytes are readable -> word reads
};
// n (1..64) bits at bit position pos, MSB first; bits past `end` (a whole number of bytes) read as 0.
// Out of line (one copy per kernel: the byte loop is ~1 KB of SASS per call site) and pure, so the
// cursor struct of the caller stays in registers.
static __device__ __noinline__ unsigned long long br_peek(const unsigned char* p, long long pos, long long end, int n) {
    const long long byte = pos >> 3, endByte = end >> 3;
    const int off = (int)(pos & 7);
    const int nb = (off + n + 7) >> 3;               // bytes touched, <= 9
    unsigned long long hi = 0ull;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        unsigned b = (i < nb && byte + i < endByte) ? p[byte + i] : 0u;
        hi = (hi << 8) | b;
    }
    unsigned long long v = hi << off;
    if (nb == 9) {
        unsigned b8 = (byte + 8 < endByte) ? p[byte + 8] : 0u;
        v |= (unsigned long long)(b8 >> (8 - off));
    }
    return n == 64 ? v : v >> (64 - n);
}
// Same result from two aligned 8-byte words (p 8-byte aligned, cap readable bytes, whatever lies
// between end and cap is ignored): ~25 instructions instead of a 9-step byte loop.
__device__ __forceinline__ unsigned long long br_bswap64(unsigned long long w) {
    const unsigned lo = (unsigned)w, hi = (unsigned)(w >> 32);
    return ((unsigned long long)__byte_perm(lo, 0, 0x0123) << 32) | (unsigned long long)__byte_perm(hi, 0, 0x0123);
}
__device__ __forceinline__ unsigned long long br_peek_words(const unsigned char* p, int cap, long long pos, long long end,
                                                            int n) {
    const long long avail = end - pos;
    if (avail <= 0) return 0ull;
    const unsigned long long* q = reinterpret_cast<const unsigned long long*>(p);
    const int wi = (int)(pos >> 6), off = (int)(pos & 63);
    const unsigned long long w0 = (wi + 1) * 8 <= cap ? br_bswap64(q[wi]) : 0ull;
    const unsigned long long w1 = (wi + 2) * 8 <= cap ? br_bswap64(q[wi + 1]) : 0ull;
    unsigned long long v = off ? (w0 << off) | (w1 >> (64 - off)) : w0;
    if (n < 64) v >>= (64 - n);
    if (avail < n) v &= ~((1ull << (int)(n - avail)) - 1ull);
    return v;
}
// Next n (<= 64) bits of the stream.
__device__ __forceinline__ unsigned long long br_get(BitR& r, int n) {
    if (n <= 0) return 0ull;
    const long long pos = r.pos;
    r.pos = pos + n;
    return r.cap ? br_peek_words(r.p, r.cap, pos, r.end, n) : br_peek(r.p, pos, r.end, n);
}
