// Gain-shape vector quantiser, encode side: mu-law gain + pyramid VQ shape with
// the recursive mid/side split flattened onto an explicit stack.  One warp per
// scale-factor band.  Functional spec: coder/gain_shape_quantize.py:30-124 and
// :249-403, :476-512 (pvq_search, pvq_codebook_size, encode_pvq_vector,
// pvq_compute_k_for_R, gain_shape_alloc, bit_allocation_ms, split_band_encode,
// quantize_gain_shape).
#pragma once
#include "quant.cuh"

#define PVQ_SPLIT_BITS 32
#define PVQ_ERR_FIELD_TOO_WIDE 1   // a theta field wider than 64 bits
#define PVQ_ERR_DEGENERATE 2       // L<=1 leaf with bits, or an all-zero shape (undefined in the reference)
#define PVQ_ERR_STACK 4

struct PvqTab {
    const unsigned long long* N;
    const int* off;
    const double* log2tab;   // log2(L), L <= PAC_PVQ_LMAX
    const int2* KW;          // {K, width}[L][bits], bits 0..32
    const double* theta;     // PacDev::thetaTab or nullptr (direct evaluation)
};

// N(l,k): rows 0..3 closed form, others from the table; entries past a row's end
// are "more than 2^32".
__device__ __forceinline__ unsigned long long pvq_N(const PvqTab& t, int l, long long k) {
    if (k <= 0) return 1ull;
    if (l <= 0) return 0ull;
    if (l == 1) return 2ull;
    if (l == 2) return 4ull * (unsigned long long)k;
    if (l == 3) return 4ull * (unsigned long long)k * (unsigned long long)k + 2ull;
    int o = t.off[l], len = t.off[l + 1] - o;
    if (k >= len) return 1ull << 40;
    return t.N[o + k];
}

// pvq_compute_k_for_R: largest K with N(L,K) <= 2^bits, and the field width
// ceil(log2(N(L,K) + eps)) -- which is 1 (not 0) for N == 1.  bits <= 32.
__device__ __forceinline__ void pvq_k_for_bits(const PvqTab& t, int L, int bits, long long* K, int* used, int* err) {
    if (L <= 1 && bits >= 1) *err |= PVQ_ERR_DEGENERATE;      // the reference never terminates here
    const int2 kw = t.KW[L * 33 + bits];
    *K = kw.x;
    *used = kw.y;
}

// gain_shape_alloc(R, L, 0): coder/gain_shape_quantize.py:57-62
__device__ __forceinline__ void gs_alloc(const PvqTab& t, int R, int L, int* rg, int* rs) {
    double g = floor((double)R / (double)L + 0.5 * t.log2tab[L]);
    *rg = (int)g;
    double s = (double)R - g;
    *rs = s > 0.0 ? (int)s : 0;
}

// bit_allocation_ms: coder/gain_shape_quantize.py:302-312
__device__ __forceinline__ void ms_alloc(int aMs, double theta, int len, int* aMid, int* aSide) {
    long long m = 0;
    if (theta != 0.0) {
        double v = floor(((double)aMs - (double)(len - 1) * log2(tan(fabs(theta)) + PAC_EPS)) / 2.0);
        m = v > 1e15 ? (long long)1e15 : (v < -1e15 ? (long long)-1e15 : (long long)v);
    }
    if (m < 0) m = 0;
    if (m > aMs) m = aMs;
    *aMid = (int)m;
    *aSide = aMs - (int)m > 0 ? aMs - (int)m : 0;
}

// The quantised angle of a split node (aT bits, code ti) and log2(tan|theta| + eps), from the
// table when it covers aT (entry = its index, for the decoder's cos / sin lookup), else evaluated.
__device__ __forceinline__ void theta_terms(const PvqTab& t, unsigned long long ti, int aT, double* theta, double* l2t,
                                            int* entry) {
    if (t.theta && aT >= 1 && aT <= PAC_THETA_BITS) {
        const int e = (1 << aT) + (int)ti;
        *theta = t.theta[4 * e];
        *l2t = t.theta[4 * e + 1];
        *entry = e;
    } else {
        const double th = dequantize_uniform(ti, aT) * 1.5707963267948966;
        *theta = th;
        *l2t = th != 0.0 ? pm_log2(pm_tan(fabs(th)) + PAC_EPS) : 0.0;
        *entry = -1;
    }
}
// bit_allocation_ms with the logarithm already in hand
__device__ __forceinline__ void ms_alloc_l2t(int aMs, double theta, double l2t, int len, int* aMid, int* aSide) {
    long long m = 0;
    if (theta != 0.0) {
        double v = floor(((double)aMs - (double)(len - 1) * l2t) / 2.0);
        m = v > 1e15 ? (long long)1e15 : (v < -1e15 ? (long long)-1e15 : (long long)v);
    }
    if (m < 0) m = 0;
    if (m > aMs) m = aMs;
    *aMid = (int)m;
    *aSide = aMs - (int)m > 0 ? aMs - (int)m : 0;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// quantize_pvq on one warp: x = vec[0..n) (unit L2 norm), residual scratch res[0..n),
// pulse magnitudes y[0..n).  Returns the Fischer index; *used = field width.
__device__ __forceinline__ unsigned long long warp_quantize_pvq(const PvqTab& t, const double* vec, double* res,
                                                                int* y, int n, int bits, int* used, int* err) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    long long K;
    pvq_k_for_bits(t, n, bits, &K, used, err);
    if (K == 0) return 0ull;
    // pvq_search (coder/gain_shape_quantize.py:41-54)
    double l1 = 0.0;
    for (int i = lane; i < n; i += 32) l1 += fabs(vec[i]);
    l1 = warp_sum(l1);
    if (l1 == 0.0) {
        *err |= PVQ_ERR_DEGENERATE;
        return 0ull;
    }
    double ps = 0.0;
    for (int i = lane; i < n; i += 32) {
        double xh = fabs(((double)K * vec[i]) / l1);
        double p = floor(xh);
        res[i] = xh - p;
        y[i] = (int)p;
        ps += p;
    }
    ps = warp_sum(ps);
    __syncwarp();
    long long rem = K - (long long)ps;
    for (; rem > 0; --rem) {
        double bv = -1e300;
        int bi = 0x7fffffff;
        for (int i = lane; i < n; i += 32) {
            double r = res[i];
            if (r > bv) { bv = r; bi = i; }           // strided: first (lowest) index wins per lane
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            double ov = __shfl_xor_sync(FULL, bv, o);
            int oi = __shfl_xor_sync(FULL, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) {
            res[bi] -= 1.0;
            y[bi] += 1;
        }
        __syncwarp();
    }
    // encode_pvq_vector (coder/gain_shape_quantize.py:105-124), lanes over elements
    unsigned long long acc = 0ull;
    long long carry = 0;                               // pulses consumed before this chunk
    for (int base = 0; base < n; base += 32) {
        int i = base + lane;
        int a = i < n ? y[i] : 0;
        int inc = a;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int v = __shfl_up_sync(FULL, inc, o);
            if (lane >= o) inc += v;
        }
        long long k = K - carry - (inc - a);           // pulses left when element i is reached
        if (a > 0) {
            // N(l-1,k) + 2 sum_{j<a} N(l-1,k-j) counts the vectors whose first component is smaller in
            // magnitude than a: N(l,k) - N(l,k-a) - N(l-1,k-a) (pvq_first_below)
            int l = n - i;
            bool neg = vec[i] < 0.0;
            unsigned long long below = pvq_N(t, l - 1, k - a);
            unsigned long long c = pvq_N(t, l, k) - pvq_N(t, l, k - a) - below;
            if (neg) c += below;
            acc += c;
        }
        carry += __shfl_sync(FULL, inc, 31);
    }
    return warp_sum_u64(acc);
}

struct PvqNode { short off, cap, n, buf; int bits; };

// Optional record of the (index, width) fields a band emits, in stream order: what
// quantize_gain_shape returns as two Python lists (coder/gain_shape_quantize.py:476-512).
// Written by lane 0 only; nullptr in the fused stream kernels.
struct FieldLog {
    unsigned long long* val;
    int* width;
    int* count;     // fields recorded so far
    int cap;
};
__device__ __forceinline__ void log_field(FieldLog* lg, unsigned long long v, int w) {
    if (!lg) return;
    int i = *lg->count;
    if (i < lg->cap) { lg->val[i] = v; lg->width[i] = w; }
    *lg->count = i + 1;
}

// quantize_gain_shape for one band on one warp.
//   x[0..L)   band lines (shared, read only)
//   wsA/wsB   double workspaces, this band's region of capacity cap = pow2 >= L
//   yws       int workspace, same region
//   out       bit writer positioned at this band's private bit buffer
// Returns the number of bits written (0: all-zero band, caller zeroes bitAlloc).
__device__ __forceinline__ int warp_quantize_gain_shape(const PvqTab& t, const double* x, int L, int R, double* wsA,
                                                        double* wsB, int* yws, int cap, BitW& out, int* err,
                                                        FieldLog* lg = nullptr) {
    const int lane = threadIdx.x & 31;
    int bg, bs;
    gs_alloc(t, R, L, &bg, &bs);
    double ss = 0.0;
    for (int i = lane; i < L; i += 32) ss += x[i] * x[i];
    double gain = sqrt(warp_sum(ss));
    if (gain == 0.0) return 0;
    int total = 0;
    if (bs != 0) {
        for (int i = lane; i < L; i += 32) wsA[i] = x[i] / gain;
        __syncwarp();
        PvqNode st[24];
        int sp = 0;
        st[sp++] = {0, (short)cap, (short)L, 0, bs};
        int used = 0;
        while (sp > 0) {
            PvqNode nd = st[--sp];
            double* cur = (nd.buf ? wsB : wsA) + nd.off;
            double* oth = (nd.buf ? wsA : wsB) + nd.off;
            if (nd.bits <= PVQ_SPLIT_BITS) {
                int w;
                unsigned long long idx = warp_quantize_pvq(t, cur, oth, yws + nd.off, nd.n, nd.bits, &w, err);
                if (lane == 0) { bw_put(out, idx, w); log_field(lg, idx, w); }
                used += w;
                continue;
            }
            const int n = nd.n, mid = n / 2, half = (n + 1) / 2, hc = nd.cap / 2;
            double sm = 0.0, sd = 0.0;
            for (int i = lane; i < half; i += 32) {
                double l = i < mid ? cur[i] : 0.0;       // odd n: the LEFT half is zero padded at its end
                double r = cur[mid + i];
                double M = (l + r) / 2.0, S = (l - r) / 2.0;
                oth[i] = M;
                oth[hc + i] = S;
                sm += M * M;
                sd += S * S;
            }
            double nm = sqrt(warp_sum(sm)), ns = sqrt(warp_sum(sd));
            __syncwarp();
            for (int i = lane; i < half; i += 32) {
                if (nm != 0.0) oth[i] = oth[i] / nm;
                if (ns != 0.0) oth[hc + i] = oth[hc + i] / ns;
            }
            __syncwarp();
            double theta = nm == 0.0 ? 0.0 : atan(ns / nm);
            int aT, aMs, aMid, aSide;
            gs_alloc(t, nd.bits, half, &aT, &aMs);
            const double hp = 1.5707963267948966;        // np.pi / 2
            if (aT > 64) { *err |= PVQ_ERR_FIELD_TOO_WIDE; aT = 64; }
            unsigned long long ti = quantize_uniform(theta / hp, aT);
            double tq = dequantize_uniform(ti, aT) * hp;
            ms_alloc(aMs, tq, half, &aMid, &aSide);
            if (lane == 0) { bw_put(out, ti, aT); log_field(lg, ti, aT); }
            used += aT;
            if (sp + 2 > 24) { *err |= PVQ_ERR_STACK; break; }
            if (aSide > 0) st[sp++] = {(short)(nd.off + hc), (short)hc, (short)half, (short)(1 - nd.buf), aSide};
            if (aMid > 0) st[sp++] = {nd.off, (short)hc, (short)half, (short)(1 - nd.buf), aMid};
        }
        bg += bs - used;
        total = used;
    }
    double g = pm_log(1.0 + 255.0 * (gain / (double)L)) / log(256.0);      // mu_law_fn
    if (bg < 0) bg = 0;
    if (bg > 64) { *err |= PVQ_ERR_FIELD_TOO_WIDE; bg = 64; }
    unsigned long long gi = quantize_uniform(g, bg);
    if (lane == 0) { bw_put(out, gi, bg); log_field(lg, gi, bg); }
    __syncwarp();
    return total + bg;
}

// ---------------------------------------------------------------------------------------------
// Level-synchronous form of quantize_gain_shape for the fused frame encoder.  The split tree
// (theta fields, bit allocations, leaf sizes, and therefore every field position) does not depend
// on the pulse search, and the mid and side subtrees of a node are independent.  So, per frame:
//   roots    warp per band: gain, normalised copy, root node
//   levels   all warps: the nodes of one tree depth of ALL bands at a time (a node's children are
//            appended to the pool, so each depth is a contiguous index range)
//   layout   thread per band: pre-order walk -> bit position of every field, theta and gain
//            fields written
//   leaves   thread per leaf, all threads: pulse search + Fischer index
struct TNode {                 // 32 bytes
    unsigned long long val;    // internal: quantised theta
    int bits;                  // shape bits given to the node; for a leaf, after layout: its bit position
    int K;                     // -2 not yet visited, -1 internal, >= 0 leaf with K pulses
    short off, n;              // absolute workspace offset, length
    short kid[2];              // mid / side child (-1: none)
    unsigned char capLog2, buf, band, w;    // w: field width (theta bits or index bits)
    short logSlot, pad;
};

struct BandEnc {
    double gain;
    int bg, bs;
    int root;                  // -1: no shape tree
    int L;
};

// np.add.reduce of |a[i]| in NumPy's pairwise order (loops_utils.h.src pairwise_sum): blocks of at
// most 128 elements are summed with 8 accumulators, larger arrays are halved recursively.
template <int XS = 1>   // XS: element stride (the batched quantiser interleaves 32 lanes)
static __device__ __noinline__ double np_abs_sum_block(const double* a, int n) {
    if (n < 8) {
        double r = 0.0;
        for (int i = 0; i < n; ++i) r += fabs(a[i * XS]);
        return r;
    }
    double r[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = fabs(a[j * XS]);
    int i;
    for (i = 8; i < n - (n % 8); i += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] += fabs(a[(i + j) * XS]);
    }
    double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
    for (; i < n; ++i) res += fabs(a[i * XS]);
    return res;
}
template <int XS = 1>
__device__ __forceinline__ double np_abs_sum(const double* a, int n) {
    if (n <= 128) return np_abs_sum_block<XS>(a, n);
    // explicit post-order walk of the halving tree (depth <= 4 for n <= 1024)
    struct Fr { const double* p; int n; int state; double left; };
    Fr st[8];
    int sp = 0;
    st[sp++] = {a, n, 0, 0.0};
    double ret = 0.0;
    while (sp > 0) {
        Fr& f = st[sp - 1];
        if (f.n <= 128) { ret = np_abs_sum_block<XS>(f.p, f.n); --sp; continue; }
        int n2 = f.n / 2;
        n2 -= n2 % 8;
        if (f.state == 0) { f.state = 1; st[sp++] = {f.p, n2, 0, 0.0}; continue; }
        if (f.state == 1) { f.left = ret; f.state = 2; st[sp++] = {f.p + n2 * XS, f.n - n2, 0, 0.0}; continue; }
        ret = f.left + ret;
        --sp;
    }
    return ret;
}

// quantize_pvq for one leaf on ONE thread (coder/gain_shape_quantize.py:41-54, :105-124).
template <int XS = 1>
__device__ __forceinline__ unsigned long long thread_quantize_pvq(const PvqTab& t, const double* __restrict__ vec,
                                                                  double* __restrict__ res, int* __restrict__ y,
                                                                  int n, long long K, int* err) {
    if (K == 0) return 0ull;
    const double l1 = np_abs_sum<XS>(vec, n);
    if (l1 == 0.0) { *err |= PVQ_ERR_DEGENERATE; return 0ull; }
    long long ps = 0;
    const PacDiv byL1 = pac_div_by(l1);
#pragma unroll 4
    for (int i = 0; i < n; ++i) {
        double xh = fabs(pac_div((double)K * vec[i * XS], byL1));
        double p = floor(xh);
        res[i * XS] = xh - p;
        y[i * XS] = (int)p;
        ps += (long long)p;
    }
    for (long long rem = K - ps; rem > 0; --rem) {        // np.argmax: first maximum
        double bv = res[0];
        int bi = 0;
        for (int i = 1; i < n; ++i) {
            double r = res[i * XS];
            if (r > bv) { bv = r; bi = i; }
        }
        res[bi * XS] -= 1.0;
        y[bi * XS] += 1;
    }
    unsigned long long acc = 0ull;
    long long k = K;
    for (int i = 0; i < n && k > 0; ++i) {
        const int a = y[i * XS];
        if (a > 0) {
            const int l = n - i;
            const unsigned long long below = pvq_N(t, l - 1, k - a);
            unsigned long long c = pvq_N(t, l, k) - pvq_N(t, l, k - a) - below;
            if (vec[i * XS] < 0.0) c += below;
            acc += c;
            k -= a;
        }
    }
    return acc;
}

// Levels step, part 1 - one tree node on one warp: the vector work of an internal node (mid/side
// split, norms, normalised halves).  The two norms go to nodes[i].val (mid) and side[i] for part 2.
__device__ __forceinline__ void warp_split_node(TNode* nodes, int i, double* side, double* wsA, double* wsB) {
    const int lane = threadIdx.x & 31;
    const TNode nd = nodes[i];
    if (nd.bits <= PVQ_SPLIT_BITS) return;                // leaf: nothing to split
    double* cur = (nd.buf ? wsB : wsA) + nd.off;
    double* oth = (nd.buf ? wsA : wsB) + nd.off;
    const int n = nd.n, mid = n / 2, half = (n + 1) / 2, hc = (1 << nd.capLog2) / 2;
    double sm = 0.0, sd = 0.0;
    for (int k = lane; k < half; k += 32) {
        double l = k < mid ? cur[k] : 0.0;               // odd n: the LEFT half is zero padded at its end
        double r = cur[mid + k];
        double M = (l + r) / 2.0, S = (l - r) / 2.0;
        oth[k] = M;
        oth[hc + k] = S;
        sm += M * M;
        sd += S * S;
    }
    const double nm = sqrt(warp_sum(sm)), ns = sqrt(warp_sum(sd));
    __syncwarp();
    for (int k = lane; k < half; k += 32) {
        if (nm != 0.0) oth[k] = oth[k] / nm;
        if (ns != 0.0) oth[hc + k] = oth[hc + k] / ns;
    }
    if (lane == 0) {
        nodes[i].val = (unsigned long long)__double_as_longlong(nm);
        side[i] = ns;
    }
}

// Levels step, part 2 - one tree node on ONE thread: the scalar work (angle, its quantisation, the
// bit split with its fp64 transcendentals, the children).  All nodes of a level run in parallel
// lanes, so a level pays the atan / tan / log2 latency chain once instead of once per node.
// count: pool fill; the node's children are appended.
__device__ __forceinline__ void thread_plan_node(const PvqTab& t, TNode* nodes, int i, const double* side, int* count,
                                                 int maxNodes, int* err) {
    const TNode nd = nodes[i];
    if (nd.bits <= PVQ_SPLIT_BITS) {                      // leaf: pulses and index width from the table
        long long K;
        int w;
        pvq_k_for_bits(t, nd.n, nd.bits, &K, &w, err);
        nodes[i].K = (int)K;
        nodes[i].w = (unsigned char)w;
        return;
    }
    const int half = (nd.n + 1) / 2, hc = (1 << nd.capLog2) / 2;
    const double nm = __longlong_as_double((long long)nd.val), ns = side[i];
    const double theta = nm == 0.0 ? 0.0 : pm_atan(ns / nm);
    int aT, aMs, aMid, aSide;
    gs_alloc(t, nd.bits, half, &aT, &aMs);
    const double hp = 1.5707963267948966;                // np.pi / 2
    if (aT > 64) { *err |= PVQ_ERR_FIELD_TOO_WIDE; aT = 64; }
    const unsigned long long ti = quantize_uniform(theta / hp, aT);
    double tq, l2t;
    int entry;
    theta_terms(t, ti, aT, &tq, &l2t, &entry);
    ms_alloc_l2t(aMs, tq, l2t, half, &aMid, &aSide);
    const int nk = (aMid > 0) + (aSide > 0);
    int slot = nk ? atomicAdd(count, nk) : 0;
    short k0 = -1, k1 = -1;
    if (slot + nk > maxNodes) {
        *err |= PVQ_ERR_STACK;
    } else {
        TNode c;
        c.val = 0ull; c.K = -2; c.n = (short)half; c.kid[0] = c.kid[1] = -1;
        c.capLog2 = (unsigned char)(nd.capLog2 - 1); c.buf = (unsigned char)(1 - nd.buf); c.band = nd.band; c.w = 0;
        c.logSlot = -1; c.pad = 0;
        if (aMid > 0) { c.off = nd.off; c.bits = aMid; k0 = (short)slot; nodes[slot++] = c; }
        if (aSide > 0) { c.off = (short)(nd.off + hc); c.bits = aSide; k1 = (short)slot; nodes[slot] = c; }
    }
    nodes[i].val = ti;
    nodes[i].K = -1;
    nodes[i].w = (unsigned char)aT;
    nodes[i].kid[0] = k0;
    nodes[i].kid[1] = k1;
}

// Layout step for one band on ONE thread: stream order is pre-order [theta, mid.., side..], then
// the gain, which also absorbs the unused shape bits.  words/bit0: the band's private bit buffer
// (zero-initialised).  Returns the bits the band emits.
__device__ __forceinline__ int thread_layout_band(TNode* nodes, const BandEnc& be, unsigned* words, int bit0, int* err,
                                                  FieldLog* lg) {
    int pos = bit0;
    if (be.root >= 0) {
        short st[48];
        int sp = 0;
        st[sp++] = (short)be.root;
        while (sp > 0) {
            const int i = st[--sp];
            const int K = nodes[i].K, w = nodes[i].w;
            if (K >= 0) {                                  // leaf; an empty one is index 0 in a 1-bit field
                if (K > 0 && lg) nodes[i].logSlot = (short)*lg->count;
                log_field(lg, 0ull, w);
                nodes[i].bits = pos;
            } else {
                BitW bw = {words, pos};
                bw_put(bw, nodes[i].val, w);
                log_field(lg, nodes[i].val, w);
                if (sp + 2 > 48) { *err |= PVQ_ERR_STACK; break; }
                if (nodes[i].kid[1] >= 0) st[sp++] = nodes[i].kid[1];
                if (nodes[i].kid[0] >= 0) st[sp++] = nodes[i].kid[0];
            }
            pos += w;
        }
    }
    int bg = be.bg + (be.bs != 0 ? be.bs - (pos - bit0) : 0);
    const double g = pm_log(1.0 + 255.0 * (be.gain / (double)be.L)) / log(256.0);      // mu_law_fn
    if (bg < 0) bg = 0;
    if (bg > 64) { *err |= PVQ_ERR_FIELD_TOO_WIDE; bg = 64; }
    const unsigned long long gi = quantize_uniform(g, bg);
    BitW bw = {words, pos};
    bw_put(bw, gi, bg);
    log_field(lg, gi, bg);
    return pos + bg - bit0;
}

// Leaves step: every thread of the CTA takes pool entries; leaf j -> warp j % nWarps, lane j / nWarps
// (the serial per-leaf walks diverge, so they are spread over all warps and schedulers).
__device__ __forceinline__ void cta_quantize_leaves(const PvqTab& t, const TNode* nodes, int nNodes, double* wsA,
                                                    double* wsB, int* yws, unsigned* words, unsigned long long* logVal,
                                                    int logCap, int tid, int nThreads, int* err) {
    const int nWarps = nThreads >> 5;
    for (int i = (tid & 31) * nWarps + (tid >> 5); i < nNodes; i += nThreads) {
        const TNode nd = nodes[i];
        if (nd.K <= 0) continue;
        int werr = 0;
        const unsigned long long idx = thread_quantize_pvq(t, (nd.buf ? wsB : wsA) + nd.off, (nd.buf ? wsA : wsB) + nd.off,
                                                           yws + nd.off, nd.n, nd.K, &werr);
        bw_put_at(words, nd.bits, idx, nd.w);
        if (logVal && nd.logSlot >= 0 && nd.logSlot < logCap) logVal[(size_t)nd.band * logCap + nd.logSlot] = idx;
        if (werr) atomicOr(err, werr);
    }
}

// ---------------------------------------------------------------------------------------------
// quantize_gain_shape for one band in ONE THREAD (k_pvq, encode.cu): the 32 lanes of a warp hold the
// same band of 32 different channel-blocks.  The tree is walked depth first, so the fields come out
// in stream order and go through a register bit accumulator into the band's private word buffer
// (written once per word, no read-modify-write).  Same arithmetic as warp_quantize_gain_shape with
// the level-synchronous encoder's table lookups (theta_terms / ms_alloc_l2t); sums run in index order.
// x, wsA, wsB, yws, words: per-lane arrays with element stride XS (lane-interleaved global memory).
struct TBitW { unsigned* w; unsigned long long acc; int n, widx, cap; };
template <int XS>
__device__ __forceinline__ void tbw_put32(TBitW& b, unsigned v, int w) {       // w in 1..32
    b.acc = (b.acc << w) | (unsigned long long)(w == 32 ? v : (v & ((1u << w) - 1u)));
    b.n += w;
    if (b.n >= 32) {
        if (b.widx < b.cap) b.w[b.widx * XS] = (unsigned)(b.acc >> (b.n - 32));
        ++b.widx;
        b.n -= 32;
    }
}
template <int XS>
__device__ __forceinline__ void tbw_put(TBitW& b, unsigned long long v, int w) {   // w in 0..64
    if (w > 32) { tbw_put32<XS>(b, (unsigned)(v >> 32), w - 32); w = 32; }
    if (w > 0) tbw_put32<XS>(b, (unsigned)v, w);
}
template <int XS>
__device__ __forceinline__ int tbw_flush(TBitW& b) {
    const int total = b.widx * 32 + b.n;
    if (b.n > 0 && b.widx < b.cap) b.w[b.widx * XS] = (unsigned)(b.acc << (32 - b.n));
    return total;
}

// Returns the number of bits written (0: all-zero band, caller zeroes bitAlloc).  env: the band is
// ONE value, envVal (SBR envelope, coder/codec.py:508-510), instead of x[0..L).
template <int XS>
__device__ __forceinline__ int thread_quantize_gain_shape(const PvqTab& t, const double* __restrict__ x, bool env,
                                                          double envVal, int L, int R, double* wsA, double* wsB, int* yws,
                                                          int cap, unsigned* words, int capWords, int* err) {
    int bg, bs;
    gs_alloc(t, R, L, &bg, &bs);
    double ss = 0.0;
    if (env) ss = envVal * envVal;
    else for (int i = 0; i < L; ++i) { const double v = x[i * XS]; ss += v * v; }
    const double gain = sqrt(ss);
    if (gain == 0.0) return 0;
    TBitW out = {words, 0ull, 0, 0, capWords};
    if (bs != 0) {
        // A node's vector stays UN-normalised in the workspace; its norm travels with the node and the
        // reader divides (the same operations on the same values as normalise-then-read, without the
        // extra load + store pass per level).  The root reads the lines themselves, norm = gain.
        struct TEncNode { short off, cap, n, buf; int bits; double nr; };
        TEncNode st[24];
        int sp = 0, used = 0;
        if (env) wsA[0] = envVal / gain;
        st[sp++] = {0, (short)cap, (short)L, 0, bs, env ? 0.0 : gain};
        bool root = true, fail = false;
        // Each round: (1) split nodes until a leaf is on top of the stack, (2) quantise that leaf.  The lanes
        // of a warp (same band, 32 blocks) so meet at the pulse search, the long step.
        while (sp > 0 && !fail) {
            while (sp > 0 && st[sp - 1].bits > PVQ_SPLIT_BITS) {
                const TEncNode nd = st[--sp];
                const double* cur = (root && !env) ? x : ((nd.buf ? wsB : wsA) + nd.off * XS);
                double* __restrict__ oth = (nd.buf ? wsA : wsB) + nd.off * XS;
                const double nr = nd.nr;
                const int n = nd.n, mid = n / 2, half = (n + 1) / 2, hc = nd.cap / 2;
                double sm = 0.0, sd = 0.0;
                const bool dv = nr != 0.0;
                const PacDiv byNr = pac_div_by(dv ? nr : 1.0);
                int i = 0;
                for (; i + 2 <= mid; i += 2) {                 // two elements per trip: four loads in flight
                    double l0 = cur[i * XS], l1 = cur[(i + 1) * XS], r0 = cur[(mid + i) * XS], r1 = cur[(mid + i + 1) * XS];
                    if (dv) { l0 = pac_div(l0, byNr); l1 = pac_div(l1, byNr); r0 = pac_div(r0, byNr); r1 = pac_div(r1, byNr); }
                    const double M0 = (l0 + r0) / 2.0, S0 = (l0 - r0) / 2.0, M1 = (l1 + r1) / 2.0, S1 = (l1 - r1) / 2.0;
                    oth[i * XS] = M0; oth[(hc + i) * XS] = S0; oth[(i + 1) * XS] = M1; oth[(hc + i + 1) * XS] = S1;
                    sm += M0 * M0; sd += S0 * S0;
                    sm += M1 * M1; sd += S1 * S1;
                }
                for (; i < half; ++i) {
                    double l = i < mid ? cur[i * XS] : 0.0;           // odd n: the LEFT half is zero padded at its end
                    double r = cur[(mid + i) * XS];
                    if (dv) { l = pac_div(l, byNr); r = pac_div(r, byNr); }
                    const double M = (l + r) / 2.0, S = (l - r) / 2.0;
                    oth[i * XS] = M;
                    oth[(hc + i) * XS] = S;
                    sm += M * M;
                    sd += S * S;
                }
                const double nm = sqrt(sm), ns = sqrt(sd);
                const double theta = nm == 0.0 ? 0.0 : pm_atan(ns / nm);
                int aT, aMs, aMid, aSide, entry;
                gs_alloc(t, nd.bits, half, &aT, &aMs);
                const double hp = 1.5707963267948966;        // np.pi / 2
                if (aT > 64) { *err |= PVQ_ERR_FIELD_TOO_WIDE; aT = 64; }
                const unsigned long long ti = quantize_uniform(theta / hp, aT);
                double tq, l2t;
                theta_terms(t, ti, aT, &tq, &l2t, &entry);
                ms_alloc_l2t(aMs, tq, l2t, half, &aMid, &aSide);
                tbw_put<XS>(out, ti, aT);
                used += aT;
                root = false;
                if (sp + 2 > 24) { *err |= PVQ_ERR_STACK; fail = true; break; }
                if (aSide > 0) st[sp++] = {(short)(nd.off + hc), (short)hc, (short)half, (short)(1 - nd.buf), aSide, ns};
                if (aMid > 0) st[sp++] = {nd.off, (short)hc, (short)half, (short)(1 - nd.buf), aMid, nm};
            }
            if (sp > 0 && !fail) {
                const TEncNode nd = st[--sp];
                const double* cur = (root && !env) ? x : ((nd.buf ? wsB : wsA) + nd.off * XS);
                double* __restrict__ oth = (nd.buf ? wsA : wsB) + nd.off * XS;
                const double nr = nd.nr;
                long long K;
                int w;
                pvq_k_for_bits(t, nd.n, nd.bits, &K, &w, err);
                // the leaf search reads its vector three times: normalise it once, in place (a root leaf: into wsA)
                double* vec = (nd.buf ? wsB : wsA) + nd.off * XS;
                if (nr != 0.0 && K != 0) {
                    const PacDiv byNr = pac_div_by(nr);
                    int i = 0;
                    for (; i + 4 <= nd.n; i += 4) {
                        const double v0 = cur[i * XS], v1 = cur[(i + 1) * XS], v2 = cur[(i + 2) * XS], v3 = cur[(i + 3) * XS];
                        vec[i * XS] = pac_div(v0, byNr); vec[(i + 1) * XS] = pac_div(v1, byNr);
                        vec[(i + 2) * XS] = pac_div(v2, byNr); vec[(i + 3) * XS] = pac_div(v3, byNr);
                    }
                    for (; i < nd.n; ++i) vec[i * XS] = pac_div(cur[i * XS], byNr);
                }
                const unsigned long long idx = thread_quantize_pvq<XS>(t, vec, oth, yws + nd.off * XS, nd.n, K, err);
                tbw_put<XS>(out, idx, w);
                used += w;
                root = false;
            }
        }
        bg += bs - used;
    }
    const double g = pm_log(1.0 + 255.0 * (gain / (double)L)) / log(256.0);      // mu_law_fn
    if (bg < 0) bg = 0;
    if (bg > 64) { *err |= PVQ_ERR_FIELD_TOO_WIDE; bg = 64; }
    tbw_put<XS>(out, quantize_uniform(g, bg), bg);
    const int total = tbw_flush<XS>(out);
    if (out.widx >= capWords && total > capWords * 32) *err |= PVQ_ERR_STACK;
    return total;
}
