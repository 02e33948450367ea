// Gain-shape vector dequantiser: parses a band's fields straight from the
// bitstream (field widths depend on the theta values just read), one warp per
// band.  Functional spec: coder/gain_shape_quantize.py:127-186 (decoder.step_*),
// :263-272 (dequantize_pvq), :406-473 (split_band_decode), :515-538
// (dequantize_gain_shape).
#pragma once
#include "pvq.cuh"

// Fischer index -> pulse vector: the inverse of encode_pvq_vector, i.e. what the reference's
// recursive step_1..step_5 walk produces for a valid index (coder/gain_shape_quantize.py:127-186).
// Element i with l components and k pulses left: codes [0, N(l-1,k)) have x_i = 0; then for
// j = 1..k a block of N(l-1,k-j) codes with x_i = +j and one with x_i = -j.  The +j block starts
// at A(j) = N(l,k) - N(l,k-j) - N(l-1,k-j) (the number of vectors with |x_i| < j), so j is found by
// bisection instead of the reference's linear scan.  Serial; run by ONE thread.  x[0..L) receives
// the signed pulse counts as doubles.
template <int XS = 1>   // XS: element stride of x (the batched dequantiser interleaves 32 lanes)
__device__ __forceinline__ double pvq_vector(const PvqTab& t, unsigned long long b, int L, long long K, double* x,
                                             int* err) {
    double ss = 0.0;                                      // sum of squares in index order (zeros add nothing)
    long long k = K;
    unsigned long long r = b;
    int i = 0;
    for (; i < L && k > 0; ++i) {
        const int l = L - i;
        if (r == 0ull) break;                             // step 1 + 5: the rest of the pulses sit on the last component
        const unsigned long long n0 = pvq_N(t, l - 1, k);
        if (r < n0) {
            // a run of zeros: x_i' = 0 while r < N(l'-1, k).  N(m, k) grows with m, so the end of the run
            // is the smallest m with N(m, k) > r, found by bisection (wide bands with a handful of
            // pulses are the critical path of the leaf phase otherwise)
            int lo = 1, hi = l - 1;                       // N(0, k) = 0 <= r, N(l-1, k) > r
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (pvq_N(t, mid, k) > r) hi = mid; else lo = mid + 1;
            }
            const int stop = L - lo;                      // first position whose remaining length is lo
            for (; i < stop; ++i) x[i * XS] = 0.0;
            --i;
            continue;
        }
        const unsigned long long nlk = pvq_N(t, l, k);
        long long lo = 1, hi = k;
        unsigned long long nj, nlj;                       // N(l-1, k-lo), N(l, k-lo)
        if (l >= 5) {
            // both rows come from the table: fetch their bases once instead of once per probe (same values as pvq_N)
            const int oL = t.off[l], o1 = t.off[l - 1];
            const int lenL = t.off[l + 1] - oL, len1 = oL - o1;
            const unsigned long long* rowL = t.N + oL;
            const unsigned long long* row1 = t.N + o1;
            const unsigned long long big = 1ull << 40;
            while (lo < hi) {
                const long long mid = (lo + hi + 1) >> 1, x = k - mid;        // 0 <= x < k
                const unsigned long long vL = x == 0 ? 1ull : (x >= lenL ? big : rowL[x]);
                const unsigned long long v1 = x == 0 ? 1ull : (x >= len1 ? big : row1[x]);
                if (nlk - vL - v1 <= r) lo = mid; else hi = mid - 1;
            }
            const long long x = k - lo;
            nlj = x == 0 ? 1ull : (x >= lenL ? big : rowL[x]);
            nj = x == 0 ? 1ull : (x >= len1 ? big : row1[x]);
        } else {
            while (lo < hi) {
                long long mid = (lo + hi + 1) >> 1;
                unsigned long long a = nlk - pvq_N(t, l, k - mid) - pvq_N(t, l - 1, k - mid);
                if (a <= r) lo = mid; else hi = mid - 1;
            }
            nj = pvq_N(t, l - 1, k - lo);
            nlj = pvq_N(t, l, k - lo);
        }
        r -= nlk - nlj - nj;
        bool neg = false;
        if (r >= nj) { neg = true; r -= nj; }
        if (r >= nj) { *err |= PVQ_ERR_DEGENERATE; r = 0ull; }    // index outside the codebook
        const double xv = neg ? -(double)lo : (double)lo;
        x[i * XS] = xv;
        ss += xv * xv;
        k -= lo;
    }
    for (; i < L; ++i) x[i * XS] = 0.0;
    if (k > 0) {
        if (r != 0ull) *err |= PVQ_ERR_DEGENERATE;
        x[(L - 1) * XS] = (double)k;
        ss += (double)k * (double)k;
    }
    return ss;
}

struct DecNode { short off, cap, n, buf; int bits; double theta; short state, root; };

// dequantize_gain_shape for one band on one warp.  Reads its fields from br and
// leaves the cursor after them (all lanes hold the same cursor).  out[0..L) <- gain * shape.
__device__ __forceinline__ void warp_dequantize_gain_shape(const PvqTab& t, BitR& br, int R, int L, double* wsA,
                                                           double* wsB, int cap, double* out, int* err) {
    const int lane = threadIdx.x & 31;
    int bg, bs;
    gs_alloc(t, R, L, &bg, &bs);
    if (bs != 0) {
        DecNode st[24];
        int sp = 0;
        st[sp++] = {0, (short)cap, (short)L, 0, bs, 0.0, 0, 1};
        int used = 0;
        while (sp > 0) {
            DecNode& nd = st[sp - 1];
            double* cur = (nd.buf ? wsB : wsA) + nd.off;
            double* oth = (nd.buf ? wsA : wsB) + nd.off;
            if (nd.bits <= PVQ_SPLIT_BITS) {              // leaf
                long long K;
                int w;
                pvq_k_for_bits(t, nd.n, nd.bits, &K, &w, err);
                unsigned long long idx = br_get(br, w);
                used += w;
                if (lane == 0) pvq_vector(t, idx, nd.n, K, cur, err);
                __syncwarp();
                const int reps = nd.root ? 2 : 1;         // a root leaf is normalised twice (:214-218)
                for (int r = 0; r < reps; ++r) {
                    double ss = 0.0;
                    for (int i = lane; i < nd.n; i += 32) ss += cur[i] * cur[i];
                    double nr = sqrt(warp_sum(ss));
                    if (nr != 0.0)
                        for (int i = lane; i < nd.n; i += 32) cur[i] = cur[i] / nr;
                    __syncwarp();
                }
                --sp;
                continue;
            }
            const int n = nd.n, mid = n / 2, half = (n + 1) / 2, hc = nd.cap / 2;
            if (nd.state == 0) {                          // first visit: theta, then children
                int aT, aMs, aMid, aSide;
                gs_alloc(t, nd.bits, half, &aT, &aMs);
                if (aT > 64) { *err |= PVQ_ERR_FIELD_TOO_WIDE; aT = 64; }
                unsigned long long ti = br_get(br, aT);
                used += aT;
                nd.theta = dequantize_uniform(ti, aT) * 1.5707963267948966;
                ms_alloc(aMs, nd.theta, half, &aMid, &aSide);
                nd.state = 1;
                for (int i = lane; i < half; i += 32) {   // children that get no bits stay zero
                    if (aMid <= 0) oth[i] = 0.0;
                    if (aSide <= 0) oth[hc + i] = 0.0;
                }
                __syncwarp();
                if (sp + 2 > 24) { *err |= PVQ_ERR_STACK; return; }
                const short off = nd.off, buf = nd.buf;   // nd is invalidated by the pushes below
                if (aSide > 0) st[sp++] = {(short)(off + hc), (short)hc, (short)half, (short)(1 - buf), aSide, 0.0, 0, 0};
                if (aMid > 0) st[sp++] = {off, (short)hc, (short)half, (short)(1 - buf), aMid, 0.0, 0, 0};
                continue;
            }
            // children done: recombine (coder/gain_shape_quantize.py:449-463)
            const double ct = cos(nd.theta), sn = sin(nd.theta), r2 = sqrt(2.0);
            double ss = 0.0;
            for (int i = lane; i < half; i += 32) {
                double m = oth[i], s = oth[hc + i];
                double l = (m * ct + s * sn) / r2;
                double r = (m * ct - s * sn) / r2;
                if (i < mid) { cur[i] = l; ss += l * l; }  // odd n: the padded last element of left is dropped
                cur[mid + i] = r;
                ss += r * r;
            }
            double nr = sqrt(warp_sum(ss));
            __syncwarp();
            if (nr != 0.0)
                for (int i = lane; i < n; i += 32) cur[i] = cur[i] / nr;
            __syncwarp();
            --sp;
        }
        bg += bs - used;
    } else {
        for (int i = lane; i < L; i += 32) wsA[i] = 1.0;
        __syncwarp();
    }
    if (bg > 64) { *err |= PVQ_ERR_FIELD_TOO_WIDE; bg = 64; }
    if (bg < 0) bg = 0;
    unsigned long long gi = br_get(br, bg);
    double gq = dequantize_uniform(gi, bg);
    // inv_mu_law_fn: sign(y)/mu * ((1+mu)^|y| - 1)
    double sg = gq > 0.0 ? 1.0 : (gq < 0.0 ? -1.0 : 0.0);
    double gain = sg / 255.0 * (pm_pow(256.0, fabs(gq)) - 1.0) * (double)L;
    for (int i = lane; i < L; i += 32) out[i] = gain * wsA[i];
    __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// Three-phase form of dequantize_gain_shape for the fused frame decoder.  The bitstream parse
// is serial within a band (field widths depend on the theta values just read) but cheap; the
// expensive part, index -> pulse vector, is independent per leaf.  So:
//   phase A  warp per band: walk the split tree in stream order, read every field, append one
//            DNode per tree node to a CTA-wide list (leaves carry their index and K)
//   phase B  thread per leaf, all threads of the CTA: pvq_vector
//   phase C  warp per band: walk the band's nodes in reverse (children before parents):
//            normalise leaves, recombine mid/side, then apply the gain
// Offsets in DNode are absolute positions in the CTA's ping-pong workspaces wsA / wsB.
struct DNode {
    unsigned long long u;      // leaf: Fischer index; internal: theta (double bits)
    int K;                     // leaf: pulses; internal: thetaTab entry (-1: evaluate cos / sin)
    short off, n, prev;        // workspace offset, length, previous node of the same band (-1: none)
    unsigned char capLog2, flags;
};
enum { DN_LEAF = 1, DN_ROOT = 2, DN_BUF = 4, DN_ZMID = 8, DN_ZSIDE = 16 };   // Z*: that child got no bits (reads as zeros)

struct BandDec {               // per band, written by phase A (lane 0), read by phase C
    double gain;
    int tail;                  // last node appended (-1: no shape bits, shape = ones)
    int L;
};

struct ParseNode { short off, cap, n, buf; int bits; short root; };

// Sub-warp groups of G lanes (G = 8, 16 or 32): the per-band phases are latency bound (serial tree
// walks, dependent fp64 transcendentals), so several bands per warp in flight beat one band
// spread over 32 mostly idle lanes.
template <int G> __device__ __forceinline__ unsigned grp_mask() {
    return G == 32 ? 0xffffffffu : (((1u << G) - 1u) << ((threadIdx.x & 31) & ~(G - 1)));
}
template <int G> __device__ __forceinline__ double grp_sum(double v, unsigned gmask) {
#pragma unroll
    for (int o = G / 2; o; o >>= 1) v += __shfl_xor_sync(gmask, v, o);
    return v;
}

template <int G>
__device__ __forceinline__ int dn_append(DNode* nodes, int* nNodes, int maxNodes, const DNode& nd, int* err) {
    const int lane = threadIdx.x & (G - 1);
    const unsigned gmask = grp_mask<G>();
    int slot = 0;
    if (lane == 0) slot = atomicAdd(nNodes, 1);
    slot = __shfl_sync(gmask, slot, 0, G);
    if (slot >= maxNodes) { *err |= PVQ_ERR_STACK; return -1; }
    if (lane == 0) nodes[slot] = nd;
    return slot;
}

// Phase A for one band on one group of G lanes (G = 8: four bands per warp in flight).  base = offset of the band's region (capacity cap, a power of
// two >= L) in the workspaces.  All lanes hold the same cursor and run the scalar logic redundantly.
template <int G>
__device__ __forceinline__ void grp_parse_gain_shape(const PvqTab& t, BitR& br, int R, int L, double* wsA, double* wsB,
                                                      int base, int cap, DNode* nodes, int* nNodes, int maxNodes,
                                                      BandDec* bd, int* err) {
    const int lane = threadIdx.x & (G - 1);
    const unsigned gmask = grp_mask<G>();
    int bg, bs, tail = -1;
    gs_alloc(t, R, L, &bg, &bs);
    if (bs != 0) {
        ParseNode st[24];
        int sp = 0, used = 0;
        st[sp++] = {(short)base, (short)cap, (short)L, 0, bs, 1};
        while (sp > 0) {
            const ParseNode nd = st[--sp];
            DNode dn;
            dn.off = nd.off; dn.n = nd.n; dn.prev = (short)tail;
            dn.capLog2 = (unsigned char)(31 - __clz((int)nd.cap));
            dn.flags = (unsigned char)((nd.buf ? DN_BUF : 0) | (nd.root ? DN_ROOT : 0));
            if (nd.bits <= PVQ_SPLIT_BITS) {
                long long K;
                int w;
                pvq_k_for_bits(t, nd.n, nd.bits, &K, &w, err);
                dn.u = br_get(br, w);
                dn.K = (int)K;
                dn.flags |= DN_LEAF;
                used += w;
                tail = dn_append<G>(nodes, nNodes, maxNodes, dn, err);
                if (tail < 0) break;
                continue;
            }
            const int half = (nd.n + 1) / 2, hc = nd.cap / 2;
            int aT, aMs, aMid, aSide;
            gs_alloc(t, nd.bits, half, &aT, &aMs);
            if (aT > 64) { *err |= PVQ_ERR_FIELD_TOO_WIDE; aT = 64; }
            const unsigned long long ti = br_get(br, aT);
            used += aT;
            double theta, l2t;
            int entry;
            theta_terms(t, ti, aT, &theta, &l2t, &entry);
            ms_alloc_l2t(aMs, theta, l2t, half, &aMid, &aSide);
            double* oth = (nd.buf ? wsA : wsB) + nd.off;
            for (int i = lane; i < half; i += G) {       // children that get no bits stay zero
                if (aMid <= 0) oth[i] = 0.0;
                if (aSide <= 0) oth[hc + i] = 0.0;
            }
            dn.u = (unsigned long long)__double_as_longlong(theta);
            dn.K = entry;                                   // >= 0: cos / sin come from the table in phase C
            tail = dn_append<G>(nodes, nNodes, maxNodes, dn, err);
            if (tail < 0) break;
            if (sp + 2 > 24) { *err |= PVQ_ERR_STACK; break; }
            if (aSide > 0) st[sp++] = {(short)(nd.off + hc), (short)hc, (short)half, (short)(1 - nd.buf), aSide, 0};
            if (aMid > 0) st[sp++] = {nd.off, (short)hc, (short)half, (short)(1 - nd.buf), aMid, 0};
        }
        bg += bs - used;
    }
    if (bg > 64) { *err |= PVQ_ERR_FIELD_TOO_WIDE; bg = 64; }
    if (bg < 0) bg = 0;
    const unsigned long long gi = br_get(br, bg);
    const double gq = dequantize_uniform(gi, bg);
    const double sg = gq > 0.0 ? 1.0 : (gq < 0.0 ? -1.0 : 0.0);
    if (lane == 0) {
        bd->gain = sg / 255.0 * (pm_pow(256.0, fabs(gq)) - 1.0) * (double)L;     // inv_mu_law_fn * L
        bd->tail = tail;
        bd->L = L;
    }
}

// dequantize_gain_shape for one band in ONE THREAD (k_dequant, decode.cu): the 32 lanes of a warp
// hold the same band of 32 different channel-blocks, so they run the same loops with similar trip
// counts, nothing waits at a barrier, and thousands of these serial chains (field widths depend on
// the angles just read; the index decode is sequential by nature) are in flight per SM.  Same tree
// walk as warp_dequantize_gain_shape above.  wsA / wsB / out: per-lane arrays with element stride XS
// (lane-interleaved global memory: every access of a converged warp is one coalesced line).
// A node leaves its vector UN-normalised in the workspace and hands its norm to the parent, which
// divides as it reads (the same operations on the same values as normalise-then-read, one load and one
// store pass fewer per node).
struct TDecNode { short off, cap, n, par; unsigned char buf, state, root, z; int bits, entry; double theta, nrMid, nrSide; };
static __device__ __noinline__ void pm_sincos_pair(double x, double* s, double* c) { sincos(x, s, c); }
template <int XS>
__device__ __forceinline__ void thread_dequantize_gain_shape(const PvqTab& t, BitR& br, int R, int L, int cap,
                                                             double* wsA, double* wsB, double* __restrict__ out, int* err) {
    int bg, bs;
    gs_alloc(t, R, L, &bg, &bs);
    bool ones = true;
    double rootNr = 0.0;
    if (bs != 0) {
        ones = false;
        TDecNode st[24];
        int sp = 0, used = 0;
        st[sp++] = {0, (short)cap, (short)L, -1, 0, 0, 1, 0, bs, -1, 0.0, 0.0, 0.0};
        // Each round: (1) descend - read angles and push children until a leaf is on top, (2) decode that leaf,
        // (3) recombine every node whose children are now complete.  The lanes of a warp (same band, 32
        // blocks) so meet at the leaf decode, the long step, instead of each being at a different kind of node.
        bool fail = false;
        while (sp > 0 && !fail) {
            while (st[sp - 1].bits > PVQ_SPLIT_BITS && st[sp - 1].state == 0) {      // (1)
                const TDecNode nd = st[sp - 1];
                const int half = (nd.n + 1) / 2, hc = nd.cap / 2;
                int aT, aMs, aMid, aSide, entry;
                gs_alloc(t, nd.bits, half, &aT, &aMs);
                if (aT > 64) { *err |= PVQ_ERR_FIELD_TOO_WIDE; aT = 64; }
                const unsigned long long ti = br_get(br, aT);
                used += aT;
                double theta, l2t;
                theta_terms(t, ti, aT, &theta, &l2t, &entry);
                ms_alloc_l2t(aMs, theta, l2t, half, &aMid, &aSide);
                st[sp - 1].theta = theta;
                st[sp - 1].entry = entry;
                st[sp - 1].state = 1;
                st[sp - 1].z = (unsigned char)((aMid <= 0 ? 1 : 0) | (aSide <= 0 ? 2 : 0));   // children without bits read as zeros
                if (sp + 2 > 24) { *err |= PVQ_ERR_STACK; ones = true; fail = true; break; }
                const short par = (short)(sp - 1);
                if (aSide > 0) st[sp++] = {(short)(nd.off + hc), (short)hc, (short)half, par, (unsigned char)(1 - nd.buf), 0, 0, 0, aSide, -1, 0.0, 0.0, 0.0};
                if (aMid > 0) st[sp++] = {nd.off, (short)hc, (short)half, par, (unsigned char)(1 - nd.buf), 0, 0, 0, aMid, -1, 0.0, 0.0, 0.0};
            }
            if (fail) break;
            if (st[sp - 1].bits <= PVQ_SPLIT_BITS) {                                  // (2)
                const TDecNode nd = st[sp - 1];
                double* cur = (nd.buf ? wsB : wsA) + nd.off * XS;
                long long K;
                int w;
                pvq_k_for_bits(t, nd.n, nd.bits, &K, &w, err);
                const unsigned long long idx = br_get(br, w);
                used += w;
                double nrOut = sqrt(pvq_vector<XS>(t, idx, nd.n, K, cur, err));
                if (nd.root) {                            // a root leaf is normalised twice (:214-218)
                    double ss = 0.0;
                    const PacDiv byNr = pac_div_by(nrOut != 0.0 ? nrOut : 1.0);
                    for (int i = 0; i < nd.n; ++i) {
                        const double c = nrOut != 0.0 ? pac_div(cur[i * XS], byNr) : cur[i * XS];
                        cur[i * XS] = c;
                        ss += c * c;
                    }
                    nrOut = sqrt(ss);
                }
                --sp;                                      // hand the norm to the parent (mid child: same offset)
                if (nd.par >= 0) {
                    if (nd.off == st[nd.par].off) st[nd.par].nrMid = nrOut; else st[nd.par].nrSide = nrOut;
                } else {
                    rootNr = nrOut;
                }
            }
            while (sp > 0 && st[sp - 1].bits > PVQ_SPLIT_BITS && st[sp - 1].state == 1) {   // (3) coder/gain_shape_quantize.py:449-463
                const TDecNode nd = st[sp - 1];
                double* __restrict__ cur = (nd.buf ? wsB : wsA) + nd.off * XS;
                const double* __restrict__ oth = (nd.buf ? wsA : wsB) + nd.off * XS;
                const int n = nd.n, mid = n / 2, half = (n + 1) / 2, hc = nd.cap / 2;
                double ct, sn;
                if (nd.entry >= 0) {
                    ct = t.theta[4 * nd.entry + 2];
                    sn = t.theta[4 * nd.entry + 3];
                } else {
                    pm_sincos_pair(nd.theta, &sn, &ct);
                }
                const double nm = nd.nrMid, ns = nd.nrSide;
                const bool zm = nd.z & 1, zs = nd.z & 2;
                // the three divisors of this node (pac_div: the IEEE quotient from one shared reciprocal each)
                const PacDiv byNm = pac_div_by(nm != 0.0 ? nm : 1.0), byNs = pac_div_by(ns != 0.0 ? ns : 1.0);
                const PacDiv byR2 = {__longlong_as_double(0x3FF6A09E667F3BCDLL), __longlong_as_double(0x3FE6A09E667F3BCCLL)};   // sqrt(2.0), RN(1 / sqrt(2.0))
                double ss = 0.0;
                int i = 0;
                for (; i + 2 <= mid; i += 2) {                 // two elements per trip: up to four loads in flight
                    double m0 = 0.0, m1 = 0.0, s0 = 0.0, s1 = 0.0;
                    if (!zm) { m0 = oth[i * XS]; m1 = oth[(i + 1) * XS]; }
                    if (!zs) { s0 = oth[(hc + i) * XS]; s1 = oth[(hc + i + 1) * XS]; }
                    if (nm != 0.0) { m0 = pac_div(m0, byNm); m1 = pac_div(m1, byNm); }
                    if (ns != 0.0) { s0 = pac_div(s0, byNs); s1 = pac_div(s1, byNs); }
                    const double l0 = pac_div(m0 * ct + s0 * sn, byR2), r0 = pac_div(m0 * ct - s0 * sn, byR2);
                    const double l1 = pac_div(m1 * ct + s1 * sn, byR2), r1 = pac_div(m1 * ct - s1 * sn, byR2);
                    cur[i * XS] = l0; cur[(mid + i) * XS] = r0; cur[(i + 1) * XS] = l1; cur[(mid + i + 1) * XS] = r1;
                    ss += l0 * l0; ss += r0 * r0;
                    ss += l1 * l1; ss += r1 * r1;
                }
                for (; i < half; ++i) {
                    double m = zm ? 0.0 : oth[i * XS], s = zs ? 0.0 : oth[(hc + i) * XS];
                    if (nm != 0.0) m = pac_div(m, byNm);
                    if (ns != 0.0) s = pac_div(s, byNs);
                    const double l = pac_div(m * ct + s * sn, byR2);
                    const double r = pac_div(m * ct - s * sn, byR2);
                    if (i < mid) { cur[i * XS] = l; ss += l * l; }     // odd n: the padded last element of left is dropped
                    cur[(mid + i) * XS] = r;
                    ss += r * r;
                }
                const double nrOut = sqrt(ss);
                --sp;
                if (nd.par >= 0) {
                    if (nd.off == st[nd.par].off) st[nd.par].nrMid = nrOut; else st[nd.par].nrSide = nrOut;
                } else {
                    rootNr = nrOut;
                }
            }
        }
        bg += bs - used;
    }
    if (bg > 64) { *err |= PVQ_ERR_FIELD_TOO_WIDE; bg = 64; }
    if (bg < 0) bg = 0;
    const unsigned long long gi = br_get(br, bg);
    const double gq = dequantize_uniform(gi, bg);
    const double sg = gq > 0.0 ? 1.0 : (gq < 0.0 ? -1.0 : 0.0);
    const double gain = sg / 255.0 * (pm_pow(256.0, fabs(gq)) - 1.0) * (double)L;     // inv_mu_law_fn * L
    const double* __restrict__ shape = wsA;
    if (ones) {
        for (int i = 0; i < L; ++i) out[i * XS] = gain;
    } else {
        int i = 0;
        const PacDiv byRoot = pac_div_by(rootNr != 0.0 ? rootNr : 1.0);
        for (; i + 4 <= L; i += 4) {
            double v0 = shape[i * XS], v1 = shape[(i + 1) * XS], v2 = shape[(i + 2) * XS], v3 = shape[(i + 3) * XS];
            if (rootNr != 0.0) { v0 = pac_div(v0, byRoot); v1 = pac_div(v1, byRoot); v2 = pac_div(v2, byRoot); v3 = pac_div(v3, byRoot); }
            out[i * XS] = gain * v0; out[(i + 1) * XS] = gain * v1; out[(i + 2) * XS] = gain * v2; out[(i + 3) * XS] = gain * v3;
        }
        for (; i < L; ++i) {
            double v = shape[i * XS];
            if (rootNr != 0.0) v = pac_div(v, byRoot);
            out[i * XS] = gain * v;
        }
    }
}

// Phase B: every thread of the CTA takes leaves tid, tid + nThreads, ...
__device__ __forceinline__ void cta_decode_leaves(const PvqTab& t, DNode* nodes, int nNodes, double* wsA,
                                                  double* wsB, int tid, int nThreads, int* err) {
    const int nWarps = nThreads >> 5;                      // node j -> warp j % nWarps, lane j / nWarps
    for (int i = (tid & 31) * nWarps + (tid >> 5); i < nNodes; i += nThreads) {
        const DNode nd = nodes[i];
        if (!(nd.flags & DN_LEAF)) continue;
        int werr = 0;
        pvq_vector(t, nd.u, nd.n, nd.K, ((nd.flags & DN_BUF) ? wsB : wsA) + nd.off, &werr);
        if (werr) atomicOr(err, werr);
    }
}

// Phase C for one band on one group of G lanes: out[0..L) <- gain * shape.  base as in phase A.
template <int G>
__device__ __forceinline__ void grp_finish_gain_shape(const double* thetaTab, const DNode* nodes, const BandDec& bd,
                                                      double* wsA, double* wsB, int base, double* out) {
    const int lane = threadIdx.x & (G - 1);
    const unsigned gmask = grp_mask<G>();
    const int L = bd.L;
    if (bd.tail < 0) {
        for (int i = lane; i < L; i += G) out[i] = bd.gain;                 // shape = ones
        return;
    }
    for (int idx = bd.tail; idx >= 0;) {
        const DNode nd = nodes[idx];
        idx = nd.prev;
        double* cur = ((nd.flags & DN_BUF) ? wsB : wsA) + nd.off;
        double* oth = ((nd.flags & DN_BUF) ? wsA : wsB) + nd.off;
        const int n = nd.n;
        if (nd.flags & DN_LEAF) {
            const int reps = (nd.flags & DN_ROOT) ? 2 : 1;                   // a root leaf is normalised twice
            for (int r = 0; r < reps; ++r) {
                double ss = 0.0;
                for (int i = lane; i < n; i += G) ss += cur[i] * cur[i];
                const double nr = sqrt(grp_sum<G>(ss, gmask));
                if (nr != 0.0)
                    for (int i = lane; i < n; i += G) cur[i] = cur[i] / nr;
                __syncwarp(gmask);
            }
            continue;
        }
        // recombine (coder/gain_shape_quantize.py:449-463)
        const int mid = n / 2, half = (n + 1) / 2, hc = (1 << nd.capLog2) / 2;
        double ct, sn;
        if (nd.K >= 0 && thetaTab) {
            ct = thetaTab[4 * nd.K + 2];
            sn = thetaTab[4 * nd.K + 3];
        } else {
            pm_sincos(__longlong_as_double((long long)nd.u), &sn, &ct);
        }
        const double r2 = sqrt(2.0);
        double ss = 0.0;
        for (int i = lane; i < half; i += G) {
            const double m = (nd.flags & DN_ZMID) ? 0.0 : oth[i], s = (nd.flags & DN_ZSIDE) ? 0.0 : oth[hc + i];
            const double l = (m * ct + s * sn) / r2;
            const double r = (m * ct - s * sn) / r2;
            if (i < mid) { cur[i] = l; ss += l * l; }      // odd n: the padded last element of left is dropped
            cur[mid + i] = r;
            ss += r * r;
        }
        const double nr = sqrt(grp_sum<G>(ss, gmask));
        __syncwarp(gmask);
        if (nr != 0.0)
            for (int i = lane; i < n; i += G) cur[i] = cur[i] / nr;
        __syncwarp(gmask);
    }
    for (int i = lane; i < L; i += G) out[i] = bd.gain * wsA[base + i];
    __syncwarp(gmask);
}
