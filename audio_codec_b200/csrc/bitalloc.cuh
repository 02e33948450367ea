// Water-filling bit allocation, one warp per frame, lanes = bands (<= 32).
// Exact emulation of coder/bitalloc.py:62-121 (BitAlloc): the Nflip needle,
// np.round (half to even), the `bits < 2` flagging, the 201-pass cap and
// NumPy's pairwise summation order for np.sum(nLines*SMR) over <= 32 terms.
#pragma once
#include "pac_common.cuh"

// np.add.reduce over a contiguous double array of n < 128 elements
// (numpy/core/src/umath/loops_utils.h.src pairwise_sum): n < 8 sequential, else 8
// accumulators combined as ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)) then the tail.
__device__ __forceinline__ double np_sum_small(const double* a, int n) {
    if (n < 8) {
        double r = 0.0;                // NumPy starts from -0.0; identical for any n >= 1 and for the
        for (int i = 0; i < n; ++i) r += a[i];   // empty sum the only consumer divides it by N anyway
        return r;
    }
    double r[8];
    for (int j = 0; j < 8; ++j) r[j] = a[j];
    int i;
    for (i = 8; i < n - (n % 8); i += 8)
        for (int j = 0; j < 8; ++j) r[j] += a[i + j];
    double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
    for (; i < n; ++i) res += a[i];
    return res;
}

// Called by ONE full warp.  nLines: effective lines per band (1 for SBR-omitted
// bands, coder/bitalloc.py:142-143).  tmp: >= 32 doubles of shared scratch.
// Result in bitsOut[band] (int, shared or global).
__device__ __forceinline__ void warp_bit_alloc(double budget, int maxMant, int nBands, const int* nLines,
                                               const double* smr, double* tmp, int* bitsOut) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const bool act = lane < nBands;
    const int nl = act ? nLines[lane] : 0;
    const double s = act ? smr[lane] : 0.0;
    if (maxMant > 16) maxMant = 16;
    int bits = 0;
    bool flagged = false;
    int nflip = 0, count = 0;
    // Everything up to the needle level depends only on WHICH bands are still valid, and most passes
    // (the nflip search) leave that set unchanged: keep the pass state keyed on the valid mask.
    unsigned vmCached = 0u, pm = 0u;
    bool haveCache = false, posf = false;
    double ropt = 0.0, fr = 0.0;
    int nD = 0, rank = 0;
    while (true) {
        const bool valid = act && !flagged;
        const unsigned vm = __ballot_sync(FULL, valid);
        if (!haveCache || vm != vmCached) {
            haveCache = true;
            vmCached = vm;
            int Ni = valid ? nl : 0;
#pragma unroll
            for (int o = 16; o; o >>= 1) Ni += __shfl_xor_sync(FULL, Ni, o);
            double N = Ni == 0 ? 1e-12 : (double)Ni;
            // compacted products in band order -> NumPy-ordered sum on lane 0
            int pos = __popc(vm & ((1u << lane) - 1));
            if (valid) tmp[pos] = (double)nl * s;
            __syncwarp();
            double S = 0.0;
            if (lane == 0) S = np_sum_small(tmp, __popc(vm));
            S = __shfl_sync(FULL, S, 0);
            __syncwarp();
            ropt = budget / N + (1.0 / 6.2) * (s - S / N);
            fr = ropt - floor(ropt) - 0.5;
            posf = valid && fr > 0.0;
            pm = __ballot_sync(FULL, posf);
            nD = __popc(pm);
            // rank of this lane's fraction among the positive fractions (ascending, ties by band)
            rank = 0;
            for (unsigned m = pm; m;) {
                const int j = __ffs(m) - 1;
                m &= m - 1;
                const double fj = __shfl_sync(FULL, fr, j);
                if (fj < fr || (fj == fr && j < lane)) ++rank;
            }
        }
        if (nflip > nD) {
            nflip -= 1;                                   // bits unchanged this pass
        } else {
            double level = 0.0;
            if (nflip > 0) {
                // value of rank nflip-1 among the positive fractions
                unsigned hit = __ballot_sync(FULL, posf && rank == nflip - 1);
                level = __shfl_sync(FULL, fr, __ffs(hit) - 1);
            }
            if (valid) bits = (int)rint(ropt - level);
        }
        if (bits > maxMant) bits = maxMant;
        const bool nf = act && bits < 2;
        if (nf) bits = 0;
        // np.logical_xor(flagged_new, valid_old).all()
        const bool stable = __all_sync(FULL, !act || (nf != valid));
        flagged = nf;
        int tot = act ? bits * nl : 0;
#pragma unroll
        for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(FULL, tot, o);
        if (stable && (double)tot <= budget) break;
        if (stable && (double)tot > budget) nflip += 1;
        if (++count > 200) break;
    }
    if (act) bitsOut[lane] = bits;
    __syncwarp();
}
