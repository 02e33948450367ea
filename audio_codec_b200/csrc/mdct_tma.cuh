// Streaming window+MDCT / IMDCT+window+overlap-add for long frames (N = 2048) with Blackwell bulk copies.
//
// Same pipeline as mdct_stream.cuh (persistent warps, a warp walks a run of consecutive frames of one row, the
// 512-point FFT is 16 x 16 x 2 in registers with one shared-memory exchange), but the bytes no longer pass through
// the LSU on their way in and out:
//   * every 1024-sample block arrives by ONE cp.async.bulk (the TMA engine's 1-D copy, SASS UBLKCP) into a
//     per-warp ring of three buffers in NATURAL order, completion signalled on an mbarrier (no LDG -> register ->
//     STS de-interleave); two blocks are in flight ahead of the frame being transformed;
//   * the fold reads (even, odd) PAIRS with one 8-byte LDS each: a lane transforms the point pair {m, 511 - m},
//     which together use both halves of every pair they touch (block element 2j goes to point 256 + j, element
//     2j + 1 to point 255 - j; 512 + 2m to m, 513 + 2m to 511 - m), and hands the second point to its owner (the
//     mirrored lane 31 - l) by a shuffle;
//   * the ring slot of the frame's first block is dead once the fold has read it: it becomes the FFT exchange tile and
//     then the natural-order output tile, which leaves by ONE cp.async.bulk shared -> global (no LDS -> register ->
//     STG); the slot is refilled (block j + 3's predecessor) only after that store has read it.  12.75 KB (fp32) per
//     warp: 16 warps per SM as in the LDG / STG pipeline.
// __syncwarp and mbarriers only; nothing is CTA-wide after the tables are loaded.
#pragma once
#include "mdct_stream.cuh"

__device__ __forceinline__ unsigned mt_smem(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mt_bar_init(void* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mt_smem(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mt_bar_expect(void* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mt_smem(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mt_bar_wait(void* bar, unsigned parity) {
    unsigned ok = 0;
    const unsigned a = mt_smem(bar);
    while (!ok)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
}
// global -> shared, bytes a multiple of 16, both addresses 16-byte aligned
__device__ __forceinline__ void mt_bulk_load(void* dst, const void* src, unsigned bytes, void* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(mt_smem(dst)), "l"(src), "r"(bytes), "r"(mt_smem(bar)) : "memory");
}
__device__ __forceinline__ void mt_bulk_store(void* dst, const void* src, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(mt_smem(src)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void mt_bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void mt_fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <typename T> struct MtCfg;
template <> struct MtCfg<float> { static constexpr int WARPS = 16; };
template <> struct MtCfg<double> { static constexpr int WARPS = 7; };
#define MT_RING 3
#define MT_SLOT (2 * MS_TILE)      // values per ring slot: a block (1024) or the skewed FFT exchange tile (1088)

template <typename T> __host__ __device__ constexpr size_t mt_smem_bytes() {
    // tables (2048 window + 2 * (512 + 256 + 256) complex parts) + per warp: the ring (a slot also serves as the FFT
    // exchange / output tile once its block has been folded); then the mbarriers
    return sizeof(T) * ((size_t)2048 + 2 * 1024 + (size_t)MtCfg<T>::WARPS * MT_RING * MT_SLOT) + (size_t)MtCfg<T>::WARPS * MT_RING * 8;
}

template <typename T> struct MtPair;           // two consecutive values: one 8-byte (fp32) / 16-byte (fp64) access
template <> struct MtPair<float> {
    static __device__ __forceinline__ void ld(const float* p, float& a, float& b) {
        const float2 v = *reinterpret_cast<const float2*>(p);
        a = v.x; b = v.y;
    }
};
template <> struct MtPair<double> {
    static __device__ __forceinline__ void ld(const double* p, double& a, double& b) {
        const double2 v = *reinterpret_cast<const double2*>(p);
        a = v.x; b = v.y;
    }
};
// (a0 b0, a1 b1): one packed multiply in fp32
template <typename T> __device__ __forceinline__ void mt_mul2(T a0, T a1, T b0, T b1, T& r0, T& r1) { r0 = a0 * b0; r1 = a1 * b1; }
#if PAC_F32X2
template <> __device__ __forceinline__ void mt_mul2<float>(float a0, float a1, float b0, float b1, float& r0, float& r1) {
    const float2 r = __fmul2_rn(make_float2(a0, a1), make_float2(b0, b1));
    r0 = r.x; r1 = r.y;
}
#endif
template <typename T> __device__ __forceinline__ void mt_add2(T a0, T a1, T b0, T b1, T& r0, T& r1) { r0 = a0 + b0; r1 = a1 + b1; }
#if PAC_F32X2
template <> __device__ __forceinline__ void mt_add2<float>(float a0, float a1, float b0, float b1, float& r0, float& r1) {
    const float2 r = __fadd2_rn(make_float2(a0, a1), make_float2(b0, b1));
    r0 = r.x; r1 = r.y;
}
#endif
template <typename T> __device__ __forceinline__ cplx<T> mt_shfl_mirror(cplx<T> v, int lane) {
    cplx<T> r = {__shfl_sync(0xffffffffu, v.x, 31 - lane), __shfl_sync(0xffffffffu, v.y, 31 - lane)};
    return r;
}

// ---------------------------------------------------------------- forward
// unit = (row, run of `seg` frames).  out[row][frame][1024].  Needs 16-byte aligned rows (the launcher checks).
// Ring discipline (per warp): block j of the run lives in slot (base + j) % 3.  Iteration j (frame j = blocks j, j + 1):
// wait for block j + 1; fold; then slot(j) is dead -> wait until the store of frame j - 1 has read slot(j - 1) =
// slot(j + 2) and refill it with block j + 2; FFT in slot(j); output tile in slot(j); bulk store.
template <typename T, bool SYM>
__global__ void __launch_bounds__(MtCfg<T>::WARPS * 32, 1)
k_mdct_fwd_tma(const T* __restrict__ x, long long rowStride, int nFrames, int seg, int segsPerRow, long long nUnits,
               const T* __restrict__ winG, const cplx<T>* __restrict__ twG, const cplx<T>* __restrict__ rootsG,
               T* __restrict__ out) {
    constexpr int N = 2048, M = 1024, Q = 512, W = MtCfg<T>::WARPS;
    constexpr unsigned BLK_BYTES = M * sizeof(T);
    extern __shared__ __align__(128) unsigned char smem_raw[];
    MsTables<T> tb;
    unsigned char* p = ms_load_tables<T, W * 32, false, 1>(smem_raw, winG, twG, rootsG, tb);     // window * 2 / N
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    T* ring = reinterpret_cast<T*>(p) + (size_t)warp * (MT_RING * MT_SLOT);
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(reinterpret_cast<T*>(p) + (size_t)W * (MT_RING * MT_SLOT)) +
                               warp * MT_RING;
    if (lane == 0) {
#pragma unroll
        for (int b = 0; b < MT_RING; ++b) mt_bar_init(&bars[b], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const T* W0 = tb.win;            // window over samples [0, M) of the frame, natural order
    const T* W1 = tb.win + M;        // [M, N)
    unsigned used = 0;               // blocks this warp has sent through the ring so far (slot and phase follow)
    for (long long unit = (long long)blockIdx.x * W + warp; unit < nUnits; unit += (long long)gridDim.x * W) {
        const long long row = unit / segsPerRow;
        const int f0 = (int)(unit - row * segsPerRow) * seg, f1 = min(f0 + seg, nFrames);
        const int nBlk = f1 - f0 + 1;
        const T* src = x + row * rowStride + (long long)f0 * M;
        T* dst = out + (row * nFrames + f0) * M;
        const unsigned base = used;                      // block j of the run is ring use number base + j
        if (lane == 0) {
            mt_bulk_wait_read();                         // the last frame of the previous run may still be leaving its slot
            for (int j = 0; j < 2; ++j) {
                const unsigned u = base + j;
                mt_bar_expect(&bars[u % MT_RING], BLK_BYTES);
                mt_bulk_load(ring + (u % MT_RING) * MT_SLOT, src + (long long)j * M, BLK_BYTES, &bars[u % MT_RING]);
            }
        }
        mt_bar_wait(&bars[base % MT_RING], (base / MT_RING) & 1);
        for (int j = 0; j + 1 < nBlk; ++j) {             // frame f0 + j = blocks j, j + 1
            const unsigned u0 = base + j, u1 = u0 + 1, u2 = u0 + 2;
            mt_bar_wait(&bars[u1 % MT_RING], (u1 / MT_RING) & 1);
            T* B0 = ring + (u0 % MT_RING) * MT_SLOT;
            const T* B1 = ring + (u1 % MT_RING) * MT_SLOT;
            // fold + window + pre-twiddle (mdct_fold) for the point pairs {m, 511 - m}, m = lane + 32 i
            cplx<T> u[16];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int m = lane + 32 * i;
                T b0ax, b0ay, b0bx, b0by, b1ax, b1ay, b1bx, b1by, w0ax, w0ay, w0bx, w0by, w1ax, w1ay, w1bx, w1by;
                MtPair<T>::ld(B0 + 510 - 2 * m, b0ax, b0ay);      // x[510 - 2m], x[511 - 2m]
                MtPair<T>::ld(B0 + 512 + 2 * m, b0bx, b0by);      // x[512 + 2m], x[513 + 2m]
                MtPair<T>::ld(B1 + 510 - 2 * m, b1ax, b1ay);      // x[1534 - 2m], x[1535 - 2m]
                MtPair<T>::ld(B1 + 512 + 2 * m, b1bx, b1by);      // x[1536 + 2m], x[1537 + 2m]
                MtPair<T>::ld(W0 + 510 - 2 * m, w0ax, w0ay);
                MtPair<T>::ld(W0 + 512 + 2 * m, w0bx, w0by);
                if (SYM) {               // w[2047 - i] = w[i]: sine / KBD / no window
                    w1ax = w0by; w1ay = w0bx; w1bx = w0ay; w1by = w0ax;
                } else {
                    MtPair<T>::ld(W1 + 510 - 2 * m, w1ax, w1ay);
                    MtPair<T>::ld(W1 + 512 + 2 * m, w1bx, w1by);
                }
                // point m (< 256): v[2n] = -x[i0] - x[i1], v[M-1-2n] = x[i2] - x[i3]
                // the eight products are the four pairs times their window pairs
                T a0, a1, a2, a3, c0, c1, c2, c3;
                mt_mul2<T>(b0ax, b0ay, w0ax, w0ay, c0, a2);
                mt_mul2<T>(b0bx, b0by, w0bx, w0by, a3, c1);
                mt_mul2<T>(b1ax, b1ay, w1ax, w1ay, c2, a0);
                mt_mul2<T>(b1bx, b1by, w1bx, w1by, a1, c3);
                cplx<T> t = {-a0 - a1, a2 - a3};
                u[i] = cmul(t, MsRun<T>::ldc(tb.tw + m));
                // point 511 - m (>= 256): v[2n] = x[i0] - x[i1], v[M-1-2n] = -x[i2] - x[i3]
                cplx<T> s = {c0 - c1, -c2 - c3};
                s = cmul(s, MsRun<T>::ldc(tb.tw + (Q - 1 - m)));
                u[15 - i] = mt_shfl_mirror(s, lane);      // point 511 - m' of lane 31 - l = l + 32 (15 - i)
            }
            __syncwarp();                                 // every lane is done with B0 (and with this frame's view of B1)
            if (lane == 0 && j + 2 < nBlk) {              // refill slot(j + 2) = slot(j - 1): the previous frame's output tile
                mt_bulk_wait_read();
                mt_bar_expect(&bars[u2 % MT_RING], BLK_BYTES);
                mt_bulk_load(ring + (u2 % MT_RING) * MT_SLOT, src + (long long)(j + 2) * M, BLK_BYTES, &bars[u2 % MT_RING]);
            }
            T* tile = B0;                                 // exchange tile, then the natural-order output tile
#ifndef MT_EXPERIMENT_NOFFT                              // (experiment: the pipeline without the transform = its copy ceiling)
            warp_fft512<T>(u, tile, tb, lane);
#endif
#pragma unroll
            for (int r = 0; r < 16; ++r) {
                const int k = (lane & 15) + 16 * r + 256 * (lane >> 4);
                const cplx<T> y = cmul(u[r], MsRun<T>::ldc(tb.tw + k));
                tile[2 * k] = y.x;                        // the 2 / N came in with the window
                tile[M - 1 - 2 * k] = -y.y;
            }
            mt_fence_async();            // generic-proxy writes visible to the bulk-copy engine
            __syncwarp();
            if (lane == 0) mt_bulk_store(dst + (long long)j * M, tile, BLK_BYTES);
        }
        used = base + nBlk;
    }
    if (lane == 0) mt_bulk_wait_read();
}

// ---------------------------------------------------------------- inverse + window + overlap-add
// unit = (row, run of `seg` output blocks), as k_imdct_ola_stream.  Coefficient frames arrive by bulk copy into the
// ring (two frames ahead); {X[2n], X[M-1-2n]} for the point pair {m, 511 - m} are the two pairs (X[2m], X[2m+1]) and
// (X[1022-2m], X[1023-2m]); once read, the frame's slot is the FFT exchange tile and the DCT-IV staging.  The windowed
// overlap-add output goes out as 16-byte coalesced stores straight from registers (the tail stays in registers).
template <typename T>
__global__ void __launch_bounds__(MtCfg<T>::WARPS * 32, 1)
k_imdct_ola_tma(const T* __restrict__ X, int nFrames, int seg, int segsPerRow, long long nUnits,
                const T* __restrict__ winG, const cplx<T>* __restrict__ twG, const cplx<T>* __restrict__ rootsG,
                T* __restrict__ y, long long rowStride) {
    constexpr int M = 1024, Q = 512, W = MtCfg<T>::WARPS;
    constexpr int NV = MsBlock<T>::NV, PER = MsBlock<T>::PER, L = MsRun<T>::L;
    constexpr unsigned BLK_BYTES = M * sizeof(T);
    extern __shared__ __align__(128) unsigned char smem_raw[];
    MsTables<T> tb;
    unsigned char* p = ms_load_tables<T, W * 32, false, 2>(smem_raw, winG, twG, rootsG, tb);     // window * (+-2)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    T* ring = reinterpret_cast<T*>(p) + (size_t)warp * (MT_RING * MT_SLOT);
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(reinterpret_cast<T*>(p) + (size_t)W * (MT_RING * MT_SLOT)) +
                               warp * MT_RING;
    if (lane == 0) {
#pragma unroll
        for (int b = 0; b < MT_RING; ++b) mt_bar_init(&bars[b], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const int nOutRow = nFrames + 1;
    const int kl = (lane & 15) + 256 * (lane >> 4), kE = kl + 16 * (lane >> 4), kO = (Q - 1 - kl) + 16 * (1 - (lane >> 4));
    const int ll = lane * (PER / 2);
    unsigned used = 0;
    for (long long unit = (long long)blockIdx.x * W + warp; unit < nUnits; unit += (long long)gridDim.x * W) {
        const long long row = unit / segsPerRow;
        const int b0 = (int)(unit - row * segsPerRow) * seg, b1 = min(b0 + seg, nOutRow);
        const T* src = X + row * (long long)nFrames * M;
        T* dst = y + row * rowStride;
        T tail[NV * PER];
#pragma unroll
        for (int i = 0; i < NV * PER; ++i) tail[i] = (T)0;
        // frames b0 - 1 (halo: only its tail is needed) .. b1 - 1; those inside [0, nFrames) are loaded
        const int fl = max(b0 - 1, 0), fh = min(b1 - 1, nFrames - 1);
        const unsigned base = used;
        if (lane == 0) {
            for (int f = fl; f <= fh && f < fl + 2; ++f) {
                const unsigned u = base + (f - fl);
                mt_bar_expect(&bars[u % MT_RING], BLK_BYTES);
                mt_bulk_load(ring + (u % MT_RING) * MT_SLOT, src + (long long)f * M, BLK_BYTES, &bars[u % MT_RING]);
            }
        }
        T* Ev = ring;
        T* Ov = ring + MS_TILE;
        for (int f = b0 - 1; f < b1; ++f) {
            const bool have = f >= 0 && f < nFrames;
            if (have) {
                const unsigned u0 = base + (f - fl), u2 = u0 + 2;
                if (lane == 0 && f + 2 <= fh) {           // slot(f + 2) = slot(f - 1): read for the last time before the last __syncwarp
                    mt_bar_expect(&bars[u2 % MT_RING], BLK_BYTES);
                    mt_bulk_load(ring + (u2 % MT_RING) * MT_SLOT, src + (long long)(f + 2) * M, BLK_BYTES, &bars[u2 % MT_RING]);
                }
                mt_bar_wait(&bars[u0 % MT_RING], (u0 / MT_RING) & 1);
                T* S = ring + (u0 % MT_RING) * MT_SLOT;
                cplx<T> u[16];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int m = lane + 32 * i;
                    T ax, ay, bx, by;
                    MtPair<T>::ld(S + 2 * m, ax, ay);                 // X[2m], X[2m + 1]
                    MtPair<T>::ld(S + 1022 - 2 * m, bx, by);          // X[1022 - 2m], X[1023 - 2m]
                    const cplx<T> t = {ax, by};                       // point m: {X[2n], X[M-1-2n]}
                    u[i] = cmul(t, MsRun<T>::ldc(tb.tw + m));
                    cplx<T> s = {bx, ay};                             // point 511 - m
                    s = cmul(s, MsRun<T>::ldc(tb.tw + (Q - 1 - m)));
                    u[15 - i] = mt_shfl_mirror(s, lane);
                }
                __syncwarp();                                         // the coefficients are consumed: the slot becomes the tile
                Ev = S;
                Ov = S + MS_TILE;
                warp_fft512<T>(u, Ev, tb, lane);
#pragma unroll
                for (int r = 0; r < 16; ++r) {
                    const int k = (lane & 15) + 16 * r + 256 * (lane >> 4);
                    const cplx<T> v = cmul(u[r], MsRun<T>::ldc(tb.tw + k));
                    Ev[kE + 16 * r] = v.x;                 // xv[2k] at MS_OSK(k): k >> 8 = lane >> 4
                    Ov[kO - 16 * r] = -v.y;                // xv[M - 1 - 2k] at MS_OSK(Q - 1 - k): (Q - 1 - k) >> 8 = 1 - (lane >> 4)
                }
            }
            __syncwarp();
            // emit output block f (tail of f-1 + head of f), then keep this frame's tail
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                const int n0 = (i * 32 + lane) * PER;
                T o[PER], tl[PER];
#pragma unroll
                for (int e = 0; e < PER; ++e) { o[e] = (T)0; tl[e] = (T)0; }
                if (have) {
                    T ev[L], ov[L];
                    if (i * 32 * PER < Q) {       // (n0 < Q, the same for every lane) head = 2 xv[Q + n], tail = -2 xv[Q - 1 - n] (the factors are in the window copy)
                        const int A = Q / 2 + i * 16 * PER, C = (Q - PER) / 2 - i * 16 * PER;      // a = (Q + n0) / 2 = A + ll, c = (Q - n0 - PER) / 2 = C - ll
                        const int a = A + 16 * (A >> 8) + ll, c = C + 16 * (C >> 8) - ll;          // MS_OSK: the skew is the same for all lanes
                        MsRun<T>::ld(Ev + a, ev);
                        MsRun<T>::ld(Ov + a, ov);
#pragma unroll
                        for (int q = 0; q < L; ++q) { o[2 * q] = ev[q]; o[2 * q + 1] = ov[q]; }
                        MsRun<T>::ld(Ev + c, ev);
                        MsRun<T>::ld(Ov + c, ov);
#pragma unroll
                        for (int q = 0; q < L; ++q) { tl[PER - 1 - 2 * q] = ev[q]; tl[PER - 2 - 2 * q] = ov[q]; }
                    } else {            // head = -2 xv[3Q - 1 - n], tail = -2 xv[n - Q]
                        const int C = (3 * Q - PER) / 2 - i * 16 * PER, A = i * 16 * PER - Q / 2;
                        const int c = C + 16 * (C >> 8) - ll, a = A + 16 * (A >> 8) + ll;
                        MsRun<T>::ld(Ev + c, ev);
                        MsRun<T>::ld(Ov + c, ov);
#pragma unroll
                        for (int q = 0; q < L; ++q) { o[PER - 1 - 2 * q] = ev[q]; o[PER - 2 - 2 * q] = ov[q]; }
                        MsRun<T>::ld(Ev + a, ev);
                        MsRun<T>::ld(Ov + a, ov);
#pragma unroll
                        for (int q = 0; q < L; ++q) { tl[2 * q] = ev[q]; tl[2 * q + 1] = ov[q]; }
                    }
                    const uint4 w0 = *reinterpret_cast<const uint4*>(tb.win + n0);
                    const uint4 w1 = *reinterpret_cast<const uint4*>(tb.win + M + n0);
                    const T* wh = reinterpret_cast<const T*>(&w0);
                    const T* wt = reinterpret_cast<const T*>(&w1);
#pragma unroll
                    for (int e = 0; e < PER; e += 2) {
                        mt_mul2<T>(o[e], o[e + 1], wh[e], wh[e + 1], o[e], o[e + 1]);
                        mt_mul2<T>(tl[e], tl[e + 1], wt[e], wt[e + 1], tl[e], tl[e + 1]);
                    }
                }
#pragma unroll
                for (int e = 0; e < PER; e += 2) {
                    mt_add2<T>(o[e], o[e + 1], tail[i * PER + e], tail[i * PER + e + 1], o[e], o[e + 1]);
                    tail[i * PER + e] = tl[e]; tail[i * PER + e + 1] = tl[e + 1];
                }
                if (f >= b0) *reinterpret_cast<uint4*>(dst + (long long)f * M + n0) = *reinterpret_cast<uint4*>(o);
            }
            __syncwarp();
        }
        used = base + (unsigned)(fh - fl + 1);
    }
}
