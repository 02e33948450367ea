// Streaming window+MDCT / IMDCT+window+overlap-add for long frames (N = 2048): persistent
// warp pipelines.  Each warp walks a run of consecutive frames of one (stream, channel) row and
// keeps everything it needs on chip:
//   * a frame is two 1024-sample blocks; consecutive frames share one, so a warp fetches every
//     block from HBM exactly once (one halo block per run), with the NEXT block already in flight
//     in registers while the current frame is transformed (16-byte coalesced loads);
//   * the N/4 = 512-point complex FFT is 16 x 16 x 2: a lane owns 16 points of stride 32, so the
//     fold / pre-twiddle output feeds the first radix-16 butterfly straight from registers, ONE
//     shared-memory exchange (skewed, conflict-free both ways) feeds the second, and the last
//     radix-2 pairs lane l with lane l^16 through shuffles.  __syncwarp only - no CTA barrier in
//     the steady state, so warps drift apart and hide each other's latency;
//   * samples and window are kept de-interleaved (even / odd index arrays) in shared memory:
//     every fold, post-twiddle and read-out access is then unit-stride across the warp;
//   * results leave as 16-byte coalesced stores.
// Algorithmic HBM traffic: one read and one write of 1024 values per channel-frame.
// Shared-memory wavefronts per frame (fp32): ~450 (previous radix-8 x 3 tile version: 1025).
#pragma once
#include "mdct.cuh"

#define MS_SKEW(p) ((p) + ((p) >> 4))            // exchange tile of complex values: p = 16 l + r (write) and l + 32 r (read)
                                                 // are conflict-free per half warp (8-byte accesses: 16 lanes per wavefront)
#define MS_OSK(k) ((k) + (((k) >> 8) << 4))      // staging arrays: k = (l & 15) + 16 r + 256 (l >> 4) conflict-free
#define MS_TILE 544                              // values per half of the tile: 512 + skew (complex view: 544 entries)

template <typename T> struct MsCfg;
#ifndef MS_INV_WARPS_F64
#define MS_INV_WARPS_F64 8
#endif
// warps per CTA (one CTA per SM): bounded by shared memory (forward, fp64 inverse) and registers
template <> struct MsCfg<float> { static constexpr int WARPS = 16, WARPS_INV = 16; };
template <> struct MsCfg<double> { static constexpr int WARPS = 7, WARPS_INV = MS_INV_WARPS_F64; };

// 16-point DFT in registers, natural order in and out (4 x 4 Cooley-Tukey).
template <typename T> __device__ __forceinline__ void bfly16(cplx<T>* u) {
    const T h = (T)0.70710678118654752440, c1 = (T)0.92387953251128675613, s1 = (T)0.38268343236508977173;
    cplx<T> a[4][4];
#pragma unroll
    for (int r0 = 0; r0 < 4; ++r0) {
        cplx<T> t[4] = {u[r0], u[4 + r0], u[8 + r0], u[12 + r0]};
        bfly4(t);
#pragma unroll
        for (int k1 = 0; k1 < 4; ++k1) a[r0][k1] = t[k1];
    }
    const cplx<T> w1 = {c1, -s1}, w2 = {h, -h}, w3 = {s1, -c1}, w6 = {-h, -h}, w9 = {-c1, s1};
    a[1][1] = cmul(a[1][1], w1); a[1][2] = cmul(a[1][2], w2); a[1][3] = cmul(a[1][3], w3);
    a[2][1] = cmul(a[2][1], w2); a[2][2] = cmnj(a[2][2]);     a[2][3] = cmul(a[2][3], w6);
    a[3][1] = cmul(a[3][1], w3); a[3][2] = cmul(a[3][2], w6); a[3][3] = cmul(a[3][3], w9);
#pragma unroll
    for (int k1 = 0; k1 < 4; ++k1) {
        cplx<T> t[4] = {a[0][k1], a[1][k1], a[2][k1], a[3][k1]};
        bfly4(t);
#pragma unroll
        for (int k2 = 0; k2 < 4; ++k2) u[k1 + 4 * k2] = t[k2];
    }
}

// PER/2 consecutive values of an array (one 8-byte access for either type).
template <typename T> struct MsRun;
template <> struct MsRun<float> {
    static constexpr int L = 2;
    static __device__ __forceinline__ void ld(const float* p, float* o) {
        float2 v = *reinterpret_cast<const float2*>(p);
        o[0] = v.x; o[1] = v.y;
    }
    static __device__ __forceinline__ void st(float* p, const float* o) { *reinterpret_cast<float2*>(p) = make_float2(o[0], o[1]); }
    static __device__ __forceinline__ cplx<float> ldc(const cplx<float>* p) {
        float2 v = *reinterpret_cast<const float2*>(p);
        return {v.x, v.y};
    }
    static __device__ __forceinline__ void stc(cplx<float>* p, cplx<float> c) { *reinterpret_cast<float2*>(p) = make_float2(c.x, c.y); }
};
template <> struct MsRun<double> {
    static constexpr int L = 1;
    static __device__ __forceinline__ void ld(const double* p, double* o) { o[0] = *p; }
    static __device__ __forceinline__ void st(double* p, const double* o) { *p = o[0]; }
    static __device__ __forceinline__ cplx<double> ldc(const cplx<double>* p) {
        double2 v = *reinterpret_cast<const double2*>(p);
        return {v.x, v.y};
    }
    static __device__ __forceinline__ void stc(cplx<double>* p, cplx<double> c) { *reinterpret_cast<double2*>(p) = make_double2(c.x, c.y); }
};

template <typename T> struct MsBlock {        // one 1024-value block in registers, 16 bytes per access
    static constexpr int NV = 1024 * sizeof(T) / 16 / 32;
    static constexpr int PER = 16 / sizeof(T);
    uint4 v[NV];
    __device__ __forceinline__ void load(const T* __restrict__ src, int lane, bool vec) {
        if (vec) {
            const uint4* p = reinterpret_cast<const uint4*>(src);
#pragma unroll
            for (int i = 0; i < NV; ++i) v[i] = __ldg(p + i * 32 + lane);
        } else {
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                T t[PER];
#pragma unroll
                for (int e = 0; e < PER; ++e) t[e] = src[(i * 32 + lane) * PER + e];
                v[i] = *reinterpret_cast<uint4*>(t);
            }
        }
    }
    // de-interleave into E[512] (even indices) and O[512] = E + 512
    __device__ __forceinline__ void to_smem_eo(T* E, int lane) const {
        constexpr int L = MsRun<T>::L;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const T* t = reinterpret_cast<const T*>(&v[i]);
            T ev[L], ov[L];
#pragma unroll
            for (int q = 0; q < L; ++q) { ev[q] = t[2 * q]; ov[q] = t[2 * q + 1]; }
            MsRun<T>::st(E + (i * 32 + lane) * L, ev);
            MsRun<T>::st(E + 512 + (i * 32 + lane) * L, ov);
        }
    }
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int i = 0; i < NV; ++i) v[i] = make_uint4(0, 0, 0, 0);
    }
};

// CTA-shared tables.  win: [2048] either natural order (inverse) or de-interleaved per block
// (forward: [(block * 2 + odd) * 512 + i] = w[block * 1024 + 2 i + odd]); tw: MDCT twiddles [512];
// r1: second-pass roots [r * 16 + k] = exp(-2 pi j 2 r k / 512); roots: exp(-2 pi j k / 512).
template <typename T> struct MsTables {
    T* win;
    cplx<T>* tw;
    cplx<T>* r1;
    cplx<T>* roots;
};
// MODE folds a power-of-two factor of the transform into the kernel's private copy of the window (exact: the
// products, and everything linear after them, scale without a rounding): 0 none; 1 forward, w * 2/N (the MDCT's
// output scale); 2 inverse, the unfolding factors of imdct_sample: +2 w on [0, N/4), -2 w on [N/4, N).
template <typename T, int NT, bool EO, int MODE = 0>
__device__ __forceinline__ unsigned char* ms_load_tables(unsigned char* raw, const T* __restrict__ winG,
                                                         const cplx<T>* __restrict__ twG,
                                                         const cplx<T>* __restrict__ rootsG, MsTables<T>& t) {
    constexpr int N = 2048, Q = 512;
    t.win = reinterpret_cast<T*>(raw);
    t.tw = reinterpret_cast<cplx<T>*>(t.win + N);
    t.r1 = t.tw + Q;
    t.roots = t.r1 + 256;
    for (int i = threadIdx.x; i < N; i += NT) {
        T w = winG ? winG[i] : (T)1;
        if (MODE == 1) w *= (T)2 / (T)N;
        if (MODE == 2) w *= i < 512 ? (T)2 : (T)-2;
        if (EO) t.win[((i >> 10) * 2 + (i & 1)) * 512 + ((i & 1023) >> 1)] = w;
        else t.win[i] = w;
    }
    for (int i = threadIdx.x; i < Q; i += NT) t.tw[i] = twG[i];
    for (int i = threadIdx.x; i < 256; i += NT) {
        t.roots[i] = rootsG[i];
        t.r1[i] = rootsG[(2 * (i >> 4) * (i & 15)) & (Q - 1)];
    }
    __syncthreads();
    return reinterpret_cast<unsigned char*>(t.roots + 256);
}

template <typename T> __host__ __device__ constexpr size_t ms_smem_bytes() {
    // tables (2048 + 2 * (512 + 256 + 256) values) + per warp: two blocks (2048) + exchange tile (2 * MS_TILE)
    return sizeof(T) * ((size_t)2048 + 2 * 1024 + (size_t)MsCfg<T>::WARPS * (2048 + 2 * MS_TILE));
}

// 512-point FFT of z[lane + 32 i] = u[i]; on return u[r] = Z[(lane & 15) + 16 r + 256 (lane >> 4)].
// The tile is free again when the function returns.
template <typename T>
__device__ __forceinline__ void warp_fft512(cplx<T>* u, T* tileRe, const MsTables<T>& tb, int lane) {
    bfly16(u);
    cplx<T>* tile = reinterpret_cast<cplx<T>*>(tileRe);     // [MS_TILE] complex = both halves of the tile
#pragma unroll
    for (int r = 0; r < 16; ++r) MsRun<T>::stc(tile + MS_SKEW(16 * lane + r), u[r]);
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 16; ++r) u[r] = MsRun<T>::ldc(tile + MS_SKEW(lane + 32 * r));
    __syncwarp();
    const int k = lane & 15;
#pragma unroll
    for (int r = 1; r < 16; ++r) u[r] = cmul(u[r], tb.r1[r * 16 + k]);
    bfly16(u);
    const bool lo = lane < 16;
#pragma unroll
    for (int r = 0; r < 16; ++r) {
        cplx<T> o = {__shfl_xor_sync(0xffffffffu, u[r].x, 16), __shfl_xor_sync(0xffffffffu, u[r].y, 16)};
        const cplx<T> a = lo ? u[r] : o, b = lo ? o : u[r];
        const cplx<T> wb = cmul(b, tb.roots[k + 16 * r]);
        u[r] = lo ? cadd(a, wb) : csub(a, wb);
    }
}

// ---------------------------------------------------------------- forward
// unit = (row, run of `seg` frames).  out[row][frame][1024].
template <typename T, bool SYM>
__global__ void __launch_bounds__(MsCfg<T>::WARPS * 32, 1)
k_mdct_fwd_stream(const T* __restrict__ x, long long rowStride, int nFrames, int seg, int segsPerRow, long long nUnits,
                  const T* __restrict__ winG, const cplx<T>* __restrict__ twG, const cplx<T>* __restrict__ rootsG,
                  T* __restrict__ out, int vec) {
    constexpr int N = 2048, M = 1024, Q = 512, H = Q / 2, W = MsCfg<T>::WARPS;
    constexpr int NV = MsBlock<T>::NV, PER = MsBlock<T>::PER, L = MsRun<T>::L;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    MsTables<T> tb;
    unsigned char* p = ms_load_tables<T, W * 32, true>(smem_raw, winG, twG, rootsG, tb);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    T* blk = reinterpret_cast<T*>(p) + (size_t)warp * (2 * M + 2 * MS_TILE);   // [2][E 512 | O 512]
    T* tileRe = blk + 2 * M;
    T* tileIm = tileRe + MS_TILE;
    const T scale = (T)2 / (T)N;
    for (long long unit = (long long)blockIdx.x * W + warp; unit < nUnits; unit += (long long)gridDim.x * W) {
        const long long row = unit / segsPerRow;
        const int f0 = (int)(unit - row * segsPerRow) * seg, f1 = min(f0 + seg, nFrames);
        const T* src = x + row * rowStride + (long long)f0 * M;
        T* dst = out + (row * nFrames + f0) * M;
        MsBlock<T> pre;
        pre.load(src, lane, vec);
        pre.to_smem_eo(blk, lane);
        pre.load(src + M, lane, vec);
        for (int f = f0; f < f1; ++f) {
            const int cur = (f - f0) & 1;
            const T* E0 = blk + cur * M;         // samples [0, M) of the frame, even / odd indices
            const T* E1 = blk + (cur ^ 1) * M;   // samples [M, N)
            const T* W0 = tb.win;
            const T* W1 = tb.win + M;
            pre.to_smem_eo(blk + (cur ^ 1) * M, lane);
            __syncwarp();
            if (f + 1 < f1) pre.load(src + (long long)(f - f0 + 2) * M, lane, vec);
            // fold + window + pre-twiddle (mdct_fold): every access is unit stride across the warp
            cplx<T> u[16];
#pragma unroll
            for (int it = 0; it < 16; ++it) {
                const int n = lane + 32 * it;
                cplx<T> t;
                if (it < 8) {
                    const int a = H - 1 - n, b = H + n;
                    // symmetric window (sine / KBD): w[i1] = w[i2], w[i3] = w[i0] - two loads instead of four
                    const T w0 = W1[512 + a], w2 = W0[512 + a];
                    const T a0 = E1[512 + a] * w0, a1 = E1[b] * (SYM ? w2 : W1[b]);
                    const T a2 = E0[512 + a] * w2, a3 = E0[b] * (SYM ? w0 : W0[b]);
                    t.x = -a0 - a1; t.y = a2 - a3;
                } else {
                    const int a = n - H, b = 3 * H - 1 - n;
                    const T w0 = W0[a], w1 = W0[512 + b];
                    const T a0 = E0[a] * w0, a1 = E0[512 + b] * w1;
                    const T a2 = E1[a] * (SYM ? w1 : W1[a]), a3 = E1[512 + b] * (SYM ? w0 : W1[512 + b]);
                    t.x = a0 - a1; t.y = -a2 - a3;
                }
                u[it] = cmul(t, MsRun<T>::ldc(tb.tw + n));
            }
            warp_fft512<T>(u, tileRe, tb, lane);
            // post-twiddle into the staging arrays (even / odd output indices)
#pragma unroll
            for (int r = 0; r < 16; ++r) {
                const int k = (lane & 15) + 16 * r + 256 * (lane >> 4);
                const cplx<T> y = cmul(u[r], MsRun<T>::ldc(tb.tw + k));
                tileRe[MS_OSK(k)] = scale * y.x;               // out[2k]
                tileIm[MS_OSK(Q - 1 - k)] = -scale * y.y;      // out[M - 1 - 2k]
            }
            __syncwarp();
            T* d = dst + (long long)(f - f0) * M;
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                const int v = i * 32 + lane;
                T ev[L], ov[L], o[PER];
                MsRun<T>::ld(tileRe + MS_OSK(v * L), ev);
                MsRun<T>::ld(tileIm + MS_OSK(v * L), ov);
#pragma unroll
                for (int q = 0; q < L; ++q) { o[2 * q] = ev[q]; o[2 * q + 1] = ov[q]; }
                if (vec) *reinterpret_cast<uint4*>(d + v * PER) = *reinterpret_cast<uint4*>(o);
                else
                    for (int e = 0; e < PER; ++e) d[v * PER + e] = o[e];
            }
            __syncwarp();
        }
    }
}

// ---------------------------------------------------------------- inverse + window + overlap-add
// unit = (row, run of `seg` output blocks); a row has nFrames + 1 output blocks:
// block b = w[M:] * imdct(frame b-1)[M:] + w[:M] * imdct(frame b)[:M]  (coder/pacfile.py:289-295).
// The windowed second half of the previous frame stays in registers (each lane re-reads exactly
// the positions it wrote).
template <typename T> __host__ __device__ constexpr size_t ms_inv_smem_bytes() {
    return sizeof(T) * ((size_t)2048 + 2 * 1024 + (size_t)MsCfg<T>::WARPS_INV * (1024 + 2 * MS_TILE));
}

template <typename T>
__global__ void __launch_bounds__(MsCfg<T>::WARPS_INV * 32, 1)
k_imdct_ola_stream(const T* __restrict__ X, int nFrames, int seg, int segsPerRow, long long nUnits,
                   const T* __restrict__ winG, const cplx<T>* __restrict__ twG, const cplx<T>* __restrict__ rootsG,
                   T* __restrict__ y, long long rowStride, int vec) {
    constexpr int M = 1024, Q = 512, W = MsCfg<T>::WARPS_INV;
    constexpr int NV = MsBlock<T>::NV, PER = MsBlock<T>::PER, L = MsRun<T>::L;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    MsTables<T> tb;
    unsigned char* p = ms_load_tables<T, W * 32, false>(smem_raw, winG, twG, rootsG, tb);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    T* Ei = reinterpret_cast<T*>(p) + (size_t)warp * (M + 2 * MS_TILE);   // coefficients: even | odd indices
    T* Ev = Ei + M;                                                        // DCT-IV output, even indices (skewed)
    T* Ov = Ev + MS_TILE;                                                  // odd indices
    const int nOutRow = nFrames + 1;
    for (long long unit = (long long)blockIdx.x * W + warp; unit < nUnits; unit += (long long)gridDim.x * W) {
        const long long row = unit / segsPerRow;
        const int b0 = (int)(unit - row * segsPerRow) * seg, b1 = min(b0 + seg, nOutRow);
        const T* src = X + row * (long long)nFrames * M;
        T* dst = y + row * rowStride;
        MsBlock<T> pre;
        T tail[NV * PER];
#pragma unroll
        for (int i = 0; i < NV * PER; ++i) tail[i] = (T)0;
        // frames b0-1 (halo: only its tail is needed) .. b1-1
        int f = b0 - 1;
        if (f >= 0) pre.load(src + (long long)f * M, lane, vec); else pre.zero();
        for (; f < b1; ++f) {
            const bool have = f >= 0 && f < nFrames;
            pre.to_smem_eo(Ei, lane);
            __syncwarp();
            if (f + 1 < b1 && f + 1 < nFrames) pre.load(src + (long long)(f + 1) * M, lane, vec); else pre.zero();
            if (have) {
                cplx<T> u[16];
#pragma unroll
                for (int it = 0; it < 16; ++it) {
                    const int n = lane + 32 * it;
                    const cplx<T> t = {Ei[n], Ei[512 + Q - 1 - n]};      // {X[2n], X[M-1-2n]}
                    u[it] = cmul(t, MsRun<T>::ldc(tb.tw + n));
                }
                warp_fft512<T>(u, Ev, tb, lane);
#pragma unroll
                for (int r = 0; r < 16; ++r) {
                    const int k = (lane & 15) + 16 * r + 256 * (lane >> 4);
                    const cplx<T> v = cmul(u[r], MsRun<T>::ldc(tb.tw + k));
                    Ev[MS_OSK(k)] = v.x;                   // xv[2k]
                    Ov[MS_OSK(Q - 1 - k)] = -v.y;          // xv[M - 1 - 2k]
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
                    if (n0 < Q) {       // head = 2 xv[Q + n], tail = -2 xv[Q - 1 - n]
                        const int a = (Q + n0) / 2, c = (Q - n0 - PER) / 2;
                        MsRun<T>::ld(Ev + MS_OSK(a), ev);
                        MsRun<T>::ld(Ov + MS_OSK(a), ov);
#pragma unroll
                        for (int q = 0; q < L; ++q) { o[2 * q] = (T)2 * ev[q]; o[2 * q + 1] = (T)2 * ov[q]; }
                        MsRun<T>::ld(Ev + MS_OSK(c), ev);
                        MsRun<T>::ld(Ov + MS_OSK(c), ov);
#pragma unroll
                        for (int q = 0; q < L; ++q) { tl[PER - 1 - 2 * q] = (T)-2 * ev[q]; tl[PER - 2 - 2 * q] = (T)-2 * ov[q]; }
                    } else {            // head = -2 xv[3Q - 1 - n], tail = -2 xv[n - Q]
                        const int c = (3 * Q - n0 - PER) / 2, a = (n0 - Q) / 2;
                        MsRun<T>::ld(Ev + MS_OSK(c), ev);
                        MsRun<T>::ld(Ov + MS_OSK(c), ov);
#pragma unroll
                        for (int q = 0; q < L; ++q) { o[PER - 1 - 2 * q] = (T)-2 * ev[q]; o[PER - 2 - 2 * q] = (T)-2 * ov[q]; }
                        MsRun<T>::ld(Ev + MS_OSK(a), ev);
                        MsRun<T>::ld(Ov + MS_OSK(a), ov);
#pragma unroll
                        for (int q = 0; q < L; ++q) { tl[2 * q] = (T)-2 * ev[q]; tl[2 * q + 1] = (T)-2 * ov[q]; }
                    }
                    const uint4 w0 = *reinterpret_cast<const uint4*>(tb.win + n0);
                    const uint4 w1 = *reinterpret_cast<const uint4*>(tb.win + M + n0);
                    const T* wh = reinterpret_cast<const T*>(&w0);
                    const T* wt = reinterpret_cast<const T*>(&w1);
#pragma unroll
                    for (int e = 0; e < PER; ++e) { o[e] *= wh[e]; tl[e] *= wt[e]; }
                }
#pragma unroll
                for (int e = 0; e < PER; ++e) { o[e] += tail[i * PER + e]; tail[i * PER + e] = tl[e]; }
                if (f >= b0) {
                    T* d = dst + (long long)f * M + n0;
                    if (vec) *reinterpret_cast<uint4*>(d) = *reinterpret_cast<uint4*>(o);
                    else
                        for (int e = 0; e < PER; ++e) d[e] = o[e];
                }
            }
            __syncwarp();
        }
    }
}
