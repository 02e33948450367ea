// Streaming window+MDCT / IMDCT+window+overlap-add for long frames (N = 2048): persistent
// warp pipelines.  Each warp walks a run of consecutive frames of one (stream, channel) row and
// keeps everything it needs on chip:
//   * a frame is two 1024-sample blocks; consecutive frames share one, so a warp fetches every
//     PCM block from HBM exactly once (one halo block per run), with the NEXT block already in
//     flight in registers while the current frame is transformed (16-byte coalesced loads);
//   * the N/4 = 512-point complex FFT is three radix-8 Stockham passes inside the warp's own
//     shared-memory tile, synchronised with __syncwarp only - there is no CTA-wide barrier in
//     the steady state, so warps drift apart and hide each other's latency;
//   * window, twiddles and FFT roots sit in shared memory once per CTA;
//   * results leave as 16-byte coalesced stores.
// Algorithmic HBM traffic: one read and one write of 1024 values per channel-frame.
#pragma once
#include "mdct.cuh"

#define MS_PAD(i) ((i) + ((i) >> 3))        // skew against bank conflicts of the strided Stockham writes

template <typename T> struct MsCfg;
template <> struct MsCfg<float> { static constexpr int WARPS = 16; };
template <> struct MsCfg<double> { static constexpr int WARPS = 7; };   // 32 KB tables + 7 x 25.6 KB <= 227 KB

// Q-point FFT of the warp's padded tile (Q = 512: three radix-8 passes).
template <typename T, int Q>
__device__ __forceinline__ void warp_fft512(cplx<T>* fb, const cplx<T>* __restrict__ roots, int lane) {
    static_assert(Q == 512, "three radix-8 passes");
    constexpr int J = Q / 8;                 // butterflies per pass
#pragma unroll
    for (int pass = 0; pass < 3; ++pass) {
        const int Ns = pass == 0 ? 1 : (pass == 1 ? 8 : 64);
        cplx<T> u[2][8];
#pragma unroll
        for (int it = 0; it < 2; ++it) {
            const int j = lane + it * 32;
            const int k = j & (Ns - 1);
            const int step = k * (Q / (Ns * 8));
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                cplx<T> v = fb[MS_PAD(r * J + j)];
                if (r > 0 && pass > 0) v = cmul(v, roots[r * step]);
                u[it][r] = v;
            }
            bfly8(u[it]);
        }
        __syncwarp();
#pragma unroll
        for (int it = 0; it < 2; ++it) {
            const int j = lane + it * 32;
            const int k = j & (Ns - 1);
            const int base = (j - k) * 8 + k;
#pragma unroll
            for (int r = 0; r < 8; ++r) fb[MS_PAD(base + r * Ns)] = u[it][r];
        }
        __syncwarp();
    }
    (void)J;
}

template <typename T> struct MsBlock {        // one 1024-value block in registers, 16 bytes per access
    static constexpr int NV = 1024 * sizeof(T) / 16 / 32;
    uint4 v[NV];
    __device__ __forceinline__ void load(const T* __restrict__ src, int lane, bool vec) {
        if (vec) {
            const uint4* p = reinterpret_cast<const uint4*>(src);
#pragma unroll
            for (int i = 0; i < NV; ++i) v[i] = __ldg(p + i * 32 + lane);
        } else {
            constexpr int PER = 16 / sizeof(T);
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                T t[PER];
#pragma unroll
                for (int e = 0; e < PER; ++e) t[e] = src[(i * 32 + lane) * PER + e];
                v[i] = *reinterpret_cast<uint4*>(t);
            }
        }
    }
    __device__ __forceinline__ void to_smem(T* dst, int lane) const {
        uint4* p = reinterpret_cast<uint4*>(dst);
#pragma unroll
        for (int i = 0; i < NV; ++i) p[i * 32 + lane] = v[i];
    }
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int i = 0; i < NV; ++i) v[i] = make_uint4(0, 0, 0, 0);
    }
};

template <typename T>
__device__ __forceinline__ void ms_store_block(T* __restrict__ dst, const T* src, int lane, bool vec) {
    constexpr int NV = 1024 * sizeof(T) / 16 / 32;
    if (vec) {
        const uint4* s = reinterpret_cast<const uint4*>(src);
        uint4* d = reinterpret_cast<uint4*>(dst);
#pragma unroll
        for (int i = 0; i < NV; ++i) d[i * 32 + lane] = s[i * 32 + lane];
    } else {
        for (int i = lane; i < 1024; i += 32) dst[i] = src[i];
    }
}

// CTA-shared tables: window [N] (if any), MDCT twiddles [Q], FFT roots [Q].
template <typename T> struct MsTables {
    T* win;
    cplx<T>* tw;
    cplx<T>* roots;
};
template <typename T, int N, int NT>
__device__ __forceinline__ unsigned char* ms_load_tables(unsigned char* raw, const T* __restrict__ winG,
                                                         const cplx<T>* __restrict__ twG,
                                                         const cplx<T>* __restrict__ rootsG, MsTables<T>& t) {
    constexpr int Q = N / 4;
    t.win = reinterpret_cast<T*>(raw);
    t.tw = reinterpret_cast<cplx<T>*>(t.win + N);
    t.roots = t.tw + Q;
    for (int i = threadIdx.x; i < N; i += NT) t.win[i] = winG ? winG[i] : (T)1;
    for (int i = threadIdx.x; i < Q; i += NT) { t.tw[i] = twG[i]; t.roots[i] = rootsG[i]; }
    __syncthreads();
    return reinterpret_cast<unsigned char*>(t.roots + Q);
}

template <typename T> __host__ __device__ constexpr size_t ms_smem_bytes() {
    // tables (N + 2*2*Q values) + per warp: two blocks (2*M) + padded FFT tile (2*(Q + Q/8))
    return sizeof(T) * ((size_t)2048 + 4 * 512 + (size_t)MsCfg<T>::WARPS * (2 * 1024 + 2 * (512 + 64)));
}

// ---------------------------------------------------------------- forward
// unit = (row, run of `seg` frames).  out[row][frame][1024].
template <typename T>
__global__ void __launch_bounds__(MsCfg<T>::WARPS * 32, 1)
k_mdct_fwd_stream(const T* __restrict__ x, long long rowStride, int nFrames, int seg, int segsPerRow, long long nUnits,
                  const T* __restrict__ winG, const cplx<T>* __restrict__ twG, const cplx<T>* __restrict__ rootsG,
                  T* __restrict__ out, int vec) {
    constexpr int N = 2048, M = 1024, Q = 512, QP = Q + Q / 8, W = MsCfg<T>::WARPS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    MsTables<T> tb;
    unsigned char* p = ms_load_tables<T, N, W * 32>(smem_raw, winG, twG, rootsG, tb);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    T* blk = reinterpret_cast<T*>(p) + (size_t)warp * (2 * M + 2 * QP);
    cplx<T>* fb = reinterpret_cast<cplx<T>*>(blk + 2 * M);
    const bool useWin = winG != nullptr;
    const T scale = (T)2 / (T)N;
    for (long long unit = (long long)blockIdx.x * W + warp; unit < nUnits; unit += (long long)gridDim.x * W) {
        const long long row = unit / segsPerRow;
        const int f0 = (int)(unit - row * segsPerRow) * seg, f1 = min(f0 + seg, nFrames);
        const T* src = x + row * rowStride + (long long)f0 * M;
        T* dst = out + (row * nFrames + f0) * M;
        MsBlock<T> pre;
        pre.load(src, lane, vec);
        pre.to_smem(blk, lane);
        pre.load(src + M, lane, vec);
        for (int f = f0; f < f1; ++f) {
            const int cur = (f - f0) & 1;
            T* b0 = blk + cur * M;           // samples [0, M) of the frame
            T* b1 = blk + (cur ^ 1) * M;     // samples [M, N)
            pre.to_smem(b1, lane);
            __syncwarp();
            if (f + 1 < f1) pre.load(src + (long long)(f - f0 + 2) * M, lane, vec);
            // fold + window + pre-twiddle (mdct_fold, one frame)
#pragma unroll 4
            for (int n = lane; n < Q; n += 32) {
                int i0, i1, i2, i3;
                T re, im;
                if (2 * n < Q) {
                    i0 = 3 * Q - 1 - 2 * n; i1 = 3 * Q + 2 * n; i2 = Q - 1 - 2 * n; i3 = Q + 2 * n;
                } else {
                    i0 = 2 * n - Q; i1 = 3 * Q - 1 - 2 * n; i2 = Q + 2 * n; i3 = 5 * Q - 1 - 2 * n;
                }
                T a0 = i0 < M ? b0[i0] : b1[i0 - M], a1 = i1 < M ? b0[i1] : b1[i1 - M];
                T a2 = i2 < M ? b0[i2] : b1[i2 - M], a3 = i3 < M ? b0[i3] : b1[i3 - M];
                if (useWin) { a0 *= tb.win[i0]; a1 *= tb.win[i1]; a2 *= tb.win[i2]; a3 *= tb.win[i3]; }
                if (2 * n < Q) { re = -a0 - a1; im = a2 - a3; } else { re = a0 - a1; im = -a2 - a3; }
                cplx<T> t = {re, im};
                fb[MS_PAD(n)] = cmul(t, tb.tw[n]);
            }
            __syncwarp();
            warp_fft512<T, Q>(fb, tb.roots, lane);
            // post-twiddle into the (now dead) first block, natural order
#pragma unroll 4
            for (int k = lane; k < Q; k += 32) {
                cplx<T> y = cmul(fb[MS_PAD(k)], tb.tw[k]);
                b0[2 * k] = scale * y.x;
                b0[M - 1 - 2 * k] = -scale * y.y;
            }
            __syncwarp();
            ms_store_block<T>(dst + (long long)(f - f0) * M, b0, lane, vec);
            __syncwarp();
        }
    }
}

// ---------------------------------------------------------------- inverse + window + overlap-add
// unit = (row, run of `seg` output blocks); a row has nFrames + 1 output blocks:
// block b = w[M:] * imdct(frame b-1)[M:] + w[:M] * imdct(frame b)[:M]  (coder/pacfile.py:289-295).
template <typename T>
__global__ void __launch_bounds__(MsCfg<T>::WARPS * 32, 1)
k_imdct_ola_stream(const T* __restrict__ X, int nFrames, int seg, int segsPerRow, long long nUnits,
                   const T* __restrict__ winG, const cplx<T>* __restrict__ twG, const cplx<T>* __restrict__ rootsG,
                   T* __restrict__ y, long long rowStride, int vec) {
    constexpr int N = 2048, M = 1024, Q = 512, QP = Q + Q / 8, W = MsCfg<T>::WARPS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    MsTables<T> tb;
    unsigned char* p = ms_load_tables<T, N, W * 32>(smem_raw, winG, twG, rootsG, tb);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    T* xv = reinterpret_cast<T*>(p) + (size_t)warp * (2 * M + 2 * QP);   // coefficients, then DCT-IV output
    T* tail = xv + M;                                                     // windowed second half of the previous frame
    cplx<T>* fb = reinterpret_cast<cplx<T>*>(xv + 2 * M);
    const int nOutRow = nFrames + 1;
    for (long long unit = (long long)blockIdx.x * W + warp; unit < nUnits; unit += (long long)gridDim.x * W) {
        const long long row = unit / segsPerRow;
        const int b0 = (int)(unit - row * segsPerRow) * seg, b1 = min(b0 + seg, nOutRow);
        const T* src = X + row * (long long)nFrames * M;
        T* dst = y + row * rowStride;
        MsBlock<T> pre;
        // frames b0-1 (halo: only its tail is needed) .. b1-1
        int f = b0 - 1;
        if (f >= 0) pre.load(src + (long long)f * M, lane, vec); else pre.zero();
        for (; f < b1; ++f) {
            const bool have = f >= 0 && f < nFrames;
            pre.to_smem(xv, lane);
            __syncwarp();
            if (f + 1 < b1 && f + 1 < nFrames) pre.load(src + (long long)(f + 1) * M, lane, vec); else pre.zero();
            if (have) {
#pragma unroll 4
                for (int n = lane; n < Q; n += 32) {
                    cplx<T> t = {xv[2 * n], xv[M - 1 - 2 * n]};
                    fb[MS_PAD(n)] = cmul(t, tb.tw[n]);
                }
                __syncwarp();
                warp_fft512<T, Q>(fb, tb.roots, lane);
#pragma unroll 4
                for (int k = lane; k < Q; k += 32) {
                    cplx<T> v = cmul(fb[MS_PAD(k)], tb.tw[k]);
                    xv[2 * k] = v.x;
                    xv[M - 1 - 2 * k] = -v.y;
                }
                __syncwarp();
            }
            // emit output block f (needs tail of f-1 and head of f), then keep this frame's tail
            constexpr int PER = 16 / sizeof(T), NV = MsBlock<T>::NV;
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                const int n0 = (i * 32 + lane) * PER;
                T o[PER], tl[PER];
#pragma unroll
                for (int e = 0; e < PER; ++e) {
                    const int n = n0 + e;
                    T head = (T)0, tn = (T)0;
                    if (have) {
                        head = n < Q ? (T)2 * xv[Q + n] : (T)-2 * xv[3 * Q - 1 - n];          // imdct_sample(v, n)
                        tn = n < Q ? (T)-2 * xv[Q - 1 - n] : (T)-2 * xv[n - Q];               // imdct_sample(v, M + n)
                        head *= tb.win[n];
                        tn *= tb.win[M + n];
                    }
                    o[e] = (f > b0 - 1 ? tail[n] : (T)0) + head;
                    tl[e] = tn;
                }
                if (f >= b0) {
                    T* d = dst + (long long)f * M + n0;
                    if (vec) *reinterpret_cast<uint4*>(d) = *reinterpret_cast<uint4*>(o);
                    else
                        for (int e = 0; e < PER; ++e) d[e] = o[e];
                }
#pragma unroll
                for (int e = 0; e < PER; ++e) tail[n0 + e] = tl[e];
            }
            __syncwarp();
        }
    }
}
