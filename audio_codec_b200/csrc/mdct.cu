// Batched window+MDCT and IMDCT+window+overlap-add kernels (fp32 and fp64).
// Replaces coder/mdct.py:43-76 + coder/window.py for whole streams at once:
// rows are (stream, channel) pairs, frames are hop-N/2 views into each row,
// so no overlapped copy of the PCM is ever materialised.
#include <algorithm>
#include "mdct.cuh"
#include "mdct_stream.cuh"
#include "mdct_tma.cuh"
#include "pac_api.h"

// ---------------------------------------------------------------- forward
// One CTA transforms FPC consecutive frames of one row: it stages (FPC+1)*M
// samples once (each sample is fetched from HBM once per CTA, the 1-block halo
// comes from L2), folds with the window, runs FPC N/4-point FFTs side by side
// in shared memory and writes M coefficients per frame, coalesced.
template <typename T, int N, int FPC, int NT>
__global__ void __launch_bounds__(NT)
k_mdct_fwd(const T* __restrict__ x, long long rowStride, int nFrames, int groupsPerRow,
           const T* __restrict__ winTab, int shape, const cplx<T>* __restrict__ tw,
           const cplx<T>* __restrict__ roots, T* __restrict__ out) {
    constexpr int M = N / 2, Q = N / 4;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* xs = reinterpret_cast<T*>(smem_raw);                       // [(FPC+1)*M]
    cplx<T>* fbuf = reinterpret_cast<cplx<T>*>(xs + (FPC + 1) * M); // [FPC][Q]
    const int tid = threadIdx.x;
    const int row = blockIdx.x / groupsPerRow;
    const int f0 = (blockIdx.x - row * groupsPerRow) * FPC;
    const int nf = min(FPC, nFrames - f0);
    const T* src = x + (long long)row * rowStride + (long long)f0 * M;
    const int nS = (nf + 1) * M;
    for (int i = tid; i < (FPC + 1) * M; i += NT) xs[i] = i < nS ? src[i] : (T)0;
    __syncthreads();
    mdct_fold<T, N, FPC, NT>(xs, M, winTab, nullptr, shape, tw, fbuf, tid);
    __syncthreads();
    cta_fft<T, Q, FPC, NT>(fbuf, roots, tid);
    mdct_post<T, N, FPC, NT>(fbuf, tw, (T)2 / (T)N, xs, tid);
    __syncthreads();
    T* dst = out + ((long long)row * nFrames + f0) * M;
    for (int i = tid; i < nf * M; i += NT) dst[i] = xs[i];
}

// ---------------------------------------------------------------- inverse
// One CTA produces FPC consecutive output blocks (M samples each) of one row.
// Output block b = windowed IMDCT(frame b-1)[M:] + windowed IMDCT(frame b)[:M],
// so the CTA inverts frames b0-1 .. b0+FPC-1 (FPC+1 transforms; the extra one is
// an L2 hit) and overlap-adds in shared memory (coder/pacfile.py:289-295).
template <typename T, int N, int FPC, int NT>
__global__ void __launch_bounds__(NT)
k_imdct_ola(const T* __restrict__ X, int nFrames, int groupsPerRow, const T* __restrict__ winTab, int shape,
            const cplx<T>* __restrict__ tw, const cplx<T>* __restrict__ roots, T* __restrict__ y,
            long long rowStride) {
    constexpr int M = N / 2, Q = N / 4, B = FPC + 1;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* v = reinterpret_cast<T*>(smem_raw);                 // [B][M] coefficients, then DCT-IV output
    cplx<T>* fbuf = reinterpret_cast<cplx<T>*>(v + B * M);  // [B][Q]
    const int tid = threadIdx.x;
    const int row = blockIdx.x / groupsPerRow;
    const int b0 = (blockIdx.x - row * groupsPerRow) * FPC;   // first output block
    const int nOut = min(FPC, nFrames + 1 - b0);
    // slot s holds frame b0-1+s
    const T* src = X + (long long)row * nFrames * M;
    for (int i = tid; i < B * M; i += NT) {
        int s = i / M, f = b0 - 1 + s;
        v[i] = (f >= 0 && f < nFrames) ? src[(long long)f * M + (i - s * M)] : (T)0;
    }
    __syncthreads();
    imdct_pre<T, N, B, NT>(v, (T)1, tw, fbuf, tid);
    __syncthreads();
    cta_fft<T, Q, B, NT>(fbuf, roots, tid);
    mdct_post<T, N, B, NT>(fbuf, tw, (T)1, v, tid);
    __syncthreads();
    const T* w = winTab ? winTab + shape * N : nullptr;
    T* dst = y + (long long)row * rowStride + (long long)b0 * M;
    for (int i = tid; i < nOut * M; i += NT) {
        int s = i / M, n = i - s * M;
        T a = imdct_sample<T, N>(v + s * M, M + n);        // tail of frame b0-1+s
        T c = imdct_sample<T, N>(v + (s + 1) * M, n);      // head of frame b0+s
        if (w) { a *= w[M + n]; c *= w[n]; }
        dst[i] = a + c;
    }
}

template <typename T> static const PacTables<T>& tabs(const pac_ctx* c);
template <> const PacTables<float>& tabs<float>(const pac_ctx* c) { return c->host.f; }
template <> const PacTables<double>& tabs<double>(const pac_ctx* c) { return c->host.d; }

template <typename T, int N>
static int launch_fwd(pac_ctx* ctx, const void* x, long long rowStride, int nRows, int nFrames, int window,
                      void* coefs, cudaStream_t st) {
    constexpr int FPC = (N == 2048) ? 4 : 16, NT = 256, M = N / 2, Q = N / 4;
    const PacTables<T>& t = tabs<T>(ctx);
    const T* win = window == PAC_WIN_NONE ? nullptr : (N == 2048 ? t.win_long : t.win_short);
    const cplx<T>* tw = reinterpret_cast<const cplx<T>*>(N == 2048 ? t.tw_mdct_long : t.tw_mdct_short);
    const cplx<T>* roots = reinterpret_cast<const cplx<T>*>(t.tw_fft[N == 2048 ? 2 : 0]);
    int groups = (nFrames + FPC - 1) / FPC;
    size_t smem = sizeof(T) * ((FPC + 1) * M + 2 * FPC * Q);
    auto kern = k_mdct_fwd<T, N, FPC, NT>;
    PAC_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<nRows * groups, NT, smem, st>>>((const T*)x, rowStride, nFrames, groups, win, window, tw, roots,
                                             (T*)coefs);
    PAC_CHECK(cudaGetLastError());
    return 0;
}

template <typename T, int N>
static int launch_inv(pac_ctx* ctx, const void* X, int nRows, int nFrames, int window, void* y,
                      long long rowStride, cudaStream_t st) {
    constexpr int FPC = (N == 2048) ? 4 : 16, NT = 256, M = N / 2, Q = N / 4, B = FPC + 1;
    const PacTables<T>& t = tabs<T>(ctx);
    const T* win = window == PAC_WIN_NONE ? nullptr : (N == 2048 ? t.win_long : t.win_short);
    const cplx<T>* tw = reinterpret_cast<const cplx<T>*>(N == 2048 ? t.tw_mdct_long : t.tw_mdct_short);
    const cplx<T>* roots = reinterpret_cast<const cplx<T>*>(t.tw_fft[N == 2048 ? 2 : 0]);
    int groups = (nFrames + 1 + FPC - 1) / FPC;
    size_t smem = sizeof(T) * (B * M + 2 * B * Q);
    auto kern = k_imdct_ola<T, N, FPC, NT>;
    PAC_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<nRows * groups, NT, smem, st>>>((const T*)X, nFrames, groups, win, window, tw, roots, (T*)y, rowStride);
    PAC_CHECK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------- any size
// Direct evaluation of the defining sums (coder/mdct.py:11-39 MDCTslow; MDCT/IMDCT :43-76 compute the
// same values through an N-point FFT) for block sizes the N/4-FFT kernels do not cover: any a, b with
// a + b even.  One thread per output value; the phase (2n + b + 1)(2k + 1) is reduced modulo 4N in
// integers, so the cosine argument is exact.  O(N^2 / 2) per frame: meant for the small sizes of the
// reference's self-tests and for odd window splits, not for the codec's 2048 / 256.
template <typename T>
__global__ void k_mdct_direct(const T* __restrict__ in, long long nRows, int a, int b, int inverse, T* __restrict__ out) {
    const int N = a + b, H = N / 2;
    const int nOut = inverse ? N : H, nIn = inverse ? H : N;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nRows * nOut) return;
    const long long row = idx / nOut;
    const int o = (int)(idx - row * nOut);
    const T* x = in + row * nIn;
    const long long per = 4LL * N;
    double acc = 0.0;
    for (int i = 0; i < nIn; ++i) {
        const int n = inverse ? o : i, k = inverse ? i : o;
        const long long ph = ((long long)(2 * n + b + 1) * (long long)(2 * k + 1)) % per;
        acc += (double)x[i] * cospi((double)ph / (double)(2 * N));
    }
    out[idx] = (T)(inverse ? 2.0 * acc : (2.0 / (double)N) * acc);
}

extern "C" int pac_mdct_direct(pac_ctx* ctx, int dtype, const void* d_in, long long n_rows, int a, int b, int inverse,
                               void* d_out, void* stream) {
    if (!ctx || a < 0 || b < 0 || a + b < 2 || ((a + b) & 1)) return PAC_ERR_ARG;
    if (n_rows <= 0) return 0;
    if (!d_in || !d_out) return PAC_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    PAC_CHECK(cudaSetDevice(ctx->device));
    const long long tot = n_rows * (inverse ? a + b : (a + b) / 2);
    const unsigned grid = (unsigned)((tot + 127) / 128);
    if (dtype == PAC_F32) k_mdct_direct<float><<<grid, 128, 0, st>>>((const float*)d_in, n_rows, a, b, inverse, (float*)d_out);
    else if (dtype == PAC_F64) k_mdct_direct<double><<<grid, 128, 0, st>>>((const double*)d_in, n_rows, a, b, inverse, (double*)d_out);
    else return PAC_ERR_ARG;
    PAC_CHECK(cudaGetLastError());
    return 0;
}

static bool aligned16(const void* p) { return ((size_t)p & 15) == 0; }

// run length of a warp's walk: long enough to amortise the one halo block, short enough that
// every warp of the grid gets several runs
static int stream_seg(long long nRows, int nPerRow, int warps) {
    long long total = nRows * (long long)nPerRow, want = total / ((long long)148 * warps * 4);
    if (want < 1) want = 1;
    if (want > 128) want = 128;
    if (want > nPerRow) want = nPerRow;
    return (int)want;
}

template <typename T>
static int launch_fwd_stream(pac_ctx* ctx, const void* x, long long rowStride, int nRows, int nFrames, int window,
                             void* coefs, cudaStream_t st) {
    const PacTables<T>& t = tabs<T>(ctx);
    constexpr int W = MsCfg<T>::WARPS;
    const T* win = window == PAC_WIN_NONE ? nullptr : t.win_long + (size_t)window * 2048;
    const int seg = stream_seg(nRows, nFrames, W), segs = (nFrames + seg - 1) / seg;
    const long long units = (long long)nRows * segs;
    const int grid = (int)std::min<long long>(148, (units + W - 1) / W);
    const int vec = aligned16(x) && aligned16(coefs) && (rowStride * sizeof(T)) % 16 == 0;
    const bool sym = window == PAC_WIN_SINE || window == PAC_WIN_KAISER || window == PAC_WIN_NONE;
    if (vec && !(ctx->optVariant & 8)) {             // bulk-copy (TMA) pipeline; pac_set_option variant bit 3: the LDG / STG one
        constexpr int WT = MtCfg<T>::WARPS;
        const int segT = stream_seg(nRows, nFrames, WT), segsT = (nFrames + segT - 1) / segT;
        const long long unitsT = (long long)nRows * segsT;
        const int gridT = (int)std::min<long long>(148, (unitsT + WT - 1) / WT);
        auto kt = sym ? k_mdct_fwd_tma<T, true> : k_mdct_fwd_tma<T, false>;
        const size_t smemT = mt_smem_bytes<T>();
        PAC_CHECK(cudaFuncSetAttribute(kt, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemT));
        kt<<<gridT, WT * 32, smemT, st>>>((const T*)x, rowStride, nFrames, segT, segsT, unitsT, win,
                                           reinterpret_cast<const cplx<T>*>(t.tw_mdct_long),
                                           reinterpret_cast<const cplx<T>*>(t.tw_fft[2]), (T*)coefs);
        PAC_CHECK(cudaGetLastError());
        return 0;
    }
    auto kern = sym ? k_mdct_fwd_stream<T, true> : k_mdct_fwd_stream<T, false>;
    const size_t smem = ms_smem_bytes<T>();
    PAC_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, W * 32, smem, st>>>((const T*)x, rowStride, nFrames, seg, segs, units, win,
                                      reinterpret_cast<const cplx<T>*>(t.tw_mdct_long),
                                      reinterpret_cast<const cplx<T>*>(t.tw_fft[2]), (T*)coefs, vec);
    PAC_CHECK(cudaGetLastError());
    return 0;
}

template <typename T>
static int launch_inv_stream(pac_ctx* ctx, const void* X, int nRows, int nFrames, int window, void* y,
                             long long rowStride, cudaStream_t st) {
    const PacTables<T>& t = tabs<T>(ctx);
    constexpr int W = MsCfg<T>::WARPS_INV;
    const T* win = window == PAC_WIN_NONE ? nullptr : t.win_long + (size_t)window * 2048;
    const int seg = stream_seg(nRows, nFrames + 1, W), segs = (nFrames + 1 + seg - 1) / seg;
    const long long units = (long long)nRows * segs;
    const int grid = (int)std::min<long long>(148, (units + W - 1) / W);
    const int vec = aligned16(X) && aligned16(y) && (rowStride * sizeof(T)) % 16 == 0;
    if (vec && !(ctx->optVariant & 8)) {             // bulk-copy (TMA) input pipeline
        constexpr int WT = MtCfg<T>::WARPS;
        const int segT = stream_seg(nRows, nFrames + 1, WT), segsT = (nFrames + 1 + segT - 1) / segT;
        const long long unitsT = (long long)nRows * segsT;
        const int gridT = (int)std::min<long long>(148, (unitsT + WT - 1) / WT);
        auto kt = k_imdct_ola_tma<T>;
        const size_t smemT = mt_smem_bytes<T>();
        PAC_CHECK(cudaFuncSetAttribute(kt, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemT));
        kt<<<gridT, WT * 32, smemT, st>>>((const T*)X, nFrames, segT, segsT, unitsT, win,
                                           reinterpret_cast<const cplx<T>*>(t.tw_mdct_long),
                                           reinterpret_cast<const cplx<T>*>(t.tw_fft[2]), (T*)y, rowStride);
        PAC_CHECK(cudaGetLastError());
        return 0;
    }
    auto kern = k_imdct_ola_stream<T>;
    const size_t smem = ms_inv_smem_bytes<T>();
    PAC_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, W * 32, smem, st>>>((const T*)X, nFrames, seg, segs, units, win,
                                      reinterpret_cast<const cplx<T>*>(t.tw_mdct_long),
                                      reinterpret_cast<const cplx<T>*>(t.tw_fft[2]), (T*)y, rowStride, vec);
    PAC_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int pac_mdct_forward(pac_ctx* ctx, int dtype, const void* x, long long row_stride, int n_rows,
                                int n_frames, int N, int window, void* coefs, void* stream) {
    if (!ctx) return PAC_ERR_ARG;
    if (n_rows <= 0 || n_frames <= 0) return 0;
    if (!x || !coefs) return PAC_ERR_ARG;
    if (window < 0 || window > PAC_WIN_NONE) return PAC_ERR_ARG;
    if (N == 256 && !(window == PAC_WIN_SINE || window == PAC_WIN_KAISER || window == PAC_WIN_NONE)) return PAC_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    PAC_CHECK(cudaSetDevice(ctx->device));
    if (N == 2048 && dtype == PAC_F32) return launch_fwd_stream<float>(ctx, x, row_stride, n_rows, n_frames, window, coefs, st);
    if (N == 2048 && dtype == PAC_F64) return launch_fwd_stream<double>(ctx, x, row_stride, n_rows, n_frames, window, coefs, st);
    if (N == 256 && dtype == PAC_F32) return launch_fwd<float, 256>(ctx, x, row_stride, n_rows, n_frames, window, coefs, st);
    if (N == 256 && dtype == PAC_F64) return launch_fwd<double, 256>(ctx, x, row_stride, n_rows, n_frames, window, coefs, st);
    return PAC_ERR_ARG;
}

extern "C" int pac_imdct_ola(pac_ctx* ctx, int dtype, const void* coefs, int n_rows, int n_frames, int N,
                             int window, void* y, long long row_stride, void* stream) {
    if (!ctx) return PAC_ERR_ARG;
    if (n_rows <= 0 || n_frames <= 0) return 0;
    if (!y || !coefs) return PAC_ERR_ARG;
    if (window < 0 || window > PAC_WIN_NONE) return PAC_ERR_ARG;
    if (N == 256 && !(window == PAC_WIN_SINE || window == PAC_WIN_KAISER || window == PAC_WIN_NONE)) return PAC_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    PAC_CHECK(cudaSetDevice(ctx->device));
    if (N == 2048 && dtype == PAC_F32) return launch_inv_stream<float>(ctx, coefs, n_rows, n_frames, window, y, row_stride, st);
    if (N == 2048 && dtype == PAC_F64) return launch_inv_stream<double>(ctx, coefs, n_rows, n_frames, window, y, row_stride, st);
    if (N == 256 && dtype == PAC_F32) return launch_inv<float, 256>(ctx, coefs, n_rows, n_frames, window, y, row_stride, st);
    if (N == 256 && dtype == PAC_F64) return launch_inv<double, 256>(ctx, coefs, n_rows, n_frames, window, y, row_stride, st);
    return PAC_ERR_ARG;
}
