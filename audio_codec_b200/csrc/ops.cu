// Per-function device ops: the reference's Python function surface, one batched
// kernel per function.  The fused stream kernels in encode.cu /
// decode.cu are built from the same device routines; these entry points exist so
// that window.*, detect_transients.parTransientDetect, psychoac.CalcSMRs,
// bitalloc.BitAlloc, quantize.*, gain_shape_quantize.*, bitpack.PackedBits and
// codec.EncodeSingleChannel / Decode can each be called on their own.
#include "decode_frame.cuh"
#include "encode_frame.cuh"
#include "pac_api.h"

#define OPS_LAUNCH_CHECK() PAC_CHECK(cudaGetLastError())

// ---------------------------------------------------------------- window.*
template <typename T>
__global__ void k_window(const T* __restrict__ x, const T* __restrict__ w, long long total, int N, T* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long step = (long long)gridDim.x * blockDim.x;
    for (; i < total; i += step) out[i] = w[i % N] * x[i];
}

extern "C" int pac_window(pac_ctx* ctx, int dtype, const void* d_x, long long n_rows, int N, int window, int n_short,
                          double alpha, void* d_out, void* stream) {
    if (!ctx || N < 1) return PAC_ERR_ARG;
    if (n_rows <= 0) return 0;
    if (!d_x || !d_out) return PAC_ERR_ARG;
    PAC_CHECK(cudaSetDevice(ctx->device));
    const double* wd;
    const float* wf;
    int rc = pac_window_table(ctx, window, N, n_short, alpha, &wd, &wf);
    if (rc) return rc;
    const long long total = n_rows * N;
    const int grid = (int)std::min<long long>((total + 255) / 256, 148 * 16);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == PAC_F32) k_window<float><<<grid, 256, 0, st>>>((const float*)d_x, wf, total, N, (float*)d_out);
    else if (dtype == PAC_F64) k_window<double><<<grid, 256, 0, st>>>((const double*)d_x, wd, total, N, (double*)d_out);
    else return PAC_ERR_ARG;
    OPS_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------- detect_transients.parTransientDetect
// One CTA per block of [C][len] doubles (coder/detect_transients.py:5-23).
__global__ void __launch_bounds__(128)
k_transient_f64(const double* __restrict__ x, int C, int len, double thresh, unsigned char* __restrict__ out) {
    __shared__ double s_v[4];
    __shared__ int s_i[4];
    __shared__ double s_max[16];
    __shared__ int s_n;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const double* blk = x + (long long)blockIdx.x * C * len;
    int nmax = 0;
    for (int c = 0; c < C; ++c) {
        const double* p = blk + (long long)c * len;
        double best = -1.0;
        int arg = 0;
        for (int i = tid; i < len; i += 128) {
            double v = fabs(p[i]);
            if (v > best) { best = v; arg = i; }
        }
        for (int o = 16; o; o >>= 1) {
            double ob = __shfl_xor_sync(0xffffffffu, best, o);
            int oa = __shfl_xor_sync(0xffffffffu, arg, o);
            if (ob > best || (ob == best && oa < arg)) { best = ob; arg = oa; }
        }
        if (lane == 0) { s_v[warp] = best; s_i[warp] = arg; }
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < 4; ++w)
                if (s_v[w] > best || (s_v[w] == best && s_i[w] < arg)) { best = s_v[w]; arg = s_i[w]; }
            if (c < 16) s_max[c] = best;
            nmax = max(nmax, arg + 500);
            if (c == C - 1) s_n = min(nmax, len);
        }
        __syncthreads();
    }
    const int n = s_n > 0 ? s_n : len;
    double sum = 0.0;
    for (int c = 0; c < C; ++c)
        for (int i = tid; i < n; i += 128) sum += fabs(blk[(long long)c * len + i]);
    sum = warp_sum(sum);
    if (lane == 0) s_v[warp] = sum;
    __syncthreads();
    if (tid == 0) {
        double avg = ((s_v[0] + s_v[1]) + (s_v[2] + s_v[3])) / ((double)C * (double)n);
        bool t = false;
        if (avg != 0.0)
            for (int c = 0; c < C && c < 16; ++c) t |= s_max[c] / avg > thresh;
        out[blockIdx.x] = t ? 1 : 0;
    }
}

extern "C" int pac_transient_detect(pac_ctx* ctx, const double* d_blocks, long long n_blocks, int n_channels, int len,
                                    double thresh, unsigned char* d_flags, void* stream) {
    if (!ctx || n_channels < 1 || n_channels > 16 || len < 1) return PAC_ERR_ARG;
    if (n_blocks <= 0) return 0;
    if (!d_blocks || !d_flags) return PAC_ERR_ARG;
    PAC_CHECK(cudaSetDevice(ctx->device));
    k_transient_f64<<<(unsigned)n_blocks, 128, 0, (cudaStream_t)stream>>>(d_blocks, n_channels, len, thresh, d_flags);
    OPS_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------- psychoac.CalcSMRs / getMaskedThreshold
template <int N, typename T>
__global__ void __launch_bounds__(ENC_NT)
k_calc_smr(const PacDev* __restrict__ dev, const double* __restrict__ data, const double* __restrict__ mdct,
           const int* __restrict__ scale, PacBands bands, double* __restrict__ smr, double* __restrict__ maskOut) {
    constexpr int H = N / 2;
    constexpr int FB = N * (int)sizeof(T) > PSY_LIST_BYTES ? N * (int)sizeof(T) : PSY_LIST_BYTES;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* lines = reinterpret_cast<double*>(smem_raw);  // [H]
    double* smrS = lines + H;                             // [32]
    T* fb = reinterpret_cast<T*>(smrS + 32);              // FB bytes: FFT buffer, then the masker list
    T* xs = reinterpret_cast<T*>(smem_raw + sizeof(double) * (H + 32) + FB);   // [N]
    T* mk = xs + N;                                       // [H + 16]
    const int tid = threadIdx.x;
    const long long f = blockIdx.x;
    const PsyConst<T> pc = psy_const<N, T>(dev);
    for (int i = tid; i < N; i += ENC_NT) xs[i] = (T)data[f * N + i];
    if (mdct) for (int i = tid; i < H; i += ENC_NT) lines[i] = mdct[f * H + i];
    __syncthreads();
    hann_rfft_mag<N, ENC_NT, T>(xs, pc, reinterpret_cast<cplx<T>*>(fb), mk, tid);
    for (int k = tid; k <= H; k += ENC_NT) {
        T m = mk[k];
        mk[k] = pc.norm * (m * m);
    }
    __syncthreads();
    const MaskerList<T> ml = masker_list<T>(fb);
    const int nMask = pick_maskers<N, ENC_NT, T>(mk, pc, ml, tid);
    masked_threshold<N, ENC_NT, T>(pc, ml, nMask, mk, tid);
    if (maskOut) for (int i = tid; i < H; i += ENC_NT) maskOut[f * H + i] = (double)mk[i];
    if (smr) {
        band_smr<ENC_NT, T>(lines, scale[f], mk, bands, smrS, tid);
        if (tid < bands.nBands) smr[f * bands.nBands + tid] = smrS[tid];
    }
}

static int fill_bands(PacBands* b, int n_bands, const int* h_lower, const int* h_nlines, int H) {
    if (n_bands < 1 || n_bands > PAC_MAX_BANDS || !h_lower || !h_nlines) return PAC_ERR_ARG;
    memset(b, 0, sizeof(*b));
    b->nBands = n_bands;
    for (int k = 0; k < n_bands; ++k) {
        if (h_lower[k] < 0 || h_nlines[k] < 0 || h_lower[k] + h_nlines[k] > H) return PAC_ERR_ARG;
        b->lower[k] = h_lower[k];
        b->nLines[k] = h_nlines[k];
    }
    return 0;
}

extern "C" int pac_calc_smr(pac_ctx* ctx, int dtype, const double* d_data, const double* d_mdct, const int* d_scale,
                            long long n_frames, int N, int n_bands, const int* h_lower, const int* h_nlines,
                            double* d_smr, double* d_mask, void* stream) {
    if (!ctx || (N != PAC_NLONG && N != PAC_NSHORT)) return PAC_ERR_ARG;
    if (n_frames <= 0) return 0;
    if (!d_data || (!d_smr && !d_mask)) return PAC_ERR_ARG;
    if (d_smr && (!d_mdct || !d_scale)) return PAC_ERR_ARG;
    PacBands b;
    memset(&b, 0, sizeof(b));
    if (d_smr) {
        int rc = fill_bands(&b, n_bands, h_lower, h_nlines, N / 2);
        if (rc) return rc;
    }
    PAC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int H = N / 2;
    if (dtype != PAC_F32 && dtype != PAC_F64) return PAC_ERR_ARG;
    const size_t eb = dtype == PAC_F32 ? sizeof(float) : sizeof(double);
    const size_t fbB = (size_t)N * eb > PSY_LIST_BYTES ? (size_t)N * eb : PSY_LIST_BYTES;
    const size_t smem = sizeof(double) * (H + 32) + fbB + eb * ((size_t)N + H + 16);
    auto go = [&](auto kern) -> int {
        PAC_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<(unsigned)n_frames, ENC_NT, smem, st>>>(ctx->dev, d_data, d_mdct, d_scale, b, d_smr, d_mask);
        return 0;
    };
    int rc2;
    if (dtype == PAC_F32) rc2 = N == PAC_NLONG ? go(k_calc_smr<PAC_NLONG, float>) : go(k_calc_smr<PAC_NSHORT, float>);
    else rc2 = N == PAC_NLONG ? go(k_calc_smr<PAC_NLONG, double>) : go(k_calc_smr<PAC_NSHORT, double>);
    if (rc2) return rc2;
    OPS_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------- bitalloc.BitAlloc
__global__ void __launch_bounds__(128)
k_bit_alloc(const double* __restrict__ budget, int maxMant, PacBands bands, const double* __restrict__ smr,
            long long nFrames, int* __restrict__ bits) {
    __shared__ double s_tmp[4][32];
    __shared__ double s_smr[4][32];
    __shared__ int s_nl[32];
    __shared__ int s_out[4][32];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid < 32) s_nl[tid] = tid < bands.nBands ? bands.nLines[tid] : 0;
    const long long f = (long long)blockIdx.x * 4 + warp;
    if (f < nFrames && lane < bands.nBands) s_smr[warp][lane] = smr[f * bands.nBands + lane];
    __syncthreads();
    if (f >= nFrames) return;
    warp_bit_alloc(budget[f], maxMant, bands.nBands, s_nl, s_smr[warp], s_tmp[warp], s_out[warp]);
    if (lane < bands.nBands) bits[f * bands.nBands + lane] = s_out[warp][lane];
}

extern "C" int pac_bit_alloc(pac_ctx* ctx, const double* d_budget, int max_mant_bits, int n_bands, const int* h_nlines,
                             const double* d_smr, long long n_frames, int* d_bits, void* stream) {
    if (!ctx || n_bands < 1 || n_bands > PAC_MAX_BANDS || !h_nlines) return PAC_ERR_ARG;
    if (n_frames <= 0) return 0;
    if (!d_budget || !d_smr || !d_bits) return PAC_ERR_ARG;
    PacBands b;
    memset(&b, 0, sizeof(b));
    b.nBands = n_bands;
    for (int k = 0; k < n_bands; ++k) b.nLines[k] = h_nlines[k];
    PAC_CHECK(cudaSetDevice(ctx->device));
    k_bit_alloc<<<(unsigned)((n_frames + 3) / 4), 128, 0, (cudaStream_t)stream>>>(d_budget, max_mant_bits, b, d_smr,
                                                                                  n_frames, d_bits);
    OPS_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------- quantize.*
enum { Q_UNIFORM = 0, Q_DEUNIFORM = 1, Q_SCALE = 2, Q_MANT = 3, Q_DEMANT = 4 };

__global__ void k_quant(int op, const double* __restrict__ xin, const long long* __restrict__ cin, long long n,
                        int scale, int nScale, int nBits, double* __restrict__ xout, long long* __restrict__ cout) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long step = (long long)gridDim.x * blockDim.x;
    for (; i < n; i += step) {
        switch (op) {
            case Q_UNIFORM: cout[i] = (long long)quantize_uniform(xin[i], nBits); break;
            case Q_DEUNIFORM: xout[i] = dequantize_uniform((unsigned long long)cin[i], nBits); break;
            case Q_SCALE: cout[i] = scale_factor(xin[i], nScale, nBits); break;
            case Q_MANT: cout[i] = mantissa_bfp(xin[i], scale, nScale, nBits); break;
            case Q_DEMANT: xout[i] = dequantize_bfp(scale, (unsigned)cin[i], nScale, nBits); break;
        }
    }
}

static int launch_quant(pac_ctx* ctx, int op, const double* xin, const long long* cin, long long n, int scale,
                        int nScale, int nBits, double* xout, long long* cout, void* stream) {
    if (!ctx) return PAC_ERR_ARG;
    if (n <= 0) return 0;
    PAC_CHECK(cudaSetDevice(ctx->device));
    const int grid = (int)std::min<long long>((n + 255) / 256, 148 * 16);
    k_quant<<<grid, 256, 0, (cudaStream_t)stream>>>(op, xin, cin, n, scale, nScale, nBits, xout, cout);
    OPS_LAUNCH_CHECK();
    return 0;
}

extern "C" int pac_quantize_uniform(pac_ctx* ctx, const double* d_x, long long n, int n_bits, long long* d_code,
                                    void* stream) {
    if (n > 0 && (!d_x || !d_code)) return PAC_ERR_ARG;
    if (n_bits > 64) return PAC_ERR_ARG;
    return launch_quant(ctx, Q_UNIFORM, d_x, nullptr, n, 0, 0, n_bits, nullptr, d_code, stream);
}
extern "C" int pac_dequantize_uniform(pac_ctx* ctx, const long long* d_code, long long n, int n_bits, double* d_x,
                                      void* stream) {
    if (n > 0 && (!d_x || !d_code)) return PAC_ERR_ARG;
    if (n_bits > 64) return PAC_ERR_ARG;
    return launch_quant(ctx, Q_DEUNIFORM, nullptr, d_code, n, 0, 0, n_bits, d_x, nullptr, stream);
}
extern "C" int pac_scale_factor(pac_ctx* ctx, const double* d_x, long long n, int n_scale_bits, int n_mant_bits,
                                long long* d_scale, void* stream) {
    if (n > 0 && (!d_x || !d_scale)) return PAC_ERR_ARG;
    if (n_scale_bits < 1 || n_scale_bits > 5 || n_mant_bits < 0 || (1 << n_scale_bits) - 1 + n_mant_bits > 62) return PAC_ERR_ARG;
    return launch_quant(ctx, Q_SCALE, d_x, nullptr, n, 0, n_scale_bits, n_mant_bits, nullptr, d_scale, stream);
}
extern "C" int pac_mantissa(pac_ctx* ctx, const double* d_x, long long n, int scale, int n_scale_bits, int n_mant_bits,
                            long long* d_mant, void* stream) {
    if (n > 0 && (!d_x || !d_mant)) return PAC_ERR_ARG;
    if (n_scale_bits < 1 || n_scale_bits > 5 || n_mant_bits < 1 || n_mant_bits > 31 || scale < 0 ||
        scale > (1 << n_scale_bits) - 1)
        return PAC_ERR_ARG;
    return launch_quant(ctx, Q_MANT, d_x, nullptr, n, scale, n_scale_bits, n_mant_bits, nullptr, d_mant, stream);
}
extern "C" int pac_dequantize(pac_ctx* ctx, int scale, const long long* d_mant, long long n, int n_scale_bits,
                              int n_mant_bits, double* d_x, void* stream) {
    if (n > 0 && (!d_x || !d_mant)) return PAC_ERR_ARG;
    if (n_scale_bits < 1 || n_scale_bits > 5 || n_mant_bits < 1 || n_mant_bits > 31 || scale < 0 ||
        scale > (1 << n_scale_bits) - 1)
        return PAC_ERR_ARG;
    return launch_quant(ctx, Q_DEMANT, nullptr, d_mant, n, scale, n_scale_bits, n_mant_bits, d_x, nullptr, stream);
}

// ---------------------------------------------------------------- psychoac.SPL / Intensity / Bark / Thresh
enum { PSY_SPL = 0, PSY_INTENSITY = 1, PSY_BARK = 2, PSY_THRESH = 3,
       PSY_SPL_SL = 4, PSY_BARK_SL = 5 };      // the straight-line forms the stream encoder uses (must give the same bits)

__global__ void k_psy_map(int op, const double* __restrict__ x, long long n, double* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long step = (long long)gridDim.x * blockDim.x;
    for (; i < n; i += step) {
        double v = x[i], r;
        switch (op) {
            case PSY_SPL: r = spl_of(v); break;                               // coder/psychoac.py:10-25
            case PSY_INTENSITY: r = pow(10.0, (v - 96.0) / 10.0); break;      // :28-32
            case PSY_BARK: r = bark_of(v); break;                             // :43-48
            case PSY_SPL_SL: r = spl_pn<double>(v); break;
            case PSY_BARK_SL: r = bark_sl(v); break;
            default: {                                                        // :35-40
                double f = v < 10.0 ? 10.0 : v;
                double u = f / 1000.0, d = u - 3.3;
                r = 3.64 * pow(u, -0.8) - 6.5 * exp(-0.6 * (d * d)) + 0.001 * pow(u, 4.0);
            }
        }
        out[i] = r;
    }
}

extern "C" int pac_psy_map(pac_ctx* ctx, int op, const double* d_x, long long n, double* d_out, void* stream) {
    if (!ctx || op < PSY_SPL || op > PSY_BARK_SL) return PAC_ERR_ARG;
    if (n <= 0) return 0;
    if (!d_x || !d_out) return PAC_ERR_ARG;
    PAC_CHECK(cudaSetDevice(ctx->device));
    const int grid = (int)std::min<long long>((n + 255) / 256, 148 * 16);
    k_psy_map<<<grid, 256, 0, (cudaStream_t)stream>>>(op, d_x, n, d_out);
    OPS_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------- pcmfile.PCMFile.ReadDataBlock's sample mapping
// int16 code -> signed fraction (coder/pcmfile.py:89-98): the division-free form the stream encoder uses
// (pcm_to_double_fast), exposed so that it can be checked for every code against the division.
__global__ void k_pcm_to_float(const short* __restrict__ x, long long n, double* __restrict__ out) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = pcm_to_double_fast(x[i]);
}

extern "C" int pac_pcm_to_float(pac_ctx* ctx, const short* d_pcm, long long n, double* d_out, void* stream) {
    if (!ctx) return PAC_ERR_ARG;
    if (n <= 0) return 0;
    if (!d_pcm || !d_out) return PAC_ERR_ARG;
    PAC_CHECK(cudaSetDevice(ctx->device));
    const int grid = (int)std::min<long long>((n + 255) / 256, 148 * 16);
    k_pcm_to_float<<<grid, 256, 0, (cudaStream_t)stream>>>(d_pcm, n, d_out);
    OPS_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------- self-test of pac_div (pac_common.cuh)
// n pseudo-random (x, d) pairs per magnitude class: quotients of the kinds the quantisers form (|x| <= d, d a norm in
// (0, 4]; also wide exponent ranges).  Counts the pairs for which pac_div differs from the IEEE division.
__global__ void k_div_selftest(long long n, unsigned long long seed, unsigned long long* __restrict__ bad) {
    unsigned long long cnt = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        unsigned long long h = (unsigned long long)i * 0x9E3779B97F4A7C15ull + seed;
        h ^= h >> 31; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 29; h *= 0x94D049BB133111EBull; h ^= h >> 32;
        unsigned long long g = h * 0xD6E8FEB86659FD93ull + 0x2545F4914F6CDD1Dull;
        g ^= g >> 33;
        const int cls = (int)(i & 3);
        // mantissas from the hash; exponents by class
        double d = __longlong_as_double((long long)((h & 0x000FFFFFFFFFFFFFull) | 0x3FF0000000000000ull));      // [1, 2)
        double x = __longlong_as_double((long long)((g & 0x000FFFFFFFFFFFFFull) | 0x3FF0000000000000ull));
        const int ed = cls == 0 ? -(int)((h >> 52) & 15) : cls == 1 ? 1 - (int)((h >> 52) & 63) : cls == 2 ? 0 : 200 - (int)((h >> 52) & 511);
        const int ex = cls == 3 ? 100 - (int)((g >> 52) & 255) : ed - (int)((g >> 52) & 63);
        d = ldexp(d, ed);
        x = ldexp(x, ex);
        if (g >> 63) x = -x;
        if (cls == 2 && ((g >> 58) & 31) == 0) x = d;                 // the exact quotient 1
        const PacDiv b = pac_div_by(d);
        if (__double_as_longlong(pac_div(x, b)) != __double_as_longlong(x / d)) ++cnt;
        const PacDiv r2 = {__longlong_as_double(0x3FF6A09E667F3BCDLL), __longlong_as_double(0x3FE6A09E667F3BCCLL)};
        if (__double_as_longlong(pac_div(x, r2)) != __double_as_longlong(x / sqrt(2.0))) ++cnt;
    }
    if (cnt) atomicAdd(bad, cnt);
}

extern "C" long long pac_selftest_division(pac_ctx* ctx, long long n_pairs, unsigned long long seed, void* stream) {
    if (!ctx || n_pairs <= 0) return PAC_ERR_ARG;
    PAC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    void* scr = nullptr;
    int rc = pac_scratch(ctx, 256, &scr);
    if (rc) return rc;
    PAC_CHECK(cudaMemsetAsync(scr, 0, 8, st));
    k_div_selftest<<<148 * 8, 256, 0, st>>>(n_pairs, seed, (unsigned long long*)scr);
    OPS_LAUNCH_CHECK();
    unsigned long long bad = 0;
    PAC_CHECK(cudaMemcpyAsync(&bad, scr, 8, cudaMemcpyDeviceToHost, st));
    PAC_CHECK(cudaStreamSynchronize(st));
    return (long long)bad;
}

// ---------------------------------------------------------------- gain_shape_quantize.*
static int pow2_cap(int L) {
    int c = 1;
    while (c < L) c <<= 1;
    return c;
}

// one warp (= one CTA) per vector
__global__ void __launch_bounds__(32)
k_gs_quantize(const PacDev* __restrict__ dev, const double* __restrict__ x, int L, int cap, const int* __restrict__ budget,
              unsigned long long* __restrict__ val, int* __restrict__ width, int* __restrict__ count, int maxF,
              int* __restrict__ err) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* xs = reinterpret_cast<double*>(smem_raw);   // [cap]
    double* wsA = xs + cap;
    double* wsB = wsA + cap;
    int* yws = reinterpret_cast<int*>(wsB + cap);       // [cap]
    unsigned* bits = reinterpret_cast<unsigned*>(yws + cap);   // scratch bit buffer
    const long long v = blockIdx.x;
    const int lane = threadIdx.x;
    const int R = budget[v];
    for (int i = lane; i < L; i += 32) xs[i] = x[v * L + i];
    const int nW = (max(R, 0) + 31) / 32 + 4;
    for (int i = lane; i < nW; i += 32) bits[i] = 0u;
    if (lane == 0) count[v] = 0;
    __syncwarp();
    PvqTab pt = {dev->pvqN, dev->pvqRowOff, dev->log2tab, dev->pvqKW, nullptr};
    BitW bw = {bits, 0};
    FieldLog lg = {val + v * maxF, width + v * maxF, count + v, maxF};
    int werr = 0;
    int used = warp_quantize_gain_shape(pt, xs, L, R, wsA, wsB, yws, cap, bw, &werr, &lg);
    if (lane == 0) {
        if (used == 0) {               // all-zero input: ([0], [0]) in the reference (:485-487)
            val[v * maxF] = 0ull;
            width[v * maxF] = 0;
            count[v] = 1;
        }
        if (werr) atomicOr(err, werr);
        if (count[v] > maxF) atomicOr(err, PVQ_ERR_STACK);
    }
}

extern "C" int pac_gain_shape_quantize(pac_ctx* ctx, const double* d_x, int L, long long n_vec, const int* d_bit_alloc,
                                       int max_bit_alloc, unsigned long long* d_idx, int* d_bits, int* d_nfields,
                                       int max_fields, int* h_err, void* stream) {
    if (!ctx || L < 1 || L > PAC_PVQ_LMAX || max_fields < 1 || max_bit_alloc < 0) return PAC_ERR_ARG;
    if (n_vec <= 0) return 0;
    if (!d_x || !d_bit_alloc || !d_idx || !d_bits || !d_nfields) return PAC_ERR_ARG;
    PAC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int cap = pow2_cap(L);
    size_t smem = (size_t)cap * (3 * sizeof(double) + sizeof(int)) + ((size_t)max_bit_alloc / 32 + 8) * sizeof(unsigned);
    if (smem > 200 * 1024) return PAC_ERR_ARG;
    void* scr = nullptr;
    int rc = pac_scratch(ctx, 256, &scr);
    if (rc) return rc;
    PAC_CHECK(cudaMemsetAsync(scr, 0, sizeof(int), st));
    PAC_CHECK(cudaFuncSetAttribute(k_gs_quantize, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_gs_quantize<<<(unsigned)n_vec, 32, smem, st>>>(ctx->dev, d_x, L, cap, d_bit_alloc, d_idx, d_bits, d_nfields,
                                                     max_fields, (int*)scr);
    OPS_LAUNCH_CHECK();
    int err = 0;
    PAC_CHECK(cudaMemcpyAsync(&err, scr, sizeof(int), cudaMemcpyDeviceToHost, st));
    PAC_CHECK(cudaStreamSynchronize(st));
    if (h_err) *h_err = err;
    return err ? PAC_ERR_SOFT : 0;
}

__global__ void __launch_bounds__(32)
k_gs_dequantize(const PacDev* __restrict__ dev, const unsigned char* __restrict__ bytes, long long nBits,
                const long long* __restrict__ bitpos, const int* __restrict__ budget, int L, int cap,
                double* __restrict__ out, long long* __restrict__ posOut, int* __restrict__ err) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* wsA = reinterpret_cast<double*>(smem_raw);
    double* wsB = wsA + cap;
    double* res = wsB + cap;
    const long long v = blockIdx.x;
    const int lane = threadIdx.x;
    PvqTab pt = {dev->pvqN, dev->pvqRowOff, dev->log2tab, dev->pvqKW, nullptr};
    BitR br = {bytes, bitpos[v], nBits, 0};
    int werr = 0;
    warp_dequantize_gain_shape(pt, br, budget[v], L, wsA, wsB, cap, res, &werr);
    for (int i = lane; i < L; i += 32) out[v * L + i] = res[i];
    if (lane == 0) {
        posOut[v] = br.pos;
        if (br.pos > nBits) werr |= PVQ_ERR_STACK;
        if (werr) atomicOr(err, werr);
    }
}

extern "C" int pac_gain_shape_dequantize(pac_ctx* ctx, const unsigned char* d_bytes, long long n_bytes,
                                         const long long* d_bitpos, const int* d_bit_alloc, int L, long long n_vec,
                                         double* d_out, long long* d_bitpos_out, int* h_err, void* stream) {
    if (!ctx || L < 1 || L > PAC_PVQ_LMAX) return PAC_ERR_ARG;
    if (n_vec <= 0) return 0;
    if (!d_bytes || !d_bitpos || !d_bit_alloc || !d_out || !d_bitpos_out) return PAC_ERR_ARG;
    PAC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int cap = pow2_cap(L);
    size_t smem = (size_t)cap * 3 * sizeof(double);
    void* scr = nullptr;
    int rc = pac_scratch(ctx, 256, &scr);
    if (rc) return rc;
    PAC_CHECK(cudaMemsetAsync(scr, 0, sizeof(int), st));
    PAC_CHECK(cudaFuncSetAttribute(k_gs_dequantize, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_gs_dequantize<<<(unsigned)n_vec, 32, smem, st>>>(ctx->dev, d_bytes, n_bytes * 8, d_bitpos, d_bit_alloc, L, cap,
                                                       d_out, d_bitpos_out, (int*)scr);
    OPS_LAUNCH_CHECK();
    int err = 0;
    PAC_CHECK(cudaMemcpyAsync(&err, scr, sizeof(int), cudaMemcpyDeviceToHost, st));
    PAC_CHECK(cudaStreamSynchronize(st));
    if (h_err) *h_err = err;
    return err ? PAC_ERR_SOFT : 0;
}

// ---- no-allocation variant of pac_gain_shape_dequantize (no counterpart in the reference: added
// for callers that cannot call cudaMalloc - CUDA graph capture, tight serving loops).
// k_gs_dequantize keeps its three per-vector workspaces in dynamic shared memory and its soft-error word in the
// context's cudaMalloc'ed scratch; this variant keeps both in ONE file-scope device array of fixed size, so the
// entry point allocates nothing.  Every GPU that loads the library holds the array (4 MiB + 256 B) until exit.
// Layout: [0, 256) the soft-error word (cleared by every call), then n_vec x 3 x cap doubles of workspace
// (written before they are read: nothing else needs clearing).  Every call shares the array: calls that overlap
// on the GPU (different streams, threads or contexts of one device) race.
#define PAC_GS_STATIC_WS_BYTES (4u << 20)
#define PAC_STATIC_MAX_DEVICES 64
static __device__ __align__(256) unsigned char g_gs_static[256 + PAC_GS_STATIC_WS_BYTES];
// In host code `g_gs_static` is a host-side stand-in, not a device address: the device address of each GPU's copy
// is looked up once per device, with that device current, and that pointer is what the kernel receives.
static unsigned char* s_gs_static_ptr[PAC_STATIC_MAX_DEVICES];

int pac_ops_static_setup(int device) {          // pac_create calls this with `device` current
    if (device < 0 || device >= PAC_STATIC_MAX_DEVICES) return 0;      // such a device has no no-allocation path
    if (s_gs_static_ptr[device]) return 0;
    void* p = nullptr;
    PAC_CHECK(cudaGetSymbolAddress(&p, g_gs_static));
    s_gs_static_ptr[device] = (unsigned char*)p;
    return 0;
}

__global__ void __launch_bounds__(32)
k_gs_dequantize_static(const PacDev* __restrict__ dev, const unsigned char* __restrict__ bytes, long long nBits,
                       const long long* __restrict__ bitpos, const int* __restrict__ budget, int L, int cap,
                       double* __restrict__ out, long long* __restrict__ posOut, unsigned char* __restrict__ scratch) {
    int* err = reinterpret_cast<int*>(scratch);
    double* wsA = reinterpret_cast<double*>(scratch + 256) + (size_t)blockIdx.x * 3 * cap;
    double* wsB = wsA + cap;
    double* res = wsB + cap;
    const long long v = blockIdx.x;
    const int lane = threadIdx.x;
    PvqTab pt = {dev->pvqN, dev->pvqRowOff, dev->log2tab, dev->pvqKW, nullptr};
    BitR br = {bytes, bitpos[v], nBits, 0};
    int werr = 0;
    warp_dequantize_gain_shape(pt, br, budget[v], L, wsA, wsB, cap, res, &werr);
    __syncwarp();
    for (int i = lane; i < L; i += 32) out[v * L + i] = res[i];
    if (lane == 0) {
        posOut[v] = br.pos;
        if (br.pos > nBits) werr |= PVQ_ERR_STACK;
        if (werr) atomicOr(err, werr);
    }
}

extern "C" long long pac_gain_shape_dequantize_static_max_vectors(int L) {
    if (L < 1 || L > PAC_PVQ_LMAX) return 0;
    return (long long)(PAC_GS_STATIC_WS_BYTES / ((size_t)3 * pow2_cap(L) * sizeof(double)));
}

extern "C" int pac_gain_shape_dequantize_static(pac_ctx* ctx, const unsigned char* d_bytes, long long n_bytes,
                                                const long long* d_bitpos, const int* d_bit_alloc, int L, long long n_vec,
                                                double* d_out, long long* d_bitpos_out, int* h_err, void* stream) {
    if (!ctx || L < 1 || L > PAC_PVQ_LMAX) return PAC_ERR_ARG;
    if (n_vec <= 0) return 0;
    if (!d_bytes || !d_bitpos || !d_bit_alloc || !d_out || !d_bitpos_out) return PAC_ERR_ARG;
    if (n_vec > pac_gain_shape_dequantize_static_max_vectors(L)) return PAC_ERR_TOO_LARGE;
    if (ctx->device < 0 || ctx->device >= PAC_STATIC_MAX_DEVICES || !s_gs_static_ptr[ctx->device]) return PAC_ERR_UNSUPPORTED;
    PAC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* scr = s_gs_static_ptr[ctx->device];
    PAC_CHECK(cudaMemsetAsync(scr, 0, sizeof(int), st));
    k_gs_dequantize_static<<<(unsigned)n_vec, 32, 0, st>>>(ctx->dev, d_bytes, n_bytes * 8, d_bitpos, d_bit_alloc, L,
                                                           pow2_cap(L), d_out, d_bitpos_out, scr);
    OPS_LAUNCH_CHECK();
    int err = 0;
    PAC_CHECK(cudaMemcpyAsync(&err, scr, sizeof(int), cudaMemcpyDeviceToHost, st));
    PAC_CHECK(cudaStreamSynchronize(st));
    if (h_err) *h_err = err;
    return err ? PAC_ERR_SOFT : 0;
}

// ---------------------------------------------------------------- bitpack.PackedBits
// One warp per row: exclusive scan of the field widths, then every lane writes its
// fields MSB-first at their bit offsets (coder/bitpack.py:37-102).  Rows are
// rowBytes long (a multiple of 4), zero filled beyond the last field.
__global__ void __launch_bounds__(128)
k_pack_bits(const unsigned long long* __restrict__ val, const int* __restrict__ width, const int* __restrict__ count,
            int maxF, long long nRows, int rowBytes, unsigned char* __restrict__ out, int* __restrict__ err) {
    const int lane = threadIdx.x & 31;
    const long long row = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= nRows) return;
    unsigned* w = reinterpret_cast<unsigned*>(out + row * rowBytes);
    const int nW = rowBytes / 4;
    for (int i = lane; i < nW; i += 32) w[i] = 0u;
    __syncwarp();
    const int nf = min(count[row], maxF);
    long long base = 0;
    for (int f0 = 0; f0 < nf; f0 += 32) {
        const int f = f0 + lane;
        int wd = f < nf ? width[row * maxF + f] : 0;
        if (wd < 0) wd = 0;
        if (wd > 64) { atomicOr(err, PVQ_ERR_FIELD_TOO_WIDE); wd = 64; }
        long long inc = wd;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            long long t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        const long long pos = base + inc - wd;
        if (wd > 0) {
            if (pos + wd > 8LL * rowBytes) atomicOr(err, PVQ_ERR_STACK);
            else {
                unsigned long long v = val[row * maxF + f];
                if (wd < 64) v &= (1ull << wd) - 1ull;
                bw_put_at(w, (int)pos, v, wd);
            }
        }
        base += __shfl_sync(0xffffffffu, inc, 31);
    }
    __syncwarp();
    for (int i = lane; i < nW; i += 32) w[i] = __byte_perm(w[i], 0, 0x0123);   // big-endian words -> byte stream
}

extern "C" int pac_pack_bits(pac_ctx* ctx, const unsigned long long* d_val, const int* d_width, const int* d_count,
                             int max_fields, long long n_rows, int row_bytes, unsigned char* d_out, int* h_err,
                             void* stream) {
    if (!ctx || max_fields < 1 || row_bytes < 4 || (row_bytes & 3) || (long long)row_bytes * 8 > 0x7fffffffLL) return PAC_ERR_ARG;
    if (n_rows <= 0) return 0;
    if (!d_val || !d_width || !d_count || !d_out) return PAC_ERR_ARG;
    PAC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    void* scr = nullptr;
    int rc = pac_scratch(ctx, 256, &scr);
    if (rc) return rc;
    PAC_CHECK(cudaMemsetAsync(scr, 0, sizeof(int), st));
    k_pack_bits<<<(unsigned)((n_rows * 32 + 127) / 128), 128, 0, st>>>(d_val, d_width, d_count, max_fields, n_rows,
                                                                      row_bytes, d_out, (int*)scr);
    OPS_LAUNCH_CHECK();
    int err = 0;
    PAC_CHECK(cudaMemcpyAsync(&err, scr, sizeof(int), cudaMemcpyDeviceToHost, st));
    PAC_CHECK(cudaStreamSynchronize(st));
    if (h_err) *h_err = err;
    return err ? PAC_ERR_SOFT : 0;
}

// ReadBits for a list of fields per row (coder/bitpack.py:105-175): field f of a row
// starts at startBit + sum of the earlier widths.
__global__ void __launch_bounds__(128)
k_unpack_bits(const unsigned char* __restrict__ in, int rowBytes, const long long* __restrict__ startBit,
              const int* __restrict__ width, const int* __restrict__ count, int maxF, long long nRows,
              unsigned long long* __restrict__ val) {
    const int lane = threadIdx.x & 31;
    const long long row = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= nRows) return;
    const int nf = min(count[row], maxF);
    long long base = startBit ? startBit[row] : 0;
    for (int f0 = 0; f0 < nf; f0 += 32) {
        const int f = f0 + lane;
        int wd = f < nf ? width[row * maxF + f] : 0;
        wd = wd < 0 ? 0 : (wd > 64 ? 64 : wd);
        long long inc = wd;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            long long t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (f < nf) {
            BitR br = {in + row * rowBytes, base + inc - wd, 8LL * rowBytes, 0};
            val[row * maxF + f] = br_get(br, wd);
        }
        base += __shfl_sync(0xffffffffu, inc, 31);
    }
}

extern "C" int pac_unpack_bits(pac_ctx* ctx, const unsigned char* d_in, int row_bytes, const long long* d_start_bit,
                               const int* d_width, const int* d_count, int max_fields, long long n_rows,
                               unsigned long long* d_val, void* stream) {
    if (!ctx || max_fields < 1 || row_bytes < 1) return PAC_ERR_ARG;
    if (n_rows <= 0) return 0;
    if (!d_in || !d_width || !d_count || !d_val) return PAC_ERR_ARG;
    PAC_CHECK(cudaSetDevice(ctx->device));
    k_unpack_bits<<<(unsigned)((n_rows * 32 + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        d_in, row_bytes, d_start_bit, d_width, d_count, max_fields, n_rows, d_val);
    OPS_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------- codec.Encode* / PACFile.WriteDataBlock
// mode 0: a whole channel-block (2048 samples = prior || data) with its three
//         transient flags -> the packed record PACFile.WriteDataBlock would write
//         (coder/pacfile.py:449-610): 3 flag bits, one long frame or eight short ones.
// mode 1: one frame of N samples -> EncodeSingleChannel[_SBR]'s structured result
//         (coder/codec.py:266-380, :426-555): overall scale, bit allocation, BFP scale
//         factors, per-band field log; the frame's packed bits (no flag bits) go to rec.
struct FrameEncArgs {
    const double* x;          // [n][2048] (mode 0) or [n][N] (mode 1)
    const unsigned char* flags;   // [n] bit0 last, bit1 cur, bit2 next
    int mode, N, sbr;
    unsigned char* rec;       // [n][REC_STRIDE]
    int* recBits;             // [n] mode 0: declared bytes; mode 1: bits written
    unsigned char* zeroSub;   // [n] (mode 0)
    int* scale;               // [n]        (mode 1)
    int* ba;                  // [n][32]    (mode 1)
    int* sf;                  // [n][32]    (mode 1)
    int* used;                // [n][32]    (mode 1) bits emitted per band
    unsigned long long* logVal;   // [n][32][logCap] (mode 1, VQ)
    int* logW;
    int* logCnt;              // [n][32]
    int logCap;
    int* err;
};

template <typename T>
__global__ void __launch_bounds__(ENC_NT)
k_encode_frames(const PacDev* __restrict__ dev, FrameEncArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    EncSmem S;
    enc_smem_carve(smem_raw, S);
    const int tid = threadIdx.x;
    const long long f = blockIdx.x;
    const bool last = a.flags[f] & 1, cur = (a.flags[f] >> 1) & 1, nxt = (a.flags[f] >> 2) & 1;
    const int nIn = a.mode == 0 ? 2048 : a.N;
    T* xs = reinterpret_cast<T*>(S.xs);
    for (int i = tid; i < 2048; i += ENC_NT) xs[i] = i < nIn ? (T)a.x[f * nIn + i] : (T)0;
    for (int i = tid; i < REC_WORDS; i += ENC_NT) S.blockbuf[i] = 0u;
    if (a.mode == 1 && a.logVal) {
        S.logVal = a.logVal + (size_t)f * 32 * a.logCap;
        S.logW = a.logW + (size_t)f * 32 * a.logCap;
        S.logCnt = a.logCnt + f * 32;
        S.logCap = a.logCap;
        if (tid < 32) S.logCnt[tid] = 0;
    } else {
        S.logVal = nullptr; S.logW = nullptr; S.logCnt = nullptr; S.logCap = 0;
    }
    if (tid == 0) {
        S.misc[3] = 0;
        S.misc[4] = 0;
    }
    __syncthreads();
    int declared = 0;
    if (a.mode == 0) {
        if (tid == 0) {
            BitW bw = {S.blockbuf, 0};
            bw_put(bw, last, 1);
            bw_put(bw, cur, 1);
            bw_put(bw, nxt, 1);
            S.misc[3] = bw.pos;
        }
        __syncthreads();
        declared = 4;
        if (!cur) {
            int shape = (last && nxt) ? PAC_WIN_STARTSTOP : last ? PAC_WIN_STOP : nxt ? PAC_WIN_START : PAC_WIN_SINE;
            declared += encode_frame<PAC_NLONG, T>(dev, S, xs, shape, false, last || nxt, a.sbr != 0, &S.misc[3], a.err);
        } else {
            for (int sf = 0; sf < PAC_SUBFRAMES; ++sf) {
                bool nz = false;
                for (int i = tid; i < PAC_NSHORT; i += ENC_NT) nz |= xs[448 + 128 * sf + i] != (T)0;
                if (!__syncthreads_or(nz) && tid == 0) S.misc[4] = 1;
            }
            for (int sf = 0; sf < PAC_SUBFRAMES; ++sf)
                declared += encode_frame<PAC_NSHORT, T>(dev, S, xs + 448 + 128 * sf, PAC_WIN_SINE, true, last || nxt,
                                                     false, &S.misc[3], a.err);
        }
    } else if (a.N == PAC_NLONG) {
        // codec.getCorrectWindow (coder/codec.py:30-44); cur on a 2048 frame still selects the sine window
        int shape = cur ? PAC_WIN_SINE : (last && nxt) ? PAC_WIN_STARTSTOP : last ? PAC_WIN_STOP : nxt ? PAC_WIN_START : PAC_WIN_SINE;
        declared = encode_frame<PAC_NLONG, T>(dev, S, xs, shape, cur, last || nxt, a.sbr != 0, &S.misc[3], a.err);
    } else {
        declared = encode_frame<PAC_NSHORT, T>(dev, S, xs, PAC_WIN_SINE, cur, last || nxt, false, &S.misc[3], a.err);
    }
    __syncthreads();
    const int nBytes = (declared + 7) / 8;
    unsigned* out = reinterpret_cast<unsigned*>(a.rec + f * REC_STRIDE);
    const int nW = min(REC_WORDS, (max(nBytes, (S.misc[3] + 7) / 8) + 3) / 4);
    for (int i = tid; i < nW; i += ENC_NT) out[i] = __byte_perm(S.blockbuf[i], 0, 0x0123);
    if (a.mode == 1 && tid < 32) {
        a.ba[f * 32 + tid] = S.ba[tid];
        a.sf[f * 32 + tid] = S.sfb[tid];
        a.used[f * 32 + tid] = S.used[tid];
    }
    if (tid == 0) {
        a.recBits[f] = a.mode == 0 ? nBytes : S.misc[3];
        if (a.mode == 0) a.zeroSub[f] = (unsigned char)S.misc[4];
        else a.scale[f] = S.misc[5];
        if (nBytes > REC_STRIDE || (a.mode == 0 && S.misc[3] > declared)) atomicOr(a.err, PVQ_ERR_STACK);
    }
}

extern "C" int pac_record_stride(void) { return REC_STRIDE; }


static int run_encode_frames(pac_ctx* ctx, FrameEncArgs& a, long long n, int* h_err, cudaStream_t st) {
    void* scr = nullptr;
    int rc = pac_scratch(ctx, 256, &scr);
    if (rc) return rc;
    PAC_CHECK(cudaMemsetAsync(scr, 0, sizeof(int), st));
    a.err = (int*)scr;
    const size_t smem = enc_smem_bytes();
    if (ctx->cfg.precision == PAC_F32) {
        PAC_CHECK(cudaFuncSetAttribute(k_encode_frames<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_encode_frames<float><<<(unsigned)n, ENC_NT, smem, st>>>(ctx->dev, a);
    } else {
        PAC_CHECK(cudaFuncSetAttribute(k_encode_frames<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_encode_frames<double><<<(unsigned)n, ENC_NT, smem, st>>>(ctx->dev, a);
    }
    OPS_LAUNCH_CHECK();
    int err = 0;
    PAC_CHECK(cudaMemcpyAsync(&err, scr, sizeof(int), cudaMemcpyDeviceToHost, st));
    PAC_CHECK(cudaStreamSynchronize(st));
    if (h_err) *h_err = err;
    return err ? PAC_ERR_SOFT : 0;
}

extern "C" int pac_encode_blocks(pac_ctx* ctx, const double* d_frames, const unsigned char* d_flags, long long n,
                                 unsigned char* d_rec, int* d_rec_bytes, unsigned char* d_zero_sub, int* h_err,
                                 void* stream) {
    if (!ctx) return PAC_ERR_ARG;
    if (n <= 0) return 0;
    if (!d_frames || !d_flags || !d_rec || !d_rec_bytes || !d_zero_sub) return PAC_ERR_ARG;
    if (ctx->cfg.use_sbr && !ctx->cfg.use_vq) return PAC_ERR_UNSUPPORTED;
    PAC_CHECK(cudaSetDevice(ctx->device));
    FrameEncArgs a;
    memset(&a, 0, sizeof(a));
    a.x = d_frames; a.flags = d_flags; a.mode = 0; a.N = PAC_NLONG; a.sbr = ctx->cfg.use_sbr;
    a.rec = d_rec; a.recBits = d_rec_bytes; a.zeroSub = d_zero_sub;
    return run_encode_frames(ctx, a, n, h_err, (cudaStream_t)stream);
}

extern "C" int pac_encode_frames(pac_ctx* ctx, const double* d_data, int N, const unsigned char* d_flags, int sbr,
                                 long long n, int* d_scale, int* d_bit_alloc, int* d_scale_factor, int* d_band_bits,
                                 unsigned char* d_rec, int* d_rec_bits, unsigned long long* d_log_val, int* d_log_width,
                                 int* d_log_count, int log_cap, int* h_err, void* stream) {
    if (!ctx || (N != PAC_NLONG && N != PAC_NSHORT)) return PAC_ERR_ARG;
    if (n <= 0) return 0;
    if (!d_data || !d_flags || !d_scale || !d_bit_alloc || !d_scale_factor || !d_band_bits || !d_rec || !d_rec_bits)
        return PAC_ERR_ARG;
    if (sbr && (!ctx->cfg.use_vq || !ctx->cfg.use_sbr || N != PAC_NLONG)) return PAC_ERR_UNSUPPORTED;   // BFP+SBR crashes in the reference
    if (d_log_val && (!d_log_width || !d_log_count || log_cap < 1)) return PAC_ERR_ARG;
    PAC_CHECK(cudaSetDevice(ctx->device));
    FrameEncArgs a;
    memset(&a, 0, sizeof(a));
    a.x = d_data; a.flags = d_flags; a.mode = 1; a.N = N; a.sbr = sbr;
    a.rec = d_rec; a.recBits = d_rec_bits; a.scale = d_scale; a.ba = d_bit_alloc; a.sf = d_scale_factor;
    a.used = d_band_bits; a.logVal = d_log_val; a.logW = d_log_width; a.logCnt = d_log_count; a.logCap = log_cap;
    return run_encode_frames(ctx, a, n, h_err, (cudaStream_t)stream);
}

// ---------------------------------------------------------------- codec.Decode* / PACFile.ReadDataBlock
// mode 0: a channel-block record (flags + frames) -> 2048 windowed samples before the
//         block overlap-add (coder/pacfile.py:256-287).
// mode 1: one frame whose fields (overall scale, bit allocations, [scale factors,]
//         band payloads) start at bit 0 of the record -> N windowed samples
//         (coder/codec.py:47-92, :95-222); bit cursor after the frame in bitsOut.
struct FrameDecArgs {
    const unsigned char* rec;   // [n][stride]
    const int* recBytes;        // [n]
    int stride, mode, N;
    const unsigned char* flags; // [n] (mode 1)
    double* out;                // [n][2048] or [n][N]
    int* bitsOut;               // [n]
    int* err;
};

template <typename T>
__global__ void __launch_bounds__(DEC_NT)
k_decode_frames(const PacDev* __restrict__ dev, FrameDecArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    DecSmem S;
    dec_smem_carve(smem_raw, S);
    const int tid = threadIdx.x;
    const long long f = blockIdx.x;
    int nBytes = a.recBytes[f];
    if (nBytes > DEC_REC_MAX || nBytes > a.stride) {
        if (tid == 0) atomicOr(a.err, PVQ_ERR_STACK);
        nBytes = min(DEC_REC_MAX, a.stride);
    }
    const unsigned char* src = a.rec + f * a.stride;
    for (int i = tid; i < nBytes; i += DEC_NT) S.rec[i] = src[i];
    for (int i = tid; i < 2048; i += DEC_NT) S.y[i] = 0.0;
    __syncthreads();
    const long long recBits = 8LL * nBytes;
    int nOut = 2048;
    if (a.mode == 0) {
        const bool last = (S.rec[0] >> 7) & 1, cur = (S.rec[0] >> 6) & 1, nxt = (S.rec[0] >> 5) & 1;
        if (tid == 0) S.misc[3] = 3;
        __syncthreads();
        if (nBytes > 0) {
            if (!cur) {
                int shape = (last && nxt) ? PAC_WIN_STARTSTOP : last ? PAC_WIN_STOP : nxt ? PAC_WIN_START : PAC_WIN_SINE;
                decode_frame<PAC_NLONG, T>(dev, S, recBits, shape, 0, a.err);
            } else {
                for (int sf = 0; sf < PAC_SUBFRAMES; ++sf)
                    decode_frame<PAC_NSHORT, T>(dev, S, recBits, PAC_WIN_SINE, 448 + 128 * sf, a.err);
            }
        }
    } else {
        const bool last = a.flags[f] & 1, cur = (a.flags[f] >> 1) & 1, nxt = (a.flags[f] >> 2) & 1;
        if (tid == 0) S.misc[3] = 0;
        __syncthreads();
        nOut = a.N;
        if (a.N == PAC_NLONG) {
            int shape = cur ? PAC_WIN_SINE : (last && nxt) ? PAC_WIN_STARTSTOP : last ? PAC_WIN_STOP : nxt ? PAC_WIN_START : PAC_WIN_SINE;
            decode_frame<PAC_NLONG, T>(dev, S, recBits, shape, 0, a.err);
        } else {
            decode_frame<PAC_NSHORT, T>(dev, S, recBits, PAC_WIN_SINE, 0, a.err);
        }
    }
    __syncthreads();
    double* dst = a.out + f * nOut;
    for (int i = tid; i < nOut; i += DEC_NT) dst[i] = S.y[i];
    if (tid == 0 && a.bitsOut) a.bitsOut[f] = S.misc[3];
}

static int run_decode_frames(pac_ctx* ctx, FrameDecArgs& a, long long n, int* h_err, cudaStream_t st) {
    void* scr = nullptr;
    int rc = pac_scratch(ctx, 256, &scr);
    if (rc) return rc;
    PAC_CHECK(cudaMemsetAsync(scr, 0, sizeof(int), st));
    a.err = (int*)scr;
    size_t smem = dec_smem_bytes();
    PAC_CHECK(cudaFuncSetAttribute(k_decode_frames<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    PAC_CHECK(cudaFuncSetAttribute(k_decode_frames<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (ctx->cfg.precision == PAC_F32) k_decode_frames<float><<<(unsigned)n, DEC_NT, smem, st>>>(ctx->dev, a);
    else k_decode_frames<double><<<(unsigned)n, DEC_NT, smem, st>>>(ctx->dev, a);
    OPS_LAUNCH_CHECK();
    int err = 0;
    PAC_CHECK(cudaMemcpyAsync(&err, scr, sizeof(int), cudaMemcpyDeviceToHost, st));
    PAC_CHECK(cudaStreamSynchronize(st));
    if (h_err) *h_err = err;
    return err ? PAC_ERR_SOFT : 0;
}

extern "C" int pac_decode_blocks(pac_ctx* ctx, const unsigned char* d_rec, int rec_stride, const int* d_rec_bytes,
                                 long long n, double* d_out, int* h_err, void* stream) {
    if (!ctx || rec_stride < 1) return PAC_ERR_ARG;
    if (n <= 0) return 0;
    if (!d_rec || !d_rec_bytes || !d_out) return PAC_ERR_ARG;
    if (ctx->cfg.use_sbr && !ctx->cfg.use_vq) return PAC_ERR_UNSUPPORTED;
    PAC_CHECK(cudaSetDevice(ctx->device));
    FrameDecArgs a;
    memset(&a, 0, sizeof(a));
    a.rec = d_rec; a.recBytes = d_rec_bytes; a.stride = rec_stride; a.mode = 0; a.N = PAC_NLONG; a.out = d_out;
    return run_decode_frames(ctx, a, n, h_err, (cudaStream_t)stream);
}

extern "C" int pac_decode_frames(pac_ctx* ctx, const unsigned char* d_rec, int rec_stride, const int* d_rec_bytes,
                                 int N, const unsigned char* d_flags, long long n, double* d_out, int* d_bits_consumed,
                                 int* h_err, void* stream) {
    if (!ctx || rec_stride < 1 || (N != PAC_NLONG && N != PAC_NSHORT)) return PAC_ERR_ARG;
    if (n <= 0) return 0;
    if (!d_rec || !d_rec_bytes || !d_flags || !d_out) return PAC_ERR_ARG;
    if (ctx->cfg.use_sbr && !ctx->cfg.use_vq) return PAC_ERR_UNSUPPORTED;
    PAC_CHECK(cudaSetDevice(ctx->device));
    FrameDecArgs a;
    memset(&a, 0, sizeof(a));
    a.rec = d_rec; a.recBytes = d_rec_bytes; a.stride = rec_stride; a.mode = 1; a.N = N; a.flags = d_flags;
    a.out = d_out; a.bitsOut = d_bits_consumed;
    return run_decode_frames(ctx, a, n, h_err, (cudaStream_t)stream);
}
