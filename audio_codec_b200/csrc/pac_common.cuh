// Shared definitions for the B200 perceptual-coder kernels (sm_100a).
//
// One context per GPU holds every read-only table the kernels need (windows,
// twiddles, band maps, Bark / threshold-in-quiet per MDCT line, PVQ codebook
// sizes).  All tables are computed on the host in pac_api.cu with long-double
// libm and uploaded once; kernels take a `const PacDev*` that lives in device
// memory.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define PAC_MAX_BANDS 32
#define PAC_NLONG 2048
#define PAC_NSHORT 256
#define PAC_HALF_LONG 1024
#define PAC_HALF_SHORT 128
#define PAC_SUBFRAMES 8
#define PAC_PVQ_LMAX 1024
#define PAC_THETA_BITS 16
#define PAC_EPS 2.220446049250313e-16   // np.finfo(float).eps

// window shapes (coder/window.py:14-92, selected by coder/codec.py:30-44)
enum { PAC_WIN_SINE = 0, PAC_WIN_START = 1, PAC_WIN_STOP = 2, PAC_WIN_STARTSTOP = 3,
       PAC_WIN_KAISER = 4, PAC_WIN_NONE = 5, PAC_WIN_COUNT = 5 };

struct PacBands {
    int nBands;
    int nLines[PAC_MAX_BANDS];
    int lower[PAC_MAX_BANDS];
    int omitted[PAC_MAX_BANDS];   // 1 if reconstructed by SBR (long bands only)
};

template <typename T> struct PacTables {
    const T* win_long;     // [PAC_WIN_COUNT][2048]
    const T* win_short;    // [PAC_WIN_COUNT][256]   (sine, kaiser used)
    const T* hann_long;    // [2048]  0.5(1-cos(2pi(n+.5)/N))
    const T* hann_short;   // [256]
    const T* tw_mdct_long;   // [512]  complex exp(-j pi (8n+1)/(8M)), interleaved re,im
    const T* tw_mdct_short;  // [64]
    const T* tw_fft[4];      // roots exp(-2 pi j k/n) for n = 64,128,512,1024 (k < n), interleaved
    const T* rtw_long;       // [1024+1] exp(-2 pi j k/2048) for the real-FFT split
    const T* rtw_short;      // [128+1]
};

struct PacDev {
    int fs;
    int nChannels;
    int nScaleBits;
    int nMantSizeBits;
    int useSBR;
    int useVQ;
    int maxMantBits;
    double targetBitsPerSample;
    PacBands longBands;
    PacBands shortBands;
    int sbrCutoff;            // first omitted line (long)
    PacTables<double> d;
    PacTables<float> f;
    // psychoacoustic per-line constants (fp64), long then short
    const double* barkLong;    // [1024] Bark((k+.5) fs/2048)
    const double* barkShort;   // [128]
    const double* threshLong;  // [1024] threshold in quiet, dB SPL
    const double* threshShort; // [128]
    const double* fftFreqLong; // [1025] rfftfreq
    const double* fftFreqShort;// [129]
    const unsigned char* lineBandLong;   // [1024] band of each MDCT line
    const unsigned char* lineBandShort;  // [128]
    double fftNormLong;        // 4/(N^2 mean(hanning(N)^2))
    double fftNormShort;
    double fftValLong;         // rfftfreq spacing 1 / (N (1 / fs)): fftFreq[k] = k * val (the same product the table holds)
    double fftValShort;
    // PVQ codebook sizes N(L,K), rows L = 0..PAC_PVQ_LMAX, each row up to the first
    // entry exceeding 2^32 (inclusive); rows 0..3 are closed-form and empty here.
    const unsigned long long* pvqN;
    const int* pvqRowOff;      // [PAC_PVQ_LMAX+2]
    const double* log2tab;     // [PAC_PVQ_LMAX+1] log2(L) (entry 0 unused)
    // pvq_compute_k_for_R as a table: [(PAC_PVQ_LMAX+1)][33] of {K, field width} for bits = 0..32
    const int2* pvqKW;
    // quantised mid/side angle: [(1 << aT) + code][4] = {theta, log2(tan|theta| + eps), cos, sin} for
    // aT = 1..PAC_THETA_BITS, filled on the device at create time with the same device functions the
    // direct evaluation uses (bit-identical), so the per-node fp64 tan / log2 / sin / cos become loads
    const double* thetaTab;
    // SBR decode: response of gaussian_filter1d(sigma=200, reflect) to unit impulses
    const double* sbrGauss;    // [nOmitted][2*(1024-cutoff)] (only first half used)
    int sbrNOmitted;
};

static __device__ __forceinline__ int pac_fft_index(int n) {
    return n == 64 ? 0 : n == 128 ? 1 : n == 512 ? 2 : 3;
}

#define PAC_CHECK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return -(int)e_ - 1000; } while (0)

// fp64 transcendentals, one out-of-line copy each: CUDA's double-precision log10 / atan / pow ...
// expand to 1-3 KB of SASS per call site and the frame kernels are instruction-cache bound.
// Same code as the inlined form, so results are bit-identical.  The float overloads stay inline.
#ifdef __CUDACC__
static __device__ __noinline__ double pm_log10(double x) { return log10(x); }
static __device__ __noinline__ double pm_log2(double x) { return log2(x); }
static __device__ __noinline__ double pm_log(double x) { return log(x); }
static __device__ __noinline__ double pm_atan(double x) { return atan(x); }
static __device__ __noinline__ double pm_tan(double x) { return tan(x); }
static __device__ __noinline__ double pm_pow(double x, double y) { return pow(x, y); }
static __device__ __noinline__ void pm_sincos(double x, double* s, double* c) { *s = sin(x); *c = cos(x); }
// x / d for many x and one divisor: the correctly rounded reciprocal y = RN(1 / d) once (__drcp_rn), then per
// quotient q0 = RN(x y), r = x - q0 d (exact, one FMA), q = RN(q0 + r y).  By Markstein's theorem q = RN(x / d), the
// IEEE quotient, bit for bit (no over / underflow: here d is a vector norm, 0 or >= 2^-537, and |x| <= ~d) - three
// instructions per quotient instead of the ~11 of a full division.  The sign of a zero quotient is not preserved
// (-0 / d gives +0): nothing on the path can observe it.  tests/test_ops_gpu.py checks 2^27 random pairs.
struct PacDiv { double d, y; };
static __device__ __forceinline__ PacDiv pac_div_by(double d) { PacDiv b = {d, __drcp_rn(d)}; return b; }
static __device__ __forceinline__ double pac_div(double x, const PacDiv& b) {
    const double q0 = __dmul_rn(x, b.y);
    const double r = __fma_rn(-q0, b.d, x);
    return __fma_rn(r, b.y, q0);
}
static __device__ __forceinline__ float pm_log10(float x) { return log10f(x); }
static __device__ __forceinline__ float pm_atan(float x) { return atanf(x); }
static __device__ __forceinline__ float pm_pow(float x, float y) { return powf(x, y); }
#endif
