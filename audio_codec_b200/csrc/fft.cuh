// CTA-cooperative complex FFT in shared memory (Stockham autosort, radix 8/4/2).
//
// B independent N-point transforms sit back to back in `buf`; NT threads of
// the CTA cooperate (every thread of the CTA must call, the passes contain
// __syncthreads()).  Each pass keeps its butterflies in registers between the
// read and the write so one buffer suffices.  Roots of unity come from a
// read-only table exp(-2 pi j k/N), k < N.
#pragma once
#include "pac_common.cuh"

template <typename T> struct cplx { T x, y; };

template <typename T> __device__ __forceinline__ cplx<T> cmul(cplx<T> a, cplx<T> b) {
    cplx<T> r;
    r.x = a.x * b.x - a.y * b.y;
    r.y = a.x * b.y + a.y * b.x;
    return r;
}
template <typename T> __device__ __forceinline__ cplx<T> cadd(cplx<T> a, cplx<T> b) { return {a.x + b.x, a.y + b.y}; }
template <typename T> __device__ __forceinline__ cplx<T> csub(cplx<T> a, cplx<T> b) { return {a.x - b.x, a.y - b.y}; }
// fp32: one packed instruction per complex add / subtract (sm_100 f32x2 arithmetic: add.rn.f32x2, and a - b as
// fma.rn.f32x2(b, -1, a): the product is exact, so both are the IEEE results of the scalar forms).  The streaming MDCT
// kernels are bound by instruction issue, and most of an FFT's instructions are these.
#ifndef PAC_F32X2
#define PAC_F32X2 1
#endif
#if PAC_F32X2
template <> __device__ __forceinline__ cplx<float> cadd<float>(cplx<float> a, cplx<float> b) {
    const float2 r = __fadd2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y));
    return {r.x, r.y};
}
template <> __device__ __forceinline__ cplx<float> csub<float>(cplx<float> a, cplx<float> b) {
    const float2 r = __ffma2_rn(make_float2(b.x, b.y), make_float2(-1.0f, -1.0f), make_float2(a.x, a.y));
    return {r.x, r.y};
}
#endif
// multiply by -j
template <typename T> __device__ __forceinline__ cplx<T> cmnj(cplx<T> a) { return {a.y, -a.x}; }

template <typename T> __device__ __forceinline__ void bfly2(cplx<T>* u) {
    cplx<T> a = u[0], b = u[1];
    u[0] = cadd(a, b);
    u[1] = csub(a, b);
}

template <typename T> __device__ __forceinline__ void bfly4(cplx<T>* u) {
    cplx<T> a = cadd(u[0], u[2]), b = csub(u[0], u[2]);
    cplx<T> c = cadd(u[1], u[3]), d = cmnj(csub(u[1], u[3]));
    u[0] = cadd(a, c);
    u[1] = cadd(b, d);
    u[2] = csub(a, c);
    u[3] = csub(b, d);
}

template <typename T> __device__ __forceinline__ void bfly8(cplx<T>* u) {
    const T h = (T)0.70710678118654752440;
    cplx<T> e[4] = {u[0], u[2], u[4], u[6]};
    cplx<T> o[4] = {u[1], u[3], u[5], u[7]};
    bfly4(e);
    bfly4(o);
    // o[r] *= exp(-2 pi j r/8)
    cplx<T> o1 = {h * (o[1].x + o[1].y), h * (o[1].y - o[1].x)};
    cplx<T> o2 = cmnj(o[2]);
    cplx<T> o3 = {h * (o[3].y - o[3].x), -h * (o[3].x + o[3].y)};
    u[0] = cadd(e[0], o[0]); u[4] = csub(e[0], o[0]);
    u[1] = cadd(e[1], o1);   u[5] = csub(e[1], o1);
    u[2] = cadd(e[2], o2);   u[6] = csub(e[2], o2);
    u[3] = cadd(e[3], o3);   u[7] = csub(e[3], o3);
}

template <typename T, int R> __device__ __forceinline__ void bfly(cplx<T>* u) {
    if (R == 2) bfly2(u);
    else if (R == 4) bfly4(u);
    else bfly8(u);
}

// exp(-2 pi j m / 1024) from its first quadrant q[0..256) (the tables are built symmetric, pac_api.cu: the value is the
// full table's entry bit for bit).  The front end keeps this quarter in shared memory: the full tables (16 KB + 8 KB,
// strided access) do not stay in the 28 KB of L1 its six CTAs leave, and every pass waited on L2 for its twiddles.
template <typename T> __device__ __forceinline__ cplx<T> tw_quarter(const cplx<T>* q, int m) {
    const cplx<T> b = q[m & 255];
    const int qd = m >> 8;                                 // times (-j)^qd
    const T x = (qd & 1) ? b.y : b.x, y = (qd & 1) ? b.x : b.y;
    cplx<T> r;
    r.x = (qd == 2 || qd == 3) ? -x : x;
    r.y = (qd == 1 || qd == 2) ? -y : y;
    return r;
}

// One Stockham pass of radix R at sub-transform size Ns (product of earlier radices).
// QSH >= 0: `roots` is the quarter table of the 1024-point roots and an index of this transform is worth 1 << QSH there.
template <typename T, int N, int R, int B, int NT, int QSH = -1>
__device__ __forceinline__ void fft_pass(cplx<T>* buf, const cplx<T>* __restrict__ roots, int Ns, int tid) {
    constexpr int J = N / R;
    constexpr int TOT = B * J;
    constexpr int IT = (TOT + NT - 1) / NT;
    cplx<T> u[IT][R];
#pragma unroll
    for (int it = 0; it < IT; ++it) {
        int idx = tid + it * NT;
        if (idx < TOT) {
            int b = idx / J, j = idx - b * J;
            int k = j & (Ns - 1);
            const cplx<T>* p = buf + b * N + j;
            int step = k * (N / (Ns * R));
#pragma unroll
            for (int r = 0; r < R; ++r) {
                cplx<T> v = p[r * J];
                if (r > 0 && Ns > 1) v = cmul(v, QSH >= 0 ? tw_quarter<T>(roots, (r * step) << (QSH >= 0 ? QSH : 0)) : roots[r * step]);
                u[it][r] = v;
            }
            bfly<T, R>(u[it]);
        }
    }
    __syncthreads();
#pragma unroll
    for (int it = 0; it < IT; ++it) {
        int idx = tid + it * NT;
        if (idx < TOT) {
            int b = idx / J, j = idx - b * J;
            int k = j & (Ns - 1);
            cplx<T>* p = buf + b * N + (j - k) * R + k;
#pragma unroll
            for (int r = 0; r < R; ++r) p[r * Ns] = u[it][r];
        }
    }
    __syncthreads();
}

// Full transform.  Radix plan: as many radix-8 passes as fit, then 4, then 2.
template <typename T, int N, int B, int NT, int QSH = -1>
__device__ __forceinline__ void cta_fft(cplx<T>* buf, const cplx<T>* __restrict__ roots, int tid) {
    int Ns = 1;
    constexpr int L2 = (N == 64) ? 6 : (N == 128) ? 7 : (N == 512) ? 9 : 10;
    constexpr int N8 = L2 / 3;
    constexpr int REM = L2 - 3 * N8;
#pragma unroll
    for (int p = 0; p < N8; ++p) {
        fft_pass<T, N, 8, B, NT, QSH>(buf, roots, Ns, tid);
        Ns *= 8;
    }
    if (REM == 2) fft_pass<T, N, 4, B, NT, QSH>(buf, roots, Ns, tid);
    if (REM == 1) fft_pass<T, N, 2, B, NT, QSH>(buf, roots, Ns, tid);
}
