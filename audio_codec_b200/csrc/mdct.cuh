// Fused window + MDCT and IMDCT + window device routines (one or several
// frames per CTA) built on an N/4-point complex FFT:
//   forward  X[k] = (2/N) sum_n w[n] x[n] cos(2pi/N (n+n0)(k+1/2)),  n0 = N/4+1/2
//   inverse  y[n] = 2 w[n] sum_k X[k] cos(2pi/N (n+n0)(k+1/2))
// which is what coder/mdct.py:43-76 computes with an N-point FFT (and
// coder/codec.py:303-305 / :88-89 apply the window around it).
// With M = N/2, Q = N/4: fold x to v[M] (TDAC), DCT-IV of v through
//   t[n] = (v[2n] + j v[M-1-2n]) e^{-j pi (8n+1)/(8M)},  T = FFT_Q(t),
//   y[k] = T[k] e^{-j pi (8k+1)/(8M)},  X[2k] = Re y[k],  X[M-1-2k] = -Im y[k].
#pragma once
#include "fft.cuh"

// Fold + window + pre-twiddle: B frames whose samples start at xs + b*hop (smem
// or global), into fft buffer [B][Q].  winTab is [PAC_WIN_COUNT][N] (nullptr =
// no window); frame b uses shape shapes[b] (shapes == nullptr: shape0 for all).
template <typename T, int N, int B, int NT>
__device__ __forceinline__ void mdct_fold(const T* xs, int hop, const T* __restrict__ winTab, const int* shapes,
                                          int shape0, const cplx<T>* __restrict__ tw, cplx<T>* fbuf, int tid) {
    constexpr int M = N / 2, Q = N / 4;
    for (int idx = tid; idx < B * Q; idx += NT) {
        int b = idx / Q, n = idx - b * Q;
        const T* x = xs + b * hop;
        const T* w = winTab ? winTab + (shapes ? shapes[b] : shape0) * N : nullptr;
        int i0, i1, i2, i3;
        T re, im;
        if (2 * n < Q) {
            i0 = 3 * Q - 1 - 2 * n; i1 = 3 * Q + 2 * n;     // v[2n]     = -x[i0] - x[i1]
            i2 = Q - 1 - 2 * n;     i3 = Q + 2 * n;         // v[M-1-2n] =  x[i2] - x[i3]
            T a0 = x[i0], a1 = x[i1], a2 = x[i2], a3 = x[i3];
            if (w) { a0 *= w[i0]; a1 *= w[i1]; a2 *= w[i2]; a3 *= w[i3]; }
            re = -a0 - a1;
            im = a2 - a3;
        } else {
            i0 = 2 * n - Q;         i1 = 3 * Q - 1 - 2 * n; // v[2n]     =  x[i0] - x[i1]
            i2 = Q + 2 * n;         i3 = 5 * Q - 1 - 2 * n; // v[M-1-2n] = -x[i2] - x[i3]
            T a0 = x[i0], a1 = x[i1], a2 = x[i2], a3 = x[i3];
            if (w) { a0 *= w[i0]; a1 *= w[i1]; a2 *= w[i2]; a3 *= w[i3]; }
            re = a0 - a1;
            im = -a2 - a3;
        }
        cplx<T> t = {re, im};
        fbuf[idx] = cmul(t, tw[n]);
    }
    (void)M;
}

// Post-twiddle of the forward transform: fbuf [B][Q] -> X [B][M] (natural order), scaled.
template <typename T, int N, int B, int NT, typename TO = T>
__device__ __forceinline__ void mdct_post(const cplx<T>* fbuf, const cplx<T>* __restrict__ tw, T scale, TO* X,
                                          int tid) {
    constexpr int M = N / 2, Q = N / 4;
    for (int idx = tid; idx < B * Q; idx += NT) {
        int b = idx / Q, k = idx - b * Q;
        cplx<T> y = cmul(fbuf[idx], tw[k]);
        X[b * M + 2 * k] = scale * y.x;
        X[b * M + M - 1 - 2 * k] = -scale * y.y;
    }
}

// Pre-twiddle for the inverse (DCT-IV is its own inverse up to scale).
template <typename T, int N, int B, int NT, typename TI = T>
__device__ __forceinline__ void imdct_pre(const TI* X, TI inScale, const cplx<T>* __restrict__ tw, cplx<T>* fbuf,
                                          int tid) {
    constexpr int M = N / 2, Q = N / 4;
    for (int idx = tid; idx < B * Q; idx += NT) {
        int b = idx / Q, n = idx - b * Q;
        cplx<T> t = {(T)(X[b * M + 2 * n] * inScale), (T)(X[b * M + M - 1 - 2 * n] * inScale)};
        fbuf[idx] = cmul(t, tw[n]);
    }
}

// Unfolded, un-windowed IMDCT sample i of a frame whose DCT-IV output is v[M]:
//   y = 2 [ v2, -rev(v2), -rev(v1), -v1 ],  v1 = v[0:Q], v2 = v[Q:M]
template <typename T, int N> __device__ __forceinline__ T imdct_sample(const T* v, int i) {
    constexpr int M = N / 2, Q = N / 4;
    if (i < Q) return (T)2 * v[Q + i];
    if (i < M + Q) return (T)-2 * v[3 * Q - 1 - i];
    return (T)-2 * v[i - 3 * Q];
}
