// Back end of the batched stream decoder for LONG blocks: scaled lines -> IMDCT -> synthesis window -> the block's
// 2048 windowed samples (coder/codec.py:85-89, coder/mdct.py:62-76, coder/window.py:14-92).
//
// k_decode<T, true> gives a channel-block a 256-thread CTA and runs the 512-point FFT in shared memory with a CTA
// barrier per stage: 27 k cycles per block, 8.4 ms per step, although the work is the streaming IMDCT's
// (mdct_stream.cuh: 1.5 ms for the same number of frames).  Here a channel-block belongs to ONE WARP:
//   * the warp gathers the block's column of the lane-interleaved line array k_dequant wrote (8 bytes out of every
//     256-byte row; the kernel walks the batch in its dense order, so the 32 blocks of a row are read by
//     neighbouring warps in the same pass and HBM sees each row once), scales and pre-twiddles in registers;
//   * warp_fft512 (16 x 16 x 2 in registers, one skewed shared-memory exchange, __syncwarp only);
//   * post-twiddle into the warp's staging tile, then the unfolded samples times the window of the block's shape
//     (read through L1 from the 5-shape table: almost every block uses the sine shape) leave as 16-byte stores.
// No CTA barrier after the tables are loaded; 16 warps per SM in both precisions.
// Blocks the kernel does not take - transient blocks (8 short frames) and long blocks with SBR reconstruction - are
// left to k_decode<T, true> (k_hdr lists them), as is the zero fill of a block without a record.
#pragma once
#include "mdct_stream.cuh"
#include "decode_frame.cuh"

#define DL_WARPS 16

template <typename T> __host__ __device__ constexpr size_t dl_smem_bytes() {
    // twiddles + second-pass roots + roots (1024 complex) + per warp the exchange / staging tile
    return sizeof(T) * (2 * 1024 + (size_t)DL_WARPS * 2 * MS_TILE);
}
// true: k_decode_long writes this block's frame; false: k_decode<T, true> does
__device__ __forceinline__ bool dl_takes(unsigned char flags, unsigned char sbr) {
    return (flags & 0x80u) && !((flags >> 1) & 1) && !sbr;     // a record, long, no SBR reconstruction
}

template <typename T>
__global__ void __launch_bounds__(DL_WARPS * 32, 1)
k_decode_long(const PacDev* __restrict__ dev, double* __restrict__ frames, const CbHdr* __restrict__ hdrs,
              const double* __restrict__ lines, const int* __restrict__ perm, const int* __restrict__ counts) {
    constexpr int M = 1024, Q = 512, W = DL_WARPS;
    constexpr int NV = MsBlock<T>::NV, PER = MsBlock<T>::PER, L = MsRun<T>::L;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const PacTables<T>& t = psy_tabs<T>(dev);
    MsTables<T> tb;
    tb.win = nullptr;
    tb.tw = reinterpret_cast<cplx<T>*>(smem_raw);
    tb.r1 = tb.tw + Q;
    tb.roots = tb.r1 + 256;
    {
        const cplx<T>* twG = reinterpret_cast<const cplx<T>*>(t.tw_mdct_long);
        const cplx<T>* rootsG = reinterpret_cast<const cplx<T>*>(t.tw_fft[2]);
        for (int i = threadIdx.x; i < Q; i += W * 32) tb.tw[i] = twG[i];
        for (int i = threadIdx.x; i < 256; i += W * 32) {
            tb.roots[i] = rootsG[i];
            tb.r1[i] = rootsG[(2 * (i >> 4) * (i & 15)) & (Q - 1)];
        }
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const T* winAll = t.win_long;
    const int kl = (lane & 15) + 256 * (lane >> 4), kE = kl + 16 * (lane >> 4), kO = (Q - 1 - kl) + 16 * (1 - (lane >> 4));
    const int ll = lane * (PER / 2);
    T* Ev = reinterpret_cast<T*>(tb.roots + 256) + (size_t)warp * 2 * MS_TILE;     // DCT-IV output, even indices (skewed)
    T* Ov = Ev + MS_TILE;                                                            // odd indices
    // in the batch's DENSE order (k_hdr: long blocks from the front): the 32 blocks of a 256-byte row of the line array
    // are then read by 32 neighbouring warps in the same pass, whatever the blocks' places in the streams (in stream
    // order a row's blocks can be far apart once transient blocks have shifted the dense order, and HBM delivered
    // 2.4 x the algorithmic bytes)
    const int nLong = counts[0];
    for (int p = blockIdx.x * W + warp; p < nLong; p += gridDim.x * W) {
        const int cb = perm[p];
        const CbHdr* h = hdrs + cb;
        const unsigned char fl = h->flags;
        if (!dl_takes(fl, h->sbr)) continue;
        double* dst = frames + (size_t)cb * 2048;
        const bool last = (fl >> 2) & 1, nxt = fl & 1;
        const int shape = (last && nxt) ? PAC_WIN_STARTSTOP : last ? PAC_WIN_STOP : nxt ? PAC_WIN_START : PAC_WIN_SINE;
        const double inv = 1.0 / (double)(1 << h->scale[0]);
        const double* src = lines + ((size_t)(p >> 5) * 1024) * 32 + (size_t)(p & 31);
        cplx<T> u[16];
#pragma unroll
        for (int it = 0; it < 16; ++it) {
            const int n = lane + 32 * it;
            const cplx<T> x = {(T)(src[(size_t)(2 * n) * 32] * inv), (T)(src[(size_t)(M - 1 - 2 * n) * 32] * inv)};
            u[it] = cmul(x, MsRun<T>::ldc(tb.tw + n));
        }
        warp_fft512<T>(u, Ev, tb, lane);
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            const int k = (lane & 15) + 16 * r + 256 * (lane >> 4);
            const cplx<T> v = cmul(u[r], MsRun<T>::ldc(tb.tw + k));
            Ev[kE + 16 * r] = v.x;                 // xv[2k] at MS_OSK(k): k >> 8 = lane >> 4
            Ov[kO - 16 * r] = -v.y;                // xv[M - 1 - 2k] at MS_OSK(Q - 1 - k): (Q - 1 - k) >> 8 = 1 - (lane >> 4)
        }
        __syncwarp();
        const T* win = winAll + (size_t)shape * 2048;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int n0 = (i * 32 + lane) * PER;
            T o[PER], tl[PER], ev[L], ov[L];
            if (i * 32 * PER < Q) {       // (n0 < Q, the same for every lane) head = 2 xv[Q + n], tail = -2 xv[Q - 1 - n]
                const int A = Q / 2 + i * 16 * PER, C = (Q - PER) / 2 - i * 16 * PER;      // a = (Q + n0) / 2 = A + ll, c = (Q - n0 - PER) / 2 = C - ll
                const int a = A + 16 * (A >> 8) + ll, c = C + 16 * (C >> 8) - ll;          // MS_OSK: the skew is the same for all lanes
                MsRun<T>::ld(Ev + a, ev);
                MsRun<T>::ld(Ov + a, ov);
#pragma unroll
                for (int q = 0; q < L; ++q) { o[2 * q] = (T)2 * ev[q]; o[2 * q + 1] = (T)2 * ov[q]; }
                MsRun<T>::ld(Ev + c, ev);
                MsRun<T>::ld(Ov + c, ov);
#pragma unroll
                for (int q = 0; q < L; ++q) { tl[PER - 1 - 2 * q] = (T)-2 * ev[q]; tl[PER - 2 - 2 * q] = (T)-2 * ov[q]; }
            } else {            // head = -2 xv[3Q - 1 - n], tail = -2 xv[n - Q]
                const int C = (3 * Q - PER) / 2 - i * 16 * PER, A = i * 16 * PER - Q / 2;
                const int c = C + 16 * (C >> 8) - ll, a = A + 16 * (A >> 8) + ll;
                MsRun<T>::ld(Ev + c, ev);
                MsRun<T>::ld(Ov + c, ov);
#pragma unroll
                for (int q = 0; q < L; ++q) { o[PER - 1 - 2 * q] = (T)-2 * ev[q]; o[PER - 2 - 2 * q] = (T)-2 * ov[q]; }
                MsRun<T>::ld(Ev + a, ev);
                MsRun<T>::ld(Ov + a, ov);
#pragma unroll
                for (int q = 0; q < L; ++q) { tl[2 * q] = (T)-2 * ev[q]; tl[2 * q + 1] = (T)-2 * ov[q]; }
            }
            const uint4 w0 = __ldg(reinterpret_cast<const uint4*>(win + n0));
            const uint4 w1 = __ldg(reinterpret_cast<const uint4*>(win + M + n0));
            const T* wh = reinterpret_cast<const T*>(&w0);
            const T* wt = reinterpret_cast<const T*>(&w1);
            double od[PER], td[PER];
#pragma unroll
            for (int e = 0; e < PER; ++e) { od[e] = (double)(wh[e] * o[e]); td[e] = (double)(wt[e] * tl[e]); }
#pragma unroll
            for (int e = 0; e < PER; e += 2) {
                *reinterpret_cast<double2*>(dst + n0 + e) = make_double2(od[e], od[e + 1]);
                *reinterpret_cast<double2*>(dst + M + n0 + e) = make_double2(td[e], td[e + 1]);
            }
        }
        __syncwarp();
    }
}
