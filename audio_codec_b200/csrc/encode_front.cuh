// Front end of the batched stream encoder: one CTA per channel-block runs window + MDCT, the Hann FFT and the
// psychoacoustic model and hands the scaled lines, the per-band SMRs, the overall scale and the SBR envelope
// sums to k_bitalloc / k_pvq / k_pack (coder/codec.py:266-311, :426-473; coder/psychoac.py:163-291).
//
// Same arithmetic as encode_frame() (every value bit-identical), organised for occupancy: the kernel is bound by
// the latency of serial chains (masker links, curve walks, fp64 transcendentals) and by barrier skew, not by issue
// slots or bandwidth, so what pays is more channel-blocks in flight per SM.  A CTA is NT = 128 threads and keeps
// the frame as the int16 codes it loaded (4 KB instead of 16 KB of doubles; the conversion is three fp64
// operations, pcm_to_double_fast): 37 KB of shared memory and <= 80 registers -> six CTAs per SM instead of four
// 256-thread ones.
#pragma once
#include "encode_frame.cuh"

#define ENF_NT 128
#define ENF_SHORT_MAXM 72          // masker capacity of a short frame (at most 65 peaks in 129 bins)
#define ENF_SHORT_DOUBLES 760      // per-warp buffers of the short-frame path (see k_encode_front); 4 x 760 <= 2048 + 1024 + 1026
#ifndef ENF_MINB
#define ENF_MINB 6
#endif

struct EnfSmem {
    short* xs;       // [2048] frame samples, int16 codes
    double* fb;      // [2048] FFT buffer; then the masker list
    double* lines;   // [1024] MDCT lines
    double* mk;      // [1026] |X| -> intensity (bins 0..1024) -> SMR bounds
    double* smr;     // [32]
    double* tmp;     // [48] band maxima keys of a short sub-frame (band_smr_pruned scratch), reductions
    double* env;     // [32] SBR: sum of |X| / H over each omitted band
    int* misc;       // [16]
};
__host__ __device__ constexpr size_t enf_smem_bytes() {
    // 37 840 bytes: six CTAs (+ 1 KB each reserved by the system) fit the 228 KB of an SM
    return 2048 * sizeof(short) + (2048 + 1024 + 1026 + 32 + 48 + 32) * sizeof(double) + 16 * sizeof(int);
}
__device__ __forceinline__ void enf_carve(unsigned char* raw, EnfSmem& S) {
    double* dp = reinterpret_cast<double*>(raw);
    S.fb = dp; dp += 2048;
    S.lines = dp; dp += 1024;
    S.mk = dp; dp += 1026;
    S.smr = dp; dp += 32;
    S.tmp = dp; dp += 48;
    S.env = dp; dp += 32;
    S.xs = reinterpret_cast<short*>(dp);
    S.misc = reinterpret_cast<int*>(S.xs + 2048);
}

template <typename T> __device__ __forceinline__ T enf_sample(short c) { return (T)pcm_to_double_fast(c); }

template <int NT> __device__ __forceinline__ double enf_block_max(double v, double* red, int tid) {
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        double t = __shfl_xor_sync(0xffffffffu, v, o);
        v = t > v ? t : v;
    }
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    double r = red[0];
#pragma unroll
    for (int w = 1; w < NT / 32; ++w) r = red[w] > r ? red[w] : r;
    return r;
}

// mdct_fold() reading int16 codes (one frame)
template <typename T, int N, int NT>
__device__ __forceinline__ void enf_fold(const short* x, const T* __restrict__ w, const cplx<T>* __restrict__ tw,
                                         cplx<T>* fbuf, int tid) {
    constexpr int Q = N / 4;
    for (int n = tid; n < Q; n += NT) {
        int i0, i1, i2, i3;
        T re, im;
        if (2 * n < Q) {
            i0 = 3 * Q - 1 - 2 * n; i1 = 3 * Q + 2 * n;
            i2 = Q - 1 - 2 * n;     i3 = Q + 2 * n;
            const T a0 = enf_sample<T>(x[i0]) * w[i0], a1 = enf_sample<T>(x[i1]) * w[i1];
            const T a2 = enf_sample<T>(x[i2]) * w[i2], a3 = enf_sample<T>(x[i3]) * w[i3];
            re = -a0 - a1;
            im = a2 - a3;
        } else {
            i0 = 2 * n - Q;         i1 = 3 * Q - 1 - 2 * n;
            i2 = Q + 2 * n;         i3 = 5 * Q - 1 - 2 * n;
            const T a0 = enf_sample<T>(x[i0]) * w[i0], a1 = enf_sample<T>(x[i1]) * w[i1];
            const T a2 = enf_sample<T>(x[i2]) * w[i2], a3 = enf_sample<T>(x[i3]) * w[i3];
            re = a0 - a1;
            im = -a2 - a3;
        }
        cplx<T> t = {re, im};
        fbuf[n] = cmul(t, tw[n]);
    }
}

// One frame of N samples (int16 codes at x): steps (1)-(3) of encode_frame() and the hand-off of encode_back().
// NT = ENF_NT: the CTA works on one frame.  NT = 32: every WARP of the CTA works on a frame of its own (S: the warp's
// buffers) - the eight short frames of a transient block, four at a time; the barriers stay CTA-wide (every warp
// passes the same unconditional barriers), so the four frames advance in step.  cand: band_smr_pruned's candidate
// list ([N / 2] ints; may be null when NT >= N / 2).
template <int N, typename T, int NT>
__device__ __forceinline__ void enf_frame(const PacDev* __restrict__ dev, const EnfSmem& S, const short* x, int shape,
                                          bool sbr, EncPre* hdr, double* linesOut, int sf, int* cand) {
    constexpr int H = N / 2, Q = N / 4;
    const int tid = NT == 32 ? (int)(threadIdx.x & 31) : (int)threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const PacBands& bands = (N == PAC_NLONG) ? dev->longBands : dev->shortBands;
    const int nB = bands.nBands;
    const PsyConst<T> pc = psy_const<N, T>(dev);
    const PacTables<T>& tabsT = psy_tabs<T>(dev);
    const T* win = ((N == PAC_NLONG) ? tabsT.win_long : tabsT.win_short) + shape * N;
    const cplx<T>* tw = reinterpret_cast<const cplx<T>*>(N == PAC_NLONG ? tabsT.tw_mdct_long : tabsT.tw_mdct_short);
    const cplx<T>* rootsQ = reinterpret_cast<const cplx<T>*>(tabsT.tw_fft[N == PAC_NLONG ? 2 : 0]);
    cplx<T>* fbuf = reinterpret_cast<cplx<T>*>(S.fb);
    T* mkT = reinterpret_cast<T*>(S.mk);
    double* red = S.tmp + 40;

    // (1) window + MDCT (coder/codec.py:301-305)
    ENF_PH_START(N == PAC_NLONG ? 0 : 32);
    // long frames: the first quadrant of the 1024-point roots (4 KB) goes to shared memory - in S.mk, which is free until
    // the Hann FFT is over - and serves both FFTs (fft.cuh: tw_quarter)
    constexpr bool QT = N == PAC_NLONG;
    cplx<T>* qtab = reinterpret_cast<cplx<T>*>(S.mk);
    if (QT) {
        const cplx<T>* r1024 = reinterpret_cast<const cplx<T>*>(tabsT.tw_fft[3]);
        for (int i = tid; i < 256; i += NT) qtab[i] = r1024[i];
    }
    enf_fold<T, N, NT>(x, win, tw, fbuf, tid);
    __syncthreads();
    ENF_PH(0);
    if (QT) cta_fft<T, Q, 1, NT, 1>(fbuf, qtab, tid);       // 512-point: an index is worth 2 in the 1024-point table
    else cta_fft<T, Q, 1, NT>(fbuf, rootsQ, tid);
    ENF_PH(1);
    mdct_post<T, N, 1, NT, double>(fbuf, tw, (T)2 / (T)N, S.lines, tid);
    __syncthreads();
    double mx = 0.0;
    for (int i = tid; i < H; i += NT) mx = fmax(mx, fabs(S.lines[i]));
    double maxLine = enf_block_max<NT>(mx, red, tid);
    ENF_PH(2);

    // (2) Hann FFT magnitude (psychoacoustic model and SBR envelope)
    for (int n = tid; n < H; n += NT) {
        cplx<T> z = {enf_sample<T>(x[2 * n]) * pc.hann[2 * n], enf_sample<T>(x[2 * n + 1]) * pc.hann[2 * n + 1]};
        fbuf[n] = z;
    }
    __syncthreads();
    ENF_PH(3);
    hann_rfft_finish<N, NT, T, QT>(pc, fbuf, mkT, tid, qtab);
    ENF_PH(4);
    if (sbr) {                                    // coder/codec.py:460-473
        double m2 = 0.0;
        for (int k = tid; k <= H; k += NT) m2 = fmax(m2, (double)mkT[k]);
        const double maxFFT = enf_block_max<NT>(m2, red, tid) / (double)H;
        maxLine = fmax(maxLine, maxFFT);
        for (int b = warp; b < nB; b += NT / 32) {
            if (!bands.omitted[b]) continue;
            double a = 0.0;
            for (int i = lane; i < bands.nLines[b]; i += 32) a += (double)mkT[bands.lower[b] + i] / (double)H;
            a = warp_sum(a);
            if (lane == 0) S.env[b] = a;
        }
        __syncthreads();
    }
    const int scale = scale_factor(maxLine, dev->nScaleBits, 5);      // codec.py:308-311
    const double up = (double)(1 << scale);
    for (int i = tid; i < H; i += NT) S.lines[i] *= up;
    for (int k = tid; k <= H; k += NT) {
        const T m = mkT[k];
        mkT[k] = pc.norm * (m * m);
    }
    __syncthreads();
    ENF_PH(5);

    // (3) psychoacoustic model -> SMR per band
    const MaskerList<T> ml = masker_list<T, (N == PAC_NLONG ? PSY_MAX_MASKERS : ENF_SHORT_MAXM)>(reinterpret_cast<T*>(S.fb));
    const int nMask = pick_maskers<N, NT, T, true>(mkT, pc, ml, tid);
    band_smr_pruned<N, NT, T>(pc, ml, nMask, S.lines, scale, mkT, bands, S.smr, S.tmp, cand, tid);

    // hand-off to k_bitalloc / k_pvq / k_pack
    for (int i = tid; i < H; i += NT) linesOut[(size_t)(sf * PAC_HALF_SHORT + i) * 32] = S.lines[i];
    if (tid < nB) {
        hdr->smr[sf * nB + tid] = S.smr[tid];
        if (sbr && bands.omitted[tid]) {
            int k = 0;
            for (int q = 0; q < tid; ++q) k += bands.omitted[q] ? 1 : 0;
            if (k < 4) hdr->env[k] = (S.env[tid] * up) / (double)bands.nLines[tid];
        }
    }
    if (tid == 0) hdr->scale[sf] = (unsigned char)scale;
    __syncthreads();
    ENF_PH(16);
}

template <typename T>
__global__ void __launch_bounds__(ENF_NT, ENF_MINB)
k_encode_front(const PacDev* __restrict__ dev, EncArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    EnfSmem S;
    enf_carve(smem_raw, S);
    const int tid = threadIdx.x;
    const int C = dev->nChannels;
    const int cb = blockIdx.x + a.cb0;        // written channel-block index
    const int wb = cb / C, c = cb - wb * C;   // written block (global), channel
    const int s = a.wbStream[wb], j = wb - a.wbOff[s], nbs = a.nb[s];
    const unsigned char* det = a.det + a.blkOff[s];
    // flags (coder/pacfile.py:717-743): written block j carries next = det[j], cur = det[j-1],
    // last = det[j-2]; the duplicate of the last block has next = False; Close()'s block has none.
    bool nxt = j < nbs && det[j];
    bool cur = j >= 1 && j - 1 < nbs && det[j - 1];
    bool last = j >= 2 && j - 2 < nbs && det[j - 2];
    if (j == nbs + 1) nxt = cur = last = false;
    const int dataBlk = j < nbs ? j : (j == nbs ? nbs - 1 : -1);
    const int priorBlk = j == 0 ? -1 : min(j - 1, nbs - 1);
    const short* base = a.pcm + ((long long)a.blkOff[s] * C + (long long)c * nbs) * 1024;
    if ((((size_t)a.pcm) & 15) == 0) {        // 16 samples (two 16-byte loads) per thread
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int i0 = (tid + h * ENF_NT) * 8;
            const int blk = i0 < 1024 ? priorBlk : dataBlk;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (blk >= 0) v = __ldg(reinterpret_cast<const uint4*>(base + (long long)blk * 1024 + (i0 & 1023)));
            *reinterpret_cast<uint4*>(S.xs + i0) = v;
        }
    } else {
        for (int i = tid; i < 2048; i += ENF_NT) {
            const int blk = i < 1024 ? priorBlk : dataBlk;
            S.xs[i] = blk < 0 ? (short)0 : base[(long long)blk * 1024 + (i & 1023)];
        }
    }
    if (tid == 0) S.misc[4] = 0;
    __syncthreads();
    EncPre* hdr = a.pre + blockIdx.x;
    if (tid == 0) {                           // dense order for k_pvq: its warps hold blocks of one kind
        const int p = cur ? a.nBatch - 1 - atomicAdd(&a.counts[1], 1) : atomicAdd(&a.counts[0], 1);
        a.perm[p] = blockIdx.x;
        S.misc[8] = p;
    }
    __syncthreads();
    double* linesOut = a.preLines + epre_line_index(S.misc[8], 0);
    if (!cur) {
        const int shape = (last && nxt) ? PAC_WIN_STARTSTOP : last ? PAC_WIN_STOP : nxt ? PAC_WIN_START : PAC_WIN_SINE;
        enf_frame<PAC_NLONG, T, ENF_NT>(dev, S, S.xs, shape, dev->useSBR != 0, hdr, linesOut, 0,
                                        reinterpret_cast<int*>(S.xs));        // the frame is dead by then
    } else {
        // zero sub-frame check (coder/pacfile.py:503-506): the block is dropped if ANY channel has one.
        // A code maps to 0.0 iff its magnitude bits are zero (0 and -32768, coder/pcmfile.py:89-98).
        for (int sf = 0; sf < PAC_SUBFRAMES; ++sf) {
            bool nz = false;
            for (int i = tid; i < PAC_NSHORT; i += ENF_NT) nz |= (S.xs[448 + 128 * sf + i] & 0x7fff) != 0;
            if (!__syncthreads_or(nz) && tid == 0) S.misc[4] = 1;
        }
        // the eight short frames: one per warp, four at a time (a short frame keeps 32 lanes busy, not 128, and a
        // transient block cost six long ones when its frames ran one after the other on the whole CTA)
        static_assert(ENF_NT == 128 && PAC_SUBFRAMES == 8, "two rounds of four short frames");
        EnfSmem W = S;
        double* wb0 = S.fb + (size_t)(tid >> 5) * ENF_SHORT_DOUBLES;
        W.fb = wb0;             // [352] FFT buffers, then the masker list
        W.lines = wb0 + 352;    // [128]
        W.mk = wb0 + 480;       // [136]
        W.smr = wb0 + 616;      // [32]
        W.tmp = wb0 + 648;      // [48]
        int* candW = reinterpret_cast<int*>(wb0 + 696);      // [128]
        for (int r = 0; r < 2; ++r) {
            const int sf = 4 * r + (tid >> 5);
            enf_frame<PAC_NSHORT, T, 32>(dev, W, S.xs + 448 + 128 * sf, PAC_WIN_SINE, false, hdr, linesOut, sf, candW);
        }
    }
    __syncthreads();
    if (tid == 0) {
        hdr->flags = (unsigned char)((last ? 4 : 0) | (cur ? 2 : 0) | (nxt ? 1 : 0));
        hdr->zeroSub = (unsigned char)S.misc[4];
        hdr->p = S.misc[8];
    }
}
