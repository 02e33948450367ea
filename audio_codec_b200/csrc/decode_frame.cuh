// Per-frame decoder (device functions shared by the stream kernel in decode.cu and the
// per-function ops in ops.cu): parse one frame's fields, dequantise every band, SBR
// reconstruction, IMDCT + synthesis window (coder/pacfile.py:177-229, coder/codec.py:47-222).
// CTA-cooperative: every thread of a DEC_NT-thread CTA calls decode_frame.
#pragma once
#include "mdct.cuh"
#include "psy.cuh"
#include "pvq_dec.cuh"

// Optional per-phase cycle accounting of the decoder (-DPAC_PHASE_TIMING; pac_debug_phases_dec).
#ifdef PAC_PHASE_TIMING
static __device__ unsigned long long g_pac_dphase[32];
#define DPHASE_T0(isLong) long long dph_t_ = clock64(); const int dph_off_ = (isLong) ? 0 : 16
#define DPHASE(i) do { __syncthreads(); if (threadIdx.x == 0) { long long n_ = clock64(); atomicAdd(&g_pac_dphase[(i) + dph_off_], (unsigned long long)(n_ - dph_t_)); dph_t_ = n_; } } while (0)
#else
#define DPHASE_T0(isLong)
#define DPHASE(i)
#endif
#define DEC_NT 256
#define DEC_REC_MAX 4096          // largest channel-block payload accepted (bytes)
#define DEC_GROUP 32              // lanes per band in the per-band phases (8 / 16 measured slower: divergent groups of a warp serialise)
#define DEC_MAX_NODES 512         // split-tree nodes per frame (a 128 kb/s long frame has ~150); sized for 3 CTAs per SM

// Batched front end of VQ streams (decode.cu): k_hdr reads the frame headers of a channel-block,
// k_dequant dequantises one band per thread into a lane-interleaved line array, k_decode then only
// runs SBR + IMDCT.  slot = sub-frame * nBands + band.
#define PRE_SLOTS 48              // 17 long bands, or 8 sub-frames x 6 short bands
#define PRE_WS_PER_CB 2048        // bound on the sum of the bands' power-of-two capacities (each < 2 nLines)
struct CbHdr {
    double env[4];                    // SBR envelope values (k_dequant), in omitted-band order
    unsigned short pos[PRE_SLOTS];    // payload bit position of the band within the record
    unsigned char ba[PRE_SLOTS];      // bit allocation
    unsigned char scale[PAC_SUBFRAMES];
    unsigned char sbr;                // long frame of an SBR stream with coded envelope bands
    unsigned char flags;              // bit 2 last, 1 cur, 0 next transient; bit 7: record present
    unsigned char pad[2];
    int p;                            // position in the batch's dense order (long blocks from the front, transient from the back)
};
// line k of sub-frame f of channel-block c (index within the batch): 32 blocks interleaved
__device__ __forceinline__ size_t pre_line_index(long long c, int k) {
    return ((size_t)(c >> 5) * 1024 + (size_t)k) * 32 + (size_t)(c & 31);
}

struct DecSmem {
    unsigned char* rec;   // [DEC_REC_MAX] payload bytes
    double* lines;        // [1024] dequantised MDCT lines of the current frame
    double* wsA;          // [2048] PVQ workspace A; later FFT buffer (first 1024) + DCT-IV output (last 1024)
    double* wsB;          // [2048] PVQ workspace B; SBR scratch
    double* y;            // [2048] windowed time samples of the block (short sub-frames overlap-added)
    int* ba;              // [32]
    int* sfb;             // [32]
    int* bandPos;         // [33] bit position of each band's payload
    int* capOff;          // [33]
    int* misc;            // [16]; [7] = number of tree nodes of the current frame
    DNode* nodes;         // [DEC_MAX_NODES] split-tree nodes of the current frame (phase A -> B -> C)
    BandDec* bd;          // [32]
    unsigned char* raw;   // base of the carve
};

__host__ __device__ constexpr size_t dec_smem_bytes() {
    return (1024 + 2048 * 3) * sizeof(double) + DEC_MAX_NODES * sizeof(DNode) + 32 * sizeof(BandDec) +
           (32 * 2 + 34 * 2 + 16) * sizeof(int) + DEC_REC_MAX;
}

// lean: the batched path (k_decode behind k_dequant) needs neither the record, the node list nor a second
// workspace (the SBR scratch and the IMDCT buffers are never live together): 41 KB instead of 75 KB.
__host__ __device__ constexpr size_t dec_lean_smem_bytes() {
    return (1024 + 2048 * 2) * sizeof(double) + (32 * 2 + 34 * 2 + 16) * sizeof(int);
}
__device__ __forceinline__ void dec_smem_carve(unsigned char* raw, DecSmem& S, bool lean = false) {
    double* dp = reinterpret_cast<double*>(raw);
    S.raw = raw;
    S.lines = dp; dp += 1024;
    S.wsA = dp; dp += 2048;
    if (lean) {
        S.wsB = S.wsA;
        S.y = dp; dp += 2048;
        S.nodes = nullptr; S.bd = nullptr; S.rec = nullptr;
        int* ip = reinterpret_cast<int*>(dp);
        S.ba = ip; ip += 32;
        S.sfb = ip; ip += 32;
        S.bandPos = ip; ip += 34;
        S.capOff = ip; ip += 34;
        S.misc = ip; ip += 16;
        return;
    }
    S.wsB = dp; dp += 2048;
    S.y = dp; dp += 2048;
    S.nodes = reinterpret_cast<DNode*>(dp);
    S.bd = reinterpret_cast<BandDec*>(S.nodes + DEC_MAX_NODES);
    int* ip = reinterpret_cast<int*>(S.bd + 32);
    S.ba = ip; ip += 32;
    S.sfb = ip; ip += 32;
    S.bandPos = ip; ip += 34;
    S.capOff = ip; ip += 34;
    S.misc = ip; ip += 16;
    S.rec = reinterpret_cast<unsigned char*>(ip);
}

// Steps (1)-(3) of a frame: header parse, dequantisation of every band, SBR reconstruction ->
// S.lines (still scaled).  One out-of-line copy shared by the long / short frames of both
// precisions (the kernel is instruction-cache bound otherwise).  Returns the overall scale.
#ifdef PAC_FRONT_INLINE
__device__ __forceinline__
#else
static __device__ __noinline__
#endif
int decode_front(const PacDev* __restrict__ dev, unsigned char* raw, long long recBits,
                                                bool isLong, int* err, const CbHdr* __restrict__ pre,
                                                const double* __restrict__ preLines, long long preCb, int sf) {
    DecSmem S;
    dec_smem_carve(raw, S, pre != nullptr);
    const int H = isLong ? PAC_HALF_LONG : PAC_HALF_SHORT;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const PacBands& bands = isLong ? dev->longBands : dev->shortBands;
    const int nB = bands.nBands;
    DPHASE_T0(isLong);
    // (1) frame header: overall scale, bit allocations (+ BFP scale factors).  VQ streams carry the
    //     allocations as fixed-width fields up front: one warp reads them in parallel (lane = band) and
    //     places the band payloads with a scan.  Block-FP interleaves them with the mantissas: thread 0.
    if (pre) {
        // the bands were dequantised by k_dequant: fetch this frame's lines (still scaled)
        const bool sbrP = isLong && pre->sbr;
        const int cut = dev->sbrCutoff;
        for (int i = tid; i < H; i += DEC_NT)
            S.lines[i] = (sbrP && i >= cut) ? 0.0 : preLines[pre_line_index(pre->p, sf * PAC_HALF_SHORT + i)];
        if (tid == 0) {
            S.misc[5] = pre->scale[sf];
            S.misc[6] = sbrP;
        }
        __syncthreads();
        if (sbrP && tid < dev->sbrNOmitted && tid < 4) S.lines[cut + tid] = pre->env[tid];   // coder/codec.py:111-132
        __syncthreads();
    } else
    if (dev->useVQ) {
        if (warp == 0) {
            const int pos0 = S.misc[3], nS = dev->nScaleBits, nM = dev->nMantSizeBits;
            const bool act = lane < nB;
            BitR br = {S.rec, (long long)pos0 + nS + (long long)lane * nM, recBits, DEC_REC_MAX};
            int bab = 0;
            if (act) {
                const int v = (int)br_get(br, nM);
                bab = v ? v + 1 : 0;
                S.ba[lane] = bab;
            }
            const bool om = act && isLong && bands.omitted[lane];
            const bool sbr = isLong && dev->useSBR && __any_sync(0xffffffffu, om && bab != 0);    // coder/pacfile.py:660-664
            const int len = bab ? bab * ((sbr && om) ? 1 : bands.nLines[lane]) : 0;
            int cap = 0;
            if (act) { cap = 1; while (cap < bands.nLines[lane]) cap <<= 1; }
            int inc = len, cinc = cap;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, inc, o), tc = __shfl_up_sync(0xffffffffu, cinc, o);
                if (lane >= o) { inc += t; cinc += tc; }
            }
            const int base = pos0 + nS + nB * nM;
            if (act) {
                S.bandPos[lane] = base + inc - len;
                S.capOff[lane] = cinc - cap;
            }
            const int total = __shfl_sync(0xffffffffu, inc, 31), ctot = __shfl_sync(0xffffffffu, cinc, 31);
            if (lane == 0) {
                BitR b0 = {S.rec, pos0, recBits, DEC_REC_MAX};
                S.misc[5] = (int)br_get(b0, nS);
                S.misc[6] = sbr;
                S.misc[7] = 0;
                S.misc[3] = base + total;
                S.capOff[nB] = ctot;
                if ((long long)base + total > recBits) atomicOr(err, PVQ_ERR_STACK);
            }
        }
    } else
    if (tid == 0) {
        BitR br = {S.rec, S.misc[3], recBits, DEC_REC_MAX};
        S.misc[7] = 0;
        S.misc[5] = (int)br_get(br, dev->nScaleBits);
        bool anyOm = false;
        if (dev->useVQ) {
            for (int b = 0; b < nB; ++b) {
                int v = (int)br_get(br, dev->nMantSizeBits);
                S.ba[b] = v ? v + 1 : 0;
            }
            for (int b = 0; b < nB; ++b) anyOm |= isLong && bands.omitted[b] && S.ba[b] != 0;
            const bool sbr = isLong && dev->useSBR && anyOm;          // coder/pacfile.py:660-664
            S.misc[6] = sbr;
            long long p = br.pos;
            for (int b = 0; b < nB; ++b) {
                S.bandPos[b] = (int)p;
                if (S.ba[b]) p += (long long)S.ba[b] * ((sbr && bands.omitted[b]) ? 1 : bands.nLines[b]);
            }
            br.pos = p;
        } else {
            S.misc[6] = 0;
            for (int b = 0; b < nB; ++b) {
                int v = (int)br_get(br, dev->nMantSizeBits);
                S.ba[b] = v ? v + 1 : 0;
                S.sfb[b] = (int)br_get(br, dev->nScaleBits);
                S.bandPos[b] = (int)br.pos;
                if (S.ba[b]) br.pos += (long long)S.ba[b] * bands.nLines[b];
            }
        }
        S.misc[3] = (int)br.pos;
        int c = 0;
        for (int b = 0; b < nB; ++b) {
            S.capOff[b] = c;
            int cap = 1;
            while (cap < bands.nLines[b]) cap <<= 1;
            c += cap;
        }
        S.capOff[nB] = c;
        if (br.pos > recBits) atomicOr(err, PVQ_ERR_STACK);
    }
    if (!pre) {
        for (int i = tid; i < H; i += DEC_NT) S.lines[i] = 0.0;
        __syncthreads();
    }
    const bool sbr = S.misc[6] != 0;
    const int scale = S.misc[5];
    DPHASE(0);
    // (2) bands: VQ in three phases (parse per band / index decode per leaf / recombine per band),
    //     block floating point directly
    PvqTab pt = {dev->pvqN, dev->pvqRowOff, dev->log2tab, dev->pvqKW, dev->thetaTab};
    const int omPos = dev->sbrCutoff;
    if (pre) {
    } else if (dev->useVQ) {
        constexpr int G = DEC_GROUP;
        const int grp = tid / G;
        // phase A is scalar: one THREAD per band, bands dealt round-robin to the warps (<= 4 per warp)
        {
            const int b = lane * (DEC_NT / 32) + warp;
            const int bb = b < nB ? S.ba[b] : 0;
            if (bb) {
            BitR br = {S.rec, S.bandPos[b], recBits, DEC_REC_MAX};
            int werr = 0;
            const int cap = S.capOff[b + 1] - S.capOff[b];
            const bool env = sbr && bands.omitted[b];       // one envelope value (coder/codec.py:111-132)
            grp_parse_gain_shape<1>(pt, br, env ? bb : bb * bands.nLines[b], env ? 1 : bands.nLines[b], S.wsA, S.wsB,
                                    S.capOff[b], cap, S.nodes, &S.misc[7], DEC_MAX_NODES, &S.bd[b], &werr);
            if (werr) atomicOr(err, werr);
            }
        }
        DPHASE(1);
        __syncthreads();
        cta_decode_leaves(pt, S.nodes, min(S.misc[7], DEC_MAX_NODES), S.wsA, S.wsB, tid, DEC_NT, err);
        DPHASE(2);
        __syncthreads();
        for (int b = grp; b < nB; b += DEC_NT / G) {
            if (!S.ba[b]) continue;
            double* out = S.lines + bands.lower[b];
            if (sbr && bands.omitted[b]) {                  // stored at consecutive lines from the cut-off
                int k = 0;
                for (int q = 0; q < b; ++q) k += bands.omitted[q] ? 1 : 0;
                out = S.lines + omPos + k;
            }
            grp_finish_gain_shape<G>(dev->thetaTab, S.nodes, S.bd[b], S.wsA, S.wsB, S.capOff[b], out);
        }
    } else {
        for (int b = warp; b < nB; b += DEC_NT / 32) {
            const int bb = S.ba[b];
            if (!bb) continue;
            const int lo = bands.lower[b], n = bands.nLines[b];
            for (int i = lane; i < n; i += 32) {
                BitR br = {S.rec, (long long)S.bandPos[b] + (long long)i * bb, recBits, DEC_REC_MAX};
                unsigned m = (unsigned)br_get(br, bb);
                S.lines[lo + i] = dequantize_bfp(S.sfb[b], m, dev->nScaleBits, bb);
            }
        }
    }
    __syncthreads();
    DPHASE(3);
    // (3) SBR: envelope smoothing, 2:1 transposition of the low band, per-band envelope
    //     scaling (coder/codec.py:134-194)
    if (isLong && sbr) {
        const int cut = dev->sbrCutoff, nLen = H - cut, nOm = dev->sbrNOmitted;
        double* hi = S.wsB;               // transposed high band [nLen]
        double* smooth = S.wsB + 1024;    // smoothed envelope [nLen]
        for (int i = tid; i < nLen; i += DEC_NT) {
            double e = 0.0;
            for (int p = 0; p < nOm; ++p) e += S.lines[cut + p] * dev->sbrGauss[(size_t)p * nLen + i];
            smooth[i] = e;
            // interp1d('slinear') of the lines at frequency (k+1/2)/2 bins: between line m-1 and m
            // (k even) or m and m+1 (k odd), k = cut+i, m = k/2; weights are exactly 1/4, 3/4
            int k = cut + i, m = k >> 1;
            double v;
            if ((k & 1) == 0) v = 0.25 * S.lines[m - 1] + 0.75 * S.lines[m];
            else v = 0.75 * S.lines[m] + 0.25 * S.lines[m + 1];
            hi[i] = v;
        }
        __syncthreads();
        for (int i = tid; i < nLen; i += DEC_NT) S.lines[cut + i] = hi[i];
        __syncthreads();
        for (int b = warp; b < nB; b += DEC_NT / 32) {
            if (!bands.omitted[b]) continue;
            const int lo = bands.lower[b], n = bands.nLines[b];
            double mx = 0.0, sm = 0.0;
            for (int i = lane; i < n; i += 32) {
                double a = fabs(S.lines[lo + i]);
                mx = fmax(mx, a);
                sm += a;
            }
#pragma unroll
            for (int o = 16; o; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            sm = warp_sum(sm);
            if (mx > 0.0) {
                double mean = sm / (double)n;
                for (int i = lane; i < n; i += 32) S.lines[lo + i] *= smooth[lo - cut + i] / mean;
            }
        }
        __syncthreads();
    }
    DPHASE(4);
    return scale;
}

// Decode one frame of H = N/2 lines whose fields start at bit S.misc[3] of the record;
// adds its windowed IMDCT into y[yOff .. yOff+N).
template <int N, typename T>
__device__ void decode_frame(const PacDev* __restrict__ dev, const DecSmem& S, long long recBits, int shape, int yOff,
                             int* err, const CbHdr* __restrict__ pre = nullptr, const double* __restrict__ preLines = nullptr,
                             long long preCb = 0, int sf = 0) {
    constexpr int Q = N / 4;
    const int tid = threadIdx.x;
    const bool isLong = N == PAC_NLONG;
    const int scale = decode_front(dev, S.raw, recBits, isLong, err, pre, preLines, preCb, sf);
    // (4) undo the overall scale, IMDCT, synthesis window (coder/codec.py:85-89)
    // T = float: the transform runs in fp32 (fast path; PCM within 1 LSB of the fp64 result).
    const PacTables<T>& t = psy_tabs<T>(dev);
    const double inv = 1.0 / (double)(1 << scale);
    const cplx<T>* tw = reinterpret_cast<const cplx<T>*>(isLong ? t.tw_mdct_long : t.tw_mdct_short);
    const cplx<T>* roots = reinterpret_cast<const cplx<T>*>(t.tw_fft[isLong ? 2 : 0]);
    cplx<T>* fbuf = reinterpret_cast<cplx<T>*>(S.wsA);
    T* v = reinterpret_cast<T*>(S.wsA + 1024);
    imdct_pre<T, N, 1, DEC_NT, double>(S.lines, inv, tw, fbuf, tid);
    __syncthreads();
    cta_fft<T, Q, 1, DEC_NT>(fbuf, roots, tid);
    mdct_post<T, N, 1, DEC_NT>(fbuf, tw, (T)1, v, tid);
    __syncthreads();
    const T* w = (isLong ? t.win_long : t.win_short) + shape * N;
    for (int i = tid; i < N; i += DEC_NT) S.y[yOff + i] += (double)(w[i] * imdct_sample<T, N>(v, i));
    __syncthreads();
}

