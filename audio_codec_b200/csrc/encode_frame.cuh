// Per-frame encoder (device functions shared by the fused stream kernel in encode.cu
// and the per-function ops in ops.cu): window + MDCT, overall scale, psychoacoustic
// SMRs, bit allocation, PVQ / block-FP quantisation and MSB-first packing of one frame
// (coder/codec.py:266-380, :426-555; coder/pacfile.py:363-447).  CTA-cooperative:
// every thread of an ENC_NT-thread CTA calls encode_frame.
#pragma once
#include "bitalloc.cuh"
#include "mdct.cuh"
#include "psy.cuh"
#include "pvq.cuh"

#define ENC_NT 256

// Optional per-phase cycle accounting (build with -DPAC_PHASE_TIMING; read with pac_debug_phases).
#ifdef PAC_PHASE_TIMING
static __device__ unsigned long long g_pac_phase[32];
#define PHASE_T0() long long ph_t_ = clock64(); const int ph_off_ = PH_IS_LONG ? 0 : 16
#define PHASE(i) do { __syncthreads(); if (threadIdx.x == 0) { long long n_ = clock64(); atomicAdd(&g_pac_phase[(i) + ph_off_], (unsigned long long)(n_ - ph_t_)); ph_t_ = n_; } } while (0)
#else
#define PHASE_T0()
#define PHASE(i)
#endif
#define REC_STRIDE 2176          // bytes per channel-block scratch record (>= worst-case declared size)
#define REC_WORDS (REC_STRIDE / 4)
#define BAND_WORDS 640
#define ENC_MAX_NODES 512        // split-tree nodes per frame (a 128 kb/s long frame has ~150)

// int16 PCM -> signed fraction (coder/pcmfile.py:89-98): 2|c|/65535 with the sign
// restored; -32768 has magnitude bits 0 and maps to 0.0.
__device__ __forceinline__ double pcm_to_double(short c) {
    int m = c < 0 ? -(int)c : (int)c;
    m &= 0x7fff;
    double v = (double)(2 * m) / 65535.0;
    return c < 0 ? -v : v;
}

// The same value without the division: q0 = x r, one Newton correction with the exact remainder
// (r = fl(1 / 65535)).  Equal to the correctly rounded x / 65535 for every even x in [0, 65534] (checked
// exhaustively in exact rational arithmetic, and on the device: tests/test_ops_gpu.py).
__device__ __forceinline__ double pcm_to_double_fast(short c) {
    int m = c < 0 ? -(int)c : (int)c;
    m &= 0x7fff;
    const double x = (double)(2 * m), r = 1.0 / 65535.0;
    const double q0 = __dmul_rn(x, r);
    const double v = fma(fma(-q0, 65535.0, x), r, q0);
    return c < 0 ? -v : v;
}

// Batched back end of VQ streams (encode.cu): k_encode stops after the bit allocation and leaves the
// scaled lines and this header; k_pvq quantises one band per thread; k_pack assembles the record.
// slot = sub-frame * nBands + band.
#define EPRE_SLOTS 48             // 17 long bands, or 8 sub-frames x 6 short bands
#define EPRE_WS_PER_CB 2048       // bound on the sum of the bands' power-of-two capacities (each < 2 nLines)
#define EPRE_WORDS_PER_CB 704     // bit buffers: nLines / 2 + 3 words per band (16 bits per line at most)
struct EncPre {
    double env[4];                    // SBR envelope value of each omitted band (coder/codec.py:508-510)
    double smr[EPRE_SLOTS];           // signal-to-mask ratio per band (k_encode -> k_bitalloc)
    unsigned short used[EPRE_SLOTS];  // bits the band emitted (k_pvq)
    unsigned char ba[EPRE_SLOTS];     // bit allocation; k_pvq zeroes it when the band emits nothing (codec.py:352-353)
    unsigned char scale[PAC_SUBFRAMES];
    unsigned char flags;              // bit 2 last, 1 cur, 0 next transient
    unsigned char zeroSub;            // a short sub-frame of this channel is all zero (coder/pacfile.py:503-506)
    unsigned char pad[2];
    int p;                            // position in the batch's dense order (long blocks from the front, transient from the back)
};
struct EncPreOut { EncPre* hdr; double* lines; int sf; int prune; };     // prune: band_smr_pruned instead of masked_threshold + band_smr     // lines: this block's column of the interleaved line array
__device__ __forceinline__ size_t epre_line_index(long long p, int k) {
    return ((size_t)(p >> 5) * 1024 + (size_t)k) * 32 + (size_t)(p & 31);
}

// ---------------------------------------------------------------- per-frame encode
struct EncSmem {
    double* xs;        // [2048] frame samples; later PVQ workspace A
    double* fb;        // [2048] FFT buffer / masker list; later PVQ workspace B
    double* lines;     // [1024] MDCT lines (scaled)
    double* mk;        // [1040] |X| -> intensity -> masked threshold; later PVQ pulse ints
    double* smr;       // [32]
    double* tmp;       // [32]
    unsigned* blockbuf;  // [REC_WORDS] the channel-block bitstream (big-endian words)
    unsigned* bandbits;  // [BAND_WORDS] per-band private bit buffers
    int* ba;           // [32] bit allocation
    int* nlEff;        // [32] lines per band as the allocator sees them (1 for SBR-omitted)
    int* used;         // [32] bits emitted per band
    int* wordOff;      // [33]
    int* capOff;       // [33] pow2 workspace offsets
    int* sfb;          // [32] BFP scale factors
    int* misc;         // [16]
    double* red;       // [16] reduction scratch
    double* env;       // [32] SBR: sum of |X|/H over each omitted band
    // optional per-band field log (global memory, [nBands][logCap]); nullptr in the stream kernel
    unsigned long long* logVal;
    int* logW;
    int* logCnt;       // [nBands]
    int logCap;
    TNode* nodes;      // [ENC_MAX_NODES] split-tree nodes of the current frame
    BandEnc* bandEnc;  // [32]
    unsigned char* raw;  // base of the carve (encode_back re-derives the pointers from it)
};
struct EncLog { unsigned long long* val; int* w; int* cnt; int cap; };

__host__ __device__ constexpr size_t enc_smem_bytes() {
    return (2048 + 2048 + 1024 + 1040 + 32 + 32 + 16 + 32) * sizeof(double) +
           (REC_WORDS + BAND_WORDS) * sizeof(unsigned) + (32 * 4 + 34 * 2 + 16) * sizeof(int) +
           ENC_MAX_NODES * sizeof(TNode) + 32 * sizeof(BandEnc);
}

// The arrays only the in-CTA quantiser and packer use come last: the batched front end (k_encode with
// EncArgs::pre) runs with enc_front_smem_bytes() and one more CTA per SM.
__host__ __device__ constexpr size_t enc_front_smem_bytes() {
    return (2048 + 2048 + 1024 + 1040 + 32 + 32 + 16 + 32) * sizeof(double) + (32 * 4 + 34 * 2 + 16) * sizeof(int);
}
__device__ __forceinline__ void enc_smem_carve(unsigned char* raw, EncSmem& S) {
    double* dp = reinterpret_cast<double*>(raw);
    S.raw = raw;
    S.xs = dp; dp += 2048;
    S.fb = dp; dp += 2048;
    S.lines = dp; dp += 1024;
    S.mk = dp; dp += 1040;
    S.smr = dp; dp += 32;
    S.tmp = dp; dp += 32;
    S.red = dp; dp += 16;
    S.env = dp; dp += 32;
    int* ip = reinterpret_cast<int*>(dp);
    S.ba = ip; ip += 32;
    S.nlEff = ip; ip += 32;
    S.used = ip; ip += 32;
    S.wordOff = ip; ip += 34;
    S.capOff = ip; ip += 34;
    S.sfb = ip; ip += 32;
    S.misc = ip; ip += 16;
    unsigned* up = reinterpret_cast<unsigned*>(ip);
    S.blockbuf = up; up += REC_WORDS;
    S.bandbits = up; up += BAND_WORDS;
    S.nodes = reinterpret_cast<TNode*>(up);
    S.bandEnc = reinterpret_cast<BandEnc*>(S.nodes + ENC_MAX_NODES);
    S.logVal = nullptr; S.logW = nullptr; S.logCnt = nullptr; S.logCap = 0;
}

__device__ __forceinline__ double block_max(double v, double* red, int tid) {
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
    for (int w = 1; w < ENC_NT / 32; ++w) r = red[w] > r ? red[w] : r;
    return r;
}

// Steps (4)-(7) of a frame: bit allocation, quantisation, header and payload merge.  One copy
// shared by the long and short instantiations of every precision (not inlined: the two encoders
// are instruction-cache bound otherwise).  flags: 1 long frame, 2 SBR, 4 transient block,
// 8 last-or-next-transient.  S.lines holds the scaled MDCT lines, S.smr the SMRs, S.env the SBR sums.
static __device__ __noinline__ int encode_back(const PacDev* __restrict__ dev, unsigned char* raw, EncLog lg, int flags,
                                               int scale, int* bpos, int* err, EncPreOut po) {
    EncSmem S;
    enc_smem_carve(raw, S);
    S.logVal = lg.val; S.logW = lg.w; S.logCnt = lg.cnt; S.logCap = lg.cap;
    const bool isLong = flags & 1, sbr = flags & 2, cur = flags & 4, lastOrNext = flags & 8;
    const int H = isLong ? PAC_HALF_LONG : PAC_HALF_SHORT;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const PacBands& bands = isLong ? dev->longBands : dev->shortBands;
    const int nB = bands.nBands;
    const double up = (double)(1 << scale);
#define PH_IS_LONG isLong
    PHASE_T0();
#undef PH_IS_LONG
    // (4) bit budget + allocation (coder/codec.py:288-299, :446-453; coder/bitalloc.py:62-145)
    if (tid < nB) S.nlEff[tid] = (sbr && bands.omitted[tid]) ? 1 : bands.nLines[tid];
    if (warp == 1) {                                       // pow2 workspace offsets of the bands (static layout)
        int c = 0;
        if (lane < nB) { c = 1; while (c < bands.nLines[lane]) c <<= 1; }
        int ci = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int tc = __shfl_up_sync(0xffffffffu, ci, o);
            if (lane >= o) ci += tc;
        }
        if (lane < nB) S.capOff[lane] = ci - c;
        if (lane == nB - 1) S.capOff[nB] = ci;
    }
    if (po.hdr) {
        // (no barrier needed before: S.lines / S.smr / S.env were completed behind barriers of steps 1-3)
        // batched path: the scaled lines, the SMRs and the envelope values leave for k_bitalloc / k_pvq
        for (int i = tid; i < H; i += ENC_NT) po.lines[(size_t)(po.sf * PAC_HALF_SHORT + i) * 32] = S.lines[i];
        if (tid < nB) {
            po.hdr->smr[po.sf * nB + tid] = S.smr[tid];
            if (sbr && bands.omitted[tid]) {
                int k = 0;
                for (int q = 0; q < tid; ++q) k += bands.omitted[q] ? 1 : 0;
                if (k < 4) po.hdr->env[k] = (S.env[tid] * up) / (double)bands.nLines[tid];
            }
        }
        if (tid == 0) po.hdr->scale[po.sf] = (unsigned char)scale;
        __syncthreads();
        return 0;
    }
    // PVQ ping-pong workspaces: a long frame reuses its dead sample / FFT buffers; a short
    // sub-frame must leave xs intact for its siblings and uses the tail of lines[]
    double* wsA = isLong ? S.xs : S.lines + 256;
    double* wsB = isLong ? S.fb : S.lines + 512;
    __syncthreads();
    if (warp != 0) {
        // while warp 0 runs the (serial) allocation: gain and normalised copy of every band - they do
        // not depend on the allocation.  SBR-omitted bands carry one value: mean |fftLines| (codec.py:508-510)
        if (dev->useVQ && S.capOff[nB] <= (isLong ? 2048 : 256)) {
            for (int b = warp - 1; b < nB; b += ENC_NT / 32 - 1) {
                const bool env = sbr && bands.omitted[b];
                const int L = env ? 1 : bands.nLines[b];
                const double* x = S.lines + bands.lower[b];
                const double envVal = env ? (S.env[b] * up) / (double)bands.nLines[b] : 0.0;
                double ss = 0.0;
                if (env) ss = lane == 0 ? envVal * envVal : 0.0;
                else for (int i = lane; i < L; i += 32) ss += x[i] * x[i];
                const double gain = sqrt(warp_sum(ss));
                if (gain != 0.0)
                    for (int i = lane; i < L; i += 32) wsA[S.capOff[b] + i] = (env ? envVal : x[i]) / gain;
                if (lane == 0) S.bandEnc[b].gain = gain;
            }
        }
    } else {
        int nfb = H;
        if (!sbr) {
            if (cur) nfb = (int)(1.45 * H);
            if (lastOrNext) nfb = (int)(0.85 * nfb);
        }
        double budget = dev->targetBitsPerSample * (double)nfb;
        if (dev->useVQ) budget -= (double)dev->nScaleBits;
        else budget -= (double)(dev->nScaleBits * (nB + 1));
        budget -= (double)(dev->nMantSizeBits * nB);
        warp_bit_alloc(budget, dev->maxMantBits, nB, S.nlEff, S.smr, S.tmp, S.ba);
        __syncwarp();
        // per-band private bit buffers (words): exclusive scan over the bands
        int w = 0;
        if (lane < nB) w = (S.ba[lane] * S.nlEff[lane] + 31) / 32 + 3;      // slack: a 64-bit put may straddle
        int wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int tw = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += tw;
        }
        if (lane < nB) S.wordOff[lane] = wi - w;
        if (lane == nB - 1) {
            S.wordOff[nB] = wi;
            if (wi > BAND_WORDS || S.capOff[nB] > (isLong ? 2048 : 256)) atomicOr(err, PVQ_ERR_STACK);
        }
        if (lane == 0) S.misc[6] = 0;                      // node pool fill
    }

    PHASE(3);
    // (5) quantise each coded band into its private bit buffer
    for (int i = tid; i < BAND_WORDS; i += ENC_NT) S.bandbits[i] = 0u;
    __syncthreads();
    if (S.wordOff[nB] > BAND_WORDS || S.capOff[nB] > (isLong ? 2048 : 256)) return 0;
    PvqTab pt = {dev->pvqN, dev->pvqRowOff, dev->log2tab, dev->pvqKW, dev->thetaTab};
    int* yws = reinterpret_cast<int*>(S.mk);              // masked threshold is dead from here on
    if (dev->useVQ) {
        // roots: one thread per band (gain and normalised copy are in place): bit split, root node
        if (tid < nB && S.ba[tid]) {
            const int b = tid;
            const bool env = sbr && bands.omitted[b];
            const int L = env ? 1 : bands.nLines[b], R = env ? S.ba[b] : S.ba[b] * bands.nLines[b];
            int bg, bs;
            gs_alloc(pt, R, L, &bg, &bs);
            const double gain = S.bandEnc[b].gain;
            int root = -1;
            if (gain != 0.0 && bs != 0) {
                root = atomicAdd(&S.misc[6], 1);
                if (root < ENC_MAX_NODES) {
                    TNode r;
                    r.val = 0ull; r.bits = bs; r.K = -2; r.off = (short)S.capOff[b]; r.n = (short)L;
                    r.kid[0] = r.kid[1] = -1;
                    r.capLog2 = (unsigned char)(31 - __clz(S.capOff[b + 1] - S.capOff[b]));
                    r.buf = 0; r.band = (unsigned char)b; r.w = 0; r.logSlot = -1; r.pad = 0;
                    S.nodes[root] = r;
                } else {
                    atomicOr(err, PVQ_ERR_STACK);
                    root = -1;
                }
            }
            BandEnc be;
            be.gain = gain; be.bg = bg; be.bs = bs; be.root = root; be.L = L;
            S.bandEnc[b] = be;
        }
        PHASE(4);
        // levels: the nodes of one tree depth of all bands - vector work on one warp per node,
        // then the scalar planning of every node of the level in parallel threads
        double* sideNorm = S.mk;                            // [ENC_MAX_NODES]; S.mk is free until the leaves step
        __syncthreads();
        for (int lvl0 = 0, lvl1 = min(S.misc[6], ENC_MAX_NODES); lvl0 < lvl1;) {
            for (int i = lvl0 + warp; i < lvl1; i += ENC_NT / 32) warp_split_node(S.nodes, i, sideNorm, wsA, wsB);
            __syncthreads();
            for (int i = lvl0 + tid; i < lvl1; i += ENC_NT) {
                int werr = 0;
                thread_plan_node(pt, S.nodes, i, sideNorm, &S.misc[6], ENC_MAX_NODES, &werr);
                if (werr) atomicOr(err, werr);
            }
            __syncthreads();
            lvl0 = lvl1;
            lvl1 = min(S.misc[6], ENC_MAX_NODES);
        }
        __syncthreads();
        // layout: one thread per band: field positions, theta and gain fields
        if (tid < nB) {
            const int b = tid;
            int used = 0;
            if (S.ba[b] && S.bandEnc[b].gain != 0.0) {
                int werr = 0;
                FieldLog flog = {S.logVal + (size_t)b * S.logCap, S.logW + (size_t)b * S.logCap, S.logCnt + b, S.logCap};
                used = thread_layout_band(S.nodes, S.bandEnc[b], S.bandbits, S.wordOff[b] * 32, &werr,
                                          S.logVal ? &flog : nullptr);
                if (werr) atomicOr(err, werr);
            }
            S.used[b] = used;
            if (used == 0) S.ba[b] = 0;                     // coder/codec.py:352-353
        }
        __syncthreads();
        PHASE(5);
        // leaves: one thread per leaf: pulse search + Fischer index
        cta_quantize_leaves(pt, S.nodes, min(S.misc[6], ENC_MAX_NODES), wsA, wsB, yws, S.bandbits, S.logVal, S.logCap,
                            tid, ENC_NT, err);
    } else {
        for (int b = warp; b < nB; b += ENC_NT / 32) {
            const int lo = bands.lower[b], n = bands.nLines[b], bb = S.ba[b];
            double m = 0.0;
            for (int i = lane; i < n; i += 32) m = fmax(m, fabs(S.lines[lo + i]));
#pragma unroll
            for (int o = 16; o; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
            const int sf = scale_factor(m, dev->nScaleBits, bb);       // coder/codec.py:369-371
            if (bb)
                for (int i = lane; i < n; i += 32)
                    bw_put_at(S.bandbits + S.wordOff[b], i * bb,
                              mantissa_bfp(S.lines[lo + i], sf, dev->nScaleBits, bb), bb);
            if (lane == 0) {
                S.sfb[b] = sf;
                S.used[b] = bb * n;
            }
        }
    }
    __syncthreads();

    PHASE(6);
    // (6) frame header + band payload positions (coder/pacfile.py:363-447): every field has a fixed
    //     width, so one warp places them with a scan over the bands (lane = band)
    if (warp == 0) {
        const int pos0 = *bpos, nS = dev->nScaleBits, nM = dev->nMantSizeBits;
        const bool vq = dev->useVQ;
        const bool act = lane < nB;
        const int bab = act ? S.ba[lane] : 0;
        const int u = bab ? S.used[lane] : 0;
        const int hdrBits = vq ? 0 : nM + nS;              // block-FP interleaves [ba, sf] with the mantissas
        int len = act ? hdrBits + u : 0, inc = len;
        int decl = act ? nM + nS + (bab ? ((sbr && bands.omitted[lane]) ? bab : bab * bands.nLines[lane]) : 0) : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) decl += __shfl_xor_sync(0xffffffffu, decl, o);
        const int base = pos0 + nS + (vq ? nB * nM : 0);
        const int mine = base + inc - len;                 // start of this band's [header,] payload
        if (lane == 0) bw_put_at(S.blockbuf, pos0, (unsigned long long)scale, nS);
        if (act) {
            const unsigned long long code = bab ? bab - 1 : 0;
            if (vq) {
                bw_put_at(S.blockbuf, pos0 + nS + lane * nM, code, nM);
            } else {
                bw_put_at(S.blockbuf, mine, code, nM);
                bw_put_at(S.blockbuf, mine + nM, (unsigned long long)S.sfb[lane], nS);
            }
            S.capOff[lane] = mine + hdrBits;              // reuse: destination bit position of band b
        }
        const int total = __shfl_sync(0xffffffffu, inc, 31);
        if (lane == 0) {
            *bpos = base + total;
            S.misc[2] = nS + decl;
            S.misc[5] = scale;
        }
    }
    __syncthreads();
    PHASE(7);
    // (7) merge the band buffers into the block bitstream (disjoint bit ranges)
    for (int b = warp; b < nB; b += ENC_NT / 32) {
        const int u = S.ba[b] ? S.used[b] : 0;
        const unsigned* src = S.bandbits + S.wordOff[b];
        for (int i = lane; i * 32 < u; i += 32) {
            int nb = min(32, u - i * 32);
            bw_put_at(S.blockbuf, S.capOff[b] + i * 32, (unsigned long long)(src[i] >> (32 - nb)), nb);
        }
    }
    __syncthreads();
    PHASE(8);
    return S.misc[2];
}


// Encode one frame of N samples at fx (shared) and append its fields to the block
// bitstream at bit position *bpos.  Returns the DECLARED size in bits
// (getNumBytesNeeded, coder/pacfile.py:342-361).  sbr: long frame of an SBR stream.
// T: arithmetic of the transform and psychoacoustic stages (double = parity path, float = fast
// path); bit allocation and quantisation are always fp64 + integer.
template <int N, typename T>
__device__ int encode_frame(const PacDev* __restrict__ dev, const EncSmem& S, const T* fx, int shape, bool cur,
                            bool lastOrNext, bool sbr, int* bpos, int* err, EncPreOut po = {nullptr, nullptr, 0, 0}) {
    constexpr int H = N / 2, Q = N / 4;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const PacBands& bands = (N == PAC_NLONG) ? dev->longBands : dev->shortBands;
    const int nB = bands.nBands;
    const PsyConst<T> pc = psy_const<N, T>(dev);
    const PacTables<T>& tabsT = psy_tabs<T>(dev);
    const T* winTab = (N == PAC_NLONG) ? tabsT.win_long : tabsT.win_short;
    const cplx<T>* tw = reinterpret_cast<const cplx<T>*>(N == PAC_NLONG ? tabsT.tw_mdct_long : tabsT.tw_mdct_short);
    const cplx<T>* rootsQ = reinterpret_cast<const cplx<T>*>(tabsT.tw_fft[N == PAC_NLONG ? 2 : 0]);
    cplx<T>* fbuf = reinterpret_cast<cplx<T>*>(S.fb);
    T* mkT = reinterpret_cast<T*>(S.mk);

#define PH_IS_LONG (N == PAC_NLONG)
    PHASE_T0();
#undef PH_IS_LONG
    // (1) window + MDCT (coder/codec.py:301-305)
    mdct_fold<T, N, 1, ENC_NT>(fx, 0, winTab, nullptr, shape, tw, fbuf, tid);
    __syncthreads();
    cta_fft<T, Q, 1, ENC_NT>(fbuf, rootsQ, tid);
    mdct_post<T, N, 1, ENC_NT, double>(fbuf, tw, (T)2 / (T)N, S.lines, tid);
    __syncthreads();
    double mx = 0.0;
    for (int i = tid; i < H; i += ENC_NT) mx = fmax(mx, fabs(S.lines[i]));
    double maxLine = block_max(mx, S.red, tid);

    PHASE(0);
    // (2) Hann FFT magnitude (shared by the psychoacoustic model and the SBR envelope)
    hann_rfft_mag<N, ENC_NT, T>(fx, pc, fbuf, mkT, tid);
    if (sbr) {                                    // coder/codec.py:460-473
        double m2 = 0.0;
        for (int k = tid; k <= H; k += ENC_NT) m2 = fmax(m2, (double)mkT[k]);
        double maxFFT = block_max(m2, S.red, tid) / (double)H;
        maxLine = fmax(maxLine, maxFFT);
        // per omitted band: sum of fftLines = |X|/H over the band (scale applied later, exact)
        for (int b = warp; b < nB; b += ENC_NT / 32) {
            if (!bands.omitted[b]) continue;
            double a = 0.0;
            for (int i = lane; i < bands.nLines[b]; i += 32) a += (double)mkT[bands.lower[b] + i] / (double)H;
            a = warp_sum(a);
            if (lane == 0) S.env[b] = a;
        }
        __syncthreads();
    }
    const int scale = scale_factor(maxLine, dev->nScaleBits, 5);      // codec.py:308-311 (default nMantBits)
    const double up = (double)(1 << scale);
    for (int i = tid; i < H; i += ENC_NT) S.lines[i] *= up;
    for (int k = tid; k <= H; k += ENC_NT) {
        T m = mkT[k];
        mkT[k] = pc.norm * (m * m);
    }
    __syncthreads();

    PHASE(1);
    // (3) psychoacoustic model -> SMR per band
    const MaskerList<T> ml = masker_list<T>(reinterpret_cast<T*>(S.fb));
    const int nMask = po.prune ? pick_maskers<N, ENC_NT, T, true>(mkT, pc, ml, tid) : pick_maskers<N, ENC_NT, T, false>(mkT, pc, ml, tid);
    PHASE(9);
    if (po.prune) {
        // candidate list: the frame samples of a long block are dead by now; a short sub-frame has every line
        // exact (no list) and must leave them to its siblings.  Band keys: tmp + red (384 bytes)
        band_smr_pruned<N, ENC_NT, T>(pc, ml, nMask, S.lines, scale, mkT, bands, S.smr, S.tmp,
                                      N == PAC_NLONG ? reinterpret_cast<int*>(S.xs) : nullptr, tid);
    } else {
        masked_threshold<N, ENC_NT, T>(pc, ml, nMask, mkT, tid);
        PHASE(10);
        band_smr<ENC_NT, T>(S.lines, scale, mkT, bands, S.smr, tid);
    }

    PHASE(2);
    const int flags = (N == PAC_NLONG ? 1 : 0) | (sbr ? 2 : 0) | (cur ? 4 : 0) | (lastOrNext ? 8 : 0);
    const EncLog lg = {S.logVal, S.logW, S.logCnt, S.logCap};
    return encode_back(dev, S.raw, lg, flags, scale, bpos, err, po);
}
