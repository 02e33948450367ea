// Fused stream encoder: PCM -> .pac bytes, all stages on device.
//
//   k_transient   per PCM block: peak-to-average transient verdict
//                 (coder/detect_transients.py:5-23 on [block || zeros], pacfile.py:717-731)
//   k_encode      one CTA per written channel-block: window + MDCT, overall scale,
//                 psychoacoustic SMRs, bit allocation, PVQ / block-FP quantisation,
//                 MSB-first bit packing into a fixed-stride scratch record
//                 (coder/codec.py:266-380, :426-555; coder/pacfile.py:363-447)
//   k_sizes/scan/k_gather  byte layout of the variable-length records with their
//                 4-byte length prefixes (coder/pacfile.py:550-608)
//
// Stream layout: stream s owns nb[s] PCM blocks of 1024 samples per channel,
// planar int16 at element offset (blkOff[s]*C + c*nb[s])*1024, zero padded to a
// whole block.  It produces nb[s]+2 written blocks: 0..nb-1, the last block AGAIN
// (pacfile.py:725-743), then Close()'s zero block (pacfile.py:612-625).
#include <cub/device/device_scan.cuh>
#include <algorithm>
#include "encode_frame.cuh"
#define ENC_BATCH_CB 131072      // channel-blocks per k_encode / k_bitalloc / k_pvq / k_pack round (bounds the scratch: ~52 KB each; larger batches amortise the long band-16 chains)
#include "pac_api.h"

// ---------------------------------------------------------------- transient detector
// One WARP per PCM block, all channels (4 blocks per CTA; no CTA barrier).  Exact integer form of
// mymax/avg > 4.5 with avg = mean(|x[:, :n]|), n = min(max_c(argmax_c)+500, 2048):
//   2*cmax_c*(C*n) > 9*sum|c|   (the second half of the look-ahead is always zero).
// A lane owns 32 consecutive samples of each channel (four 16-byte loads when the buffer allows), kept in registers
// for both passes (the first 2 channels; further ones are read again in the second pass).
#define TR_WARPS 4
__global__ void __launch_bounds__(TR_WARPS * 32)
k_transient(const short* __restrict__ pcm, const int* __restrict__ blkOff, const int* __restrict__ nb, int nStreams,
            int C, int totBlk, unsigned char* __restrict__ det) {
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * TR_WARPS + (threadIdx.x >> 5);     // global PCM block index
    if (g >= totBlk) return;
    int lo = 0, hi = nStreams;                // stream with blkOff[s] <= g < blkOff[s+1]
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (blkOff[mid] <= g) lo = mid; else hi = mid;
    }
    const int s = lo, k = g - blkOff[s], nbs = nb[s];
    const bool vec = (((size_t)pcm) & 15) == 0;
    uint4 held[2][4];
    int cmaxAll[8];
    int nmax = 0;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        if (c >= C) break;
        const short* p = pcm + ((long long)blkOff[s] * C + (long long)c * nbs + k) * 1024;
        int best = -1, arg = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint4 v4;
            if (vec) v4 = __ldg(reinterpret_cast<const uint4*>(p) + lane * 4 + j);
            else {
                short t[8];
                for (int e = 0; e < 8; ++e) t[e] = p[lane * 32 + j * 8 + e];
                v4 = *reinterpret_cast<uint4*>(t);
            }
            if (c < 2) held[c][j] = v4;
            const short* q = reinterpret_cast<const short*>(&v4);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                int v = q[e];
                v = (v < 0 ? -v : v) & 0x7fff;
                if (v > best) { best = v; arg = lane * 32 + j * 8 + e; }      // first maximum (ascending index)
            }
        }
        for (int o = 16; o; o >>= 1) {
            int ob = __shfl_xor_sync(0xffffffffu, best, o), oa = __shfl_xor_sync(0xffffffffu, arg, o);
            if (ob > best || (ob == best && oa < arg)) { best = ob; arg = oa; }
        }
        cmaxAll[c] = best;
        nmax = max(nmax, arg + 500);
    }
    const int n = min(nmax, 2048);
    const int nn = min(n, 1024);
    unsigned long long sum = 0;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        if (c >= C) break;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint4 v4;
            if (c < 2) v4 = held[c][j];
            else {
                const short* p = pcm + ((long long)blkOff[s] * C + (long long)c * nbs + k) * 1024;
                short t[8];
                for (int e = 0; e < 8; ++e) t[e] = p[lane * 32 + j * 8 + e];
                v4 = *reinterpret_cast<uint4*>(t);
            }
            const short* q = reinterpret_cast<const short*>(&v4);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                int v = q[e];
                if (lane * 32 + j * 8 + e < nn) sum += (unsigned long long)((v < 0 ? -v : v) & 0x7fff);
            }
        }
    }
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) {
        bool t = false;
        if (sum != 0) {
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                if (c >= C) break;
                if (2ull * (unsigned long long)cmaxAll[c] * (unsigned long long)(C * n) > 9ull * sum) t = true;
            }
        }
        det[g] = t ? 1 : 0;
    }
}

// ---------------------------------------------------------------- block kernel
struct EncArgs {
    const short* pcm;
    const int* blkOff;     // [S+1] PCM block prefix
    const int* nb;         // [S]
    const int* wbOff;      // [S+1] written-block prefix (nb+2 each)
    const unsigned char* det;
    int nStreams;
    unsigned char* rec;    // [nChBlocks][REC_STRIDE]
    int* recBytes;         // [nChBlocks] declared byte size
    unsigned char* zeroSub;// [nChBlocks] 1 if a short sub-frame of this channel is all zero
    int* err;              // [1] OR of soft error flags
    // batched VQ back end (nullptr: everything in k_encode)
    EncPre* pre;           // [nBatch]
    double* preLines;      // [nBatch / 32][1024][32]
    int* perm;             // [nBatch] dense position -> block of the batch
    int* counts;           // [2] long / transient blocks placed so far
    int cb0, nBatch;
    const int* wbStream;   // [totWb] stream of every written block (k_wbmap), or nullptr
    int variant;           // pac_set_option(PAC_OPT_VARIANT): 0 current kernels; bit 0: the 256-thread k_encode<T, true> front end;
                           // bit 1: with the exhaustive masked threshold and scalar loads (A/B timing, same results)
};

#include "encode_front.cuh"     // k_encode_front (needs EncArgs)

// stream index of every written block (one binary search per block here instead of one per channel-block CTA)
__global__ void k_wbmap(const int* __restrict__ wbOff, int nStreams, int totWb, int* __restrict__ wbStream) {
    const int wb = blockIdx.x * blockDim.x + threadIdx.x;
    if (wb >= totWb) return;
    int lo = 0, hi = nStreams;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (wbOff[mid] <= wb) lo = mid; else hi = mid;
    }
    wbStream[wb] = lo;
}

// FRONT: the batched path (a.pre set): stops after the psychoacoustic model, needs enc_front_smem_bytes()
// only and runs four CTAs per SM.
#ifndef ENC_FRONT_MINB
#define ENC_FRONT_MINB 4
#endif
template <typename T, bool FRONT>
__global__ void __launch_bounds__(ENC_NT, FRONT ? ENC_FRONT_MINB : 3)
k_encode(const PacDev* __restrict__ dev, EncArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    EncSmem S;
    enc_smem_carve(smem_raw, S);
    S.logVal = nullptr; S.logW = nullptr; S.logCnt = nullptr; S.logCap = 0;

    const int tid = threadIdx.x;
#ifdef PAC_PHASE_TIMING
    const long long t_k0 = clock64();
#endif
    const int C = dev->nChannels;
    const int cb = blockIdx.x + a.cb0;        // written channel-block index
    const int wb = cb / C, c = cb - wb * C;   // written block (global), channel
    int lo = 0, hi = a.nStreams;
    if (FRONT && a.wbStream) lo = a.wbStream[wb];            // k_wbmap did the search once per written block
    else
        while (hi - lo > 1) {
            int mid = (lo + hi) >> 1;
            if (a.wbOff[mid] <= wb) lo = mid; else hi = mid;
        }
    const int s = lo, j = wb - a.wbOff[s], nbs = a.nb[s];
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
    T* xs = reinterpret_cast<T*>(S.xs);
    if (FRONT && !(a.variant & 2) && (((size_t)a.pcm) & 15) == 0) {
        // eight samples (one 16-byte load) per thread, division-free exact conversion
        const int i0 = tid * 8;
        const int blk = i0 < 1024 ? priorBlk : dataBlk;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (blk >= 0) v = __ldg(reinterpret_cast<const uint4*>(base + (long long)blk * 1024 + (i0 & 1023)));
        const short* sv = reinterpret_cast<const short*>(&v);
#pragma unroll
        for (int e = 0; e < 8; ++e) xs[i0 + e] = (T)pcm_to_double_fast(sv[e]);
    } else
    for (int i = tid; i < 2048; i += ENC_NT) {
        int blk = i < 1024 ? priorBlk : dataBlk;
        xs[i] = blk < 0 ? (T)0 : (T)pcm_to_double(base[(long long)blk * 1024 + (i & 1023)]);
    }
    if (!FRONT) for (int i = tid; i < REC_WORDS; i += ENC_NT) S.blockbuf[i] = 0u;
    if (tid == 0) {
        S.misc[3] = 0;
        S.misc[4] = 0;
    }
    __syncthreads();
    if (tid == 0 && FRONT) {                  // dense order for k_pvq: its warps hold blocks of one kind
        const int p = cur ? a.nBatch - 1 - atomicAdd(&a.counts[1], 1) : atomicAdd(&a.counts[0], 1);
        a.perm[p] = blockIdx.x;
        S.misc[8] = p;
    }
    if (tid == 0 && !FRONT) {                 // 3 flag bits (coder/pacfile.py:575-577)
        BitW bw = {S.blockbuf, 0};
        bw_put(bw, last, 1);
        bw_put(bw, cur, 1);
        bw_put(bw, nxt, 1);
        S.misc[3] = bw.pos;
    }
    __syncthreads();
#ifdef PAC_PHASE_TIMING
    if (tid == 0) atomicAdd(&g_pac_phase[11 + (cur ? 16 : 0)], (unsigned long long)(clock64() - t_k0));   // prologue: stream search + PCM load
#endif
    int declared = 4;                         // "+4 for block switching" (pacfile.py:560)
    EncPreOut po = {nullptr, nullptr, 0, 0};
    if (FRONT) { po.hdr = a.pre + blockIdx.x; po.lines = a.preLines + epre_line_index(S.misc[8], 0); po.prune = !(a.variant & 2); }
    if (!cur) {
        int shape = (last && nxt) ? PAC_WIN_STARTSTOP : last ? PAC_WIN_STOP : nxt ? PAC_WIN_START : PAC_WIN_SINE;
        declared += encode_frame<PAC_NLONG, T>(dev, S, xs, shape, false, last || nxt, dev->useSBR != 0, &S.misc[3], a.err, po);
    } else {
        // zero sub-frame check (coder/pacfile.py:503-506): the block is dropped if ANY channel has one
        for (int sf = 0; sf < PAC_SUBFRAMES; ++sf) {
            bool nz = false;
            for (int i = tid; i < PAC_NSHORT; i += ENC_NT) nz |= xs[448 + 128 * sf + i] != (T)0;
            if (!__syncthreads_or(nz) && tid == 0) S.misc[4] = 1;
        }
        for (int sf = 0; sf < PAC_SUBFRAMES; ++sf) {
            po.sf = sf;
            declared += encode_frame<PAC_NSHORT, T>(dev, S, xs + 448 + 128 * sf, PAC_WIN_SINE, true, last || nxt, false,
                                                 &S.misc[3], a.err, po);
        }
    }
    __syncthreads();
    if (FRONT) {                              // the record is assembled by k_pack
        if (tid == 0) {
            po.hdr->flags = (unsigned char)((last ? 4 : 0) | (cur ? 2 : 0) | (nxt ? 1 : 0));
            po.hdr->zeroSub = (unsigned char)S.misc[4];
            po.hdr->p = S.misc[8];
        }
        return;
    }
    const int nBytes = (declared + 7) / 8;
    unsigned* out = reinterpret_cast<unsigned*>(a.rec + (long long)cb * REC_STRIDE);
    const int nW = min(REC_WORDS, (nBytes + 3) / 4);
    for (int i = tid; i < nW; i += ENC_NT) out[i] = __byte_perm(S.blockbuf[i], 0, 0x0123);
    if (tid == 0) {
        a.recBytes[cb] = nBytes;
        a.zeroSub[cb] = (unsigned char)S.misc[4];
        if (nBytes > REC_STRIDE || S.misc[3] > declared) atomicOr(a.err, PVQ_ERR_STACK);
    }
}

// ---------------------------------------------------------------- batched VQ back end
// k_bitalloc: one warp per channel-block runs the water-filling allocation of its frames
// (coder/codec.py:288-299, :446-453; coder/bitalloc.py:62-145) from the SMRs k_encode left.  Serial in
// the passes, so it runs here at full occupancy instead of on one warp of a waiting 8-warp CTA.
#define BA_WARPS 8
__global__ void __launch_bounds__(BA_WARPS * 32)
k_bitalloc(const PacDev* __restrict__ dev, EncArgs a) {
    __shared__ double s_smr[BA_WARPS][32], s_tmp[BA_WARPS][32];
    __shared__ int s_nl[BA_WARPS][32], s_ba[BA_WARPS][32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int local = blockIdx.x * BA_WARPS + warp;
    if (local >= a.nBatch) return;
    EncPre* h = a.pre + local;
    const bool last = (h->flags >> 2) & 1, cur = (h->flags >> 1) & 1, nxt = h->flags & 1;
    const PacBands& bands = cur ? dev->shortBands : dev->longBands;
    const int nB = bands.nBands, nF = cur ? PAC_SUBFRAMES : 1, H = cur ? PAC_HALF_SHORT : PAC_HALF_LONG;
    const bool sbr = !cur && dev->useSBR;
    int nfb = H;
    if (!sbr) {
        if (cur) nfb = (int)(1.45 * H);
        if (last || nxt) nfb = (int)(0.85 * nfb);
    }
    double budget = dev->targetBitsPerSample * (double)nfb;
    budget -= (double)dev->nScaleBits;
    budget -= (double)(dev->nMantSizeBits * nB);
    if (lane < nB) s_nl[warp][lane] = (sbr && bands.omitted[lane]) ? 1 : bands.nLines[lane];
    for (int f = 0; f < nF; ++f) {
        if (lane < nB) s_smr[warp][lane] = h->smr[f * nB + lane];
        __syncwarp();
        warp_bit_alloc(budget, dev->maxMantBits, nB, s_nl[warp], s_smr[warp], s_tmp[warp], s_ba[warp]);
        __syncwarp();
        if (lane < nB) h->ba[f * nB + lane] = (unsigned char)s_ba[warp][lane];
        __syncwarp();
    }
}

// k_pvq: one thread per (channel-block, band).  A CTA holds one band slot of PQ_NT consecutive blocks
// of the dense order; CTAs are ordered widest band first.  Slot q of a long block is band nB-1-q; of
// a transient block, sub-frame q / 6, band 5 - q % 6.
#ifndef PQ_NT
#define PQ_NT 256         // the tile the bit-allocation sort runs over (measured 128 / 256 / 384 / 768: 8.5 / 8.07 / 8.02 / 8.18 ms)
#endif
__device__ __forceinline__ void epre_band_layout(const PacBands& bands, int f, int b, int* base, int* cap, int* wbase) {
    int c0 = 1, bs = 0, all = 0, wb = 0, wall = 0;
    while (c0 < bands.nLines[b]) c0 <<= 1;
    for (int j = 0; j < bands.nBands; ++j) {
        int c = 1;
        while (c < bands.nLines[j]) c <<= 1;
        const int w = bands.nLines[j] / 2 + 3;
        if (j < b) { bs += c; wb += w; }
        all += c; wall += w;
    }
    *base = bs + f * all;
    *cap = c0;
    *wbase = wb + f * wall;
}
#ifndef PQ_SORT
#define PQ_SORT 1
#endif
#ifndef PQ_MINB
#define PQ_MINB 4          // 64 registers, 1024 threads per SM (the compiler takes 158 when free; measured 2 / 3 / 4 / 5 CTAs of 256 threads: 9.4 / 8.1 / 7.5 / 8.4 ms)
#endif
__global__ void __launch_bounds__(PQ_NT, PQ_MINB)
k_pvq(const PacDev* __restrict__ dev, EncArgs a, int nChunks, double* __restrict__ ws, int* __restrict__ yws,
      unsigned* __restrict__ words) {
    __shared__ int s_off[PAC_PVQ_LMAX + 2];                  // row offsets of the N(L,K) table: read by every pvq_N call
    for (int i = threadIdx.x; i < PAC_PVQ_LMAX + 2; i += PQ_NT) s_off[i] = dev->pvqRowOff[i];
    __syncthreads();
    const int q = blockIdx.x / nChunks, chunk = blockIdx.x - q * nChunks;
    const int nLong = a.counts[0], nShort = a.counts[1];
    const int nBL = dev->longBands.nBands, nBS = dev->shortBands.nBands;
    // As k_dequant: the 128 blocks of the CTA are dealt to its threads in the order of this band's bit allocation, within
    // the long / transient parts of the tile; a thread keeps the workspace column of its own place, the block it
    // quantises (its lines, its word buffer) is permuted.
#if PQ_SORT
    __shared__ int s_cnt[3][34];
    __shared__ unsigned short s_order[PQ_NT];
    {
        const int tile0 = chunk * PQ_NT, pos0 = tile0 + threadIdx.x;
        const int gap0 = min(max(nLong - tile0, 0), PQ_NT), tr0 = min(max(a.nBatch - nShort - tile0, 0), PQ_NT);
        const int cls = threadIdx.x < gap0 ? 0 : (threadIdx.x < tr0 ? 1 : 2);
        int key = 33;                                          // nothing to do: last of its part
        if (cls != 1 && pos0 < a.nBatch && !(q >= nBL && cls == 0)) {
            const EncPre* h0 = a.pre + a.perm[pos0];
            const bool cur0 = (h0->flags >> 1) & 1;
            const int nB0 = cur0 ? nBS : nBL;
            if (q < (cur0 ? PAC_SUBFRAMES * nB0 : nB0)) {
                const int f0 = q / nB0, b0 = nB0 - 1 - (q - f0 * nB0);
                key = min((int)h0->ba[f0 * nB0 + b0], 32);
            }
        }
        if (threadIdx.x < 3 * 34) (&s_cnt[0][0])[threadIdx.x] = 0;
        __syncthreads();
        const int rank = atomicAdd(&s_cnt[cls][key], 1);
        __syncthreads();
        if (threadIdx.x < 3) {
            int acc = threadIdx.x == 0 ? 0 : (threadIdx.x == 1 ? gap0 : tr0);
            for (int j = 0; j < 34; ++j) { const int c = s_cnt[threadIdx.x][j]; s_cnt[threadIdx.x][j] = acc; acc += c; }
        }
        __syncthreads();
        s_order[s_cnt[cls][key] + rank] = (unsigned short)threadIdx.x;
        __syncthreads();
    }
    const int wsPos = chunk * PQ_NT + threadIdx.x;
    const int pos = chunk * PQ_NT + s_order[threadIdx.x];
#else
    const int pos = chunk * PQ_NT + threadIdx.x, wsPos = pos;
#endif
    if (pos >= a.nBatch || (pos >= nLong && pos < a.nBatch - nShort)) return;
    if (q >= nBL && pos < nLong) return;
    EncPre* h = a.pre + a.perm[pos];
    const bool cur = (h->flags >> 1) & 1;
    const PacBands& bands = cur ? dev->shortBands : dev->longBands;
    const int nB = bands.nBands;
    if (q >= (cur ? PAC_SUBFRAMES * nB : nB)) return;
    const int f = q / nB, b = nB - 1 - (q - f * nB), slot = f * nB + b;
    const int bb = h->ba[slot];
    if (!bb) { h->used[slot] = 0; return; }
    const bool env = !cur && dev->useSBR && bands.omitted[b];
    const int nl = bands.nLines[b], L = env ? 1 : nl, R = env ? bb : bb * nl;
    int base, cap, wbase;
    epre_band_layout(bands, f, b, &base, &cap, &wbase);
    int k = 0;
    if (env) for (int j = 0; j < b; ++j) k += bands.omitted[j] ? 1 : 0;
    const size_t grp = (size_t)(pos >> 5), ln = (size_t)(pos & 31);               // the block: lines in, words out
    const size_t wgrp = (size_t)(wsPos >> 5), wln = (size_t)(wsPos & 31);         // the thread's own workspace column
    const double* x = a.preLines + epre_line_index(pos, f * PAC_HALF_SHORT + bands.lower[b]);
    double* wsA = ws + (wgrp * (2 * EPRE_WS_PER_CB) + (size_t)base) * 32 + wln;
    double* wsB = wsA + (size_t)EPRE_WS_PER_CB * 32;
    int* y = yws + (wgrp * EPRE_WS_PER_CB + (size_t)base) * 32 + wln;
    unsigned* w = words + (grp * EPRE_WORDS_PER_CB + (size_t)wbase) * 32 + ln;
    PvqTab pt = {dev->pvqN, s_off, dev->log2tab, dev->pvqKW, dev->thetaTab};
    int werr = 0;
    const int used = thread_quantize_gain_shape<32>(pt, x, env, (env && k < 4) ? h->env[k] : 0.0, L, R, wsA, wsB, y, cap, w,
                                                    nl / 2 + 3, &werr);
    if (werr) atomicOr(a.err, werr);
    h->used[slot] = (unsigned short)used;
    if (used == 0) h->ba[slot] = 0;                       // coder/codec.py:352-353
}

// k_pack: one warp per channel-block assembles the record (coder/pacfile.py:363-402, :550-608):
// 3 flag bits, then per frame the overall scale, the bit allocations and the band payloads.
#define PK_WARPS 4
__global__ void __launch_bounds__(PK_WARPS * 32)
k_pack(const PacDev* __restrict__ dev, EncArgs a, const unsigned* __restrict__ words) {
    __shared__ unsigned s_buf[PK_WARPS][REC_WORDS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int local = blockIdx.x * PK_WARPS + warp;
    if (local >= a.nBatch) return;
    unsigned* buf = s_buf[warp];
    for (int i = lane; i < REC_WORDS; i += 32) buf[i] = 0u;
    __syncwarp();
    const EncPre* h = a.pre + local;
    const int pos = h->p;
    const bool cur = (h->flags >> 1) & 1;
    const PacBands& bands = cur ? dev->shortBands : dev->longBands;
    const int nB = bands.nBands, nF = cur ? PAC_SUBFRAMES : 1, nS = dev->nScaleBits, nM = dev->nMantSizeBits;
    const bool sbr = !cur && dev->useSBR;
    if (lane == 0) bw_put_at(buf, 0, (unsigned long long)(h->flags & 7), 3);
    int p0 = 3, declared = 4;                              // "+4 for block switching" (pacfile.py:560)
    const unsigned* wcol = words + ((size_t)(pos >> 5) * EPRE_WORDS_PER_CB) * 32 + (pos & 31);
    for (int f = 0; f < nF; ++f) {
        const bool act = lane < nB;
        const int slot = f * nB + lane;
        const int bab = act ? h->ba[slot] : 0;
        const int u = bab ? h->used[slot] : 0;
        int inc = u;
        int decl = act ? nM + nS + (bab ? ((sbr && bands.omitted[lane]) ? bab : bab * bands.nLines[lane]) : 0) : 0;   // getNumBytesNeeded
        int wofs = act ? bands.nLines[lane] / 2 + 3 : 0, winc = wofs;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o), tw = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) { inc += t; winc += tw; }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) decl += __shfl_xor_sync(0xffffffffu, decl, o);
        const int wall = __shfl_sync(0xffffffffu, winc, 31);
        const int base = p0 + nS + nB * nM;
        const int mine = base + inc - u, wmine = f * wall + winc - wofs;
        if (lane == 0) bw_put_at(buf, p0, (unsigned long long)h->scale[f], nS);
        if (act) bw_put_at(buf, p0 + nS + lane * nM, (unsigned long long)(bab ? bab - 1 : 0), nM);
        for (int b = 0; b < nB; ++b) {                     // merge the band payloads (disjoint bit ranges)
            const int ub = __shfl_sync(0xffffffffu, u, b), mb = __shfl_sync(0xffffffffu, mine, b);
            const int wb = __shfl_sync(0xffffffffu, wmine, b);
            for (int i = lane; i * 32 < ub; i += 32) {
                const int nb = min(32, ub - i * 32);
                bw_put_at(buf, mb + i * 32, (unsigned long long)(wcol[(size_t)(wb + i) * 32] >> (32 - nb)), nb);
            }
        }
        p0 = base + __shfl_sync(0xffffffffu, inc, 31);
        declared += nS + decl;
    }
    __syncwarp();
    const long long cb = (long long)a.cb0 + local;
    const int nBytes = (declared + 7) / 8;
    unsigned* out = reinterpret_cast<unsigned*>(a.rec + cb * REC_STRIDE);
    const int nW = min(REC_WORDS, (nBytes + 3) / 4);
    for (int i = lane; i < nW; i += 32) out[i] = __byte_perm(buf[i], 0, 0x0123);
    if (lane == 0) {
        a.recBytes[cb] = nBytes;
        a.zeroSub[cb] = h->zeroSub;
        if (nBytes > REC_STRIDE || p0 > declared) atomicOr(a.err, PVQ_ERR_STACK);
    }
}

// ---------------------------------------------------------------- byte layout
// size of every channel-block in the output: 4 + nBytes, or 0 if its block is
// dropped (a transient block with an all-zero sub-frame in any channel); the
// first channel-block of a stream also carries the stream's header bytes.
__global__ void k_sizes(const int* __restrict__ recBytes, const unsigned char* __restrict__ zeroSub,
                        const int* __restrict__ wbOff, int nStreams, int C, int headerBytes, int nChBlocks,
                        long long* __restrict__ sizes) {
    int cb = blockIdx.x * blockDim.x + threadIdx.x;
    if (cb >= nChBlocks) return;
    int wb = cb / C;
    bool drop = false;
    for (int c = 0; c < C; ++c) drop |= zeroSub[wb * C + c] != 0;
    long long sz = drop ? 0 : 4 + recBytes[cb];
    int lo = 0, hi = nStreams;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (wbOff[mid] <= wb) lo = mid; else hi = mid;
    }
    if (cb == wbOff[lo] * C) sz += headerBytes;
    sizes[cb] = sz;
}

// one warp per channel-block: <L length prefix + payload to its final offset
__global__ void k_gather(const unsigned char* __restrict__ rec, const int* __restrict__ recBytes,
                         const long long* __restrict__ sizes, const long long* __restrict__ offs,
                         const int* __restrict__ wbOff, int nStreams, int C, int headerBytes, int nChBlocks,
                         unsigned char* __restrict__ out, long long cap, long long* __restrict__ streamOff) {
    int cb = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (cb >= nChBlocks) return;
    long long o = offs[cb], sz = sizes[cb];
    int wb = cb / C;
    int lo = 0, hi = nStreams;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (wbOff[mid] <= wb) lo = mid; else hi = mid;
    }
    if (cb == wbOff[lo] * C) {
        if (lane == 0) streamOff[lo] = o;
        o += headerBytes;
        sz -= headerBytes;
    }
    if (cb == nChBlocks - 1 && lane == 0) streamOff[nStreams] = offs[cb] + sizes[cb];
    if (sz == 0 || o + sz > cap) return;
    int nBytes = recBytes[cb];
    if (lane < 4) out[o + lane] = (unsigned char)((unsigned)nBytes >> (8 * lane));
    const unsigned char* src = rec + (long long)cb * REC_STRIDE;
    for (int i = lane; i < nBytes; i += 32) out[o + 4 + i] = src[i];
}

// .pac header (coder/pacfile.py:306-333) with the numSamples rule of :309-315
__global__ void k_headers(const unsigned char* __restrict__ tmpl, int headerBytes, const int* __restrict__ hdrSamples,
                          const long long* __restrict__ streamOff, int nStreams, unsigned char* __restrict__ out,
                          long long cap) {
    int s = blockIdx.x;
    long long o = streamOff[s];
    if (o + headerBytes > cap) return;
    for (int i = threadIdx.x; i < headerBytes; i += blockDim.x) out[o + i] = tmpl[i];
    __syncthreads();
    if (threadIdx.x < 4) {
        unsigned n = (unsigned)hdrSamples[s];
        if (n % PAC_HALF_LONG == 0) n += PAC_HALF_LONG;
        n += PAC_HALF_LONG;
        out[o + 10 + threadIdx.x] = (unsigned char)(n >> (8 * threadIdx.x));
    }
}

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

extern "C" int pac_header_bytes(const pac_ctx* ctx) {
    return ctx ? 30 + 2 * ctx->host.longBands.nBands : PAC_ERR_ARG;
}

extern "C" long long pac_encode_bound(const pac_ctx* ctx, const int* h_nblocks, int n_streams) {
    if (!ctx || !h_nblocks) return PAC_ERR_ARG;
    long long tot = 0;
    for (int s = 0; s < n_streams; ++s)
        tot += pac_header_bytes(ctx) + (long long)(h_nblocks[s] + 2) * ctx->cfg.n_channels * (4 + REC_STRIDE);
    return tot;
}

extern "C" int pac_encode_streams(pac_ctx* ctx, const short* d_pcm, const int* h_nblocks, const int* h_hdr_samples,
                                  int n_streams, unsigned char* d_out, long long out_capacity,
                                  long long* h_stream_off, int* h_err, void* stream) {
    if (!ctx || !h_nblocks || !h_hdr_samples || !h_stream_off) return PAC_ERR_ARG;
    if (n_streams <= 0) return 0;
    if (!d_pcm || !d_out) return PAC_ERR_ARG;
    if (ctx->cfg.use_sbr && !ctx->cfg.use_vq) return PAC_ERR_UNSUPPORTED;   // crashes in the reference too
    cudaStream_t st = (cudaStream_t)stream;
    PAC_CHECK(cudaSetDevice(ctx->device));
    const int C = ctx->cfg.n_channels, S = n_streams;
    const int hdrB = pac_header_bytes(ctx);
    // host-side prefix arrays
    size_t hostInts = (size_t)4 * (S + 1);
    void* pin = nullptr;
    int rc = pac_pinned(ctx, hostInts * sizeof(int) + 128 + (S + 2) * sizeof(long long), &pin);
    if (rc) return rc;
    int* hBlkOff = (int*)pin;
    int* hNb = hBlkOff + (S + 1);
    int* hWbOff = hNb + (S + 1);
    int* hHdr = hWbOff + (S + 1);
    unsigned char* hTmpl = (unsigned char*)(hHdr + (S + 1));
    long long* hOff = (long long*)(hTmpl + 128);
    long long totBlk = 0, totWb = 0;
    for (int s = 0; s < S; ++s) {
        if (h_nblocks[s] < 1) return PAC_ERR_ARG;
        hBlkOff[s] = (int)totBlk;
        hWbOff[s] = (int)totWb;
        hNb[s] = h_nblocks[s];
        hHdr[s] = h_hdr_samples[s];
        totBlk += h_nblocks[s];
        totWb += h_nblocks[s] + 2;
        if (totWb * C > 0x7ffffff0LL) return PAC_ERR_ARG;
    }
    hBlkOff[S] = (int)totBlk;
    hWbOff[S] = (int)totWb;
    hNb[S] = 0;
    hHdr[S] = 0;
    {   // header template; numSamples patched per stream on device
        unsigned char* h = hTmpl;
        memset(h, 0, 128);
        memcpy(h, "PAC ", 4);
        auto p32 = [&](int o, unsigned v) { for (int i = 0; i < 4; ++i) h[o + i] = (unsigned char)(v >> (8 * i)); };
        auto p16 = [&](int o, unsigned v) { for (int i = 0; i < 2; ++i) h[o + i] = (unsigned char)(v >> (8 * i)); };
        p32(4, (unsigned)ctx->cfg.sample_rate);
        p16(8, (unsigned)C);
        p32(10, 0);
        p32(14, PAC_HALF_LONG);
        p16(18, (unsigned)ctx->cfg.n_scale_bits);
        p16(20, (unsigned)ctx->cfg.n_mant_size_bits);
        p16(22, ctx->cfg.use_sbr ? 1 : 0);
        p16(24, ctx->cfg.use_vq ? 1 : 0);
        p32(26, (unsigned)ctx->host.longBands.nBands);
        for (int b = 0; b < ctx->host.longBands.nBands; ++b) p16(30 + 2 * b, (unsigned)ctx->host.longBands.nLines[b]);
    }
    const long long nCh = totWb * C;
    // device scratch carve-up
    size_t tempBytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tempBytes, (long long*)nullptr, (long long*)nullptr, (int)nCh, st);
    size_t o = 0;
    auto carve = [&](size_t bytes) { size_t r = o; o += align256(bytes); return r; };
    size_t oInts = carve(hostInts * sizeof(int) + 128);
    size_t oDet = carve((size_t)totBlk);
    size_t oRec = carve((size_t)nCh * REC_STRIDE);
    size_t oRecB = carve((size_t)nCh * sizeof(int));
    size_t oZero = carve((size_t)nCh);
    size_t oSizes = carve((size_t)nCh * sizeof(long long));
    size_t oOffs = carve((size_t)nCh * sizeof(long long));
    size_t oSOff = carve((size_t)(S + 2) * sizeof(long long));
    size_t oErr = carve(sizeof(int));
    size_t oTemp = carve(tempBytes);
    // VQ streams: scaled lines, workspaces and band bit buffers of one batch of channel-blocks at a time
    // (band layouts the batched back end is not sized for stay in k_encode)
    const bool vq = ctx->cfg.use_vq != 0 && ctx->host.longBands.nBands <= EPRE_SLOTS &&
                    PAC_SUBFRAMES * ctx->host.shortBands.nBands <= EPRE_SLOTS && ctx->host.sbrNOmitted <= 4;
    const long long maxB = ctx->optBatchCb > 0 ? ctx->optBatchCb : ENC_BATCH_CB;    // pac_set_option: small batches exercise the batch loop
    const long long batchCb = vq ? std::min<long long>(nCh, maxB) : 0, batchGroups = (batchCb + 31) / 32;
    size_t oPre = carve((size_t)batchCb * sizeof(EncPre));
    size_t oLines = carve((size_t)batchGroups * 32 * 1024 * sizeof(double));
    size_t oWs = carve((size_t)batchGroups * 32 * 2 * EPRE_WS_PER_CB * sizeof(double));
    size_t oY = carve((size_t)batchGroups * 32 * EPRE_WS_PER_CB * sizeof(int));
    size_t oWords = carve((size_t)batchGroups * 32 * EPRE_WORDS_PER_CB * sizeof(unsigned));
    size_t oPerm = carve((size_t)(batchCb + 2) * sizeof(int));
    size_t oWbS = carve((size_t)totWb * sizeof(int));
    void* scr = nullptr;
    if ((rc = pac_scratch(ctx, o, &scr))) return rc;
    unsigned char* sb = (unsigned char*)scr;
    int* dBlkOff = (int*)(sb + oInts);
    int* dNb = dBlkOff + (S + 1);
    int* dWbOff = dNb + (S + 1);
    int* dHdr = dWbOff + (S + 1);
    unsigned char* dTmpl = (unsigned char*)(dHdr + (S + 1));
    PAC_CHECK(cudaMemcpyAsync(dBlkOff, hBlkOff, hostInts * sizeof(int) + 128, cudaMemcpyHostToDevice, st));
    PAC_CHECK(cudaMemsetAsync(sb + oErr, 0, sizeof(int), st));
    ctx->kevN[0] = ctx->kevN[1] = 0;
    ctx->launches += 6;       // k_transient, k_sizes, the two kernels of the cub scan, k_gather, k_headers (per batch below: k_encode, k_bitalloc, k_pvq, k_pack)
    k_transient<<<(unsigned)((totBlk + TR_WARPS - 1) / TR_WARPS), TR_WARPS * 32, 0, st>>>(d_pcm, dBlkOff, dNb, S, C, (int)totBlk, sb + oDet);
    EncArgs a;
    a.pcm = d_pcm; a.blkOff = dBlkOff; a.nb = dNb; a.wbOff = dWbOff; a.det = sb + oDet; a.nStreams = S;
    a.rec = sb + oRec; a.recBytes = (int*)(sb + oRecB); a.zeroSub = sb + oZero; a.err = (int*)(sb + oErr);
    size_t smem = enc_smem_bytes() + (size_t)ctx->optSmemPad;
    a.pre = nullptr; a.preLines = nullptr; a.perm = nullptr; a.counts = nullptr; a.cb0 = 0; a.nBatch = (int)nCh;
    a.variant = ctx->optVariant;
    a.wbStream = nullptr;
    if (vq) {
        k_wbmap<<<(unsigned)((totWb + 255) / 256), 256, 0, st>>>(dWbOff, S, (int)totWb, (int*)(sb + oWbS));
        a.wbStream = (int*)(sb + oWbS);
        ++ctx->launches;
    }
    const size_t smemF = enc_front_smem_bytes();
    PAC_CHECK(cudaFuncSetAttribute(k_encode<float, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    PAC_CHECK(cudaFuncSetAttribute(k_encode<double, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    PAC_CHECK(cudaFuncSetAttribute(k_encode<float, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemF));
    PAC_CHECK(cudaFuncSetAttribute(k_encode<double, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemF));
    PAC_CHECK(cudaFuncSetAttribute(k_encode_front<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)enf_smem_bytes()));
    PAC_CHECK(cudaFuncSetAttribute(k_encode_front<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)enf_smem_bytes()));
    PAC_CHECK(cudaFuncSetAttribute(k_encode_front<float>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    PAC_CHECK(cudaFuncSetAttribute(k_encode_front<double>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    cudaEventRecord(ctx->ev[0], st);
    for (long long q0 = 0; q0 < nCh; q0 += vq ? batchCb : nCh) {
        const long long nq = vq ? std::min(batchCb, nCh - q0) : nCh;
        a.cb0 = (int)q0; a.nBatch = (int)nq;
        if (vq) {
            a.pre = (EncPre*)(sb + oPre); a.preLines = (double*)(sb + oLines);
            a.perm = (int*)(sb + oPerm); a.counts = a.perm + batchCb;
            PAC_CHECK(cudaMemsetAsync(a.counts, 0, 2 * sizeof(int), st));
        }
        pac_kev_begin(ctx, 0, st);
        if (vq && !(a.variant & 1)) {
            if (ctx->cfg.precision == PAC_F32) k_encode_front<float><<<(unsigned)nq, ENF_NT, enf_smem_bytes(), st>>>(ctx->dev, a);
            else k_encode_front<double><<<(unsigned)nq, ENF_NT, enf_smem_bytes(), st>>>(ctx->dev, a);
        } else if (vq) {
            if (ctx->cfg.precision == PAC_F32) k_encode<float, true><<<(unsigned)nq, ENC_NT, smemF, st>>>(ctx->dev, a);
            else k_encode<double, true><<<(unsigned)nq, ENC_NT, smemF, st>>>(ctx->dev, a);
        } else {
            if (ctx->cfg.precision == PAC_F32) k_encode<float, false><<<(unsigned)nq, ENC_NT, smem, st>>>(ctx->dev, a);
            else k_encode<double, false><<<(unsigned)nq, ENC_NT, smem, st>>>(ctx->dev, a);
        }
        pac_kev_end(ctx, 0, st);
        if (vq) {
            const int nChunks = (int)((nq + PQ_NT - 1) / PQ_NT);
            k_bitalloc<<<(unsigned)((nq + BA_WARPS - 1) / BA_WARPS), BA_WARPS * 32, 0, st>>>(ctx->dev, a);
            pac_kev_begin(ctx, 1, st);
            k_pvq<<<(unsigned)(nChunks * EPRE_SLOTS), PQ_NT, 0, st>>>(ctx->dev, a, nChunks, (double*)(sb + oWs), (int*)(sb + oY),
                                                                   (unsigned*)(sb + oWords));
            pac_kev_end(ctx, 1, st);
            ctx->launches += 2;
            k_pack<<<(unsigned)((nq + PK_WARPS - 1) / PK_WARPS), PK_WARPS * 32, 0, st>>>(ctx->dev, a, (unsigned*)(sb + oWords));
        }
    }
    cudaEventRecord(ctx->ev[1], st);
    ctx->evValid[0] = 1;
    PAC_CHECK(cudaGetLastError());
    k_sizes<<<(unsigned)((nCh + 255) / 256), 256, 0, st>>>(a.recBytes, a.zeroSub, dWbOff, S, C, hdrB, (int)nCh,
                                                           (long long*)(sb + oSizes));
    PAC_CHECK(cub::DeviceScan::ExclusiveSum(sb + oTemp, tempBytes, (long long*)(sb + oSizes), (long long*)(sb + oOffs),
                                            (int)nCh, st));
    k_gather<<<(unsigned)((nCh * 32 + 255) / 256), 256, 0, st>>>(a.rec, a.recBytes, (long long*)(sb + oSizes),
                                                                  (long long*)(sb + oOffs), dWbOff, S, C, hdrB, (int)nCh,
                                                                  d_out, out_capacity, (long long*)(sb + oSOff));
    k_headers<<<S, 64, 0, st>>>(dTmpl, hdrB, dHdr, (long long*)(sb + oSOff), S, d_out, out_capacity);
    PAC_CHECK(cudaGetLastError());
    PAC_CHECK(cudaMemcpyAsync(hOff, sb + oSOff, (S + 1) * sizeof(long long), cudaMemcpyDeviceToHost, st));
    PAC_CHECK(cudaMemcpyAsync(hHdr, sb + oErr, sizeof(int), cudaMemcpyDeviceToHost, st));
    PAC_CHECK(cudaStreamSynchronize(st));
    memcpy(h_stream_off, hOff, (S + 1) * sizeof(long long));
    if (h_err) *h_err = hHdr[0];
    if (hOff[S] > out_capacity) return PAC_ERR_ARG;
    return hHdr[0] ? PAC_ERR_SOFT : 0;
}

#ifdef PAC_PHASE_TIMING
// debug: copy and reset the fine-grained counters of the front end ([0..31] long frames, [32..63] short sub-frames)
extern "C" int pac_debug_phases_front(unsigned long long* h_out) {
    unsigned long long z[64] = {0};
    cudaMemcpyFromSymbol(h_out, g_enf_phase, sizeof(z));
    cudaMemcpyToSymbol(g_enf_phase, z, sizeof(z));
    return 0;
}
// debug: copy and reset the per-phase cycle counters ([0..15] long frames, [16..31] short sub-frames)
extern "C" int pac_debug_phases(unsigned long long* h_out) {
    unsigned long long z[32] = {0};
    cudaMemcpyFromSymbol(h_out, g_pac_phase, sizeof(z));
    cudaMemcpyToSymbol(g_pac_phase, z, sizeof(z));
    return 0;
}
#endif
