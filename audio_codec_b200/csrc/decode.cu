// Stream decoder: .pac bytes -> int16 PCM, all stages on device.
//
//   k_walk      per stream: follow the <L length prefixes (coder/pacfile.py:239-256)
//   k_decode    one CTA per channel-block: parse flags / scale / bit allocation,
//               dequantise every band (PVQ parse or block-FP mantissas), SBR
//               reconstruction, IMDCT + synthesis window, short-block overlap
//               (coder/pacfile.py:177-287, coder/codec.py:47-222)
//   k_ola       overlap-add consecutive blocks and quantise to 16-bit PCM
//               (coder/pacfile.py:289-295, coder/pcmfile.py:118-147)
#include <string.h>
#include <algorithm>
#include <vector>
#include "decode_frame.cuh"
#include "decode_long.cuh"
#include "pac_api.h"

extern "C" int pac_header_bytes(const pac_ctx* ctx);
#define DEC_BATCH_CB 131072      // channel-blocks per k_hdr / k_dequant / k_decode round (bounds the scratch: 40 KB each)

// pass 0: count channel-blocks; pass 1: record payload offsets (absolute, past the prefix).
// The prefix chain is serial, but its cost is the memory latency per hop: one WARP per stream
// fetches an 8 KB window (16-byte loads, all in flight together) into shared memory and lane 0
// follows the chain inside it, so DRAM latency is paid once per ~23 records instead of per record.
#define WALK_WARPS 4
#define WALK_WIN 8192
__global__ void __launch_bounds__(WALK_WARPS * 32)
k_walk(const unsigned char* __restrict__ buf, const long long* __restrict__ streamOff, int nStreams,
       int headerBytes, int* __restrict__ counts, const long long* __restrict__ cbOff,
       long long* __restrict__ payOff, int* __restrict__ payBytes) {
    __shared__ __align__(16) unsigned char win[WALK_WARPS][WALK_WIN + 16];
    __shared__ long long s_pos[WALK_WARPS];
    __shared__ int s_n[WALK_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int s = blockIdx.x * WALK_WARPS + warp;
    if (s >= nStreams) return;
    const long long end = streamOff[s + 1], bufEnd = streamOff[nStreams];
    const long long base = cbOff ? cbOff[s] : 0;
    const int limit = cbOff ? (int)(cbOff[s + 1] - cbOff[s]) : 0;
    const unsigned long long addr0 = (unsigned long long)buf;
    long long pos = streamOff[s] + headerBytes;
    int n = 0;
    bool done = false;
    while (!done && pos + 4 <= end) {
        // window of 16-byte chunks (absolute alignment) starting at the chunk that holds pos
        const long long w0 = (long long)(((addr0 + (unsigned long long)pos) & ~15ull) - addr0);
        if (w0 >= 0 && w0 + WALK_WIN <= bufEnd) {          // interior window: no per-chunk bounds checks
            const uint4* src = reinterpret_cast<const uint4*>(buf + w0) + lane;
            uint4 v[WALK_WIN / 512];
#pragma unroll
            for (int j = 0; j < WALK_WIN / 512; ++j) v[j] = __ldg(src + j * 32);
#pragma unroll
            for (int j = 0; j < WALK_WIN / 512; ++j) *reinterpret_cast<uint4*>(&win[warp][(j * 32 + lane) * 16]) = v[j];
        } else {
#pragma unroll 1
            for (int j = 0; j < WALK_WIN / 512; ++j) {
                const long long o = w0 + (j * 32 + lane) * 16;
                uint4 v = make_uint4(0, 0, 0, 0);
                if (o >= 0 && o + 16 <= bufEnd) v = __ldg(reinterpret_cast<const uint4*>(buf + o));
                else if (o + 16 > 0 && o < bufEnd) {       // first / last partial chunk of the buffer: byte loads, nothing outside [0, bufEnd)
                    __align__(16) unsigned char t[16];
#pragma unroll
                    for (int b = 0; b < 16; ++b) t[b] = (o + b >= 0 && o + b < bufEnd) ? buf[o + b] : (unsigned char)0;
                    v = *reinterpret_cast<uint4*>(t);
                }
                *reinterpret_cast<uint4*>(&win[warp][(j * 32 + lane) * 16]) = v;
            }
        }
        __syncwarp();
        if (lane == 0) {
            // the chain inside the window, window-relative; a prefix at any byte alignment is two aligned words and a
            // funnel shift (the row is padded so that the second word always exists)
            const unsigned* w32 = reinterpret_cast<const unsigned*>(&win[warp][0]);
            const long long endRel = end - w0, lim = endRel < WALK_WIN ? endRel : (long long)WALK_WIN;
            long long p = pos - w0;
            while (p + 4 <= lim) {
                const int pi = (int)p;
                const unsigned lo = w32[pi >> 2], hi = w32[(pi >> 2) + 1];
                const unsigned nb = __funnelshift_r(lo, hi, (pi & 3) * 8);
                const long long np = p + 4 + (long long)nb;
                if (np > endRel) { done = true; break; }   // truncated block
                if (payOff && n < limit) {                 // a damaged stream may end on a partial block: its
                    payOff[base + n] = w0 + p + 4;          // extra record must not land in the next stream's slot
                    payBytes[base + n] = (int)nb;
                }
                ++n;
                p = np;
            }
            pos = w0 + p;
            s_pos[warp] = pos;
            s_n[warp] = done ? -n - 1 : n;
        }
        __syncwarp();
        pos = s_pos[warp];
        const int nn = s_n[warp];
        done = nn < 0;
        n = nn < 0 ? -nn - 1 : nn;
        __syncwarp();
    }
    if (counts && lane == 0) counts[s] = n;
}

struct DecArgs {
    const unsigned char* buf;
    const long long* payOff;   // [nCb]
    const int* payBytes;       // [nCb]
    double* frames;            // [nCb][2048] windowed block output before block overlap-add
    long long bufBytes;        // size of buf
    long long cb0;             // first channel-block of this launch
    int* err;
};

// ---- batched front end of VQ streams ------------------------------------------------------------
// Bit reader over a record in global memory: 8-byte words at absolute alignment (a record may start
// at any byte); bit0 = position of the record's first bit relative to the aligned base.
__device__ __forceinline__ BitR rec_reader(const DecArgs& a, long long cb, int* nBytesOut, long long* bit0Out) {
    const int nBytes = min(a.payBytes[cb], DEC_REC_MAX);
    const unsigned char* rec = a.buf + a.payOff[cb];
    const unsigned long long ad = (unsigned long long)rec;
    const unsigned char* al = rec - (ad & 7ull);
    long long bit0 = 8LL * (long long)(ad & 7ull);
    BitR br = {al, bit0, bit0 + 8LL * nBytes, 0};
    const long long room = (a.buf + a.bufBytes) - al;         // readable bytes from the aligned base
    br.cap = (int)min((long long)(DEC_REC_MAX + 16), room & ~7LL);
    if (br.cap < (int)(ad & 7ull) + nBytes || al < a.buf) {   // the record ends inside the buffer's last partial word, or the
                                                              // aligned base lies before the caller's buffer: byte reads only
        br.p = rec; br.end = 8LL * nBytes; br.cap = 0; bit0 = 0; br.pos = 0;
    }
    *nBytesOut = nBytes;
    *bit0Out = bit0;
    return br;
}

// k_hdr: one thread per channel-block reads the block flags and every frame header (overall scale,
// bit allocations) and places the band payloads (coder/pacfile.py:177-229).
__global__ void __launch_bounds__(128)
k_hdr(const PacDev* __restrict__ dev, DecArgs a, long long cbBase, int nCb, CbHdr* __restrict__ hdrs,
      int* __restrict__ perm, int* __restrict__ counts) {
    const int local = blockIdx.x * blockDim.x + threadIdx.x;
    if (local >= nCb) return;
    int nBytes;
    long long bit0;
    BitR br = rec_reader(a, cbBase + local, &nBytes, &bit0);
    CbHdr* h = hdrs + local;
    unsigned char fl = 0, sbrF = 0;
    if (nBytes > 0) {
        const int nS = dev->nScaleBits, nM = dev->nMantSizeBits;
        fl = (unsigned char)(br_get(br, 3) | 0x80u);          // last, cur, next (coder/pacfile.py:196-198)
        const bool cur = (fl >> 1) & 1;
        const PacBands& bands = cur ? dev->shortBands : dev->longBands;
        const int nB = bands.nBands, nF = cur ? PAC_SUBFRAMES : 1;
        for (int f = 0; f < nF; ++f) {
            h->scale[f] = (unsigned char)br_get(br, nS);
            unsigned omMask = 0, baMask = 0;
            for (int b = 0; b < nB; ++b) {
                const int v = (int)br_get(br, nM);
                const int bab = v ? v + 1 : 0;
                h->ba[f * nB + b] = (unsigned char)bab;
                if (bab) baMask |= 1u << b;
                if (bands.omitted[b]) omMask |= 1u << b;
            }
            const bool sbr = !cur && dev->useSBR && (omMask & baMask);    // coder/pacfile.py:660-664
            if (sbr) sbrF = 1;
            long long p = br.pos;
            for (int b = 0; b < nB; ++b) {
                h->pos[f * nB + b] = (unsigned short)min(p - bit0, 65535LL);   // past the end: reads zeros
                if ((baMask >> b) & 1) {
                    const int bab = h->ba[f * nB + b];
                    p += (long long)bab * ((sbr && bands.omitted[b]) ? 1 : bands.nLines[b]);
                }
            }
            if (p > br.end) atomicOr(a.err, PVQ_ERR_STACK);
            br.pos = p;
        }
    }
    h->sbr = sbrF;
    h->flags = fl;
    // dense order for k_dequant: its warps then hold 32 blocks of the same kind (17 long bands against
    // 48 short slots); blocks without a record take no position
    int p = -1;
    if (fl & 0x80u) p = ((fl >> 1) & 1) ? nCb - 1 - atomicAdd(&counts[1], 1) : atomicAdd(&counts[0], 1);
    h->p = p;
    if (p >= 0) perm[p] = local;
    // transient and SBR blocks stay with the CTA-per-block kernel: a work list it pulls from (counts[2] entries)
    if (!dl_takes(fl, sbrF)) (counts + 4)[atomicAdd(&counts[2], 1)] = local;
}

// k_dequant: one thread per (channel-block, band).  A CTA holds one band slot of DQ_NT consecutive
// channel-blocks; CTAs are ordered widest band first so the long chains start first.  Slot q of a long
// block is band nB-1-q; of a transient block, sub-frame q / 6, band 5 - q % 6.  Lines go to
// lines[pre_line_index(cb, k)]; the workspaces are lane-interleaved the same way.
#ifndef DQ_NT
#define DQ_NT 128
#endif
#ifndef DQ_SORT
#define DQ_SORT 1
#endif
#ifndef DQ_MINB
#define DQ_MINB 7          // 72 registers (96 when free): seven CTAs per SM (measured 5 / 6 / 7 / 8: 32.0 / 29.7 / 29.2 / 31.3 ms decode)
#endif
__global__ void __launch_bounds__(DQ_NT, DQ_MINB)
k_dequant(const PacDev* __restrict__ dev, DecArgs a, long long cbBase, int nCb, int nChunks, CbHdr* __restrict__ hdrs,
          double* __restrict__ lines, double* __restrict__ ws, const int* __restrict__ perm,
          const int* __restrict__ counts) {
    __shared__ int s_off[PAC_PVQ_LMAX + 2];                  // row offsets of the N(L,K) table: read by every pvq_N call
    for (int i = threadIdx.x; i < PAC_PVQ_LMAX + 2; i += DQ_NT) s_off[i] = dev->pvqRowOff[i];
    __syncthreads();
    const int q = blockIdx.x / nChunks, chunk = blockIdx.x - q * nChunks;
    const int nLong = counts[0], nShort = counts[1];
    const int nBL = dev->longBands.nBands, nBS = dev->shortBands.nBands;
    // The 128 blocks of the CTA are dealt to its threads in the order of this band's bit allocation (a counting sort
    // in shared memory): the lanes of a warp then walk trees of similar size.  A thread keeps the workspace of its own
    // place; only the block it decodes - and so where its lines go - is permuted.
#if DQ_SORT
    __shared__ int s_cnt[3][34];
    __shared__ unsigned short s_order[DQ_NT];
    {
        // The dense order puts long blocks at the front of the batch and transient blocks at its back; the tile that
        // holds the boundary has both (and the unused places between them).  Blocks are permuted WITHIN their kind's
        // part of the tile only: a workspace column then serves one kind in every band's CTA (the two kinds lay out
        // their bands' rows differently).
        const int tile0 = chunk * DQ_NT, pos0 = tile0 + threadIdx.x;
        const int gap0 = min(max(nLong - tile0, 0), DQ_NT), tr0 = min(max(nCb - nShort - tile0, 0), DQ_NT);   // long [0, gap0), unused [gap0, tr0), transient [tr0, ..)
        const int cls = threadIdx.x < gap0 ? 0 : (threadIdx.x < tr0 ? 1 : 2);
        int key = 33;                                          // nothing to do: last of its part
        if (cls != 1 && pos0 < nCb && !(q >= nBL && cls == 0)) {
            const CbHdr* h0 = hdrs + perm[pos0];
            const unsigned char fl0 = h0->flags;
            if (fl0 & 0x80u) {
                const bool cur0 = (fl0 >> 1) & 1;
                const int nB0 = cur0 ? nBS : nBL;
                if (q < (cur0 ? PAC_SUBFRAMES * nB0 : nB0)) {
                    const int f0 = q / nB0, b0 = nB0 - 1 - (q - f0 * nB0);
                    key = min((int)h0->ba[f0 * nB0 + b0], 32);
                }
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
    const int wsPos = chunk * DQ_NT + threadIdx.x;             // workspace: this thread's own place
    const int pos = chunk * DQ_NT + s_order[threadIdx.x];      // the block it decodes
#else
    const int pos = chunk * DQ_NT + threadIdx.x, wsPos = pos;
#endif
    if (pos >= nCb || (pos >= nLong && pos < nCb - nShort)) return;
    if (q >= nBL && pos < nLong) return;
    const int local = perm[pos];
    CbHdr* h = hdrs + local;
    const unsigned char fl = h->flags;
    if (!(fl & 0x80u)) return;
    const bool cur = (fl >> 1) & 1;
    const PacBands& bands = cur ? dev->shortBands : dev->longBands;
    const int nB = bands.nBands;
    if (q >= (cur ? PAC_SUBFRAMES * nB : nB)) return;
    const int f = q / nB, b = nB - 1 - (q - f * nB), slot = f * nB + b;
    const bool sbr = h->sbr != 0, env = sbr && bands.omitted[b];
    const int bb = h->ba[slot], L = bands.nLines[b];
    double* out = lines + pre_line_index(pos, f * PAC_HALF_SHORT + bands.lower[b]);
    int k = 0;
    if (env) for (int j = 0; j < b; ++j) k += bands.omitted[j] ? 1 : 0;
    if (!bb) {
        if (env) { if (k < 4) h->env[k] = 0.0; }
        else for (int i = 0; i < L; ++i) out[i * 32] = 0.0;
        return;
    }
    int cap = 1, base = 0, all = 0;                            // workspace offset: power-of-two capacities of the bands before
    while (cap < L) cap <<= 1;
    for (int j = 0; j < nB; ++j) {
        int c = 1;
        while (c < bands.nLines[j]) c <<= 1;
        if (j < b) base += c;
        all += c;
    }
    base += f * all;                                           // <= 2 * 1024 in total for either frame type
    double* wsA = ws + ((size_t)(wsPos >> 5) * (2 * PRE_WS_PER_CB) + (size_t)base) * 32 + (wsPos & 31);
    double* wsB = wsA + (size_t)PRE_WS_PER_CB * 32;
    int nBytes;
    long long bit0;
    BitR br = rec_reader(a, cbBase + local, &nBytes, &bit0);
    br.pos = bit0 + h->pos[slot];
    PvqTab pt = {dev->pvqN, s_off, dev->log2tab, dev->pvqKW, dev->thetaTab};
    int werr = 0;
    if (env) {                                                 // one envelope value (coder/codec.py:111-132)
        double v;
        thread_dequantize_gain_shape<1>(pt, br, bb, 1, 1, wsA, wsB, &v, &werr);
        if (k < 4) h->env[k] = v;
    } else {
        thread_dequantize_gain_shape<32>(pt, br, bb * L, L, cap, wsA, wsB, out, &werr);
    }
    if (werr) atomicOr(a.err, werr);
}

// PRE: behind k_dequant (preAll set): SBR + IMDCT only, dec_lean_smem_bytes(), four CTAs per SM.
#ifndef DEC_PRE_MINB
#define DEC_PRE_MINB 4
#endif
// The eight short frames of a transient block behind k_dequant, ONE PER WARP, all at once (run one after the other on
// the whole CTA a transient block took 73 us: a 64-point FFT keeps 8 of 256 threads busy).  Each warp: its 128 lines,
// scale, IMDCT (the CTA FFT code with NT = 32 on the warp's own buffers; its barriers are unconditional, so the eight
// warps pass them together), sine window.  The frames overlap by half: even frames add into y first, then odd ones -
// every sample gets the same two addends as in frame order (coder/pacfile.py:258-287).
template <typename T>
__device__ __forceinline__ void decode_short_frames_pre(const PacDev* __restrict__ dev, const DecSmem& S,
                                                        const CbHdr* __restrict__ pre, const double* __restrict__ preLines) {
    constexpr int N = PAC_NSHORT, H = N / 2, Q = N / 4;
    const int sf = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const PacTables<T>& t = psy_tabs<T>(dev);
    const cplx<T>* tw = reinterpret_cast<const cplx<T>*>(t.tw_mdct_short);
    const cplx<T>* roots = reinterpret_cast<const cplx<T>*>(t.tw_fft[0]);
    double* lines = S.lines + H * sf;                                     // [128]
    cplx<T>* fbuf = reinterpret_cast<cplx<T>*>(S.wsA + 256 * sf);         // [64] complex
    T* v = reinterpret_cast<T*>(S.wsA + 256 * sf + 128);                  // [128]
    for (int i = lane; i < H; i += 32) lines[i] = preLines[pre_line_index(pre->p, sf * H + i)];
    const double inv = 1.0 / (double)(1 << pre->scale[sf]);
    __syncthreads();
    imdct_pre<T, N, 1, 32, double>(lines, inv, tw, fbuf, lane);
    __syncthreads();
    cta_fft<T, Q, 1, 32>(fbuf, roots, lane);
    mdct_post<T, N, 1, 32>(fbuf, tw, (T)1, v, lane);
    __syncthreads();
    const T* w = t.win_short + PAC_WIN_SINE * N;
    const int yOff = 448 + 128 * sf;
#pragma unroll
    for (int ph = 0; ph < 2; ++ph) {
        if ((sf & 1) == ph)
            for (int i = lane; i < N; i += 32) S.y[yOff + i] += (double)(w[i] * imdct_sample<T, N>(v, i));
        __syncthreads();
    }
}

template <typename T, bool PRE>
__global__ void __launch_bounds__(DEC_NT, PRE ? DEC_PRE_MINB : 3)
k_decode(const PacDev* __restrict__ dev, DecArgs a, const CbHdr* __restrict__ preAll, const double* __restrict__ preLines,
         int* __restrict__ work) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int s_next;
    DecSmem S;
    dec_smem_carve(smem_raw, S, PRE);
    const int tid = threadIdx.x;
    // work (behind k_decode_long): counts[2] = number of listed blocks, counts[3] = next to hand out, list at
    // counts + 4.  A fixed grid of CTAs pulls blocks until the list is empty (transient blocks come in bursts: a
    // static split leaves most CTAs idle).  Without a list: one CTA per channel-block.
    for (;;) {
    int local = (int)blockIdx.x;
    if (PRE && work) {
        if (tid == 0) s_next = atomicAdd(&work[3], 1);
        __syncthreads();
        const int j = s_next;
        __syncthreads();
        if (j >= work[2]) break;
        local = (work + 4)[j];
    }
    const long long cb = (long long)local + a.cb0;
    int nBytes = a.payBytes[cb];
    if (nBytes > DEC_REC_MAX) {
        if (tid == 0) atomicOr(a.err, PVQ_ERR_STACK);
        nBytes = DEC_REC_MAX;
    }
    const CbHdr* pre = PRE ? preAll + local : nullptr;
    const unsigned char* src = a.buf + a.payOff[cb];
    if (!pre) for (int i = tid; i < nBytes; i += DEC_NT) S.rec[i] = src[i];
    for (int i = tid; i < 2048; i += DEC_NT) S.y[i] = 0.0;
    __syncthreads();
    const long long recBits = 8LL * nBytes;
    const unsigned char b0 = pre ? (unsigned char)(pre->flags << 5) : S.rec[0];
    const bool last = (b0 >> 7) & 1, cur = (b0 >> 6) & 1, nxt = (b0 >> 5) & 1;
    if (tid == 0) S.misc[3] = 3;
    __syncthreads();
    if (nBytes > 0) {
        if (!cur) {
            int shape = (last && nxt) ? PAC_WIN_STARTSTOP : last ? PAC_WIN_STOP : nxt ? PAC_WIN_START : PAC_WIN_SINE;
            decode_frame<PAC_NLONG, T>(dev, S, recBits, shape, 0, a.err, pre, preLines, local, 0);
        } else if (PRE && DEC_NT == 32 * PAC_SUBFRAMES) {
            decode_short_frames_pre<T>(dev, S, pre, preLines);
        } else {
            for (int sf = 0; sf < PAC_SUBFRAMES; ++sf)
                decode_frame<PAC_NSHORT, T>(dev, S, recBits, PAC_WIN_SINE, 448 + 128 * sf, a.err, pre, preLines, local, sf);
        }
    }
    double* dst = a.frames + cb * 2048;
    for (int i = tid; i < 2048; i += DEC_NT) dst[i] = S.y[i];
    if (!(PRE && work)) break;
    __syncthreads();
    }
}

// PCM quantiser of the WAV writer (coder/pcmfile.py:127-134): sign-magnitude midtread, 16 bit.
__device__ __forceinline__ short double_to_pcm(double x) {
    double a = fabs(x);
    int code = a >= 1.0 ? 32767 : (int)floor((65535.0 * a + 1.0) * 0.5);
    return (short)(signbit(x) ? -code : code);
}

// Output block b of (stream, channel) = tail of written block b-1 + head of written block b;
// the final block is the overlap flush (coder/pacfile.py:247-251, :289-295).
__global__ void k_ola(const double* __restrict__ frames, const long long* __restrict__ cbOff,
                      const long long* __restrict__ pcmOff, int nStreams, int C, short* __restrict__ pcm, long long g0) {
    // grid: one CTA per (output block, channel) over all streams; find the stream by binary search
    const long long g = blockIdx.x + g0;   // index over sum_s (nBlk_s + 1) * C
    int lo = 0, hi = nStreams;
    // output blocks of stream s start at pcmOff[s]/1024 (pcmOff counts samples over all channels)
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (pcmOff[mid] / 1024 <= g) lo = mid; else hi = mid;
    }
    const int s = lo;
    const long long nCb = cbOff[s + 1] - cbOff[s];        // channel-blocks of this stream
    const long long nBlk = nCb / C;                        // written blocks
    const long long local = g - pcmOff[s] / 1024;          // (channel, block) index: c*(nBlk+1) + b
    const int c = (int)(local / (nBlk + 1));
    const long long b = local - (long long)c * (nBlk + 1);
    const double* head = b < nBlk ? frames + (cbOff[s] + b * C + c) * 2048 : nullptr;
    const double* tail = b > 0 ? frames + (cbOff[s] + (b - 1) * C + c) * 2048 + 1024 : nullptr;
    short* dst = pcm + pcmOff[s] + ((long long)c * (nBlk + 1) + b) * 1024;
    // four samples per thread: 2 x 32 bytes in, 8 bytes out (the caller's buffer may be 2-byte aligned only)
    const bool vec = ((unsigned long long)dst & 7ull) == 0ull;
    for (int i = threadIdx.x * 4; i < 1024; i += blockDim.x * 4) {
        double2 t0 = make_double2(0.0, 0.0), t1 = t0, h0 = t0, h1 = t0;
        if (tail) { t0 = *reinterpret_cast<const double2*>(tail + i); t1 = *reinterpret_cast<const double2*>(tail + i + 2); }
        if (head) { h0 = *reinterpret_cast<const double2*>(head + i); h1 = *reinterpret_cast<const double2*>(head + i + 2); }
        const short q0 = double_to_pcm(t0.x + h0.x), q1 = double_to_pcm(t0.y + h0.y);
        const short q2 = double_to_pcm(t1.x + h1.x), q3 = double_to_pcm(t1.y + h1.y);
        if (vec) *reinterpret_cast<short4*>(dst + i) = make_short4(q0, q1, q2, q3);
        else { dst[i] = q0; dst[i + 1] = q1; dst[i + 2] = q2; dst[i + 3] = q3; }
    }
}

static size_t dalign(size_t x) { return (x + 255) & ~(size_t)255; }

// Shared by plan and decode: walk the length prefixes; fills host prefix arrays.
static int walk_streams(pac_ctx* ctx, const unsigned char* d_pac, const long long* h_stream_off, int S,
                        cudaStream_t st, std::vector<long long>& cbOff, long long** dStreamOff, int** dCounts) {
    void* scr = nullptr;
    int rc = pac_scratch(ctx, dalign((S + 1) * sizeof(long long)) + dalign(S * sizeof(int)), &scr);
    if (rc) return rc;
    long long* dOff = (long long*)scr;
    int* dCnt = (int*)((unsigned char*)scr + dalign((S + 1) * sizeof(long long)));
    PAC_CHECK(cudaMemcpyAsync(dOff, h_stream_off, (S + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
    k_walk<<<(S + WALK_WARPS - 1) / WALK_WARPS, WALK_WARPS * 32, 0, st>>>(d_pac, dOff, S, pac_header_bytes(ctx), dCnt, nullptr, nullptr, nullptr);
    std::vector<int> cnt(S);
    PAC_CHECK(cudaMemcpyAsync(cnt.data(), dCnt, S * sizeof(int), cudaMemcpyDeviceToHost, st));
    PAC_CHECK(cudaStreamSynchronize(st));
    cbOff.assign(S + 1, 0);
    for (int s = 0; s < S; ++s) cbOff[s + 1] = cbOff[s] + cnt[s] - cnt[s] % ctx->cfg.n_channels;
    if (dStreamOff) *dStreamOff = dOff;
    if (dCounts) *dCounts = dCnt;
    return 0;
}

extern "C" int pac_decode_plan(pac_ctx* ctx, const unsigned char* d_pac, const long long* h_stream_off,
                               int n_streams, long long* h_pcm_off, void* stream) {
    if (!ctx || !h_stream_off || !h_pcm_off) return PAC_ERR_ARG;
    if (n_streams <= 0) return 0;
    if (!d_pac) return PAC_ERR_ARG;
    PAC_CHECK(cudaSetDevice(ctx->device));
    std::vector<long long> cbOff;
    int rc = walk_streams(ctx, d_pac, h_stream_off, n_streams, (cudaStream_t)stream, cbOff, nullptr, nullptr);
    if (rc) return rc;
    const int C = ctx->cfg.n_channels;
    h_pcm_off[0] = 0;
    for (int s = 0; s < n_streams; ++s) {
        long long nBlk = (cbOff[s + 1] - cbOff[s]) / C;
        h_pcm_off[s + 1] = h_pcm_off[s] + (nBlk + 1) * 1024 * C;
    }
    return 0;
}

// The decoder proper.  chunks: stream boundaries [c0 = 0, c1, ..., n_streams] (nullptr: one chunk); the
// kernels are launched per chunk and chunkEv[c] is recorded after chunk c, so a caller can move
// finished chunks off the device while the rest decodes; with h_copy_dst the chunk's PCM is sent to
// that host buffer on copyStream as soon as its event fires (the host-buffer entry point).
int pac_decode_streams_chunked(pac_ctx* ctx, const unsigned char* d_pac, const long long* h_stream_off,
                               int n_streams, short* d_pcm, long long pcm_capacity, long long* h_pcm_off,
                               int* h_err, void* stream, const int* chunks, int n_chunks, cudaEvent_t* chunkEv,
                               cudaStream_t copyStream, short* h_copy_dst, const int* h_counts) {
    if (!ctx || !h_stream_off || !h_pcm_off) return PAC_ERR_ARG;
    if (n_streams <= 0) return 0;
    if (!d_pac || !d_pcm) return PAC_ERR_ARG;
    if (ctx->cfg.use_sbr && !ctx->cfg.use_vq) return PAC_ERR_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    PAC_CHECK(cudaSetDevice(ctx->device));
    const int C = ctx->cfg.n_channels, S = n_streams;
    std::vector<long long> cbOff;
    int rc = 0;
    if (h_counts) {                                        // the caller has walked the streams already (pac_decode_count_async)
        cbOff.assign(S + 1, 0);
        for (int s = 0; s < S; ++s) cbOff[s + 1] = cbOff[s] + h_counts[s] - h_counts[s] % C;
    } else if ((rc = walk_streams(ctx, d_pac, h_stream_off, S, st, cbOff, nullptr, nullptr))) return rc;
    std::vector<long long> pcmOff(S + 1, 0);
    for (int s = 0; s < S; ++s) pcmOff[s + 1] = pcmOff[s] + ((cbOff[s + 1] - cbOff[s]) / C + 1) * 1024 * C;
    memcpy(h_pcm_off, pcmOff.data(), (S + 1) * sizeof(long long));
    if (pcmOff[S] > pcm_capacity) return PAC_ERR_ARG;
    const long long nCb = cbOff[S];
    size_t o = 0;
    auto carve = [&](size_t b) { size_t r = o; o += dalign(b); return r; };
    size_t oSOff = carve((S + 1) * sizeof(long long));
    size_t oCnt = carve(S * sizeof(int));
    size_t oCbOff = carve((S + 1) * sizeof(long long));
    size_t oPcmOff = carve((S + 1) * sizeof(long long));
    size_t oPay = carve((size_t)(nCb + C) * sizeof(long long));
    size_t oPayB = carve((size_t)(nCb + C) * sizeof(int));
    size_t oErr = carve(sizeof(int));
    size_t oFrames = carve((size_t)nCb * 2048 * sizeof(double));
    // VQ streams: dequantised lines + workspaces for one batch of channel-blocks at a time (reused batch after batch)
    // (band layouts the batched front end is not sized for fall back to the in-CTA parse of decode_front)
    const bool vq = ctx->cfg.use_vq != 0 && ctx->host.longBands.nBands <= PRE_SLOTS &&
                    PAC_SUBFRAMES * ctx->host.shortBands.nBands <= PRE_SLOTS && ctx->host.sbrNOmitted <= 4;
    const long long maxB = ctx->optBatchCb > 0 ? ctx->optBatchCb : DEC_BATCH_CB;    // pac_set_option: small batches exercise the batch loop
    const long long batchCb = vq ? std::min<long long>(nCb, maxB) : 0;
    const long long batchGroups = (batchCb + 31) / 32;
    size_t oPre = carve((size_t)batchCb * sizeof(CbHdr));
    size_t oLines = carve((size_t)batchGroups * 32 * 1024 * sizeof(double));
    size_t oWs = carve((size_t)batchGroups * 32 * 2 * PRE_WS_PER_CB * sizeof(double));
    size_t oPerm = carve((size_t)(2 * batchCb + 4) * sizeof(int));     // perm | counts[4] | blocks left to k_decode
    void* scr = nullptr;
    if ((rc = pac_scratch(ctx, o, &scr))) return rc;
    unsigned char* sb = (unsigned char*)scr;
    // pac_scratch may have re-allocated: upload the prefix arrays again
    PAC_CHECK(cudaMemcpyAsync(sb + oSOff, h_stream_off, (S + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
    PAC_CHECK(cudaMemcpyAsync(sb + oCbOff, cbOff.data(), (S + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
    PAC_CHECK(cudaMemcpyAsync(sb + oPcmOff, pcmOff.data(), (S + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
    PAC_CHECK(cudaMemsetAsync(sb + oErr, 0, sizeof(int), st));
    // a trailing partial block (fewer than C channel-blocks) lands in the C spare slots and is ignored
    k_walk<<<(S + WALK_WARPS - 1) / WALK_WARPS, WALK_WARPS * 32, 0, st>>>(d_pac, (long long*)(sb + oSOff), S, pac_header_bytes(ctx), nullptr,
                                            (long long*)(sb + oCbOff), (long long*)(sb + oPay), (int*)(sb + oPayB));
    ctx->kevN[2] = ctx->kevN[3] = 0;
    ctx->launches += 2;       // k_walk, twice (count, then offsets)
    const int oneChunk[2] = {0, S};
    if (!chunks) { chunks = oneChunk; n_chunks = 1; }
    for (int c = 0; c < n_chunks; ++c) {
    const int s0 = chunks[c], s1 = chunks[c + 1];
    const long long cb0 = cbOff[s0], nCbC = cbOff[s1] - cb0;
    if (nCbC > 0) {
        DecArgs a;
        a.buf = d_pac; a.payOff = (long long*)(sb + oPay); a.payBytes = (int*)(sb + oPayB);
        a.frames = (double*)(sb + oFrames); a.err = (int*)(sb + oErr);
        a.bufBytes = h_stream_off[S];
        size_t smem = dec_smem_bytes() + (size_t)ctx->optSmemPad;
        const size_t smemL = dec_lean_smem_bytes();
        PAC_CHECK(cudaFuncSetAttribute(k_decode<double, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        PAC_CHECK(cudaFuncSetAttribute(k_decode<float, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        PAC_CHECK(cudaFuncSetAttribute(k_decode<double, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemL));
        PAC_CHECK(cudaFuncSetAttribute(k_decode<float, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemL));
        PAC_CHECK(cudaFuncSetAttribute(k_decode_long<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dl_smem_bytes<double>()));
        PAC_CHECK(cudaFuncSetAttribute(k_decode_long<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dl_smem_bytes<float>()));
        if (c == 0) cudaEventRecord(ctx->ev[2], st);      // pac_last_kernel_ms: the first chunk (the only one unless chunked)
        for (long long q0 = 0; q0 < nCbC; q0 += vq ? batchCb : nCbC) {
            const long long nq = vq ? std::min(batchCb, nCbC - q0) : nCbC;
            a.cb0 = cb0 + q0;
            CbHdr* dPre = vq ? (CbHdr*)(sb + oPre) : nullptr;
            double* dLines = vq ? (double*)(sb + oLines) : nullptr;
            int* dPerm = (int*)(sb + oPerm);
            int* dCounts = dPerm + batchCb;
            if (vq) {
                const int nChunks = (int)((nq + DQ_NT - 1) / DQ_NT);
                PAC_CHECK(cudaMemsetAsync(dCounts, 0, 4 * sizeof(int), st));
                k_hdr<<<(unsigned)((nq + 127) / 128), 128, 0, st>>>(ctx->dev, a, a.cb0, (int)nq, dPre, dPerm, dCounts);
                pac_kev_begin(ctx, 2, st);
                k_dequant<<<(unsigned)(nChunks * PRE_SLOTS), DQ_NT, 0, st>>>(ctx->dev, a, a.cb0, (int)nq, nChunks, dPre, dLines,
                                                                          (double*)(sb + oWs), dPerm, dCounts);
                pac_kev_end(ctx, 2, st);
                ++ctx->launches;      // k_hdr
            }
            pac_kev_begin(ctx, 3, st);
            if (vq && !(ctx->optVariant & 16)) {
                // long blocks: one warp each (decode_long.cuh); transient and SBR blocks: the CTA-per-block kernel
                const unsigned gL = (unsigned)std::min<long long>(148, (nq + DL_WARPS - 1) / DL_WARPS);
                const unsigned gS = (unsigned)std::min<long long>(148 * DEC_PRE_MINB, nq);
                double* fr = a.frames + a.cb0 * 2048;
                if (ctx->cfg.precision == PAC_F32) {
                    k_decode_long<float><<<gL, DL_WARPS * 32, dl_smem_bytes<float>(), st>>>(ctx->dev, fr, dPre, dLines, dPerm, dCounts);
                    k_decode<float, true><<<gS, DEC_NT, smemL, st>>>(ctx->dev, a, dPre, dLines, dCounts);
                } else {
                    k_decode_long<double><<<gL, DL_WARPS * 32, dl_smem_bytes<double>(), st>>>(ctx->dev, fr, dPre, dLines, dPerm, dCounts);
                    k_decode<double, true><<<gS, DEC_NT, smemL, st>>>(ctx->dev, a, dPre, dLines, dCounts);
                }
                ++ctx->launches;
            } else if (vq) {
                if (ctx->cfg.precision == PAC_F32) k_decode<float, true><<<(unsigned)nq, DEC_NT, smemL, st>>>(ctx->dev, a, dPre, dLines, nullptr);
                else k_decode<double, true><<<(unsigned)nq, DEC_NT, smemL, st>>>(ctx->dev, a, dPre, dLines, nullptr);
            } else {
                if (ctx->cfg.precision == PAC_F32) k_decode<float, false><<<(unsigned)nq, DEC_NT, smem, st>>>(ctx->dev, a, dPre, dLines, nullptr);
                else k_decode<double, false><<<(unsigned)nq, DEC_NT, smem, st>>>(ctx->dev, a, dPre, dLines, nullptr);
            }
            pac_kev_end(ctx, 3, st);
        }
        if (c == 0) { cudaEventRecord(ctx->ev[3], st); ctx->evValid[1] = 1; }
        PAC_CHECK(cudaGetLastError());
    }
    const long long g0 = pcmOff[s0] / 1024, nOut = pcmOff[s1] / 1024 - g0;
    if (nOut > 0) ++ctx->launches;
    if (nOut > 0)
        k_ola<<<(unsigned)nOut, 256, 0, st>>>((const double*)(sb + oFrames), (long long*)(sb + oCbOff),
                                              (long long*)(sb + oPcmOff), S, C, d_pcm, g0);
    PAC_CHECK(cudaGetLastError());
    if (chunkEv) {
        PAC_CHECK(cudaEventRecord(chunkEv[c], st));
        if (h_copy_dst && pcmOff[s1] > pcmOff[s0]) {
            PAC_CHECK(cudaStreamWaitEvent(copyStream, chunkEv[c], 0));
            PAC_CHECK(cudaMemcpyAsync(h_copy_dst + pcmOff[s0], d_pcm + pcmOff[s0], (size_t)(pcmOff[s1] - pcmOff[s0]) * sizeof(short),
                                      cudaMemcpyDeviceToHost, copyStream));
        }
    }
    }
    int err = 0;
    PAC_CHECK(cudaMemcpyAsync(&err, sb + oErr, sizeof(int), cudaMemcpyDeviceToHost, st));
    PAC_CHECK(cudaStreamSynchronize(st));
    if (h_err) *h_err = err;
    return err ? PAC_ERR_SOFT : 0;
}

extern "C" int pac_decode_streams(pac_ctx* ctx, const unsigned char* d_pac, const long long* h_stream_off,
                                  int n_streams, short* d_pcm, long long pcm_capacity, long long* h_pcm_off,
                                  int* h_err, void* stream) {
    return pac_decode_streams_chunked(ctx, d_pac, h_stream_off, n_streams, d_pcm, pcm_capacity, h_pcm_off, h_err, stream,
                                      nullptr, 0, nullptr, nullptr, nullptr, nullptr);
}

#ifdef PAC_PHASE_TIMING
// debug: copy and reset the decoder's per-phase cycle counters ([0..15] long frames, [16..31] short sub-frames)
extern "C" int pac_debug_phases_dec(unsigned long long* h_out) {
    unsigned long long z[32] = {0};
    cudaMemcpyFromSymbol(h_out, g_pac_dphase, sizeof(z));
    cudaMemcpyToSymbol(g_pac_dphase, z, sizeof(z));
    return 0;
}
#endif
