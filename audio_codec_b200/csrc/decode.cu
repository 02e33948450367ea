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
#include "pac_api.h"

extern "C" int pac_header_bytes(const pac_ctx* ctx);
#define DEC_BATCH_CB 65536       // channel-blocks per k_parse / k_decode pair (bounds the parse scratch: 18 KB each)

// pass 0: count channel-blocks; pass 1: record payload offsets (absolute, past the prefix).
// The prefix chain is serial, but its cost is the memory latency per hop: one WARP per stream
// fetches a 2 KB window (16-byte loads, all in flight together) into shared memory and lane 0
// follows the chain inside it, so DRAM latency is paid once per ~6 records instead of per record.
#define WALK_WARPS 4
#define WALK_WIN 2048
__global__ void __launch_bounds__(WALK_WARPS * 32)
k_walk(const unsigned char* __restrict__ buf, const long long* __restrict__ streamOff, int nStreams,
       int headerBytes, int* __restrict__ counts, const long long* __restrict__ cbOff,
       long long* __restrict__ payOff, int* __restrict__ payBytes) {
    __shared__ __align__(16) unsigned char win[WALK_WARPS][WALK_WIN];
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
#pragma unroll
        for (int j = 0; j < WALK_WIN / 512; ++j) {
            const long long o = w0 + (j * 32 + lane) * 16;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (o < bufEnd && o >= 0) v = __ldg(reinterpret_cast<const uint4*>(buf + o));
            *reinterpret_cast<uint4*>(&win[warp][(j * 32 + lane) * 16]) = v;
        }
        __syncwarp();
        if (lane == 0) {
            const long long wEnd = w0 + WALK_WIN;
            while (pos + 4 <= end && pos + 4 <= wEnd) {
                const unsigned char* q = &win[warp][pos - w0];
                const unsigned nb = (unsigned)q[0] | ((unsigned)q[1] << 8) | ((unsigned)q[2] << 16) | ((unsigned)q[3] << 24);
                if (pos + 4 + (long long)nb > end) { done = true; break; }      // truncated block
                if (payOff && n < limit) {                 // a damaged stream may end on a partial block: its
                    payOff[base + n] = pos + 4;             // extra record must not land in the next stream's slot
                    payBytes[base + n] = (int)nb;
                }
                ++n;
                pos += 4 + nb;
            }
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

// Header + split-tree parse of VQ streams, ahead of k_decode.  A CTA takes PARSE_CBS channel-blocks;
// lane = channel-block, so a warp parses the SAME band of 32 blocks at a time (similar trees, little
// divergence) and the serial, latency-bound chains (field widths depend on the angles just read)
// run thousands at a time per SM instead of 17 per CTA.  Writes CbPre + the node lists.
#define PARSE_CBS 32
#define PARSE_NT 256
__global__ void __launch_bounds__(PARSE_NT)
k_parse(const PacDev* __restrict__ dev, DecArgs a, long long cbBase, int nCb, CbPre* __restrict__ pre,
        DNode* __restrict__ nodesG) {
    __shared__ unsigned short s_pos[PRE_SLOTS][PARSE_CBS];
    __shared__ unsigned char s_ba[PRE_SLOTS][PARSE_CBS];
    __shared__ int s_cnt[PAC_SUBFRAMES][PARSE_CBS];
    __shared__ unsigned char s_flags[PARSE_CBS], s_sbr[PARSE_CBS], s_scale[PAC_SUBFRAMES][PARSE_CBS];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int local = blockIdx.x * PARSE_CBS + lane;          // index into this launch's CbPre / node regions
    const bool have = local < nCb;
    const long long cb = cbBase + local;
    // bit reader over global memory: 8-byte words at absolute alignment (the record may start at any byte)
    BitR br = {nullptr, 0, 0, 0};
    long long bit0 = 0;
    int nBytes = 0;
    if (have) {
        nBytes = min(a.payBytes[cb], DEC_REC_MAX);
        const unsigned char* rec = a.buf + a.payOff[cb];
        const unsigned long long ad = (unsigned long long)rec;
        const unsigned char* al = rec - (ad & 7ull);
        bit0 = 8LL * (long long)(ad & 7ull);
        br.p = al; br.end = bit0 + 8LL * nBytes;
        const long long room = (a.buf + a.bufBytes) - al;     // readable bytes from the aligned base
        br.cap = (int)min((long long)(DEC_REC_MAX + 16), room & ~7LL);
        if (br.cap < (int)(ad & 7ull) + nBytes) {             // the record ends inside the buffer's last partial word
            br.p = rec; br.end = 8LL * nBytes; br.cap = 0; bit0 = 0;
        }
    }
    const int nS = dev->nScaleBits, nM = dev->nMantSizeBits;
    if (warp == 0) {
        unsigned char fl = 0, sbrF = 0;
        for (int f = 0; f < PAC_SUBFRAMES; ++f) s_cnt[f][lane] = 0;
        if (have && nBytes > 0) {
            br.pos = bit0;
            fl = (unsigned char)(br_get(br, 3) | 0x80u);      // last, cur, next (coder/pacfile.py:196-198)
            const bool cur = (fl >> 1) & 1;
            const PacBands& bands = cur ? dev->shortBands : dev->longBands;
            const int nB = bands.nBands, nF = cur ? PAC_SUBFRAMES : 1;
            for (int f = 0; f < nF; ++f) {
                s_scale[f][lane] = (unsigned char)br_get(br, nS);
                bool anyOm = false;
                for (int b = 0; b < nB; ++b) {
                    const int v = (int)br_get(br, nM);
                    const int bab = v ? v + 1 : 0;
                    s_ba[f * nB + b][lane] = (unsigned char)bab;
                    anyOm |= !cur && bands.omitted[b] && bab != 0;
                }
                const bool sbr = !cur && dev->useSBR && anyOm;            // coder/pacfile.py:660-664
                if (sbr) sbrF = 1;
                long long p = br.pos;
                for (int b = 0; b < nB; ++b) {
                    s_pos[f * nB + b][lane] = (unsigned short)min(p - bit0, 65535LL);   // past the end: reads zeros
                    const int bab = s_ba[f * nB + b][lane];
                    if (bab) p += (long long)bab * ((sbr && bands.omitted[b]) ? 1 : bands.nLines[b]);
                }
                if (p > br.end) atomicOr(a.err, PVQ_ERR_STACK);
                br.pos = p;
            }
        }
        s_flags[lane] = fl;
        s_sbr[lane] = sbrF;
    }
    __syncthreads();
    const unsigned char fl = s_flags[lane];
    if (fl & 0x80u) {
        const bool cur = (fl >> 1) & 1, sbr = s_sbr[lane] != 0;
        const PacBands& bands = cur ? dev->shortBands : dev->longBands;
        const int nB = bands.nBands, nSlots = cur ? PAC_SUBFRAMES * nB : nB;
        PvqTab pt = {dev->pvqN, dev->pvqRowOff, dev->log2tab, dev->pvqKW, dev->thetaTab};
        CbPre* pc = pre + local;
        DNode* region = nodesG + (size_t)local * PRE_NODES_PER_CB;
        // widest bands first (they are the long chains)
        for (int q = warp; q < PRE_SLOTS; q += PARSE_NT / 32) {
            if (q >= nSlots) break;
            const int f = q / nB, b = nB - 1 - (q - f * nB), slot = f * nB + b;
            const int bb = s_ba[slot][lane];
            if (!bb) continue;
            int cap = 1, base = 0;
            while (cap < bands.nLines[b]) cap <<= 1;
            for (int j = 0; j < b; ++j) { int c = 1; while (c < bands.nLines[j]) c <<= 1; base += c; }
            const bool env = sbr && bands.omitted[b];          // one envelope value (coder/codec.py:111-132)
            br.pos = bit0 + s_pos[slot][lane];
            int werr = 0;
            double gain;
            short tail;
            thread_parse_band(pt, br, env ? bb : bb * bands.nLines[b], env ? 1 : bands.nLines[b], base, cap,
                              region + (cur ? f * PRE_SHORT_NODES : 0), &s_cnt[f][lane],
                              cur ? PRE_SHORT_NODES : DEC_MAX_NODES, &gain, &tail, &werr);
            if (werr) atomicOr(a.err, werr);
            pc->gain[slot] = gain;
            pc->tail[slot] = tail;
        }
    }
    __syncthreads();
    if (warp == 0 && have) {
        CbPre* pc = pre + local;
        const bool cur = (fl >> 1) & 1;
        const int nB = cur ? dev->shortBands.nBands : dev->longBands.nBands, nSlots = (fl & 0x80u) ? (cur ? PAC_SUBFRAMES * nB : nB) : 0;
        for (int q = 0; q < nSlots; ++q) pc->ba[q] = s_ba[q][lane];
        for (int f = 0; f < PAC_SUBFRAMES; ++f) {
            pc->nNodes[f] = (unsigned short)min(s_cnt[f][lane], 65535);
            pc->scale[f] = s_scale[f][lane];
        }
        pc->sbr = s_sbr[lane];
        pc->flags = fl;
    }
}

template <typename T>
__global__ void __launch_bounds__(DEC_NT, 3)
k_decode(const PacDev* __restrict__ dev, DecArgs a, const CbPre* __restrict__ preAll, const DNode* __restrict__ nodesG) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    DecSmem S;
    dec_smem_carve(smem_raw, S);
    const int tid = threadIdx.x;
    const long long cb = (long long)blockIdx.x + a.cb0;
    int nBytes = a.payBytes[cb];
    if (nBytes > DEC_REC_MAX) {
        if (tid == 0) atomicOr(a.err, PVQ_ERR_STACK);
        nBytes = DEC_REC_MAX;
    }
    const CbPre* pre = preAll ? preAll + blockIdx.x : nullptr;
    const DNode* region = preAll ? nodesG + (size_t)blockIdx.x * PRE_NODES_PER_CB : nullptr;
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
            decode_frame<PAC_NLONG, T>(dev, S, recBits, shape, 0, a.err, pre, region, 0);
        } else {
            for (int sf = 0; sf < PAC_SUBFRAMES; ++sf)
                decode_frame<PAC_NSHORT, T>(dev, S, recBits, PAC_WIN_SINE, 448 + 128 * sf, a.err, pre,
                                            region ? region + sf * PRE_SHORT_NODES : nullptr, sf);
        }
    }
    double* dst = a.frames + cb * 2048;
    for (int i = tid; i < 2048; i += DEC_NT) dst[i] = S.y[i];
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
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) {
        double v = (tail ? tail[i] : 0.0) + (head ? head[i] : 0.0);
        dst[i] = double_to_pcm(v);
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
                               cudaStream_t copyStream, short* h_copy_dst) {
    if (!ctx || !h_stream_off || !h_pcm_off) return PAC_ERR_ARG;
    if (n_streams <= 0) return 0;
    if (!d_pac || !d_pcm) return PAC_ERR_ARG;
    if (ctx->cfg.use_sbr && !ctx->cfg.use_vq) return PAC_ERR_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    PAC_CHECK(cudaSetDevice(ctx->device));
    const int C = ctx->cfg.n_channels, S = n_streams;
    std::vector<long long> cbOff;
    int rc = walk_streams(ctx, d_pac, h_stream_off, S, st, cbOff, nullptr, nullptr);
    if (rc) return rc;
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
    // VQ streams: k_parse output for one batch of channel-blocks at a time (reused batch after batch)
    const bool vq = ctx->cfg.use_vq != 0;
    const long long batchCb = vq ? std::min<long long>(nCb, DEC_BATCH_CB) : 0;
    size_t oPre = carve((size_t)batchCb * sizeof(CbPre));
    size_t oNodes = carve((size_t)batchCb * PRE_NODES_PER_CB * sizeof(DNode));
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
        size_t smem = dec_smem_bytes() + (getenv("PAC_SMEM_PAD") ? atoi(getenv("PAC_SMEM_PAD")) : 0);
        PAC_CHECK(cudaFuncSetAttribute(k_decode<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        PAC_CHECK(cudaFuncSetAttribute(k_decode<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (c == 0) cudaEventRecord(ctx->ev[2], st);      // pac_last_kernel_ms: the first chunk (the only one unless chunked)
        for (long long q0 = 0; q0 < nCbC; q0 += vq ? batchCb : nCbC) {
            const long long nq = vq ? std::min(batchCb, nCbC - q0) : nCbC;
            a.cb0 = cb0 + q0;
            CbPre* dPre = vq ? (CbPre*)(sb + oPre) : nullptr;
            DNode* dNodes = vq ? (DNode*)(sb + oNodes) : nullptr;
            if (vq) k_parse<<<(unsigned)((nq + PARSE_CBS - 1) / PARSE_CBS), PARSE_NT, 0, st>>>(ctx->dev, a, a.cb0, (int)nq, dPre, dNodes);
            if (ctx->cfg.precision == PAC_F32) k_decode<float><<<(unsigned)nq, DEC_NT, smem, st>>>(ctx->dev, a, dPre, dNodes);
            else k_decode<double><<<(unsigned)nq, DEC_NT, smem, st>>>(ctx->dev, a, dPre, dNodes);
        }
        if (c == 0) { cudaEventRecord(ctx->ev[3], st); ctx->evValid[1] = 1; }
        PAC_CHECK(cudaGetLastError());
    }
    const long long g0 = pcmOff[s0] / 1024, nOut = pcmOff[s1] / 1024 - g0;
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
                                      nullptr, 0, nullptr, nullptr, nullptr);
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
