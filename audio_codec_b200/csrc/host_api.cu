// Host-buffer entry points: the calls a host program (the reference's PACFile
// driver, coder/pacfile.py:674-756) makes with ordinary memory.  They stage
// through context-owned device buffers.  Large batches are cut into a few chunks
// of whole streams so the copies ride under the kernels: while chunk c is being
// coded, chunk c+1's input is already on its way up (copy stream 0) and chunk
// c-1's output on its way down (copy stream 1).  Pageable host memory still
// works; the copies then just do not overlap.
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <vector>
#include "pac_api.h"

static int copy_streams(pac_ctx* ctx) {
    if (ctx->cpReady) return 0;
    for (int i = 0; i < 2; ++i) {
        PAC_CHECK(cudaStreamCreateWithFlags(&ctx->cp[i], cudaStreamNonBlocking));
        PAC_CHECK(cudaEventCreateWithFlags(&ctx->cpEv[i], cudaEventDisableTiming));
    }
    for (int i = 0; i < 8; ++i) PAC_CHECK(cudaEventCreateWithFlags(&ctx->chunkEv[i], cudaEventDisableTiming));
    ctx->cpReady = 1;
    return 0;
}

// Smallest chunk worth its per-call latency, in MB of PCM (the decoder scales it by the ~6:1
// compression); pac_set_option(PAC_OPT_HOST_CHUNK_MB) overrides it (tests use tiny chunks).
static long long chunk_mb(const pac_ctx* ctx) {
    const double v = ctx->optHostChunkMb > 0.0 ? ctx->optHostChunkMb : 24.0;
    return (long long)(v * 1048576.0);
}

// Chunk boundaries (in streams): up to 6 chunks whose sizes ramp up and down (1 2 3 3 2 1), so the
// first upload and the last download - the two copies nothing can hide - are short; one chunk for
// small batches (the per-call latency of the stream walk and the scans would dominate).
static std::vector<int> chunk_bounds(const std::vector<long long>& weight, long long minChunk) {
    const int n = (int)weight.size();
    long long tot = 0;
    for (long long w : weight) tot += w;
    int k = (int)std::min<long long>(6, std::max<long long>(1, tot / std::max<long long>(1, minChunk)));
    k = std::min(k, n);
    static const int ramp[7][6] = {{0}, {1}, {1, 1}, {1, 2, 1}, {1, 2, 2, 1}, {1, 2, 3, 2, 1}, {1, 2, 3, 3, 2, 1}};
    int rsum = 0;
    for (int c = 0; c < k; ++c) rsum += ramp[k][c];
    std::vector<int> b{0};
    long long acc = 0;
    int racc = ramp[k][0];
    for (int s = 0; s < n; ++s) {
        acc += weight[s];
        if ((int)b.size() < k && acc * rsum >= tot * racc && s + 1 < n) {
            b.push_back(s + 1);
            racc += ramp[k][b.size() - 1];
        }
    }
    b.push_back(n);
    return b;
}

extern "C" int pac_encode_host(pac_ctx* ctx, const short* h_pcm, const int* h_nblocks, const int* h_hdr_samples,
                               int n_streams, unsigned char* h_out, long long out_capacity,
                               long long* h_stream_off, int* h_err, void* stream) {
    if (!ctx || !h_nblocks || !h_hdr_samples || !h_stream_off) return PAC_ERR_ARG;   // as pac_encode_streams
    if (n_streams <= 0) return 0;
    if (!h_pcm || !h_out) return PAC_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    PAC_CHECK(cudaSetDevice(ctx->device));
    int rc;
    if ((rc = copy_streams(ctx))) return rc;
    const size_t perBlock = (size_t)ctx->cfg.n_channels * 1024 * sizeof(short);
    std::vector<long long> weight(n_streams);
    long long blocks = 0;
    for (int s = 0; s < n_streams; ++s) { weight[s] = h_nblocks[s]; blocks += h_nblocks[s]; }
    const std::vector<int> cb = chunk_bounds(weight, std::max<long long>(1, chunk_mb(ctx) / (long long)perBlock));
    const int nChunks = (int)cb.size() - 1;
    long long bound = 0;                                    // per-chunk bounds laid end to end
    for (int c = 0; c < nChunks; ++c) bound += pac_encode_bound(ctx, h_nblocks + cb[c], cb[c + 1] - cb[c]);
    void *dIn = nullptr, *dOut = nullptr;
    if ((rc = pac_io(ctx, 0, (size_t)blocks * perBlock, &dIn))) return rc;
    if ((rc = pac_io(ctx, 1, (size_t)bound, &dOut))) return rc;
    std::vector<long long> blk0(nChunks + 1, 0);
    for (int c = 0; c < nChunks; ++c) {
        blk0[c + 1] = blk0[c];
        for (int s = cb[c]; s < cb[c + 1]; ++s) blk0[c + 1] += h_nblocks[s];
    }
    auto upload = [&](int c) -> int {
        const size_t o = (size_t)blk0[c] * perBlock, n = (size_t)(blk0[c + 1] - blk0[c]) * perBlock;
        PAC_CHECK(cudaMemcpyAsync((char*)dIn + o, (const char*)h_pcm + o, n, cudaMemcpyHostToDevice, ctx->cp[0]));
        return 0;
    };
    // the caller's stream may still be producing h_pcm / using the staging buffers
    PAC_CHECK(cudaStreamSynchronize(st));
    if ((rc = upload(0))) return rc;
    int soft = 0, errAll = 0;
    long long outPos = 0;
    h_stream_off[0] = 0;
    std::vector<long long> off;
    for (int c = 0; c < nChunks; ++c) {
        PAC_CHECK(cudaEventRecord(ctx->cpEv[0], ctx->cp[0]));
        PAC_CHECK(cudaStreamWaitEvent(st, ctx->cpEv[0], 0));
        if (c + 1 < nChunks && (rc = upload(c + 1))) return rc;
        const int s0 = cb[c], ns = cb[c + 1] - s0;
        const long long cbound = pac_encode_bound(ctx, h_nblocks + s0, ns);
        if (outPos + cbound > bound) return PAC_ERR_ARG;
        off.assign(ns + 1, 0);
        int e = 0;
        rc = pac_encode_streams(ctx, (const short*)((char*)dIn + (size_t)blk0[c] * perBlock), h_nblocks + s0,
                                h_hdr_samples + s0, ns, (unsigned char*)dOut + outPos, cbound,
                                off.data(), &e, stream);
        if (rc != 0 && rc != PAC_ERR_SOFT) return rc;
        soft |= rc == PAC_ERR_SOFT;
        errAll |= e;
        const long long total = off[ns];
        if (outPos + total > out_capacity) return PAC_ERR_ARG;
        // pac_encode_streams has synchronised st: the chunk is complete, send it down behind the next one
        PAC_CHECK(cudaMemcpyAsync(h_out + outPos, (unsigned char*)dOut + outPos, (size_t)total, cudaMemcpyDeviceToHost,
                                  ctx->cp[1]));
        for (int s = 0; s < ns; ++s) h_stream_off[s0 + s + 1] = outPos + off[s + 1];
        outPos += total;
    }
    PAC_CHECK(cudaStreamSynchronize(ctx->cp[1]));
    if (h_err) *h_err = errAll;
    return soft ? PAC_ERR_SOFT : 0;
}

extern "C" int pac_decode_host(pac_ctx* ctx, const unsigned char* h_pac, const long long* h_stream_off, int n_streams,
                               short* h_pcm, long long pcm_capacity, long long* h_pcm_off, int* h_err, void* stream) {
    if (!ctx || !h_stream_off || !h_pcm_off) return PAC_ERR_ARG;
    if (n_streams <= 0) return 0;
    if (!h_pac) return PAC_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    PAC_CHECK(cudaSetDevice(ctx->device));
    int rc;
    if ((rc = copy_streams(ctx))) return rc;
    const size_t pacBytes = (size_t)(h_stream_off[n_streams] - h_stream_off[0]);
    void *dIn = nullptr, *dOut = nullptr;
    if ((rc = pac_io(ctx, 0, pacBytes, &dIn))) return rc;
    // the coded input is ~1/6 of the PCM: one copy; the PCM output is what gets pipelined
    PAC_CHECK(cudaMemcpyAsync(dIn, h_pac + h_stream_off[0], pacBytes, cudaMemcpyHostToDevice, st));
    std::vector<long long> off(n_streams + 1);
    for (int s = 0; s <= n_streams; ++s) off[s] = h_stream_off[s] - h_stream_off[0];
    if (!h_pcm)                                            // plan only: caller sizes h_pcm from h_pcm_off
        return pac_decode_plan(ctx, (const unsigned char*)dIn, off.data(), n_streams, h_pcm_off, stream);
    // sizes first (one walk over the length prefixes), so the staging buffer is allocated once
    if ((rc = pac_decode_plan(ctx, (const unsigned char*)dIn, off.data(), n_streams, h_pcm_off, stream))) return rc;
    const long long totalAll = h_pcm_off[n_streams];
    if (totalAll > pcm_capacity) return PAC_ERR_ARG;
    if ((rc = pac_io(ctx, 1, (size_t)totalAll * sizeof(short), &dOut))) return rc;
    std::vector<long long> weight(n_streams);
    for (int s = 0; s < n_streams; ++s) weight[s] = off[s + 1] - off[s];
    const std::vector<int> cb = chunk_bounds(weight, std::max<long long>(1, chunk_mb(ctx) / 6));
    const int nChunks = (int)cb.size() - 1;
    // one walk, the kernels launched per chunk with an event after each: the copy stream takes every
    // finished chunk down while the next ones decode
    PAC_CHECK(cudaStreamSynchronize(ctx->cp[1]));         // the staging buffer may still be draining from the last call
    int errAll = 0;
    // the plan above has counted the channel-blocks of every stream: hand the counts on instead of walking again
    std::vector<int> counts(n_streams);
    for (int s = 0; s < n_streams; ++s)
        counts[s] = (int)(((h_pcm_off[s + 1] - h_pcm_off[s]) / (1024LL * ctx->cfg.n_channels) - 1) * ctx->cfg.n_channels);
    rc = pac_decode_streams_chunked(ctx, (const unsigned char*)dIn, off.data(), n_streams, (short*)dOut, totalAll, h_pcm_off,
                                    &errAll, stream, cb.data(), nChunks, ctx->chunkEv, ctx->cp[1], h_pcm, counts.data());
    if (rc != 0 && rc != PAC_ERR_SOFT) return rc;
    const int soft = rc == PAC_ERR_SOFT;
    PAC_CHECK(cudaStreamSynchronize(ctx->cp[1]));
    if (h_err) *h_err = errAll;
    return soft ? PAC_ERR_SOFT : 0;
}
