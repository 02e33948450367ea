// Host-buffer entry points: the calls a host program (the reference's PACFile
// driver, coder/pacfile.py:674-756) makes with ordinary memory.  They stage
// through context-owned device buffers: H2D copy, the device pipeline, D2H copy.
#include <string.h>
#include <vector>
#include "pac_api.h"

extern "C" int pac_encode_host(pac_ctx* ctx, const short* h_pcm, const int* h_nblocks, const int* h_hdr_samples,
                               int n_streams, unsigned char* h_out, long long out_capacity,
                               long long* h_stream_off, int* h_err, void* stream) {
    if (!ctx || !h_nblocks || !h_stream_off) return PAC_ERR_ARG;
    if (n_streams <= 0) return 0;
    if (!h_pcm || !h_out) return PAC_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    PAC_CHECK(cudaSetDevice(ctx->device));
    long long blocks = 0;
    for (int s = 0; s < n_streams; ++s) blocks += h_nblocks[s];
    const size_t pcmBytes = (size_t)blocks * ctx->cfg.n_channels * 1024 * sizeof(short);
    const long long bound = pac_encode_bound(ctx, h_nblocks, n_streams);
    void *dIn = nullptr, *dOut = nullptr;
    int rc;
    if ((rc = pac_io(ctx, 0, pcmBytes, &dIn))) return rc;
    if ((rc = pac_io(ctx, 1, (size_t)bound, &dOut))) return rc;
    PAC_CHECK(cudaMemcpyAsync(dIn, h_pcm, pcmBytes, cudaMemcpyHostToDevice, st));
    rc = pac_encode_streams(ctx, (const short*)dIn, h_nblocks, h_hdr_samples, n_streams, (unsigned char*)dOut, bound,
                            h_stream_off, h_err, stream);
    if (rc != 0 && rc != PAC_ERR_SOFT) return rc;
    const long long total = h_stream_off[n_streams];
    if (total > out_capacity) return PAC_ERR_ARG;
    PAC_CHECK(cudaMemcpyAsync(h_out, dOut, (size_t)total, cudaMemcpyDeviceToHost, st));
    PAC_CHECK(cudaStreamSynchronize(st));
    return rc;
}

extern "C" int pac_decode_host(pac_ctx* ctx, const unsigned char* h_pac, const long long* h_stream_off, int n_streams,
                               short* h_pcm, long long pcm_capacity, long long* h_pcm_off, int* h_err, void* stream) {
    if (!ctx || !h_stream_off || !h_pcm_off) return PAC_ERR_ARG;
    if (n_streams <= 0) return 0;
    if (!h_pac) return PAC_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    PAC_CHECK(cudaSetDevice(ctx->device));
    const size_t pacBytes = (size_t)(h_stream_off[n_streams] - h_stream_off[0]);
    void *dIn = nullptr, *dOut = nullptr;
    int rc;
    if ((rc = pac_io(ctx, 0, pacBytes, &dIn))) return rc;
    PAC_CHECK(cudaMemcpyAsync(dIn, h_pac + h_stream_off[0], pacBytes, cudaMemcpyHostToDevice, st));
    std::vector<long long> off(n_streams + 1);
    for (int s = 0; s <= n_streams; ++s) off[s] = h_stream_off[s] - h_stream_off[0];
    if ((rc = pac_decode_plan(ctx, (const unsigned char*)dIn, off.data(), n_streams, h_pcm_off, stream))) return rc;
    const long long total = h_pcm_off[n_streams];
    if (!h_pcm) return 0;                                  // plan only: caller sizes h_pcm from h_pcm_off
    if (total > pcm_capacity) return PAC_ERR_ARG;
    if ((rc = pac_io(ctx, 1, (size_t)total * sizeof(short), &dOut))) return rc;
    rc = pac_decode_streams(ctx, (const unsigned char*)dIn, off.data(), n_streams, (short*)dOut, total, h_pcm_off,
                            h_err, stream);
    if (rc != 0 && rc != PAC_ERR_SOFT) return rc;
    PAC_CHECK(cudaMemcpyAsync(h_pcm, dOut, (size_t)total * sizeof(short), cudaMemcpyDeviceToHost, st));
    PAC_CHECK(cudaStreamSynchronize(st));
    return rc;
}
