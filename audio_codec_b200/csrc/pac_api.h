// Internal: the context behind the opaque pac_ctx of include/pac_b200.h.
#pragma once
#include <vector>
#include "../../include/pac_b200.h"
#include "pac_common.cuh"

#define PAC_KEV_MAX 64
struct pac_ctx {
    int device;
    pac_config cfg;
    PacDev host;     // table pointers are DEVICE pointers; scalars usable on host
    PacDev* dev;     // device copy of `host`
    std::vector<void*> allocs;
    // lazily grown scratch for the fused encode / decode paths
    void* scratch;
    size_t scratchBytes;
    void* pinned;
    size_t pinnedBytes;
    void* io[2];         // device staging for the host-buffer entry points
    size_t ioBytes[2];
    cudaStream_t cp[2];  // copy streams of the host-buffer entry points (created on first use)
    cudaEvent_t cpEv[2];
    cudaEvent_t chunkEv[8];
    int cpReady;
    cudaEvent_t ev[4];   // around k_encode and k_decode, for pac_last_kernel_ms
    int evValid[2];
    // per-kernel device time of the last encode / decode call: events around every launch of
    // 0 k_encode, 1 k_pvq, 2 k_dequant, 3 k_decode (one launch per batch of channel-blocks)
    cudaEvent_t kev[4][PAC_KEV_MAX][2];
    int kevN[4];
    long long launches;  // kernels launched by the library on this context
    // tuning knobs (pac_set_option): not part of the codec configuration, results do not depend on them
    long long optBatchCb;    // channel-blocks per batch of the thread-per-band kernels (0: default)
    int optSmemPad;          // extra dynamic shared memory for the frame kernels (occupancy experiments)
    double optHostChunkMb;   // smallest chunk of the host-buffer pipeline, MB of PCM (0: default)
    int optVariant;          // 0: current kernels; bit flags select earlier forms of a stage (A/B timing, same results)
    // windows of arbitrary length for pac_window, built on demand
    struct WinEntry { int kind, N, Ns; double alpha; const double* d; const float* f; };
    std::vector<WinEntry> winCache;
};

// events around one kernel launch (kind as in kev); launches beyond PAC_KEV_MAX per call go untimed
static inline void pac_kev_begin(pac_ctx* c, int kind, cudaStream_t st) {
    if (c->kevN[kind] < PAC_KEV_MAX) cudaEventRecord(c->kev[kind][c->kevN[kind]][0], st);
}
static inline void pac_kev_end(pac_ctx* c, int kind, cudaStream_t st) {
    if (c->kevN[kind] < PAC_KEV_MAX) cudaEventRecord(c->kev[kind][c->kevN[kind]][1], st);
    ++c->kevN[kind];
    ++c->launches;
}
int pac_scratch(pac_ctx* ctx, size_t bytes, void** out);
// ops.cu: device address of the static scratch of the no-allocation entry point, once per device (`device` current)
int pac_ops_static_setup(int device);
int pac_pinned(pac_ctx* ctx, size_t bytes, void** out);
int pac_io(pac_ctx* ctx, int which, size_t bytes, void** out);
// window table of length N (kind: PAC_W_*, Ns = short length for the Edler shapes)
int pac_window_table(pac_ctx* ctx, int kind, int N, int Ns, double alpha, const double** d, const float** f);

// decode.cu: pac_decode_streams launched per chunk of streams, chunkEv[c] recorded after chunk c
int pac_decode_streams_chunked(pac_ctx* ctx, const unsigned char* d_pac, const long long* h_stream_off, int n_streams,
                               short* d_pcm, long long pcm_capacity, long long* h_pcm_off, int* h_err, void* stream,
                               const int* chunks, int n_chunks, cudaEvent_t* chunkEv, cudaStream_t copyStream,
                               short* h_copy_dst, const int* h_counts);
