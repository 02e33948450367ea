// Internal: the context behind the opaque pac_ctx of include/pac_b200.h.
#pragma once
#include <vector>
#include "../../include/pac_b200.h"
#include "pac_common.cuh"

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
};

int pac_scratch(pac_ctx* ctx, size_t bytes, void** out);
int pac_pinned(pac_ctx* ctx, size_t bytes, void** out);
