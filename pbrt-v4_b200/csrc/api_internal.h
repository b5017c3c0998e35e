// api_internal.h - the handle structs behind include/b200pt.h, shared by api.cpp (geometry
// entry points) and render_api.cpp (render entry points).
#pragma once
#include <vector>
#include <cuda_runtime.h>

#include <string>

#include "../../include/b200pt.h"
#include "device_scene.h"
#include "scene_host.h"

struct b200pt_ctx {
    // host-buffer calls are pipelined in chunks: copy-in, traversal and copy-out of successive chunks overlap
    cudaStream_t sIn = nullptr, sOut = nullptr;
    std::vector<cudaEvent_t> evIn, evK;
    cudaEvent_t evStart = nullptr;
    int device = 0;
    int smCount = 0;
    cudaStream_t stream = nullptr;
    bool ownStream = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string err;
    // grow-only device staging for host-pointer calls
    void *dIn = nullptr;
    size_t dInCap = 0;
    void *dOut = nullptr;
    size_t dOutCap = 0;
    unsigned long long *dCounters = nullptr;  // [0..2] trace counters, [7] ray-fetch counter
    // scheduling knobs of the persistent kernel (B200PT_REFILL_MIN / B200PT_NODE_MIN override)
    int refillMin = 4, nodeMin = 16;
    bool simpleKernel = false;  // B200PT_TRAVERSAL=simple: the one-thread-per-ray parity kernel
    bool wideKernel = false;    // B200PT_TRAVERSAL=wide: the 8-wide walk (irregular rays still take the binary kernel)
    // [0] = count, [1..] = indices of the rays the wide kernel handed to the exact binary kernel (grow-only)
    int *dExact = nullptr;
    size_t dExactCap = 0;
};

// device scratch for `n` rays of a wide-kernel launch, or null when the wide kernel is off; B200PT_ERR_* on failure
int EnsureExactScratch(b200pt_ctx *ctx, size_t n, int **out);

struct b200pt_scene {
    b200pt_ctx *ctx = nullptr;
    b200pt::HostScene host;
    b200pt::DeviceScene dev{};
    void *dNodes = nullptr, *dPrims = nullptr, *dInst = nullptr, *dPairs = nullptr, *dWide = nullptr, *dLeafBoxes = nullptr;
    b200pt_scene_stats stats{};
    b200pt_trace_stats trace{};
};

extern std::string g_createErr;

#define B2_CUDA(ctx, call)                                                                  \
    do {                                                                                    \
        cudaError_t e_ = (call);                                                            \
        if (e_ != cudaSuccess) {                                                            \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                \
            return B200PT_ERR_CUDA;                                                         \
        }                                                                                   \
    } while (0)

