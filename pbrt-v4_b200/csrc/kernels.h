// kernels.h - launchers of the device kernels (one per .cu).
#pragma once
#include <cuda_runtime.h>

#include "device_scene.h"
struct b200pt_hit;

namespace b200pt {

cudaError_t LaunchTraverseBVH2(const DeviceScene &sc, const float *rays, long long n, b200pt_hit *hits,
                               unsigned char *occluded, bool any, unsigned long long *counters,
                               int smCount, cudaStream_t stream);

cudaError_t LaunchTraversePersist(const DeviceScene &sc, const float *rays, long long n, b200pt_hit *hits,
                                  unsigned char *occluded, bool any, int *rayCounter,
                                  unsigned long long *counters, int smCount, int stackDepth,
                                  int refillMin, int nodeMin, cudaStream_t stream, bool internalHits = false,
                                  const int *nDev = nullptr, int *exactScratch = nullptr);

}  // namespace b200pt
