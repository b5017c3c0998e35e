// traverse_bvh2.cu - closest-hit / any-hit traversal over the binary tree in the
// reference's own node order.  One thread per ray, near-child-first by
// dirIsNeg[axis], explicit stack - the control flow of BVHAggregate::Intersect /
// IntersectP (cpu/aggregates.cpp:529-624), which makes hit ids and t bit-exact
// by construction (same visit order => same tie-breaking).
// This is the parity anchor; the wide-BVH kernel (traverse_wide.cu) is the fast path.
#include <cuda_runtime.h>

#include "bvh2.h"
#include "device_scene.h"
#include "isect.h"
#include "kernels.h"
#include "../../include/b200pt.h"

namespace b200pt {

struct NodeRegs {
    float bmin[3], bmax[3];
    int offset;
    int nPrims;
    int axis;
};

__device__ __forceinline__ NodeRegs LoadNode32(const float4 *nodes, int idx) {
    float4 a = __ldg(nodes + 2 * idx), b = __ldg(nodes + 2 * idx + 1);
    NodeRegs n;
    n.bmin[0] = a.x; n.bmin[1] = a.y; n.bmin[2] = a.z;
    n.bmax[0] = a.w; n.bmax[1] = b.x; n.bmax[2] = b.y;
    n.offset = __float_as_int(b.z);
    unsigned w = __float_as_uint(b.w);
    n.nPrims = w & 0xFFFFu;
    n.axis = (w >> 16) & 0xFFu;
    return n;
}

template <bool ANY, bool COUNT>
__global__ void __launch_bounds__(128)
TraverseBVH2Kernel(DeviceScene sc, const float *__restrict__ rays, long long n,
                   b200pt_hit *__restrict__ hits, unsigned char *__restrict__ occluded,
                   unsigned long long *__restrict__ counters) {
    const float4 *nodes = reinterpret_cast<const float4 *>(sc.nodes);
    const float4 *prims = reinterpret_cast<const float4 *>(sc.prims);
    unsigned long long cNodes = 0, cLeaves = 0, cPrims = 0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const float *q = rays + 7 * i;
        RayCtx r;
        RaySetup(r, q[0], q[1], q[2], q[3], q[4], q[5]);
        float tMax = q[6];
        int stack[64];
        int sp = 0, cur = (int)sc.rootNode;
        int bestSlot = -1;
        float bestB1 = 0, bestB2 = 0;
        bool any = false;
        while (true) {
            if (COUNT) ++cNodes;
            NodeRegs nd = LoadNode32(nodes, cur);
            if (SlabTest(r, nd.bmin, nd.bmax, tMax)) {
                if (nd.nPrims > 0) {
                    if (COUNT) ++cLeaves;
                    for (int k = 0; k < nd.nPrims; ++k) {
                        int slot = nd.offset + k;
                        float4 a = __ldg(prims + 4 * slot), b = __ldg(prims + 4 * slot + 1),
                               c = __ldg(prims + 4 * slot + 2);
                        float p0[3] = {a.x, a.y, a.z}, p1[3] = {a.w, b.x, b.y}, p2[3] = {b.z, b.w, c.x};
                        float t, b0, b1, b2;
                        if (COUNT) ++cPrims;
                        bool primHit;
                        if (sc.hasPatches) {
                            float4 d4 = __ldg(prims + 4 * slot + 3);
                            float p3[3] = {d4.x, d4.y, d4.z};
                            const uint32_t kind = __float_as_uint(d4.w);
                            primHit = (kind & 1u) ? BilinearHit(r, tMax, p0, p1, p2, p3, t, b1, b2)
                                                  : TriHit(r, tMax, p0, p1, p2, t, b0, b1, b2);
                            if (primHit && (kind & 2u)) primHit = AlphaKeepsHit(r, d4.x);
                        } else
                            primHit = TriHit(r, tMax, p0, p1, p2, t, b0, b1, b2);
                        if (primHit) {
                            if (ANY) {
                                any = true;
                                break;
                            }
                            tMax = t;
                            bestSlot = slot;
                            bestB1 = b1;
                            bestB2 = b2;
                        }
                    }
                    if (ANY && any) break;
                    if (sp == 0) break;
                    cur = stack[--sp];
                } else {
                    if (r.neg[nd.axis]) {
                        stack[sp++] = cur + 1;
                        cur = nd.offset;
                    } else {
                        stack[sp++] = nd.offset;
                        cur = cur + 1;
                    }
                }
            } else {
                if (sp == 0) break;
                cur = stack[--sp];
            }
        }
        if (ANY) {
            occluded[i] = any ? 1 : 0;
        } else {
            b200pt_hit h;
            if (bestSlot < 0) {
                h.mesh = kMissTag; h.prim = 0; h.instance = 0; h.t = 0; h.b1 = 0; h.b2 = 0;
            } else {
                float4 c = __ldg(prims + 4 * bestSlot + 2);
                h.mesh = __float_as_uint(c.y);
                h.prim = __float_as_uint(c.z);
                h.instance = 0;
                h.t = tMax; h.b1 = bestB1; h.b2 = bestB2;
            }
            hits[i] = h;
        }
    }
    if (COUNT) {
        atomicAdd(counters + 0, cNodes);
        atomicAdd(counters + 1, cLeaves);
        atomicAdd(counters + 2, cPrims);
    }
}

cudaError_t LaunchTraverseBVH2(const DeviceScene &sc, const float *rays, long long n, b200pt_hit *hits,
                               unsigned char *occluded, bool any, unsigned long long *counters,
                               int smCount, cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    const int block = 128;
    long long want = (n + block - 1) / block;
    long long cap = (long long)smCount * 16;  // 16 resident CTAs of 128 threads per SM
    int grid = (int)(want < cap ? want : cap);
    if (any) {
        if (counters)
            TraverseBVH2Kernel<true, true><<<grid, block, 0, stream>>>(sc, rays, n, hits, occluded, counters);
        else
            TraverseBVH2Kernel<true, false><<<grid, block, 0, stream>>>(sc, rays, n, hits, occluded, counters);
    } else {
        if (counters)
            TraverseBVH2Kernel<false, true><<<grid, block, 0, stream>>>(sc, rays, n, hits, occluded, counters);
        else
            TraverseBVH2Kernel<false, false><<<grid, block, 0, stream>>>(sc, rays, n, hits, occluded, counters);
    }
    return cudaGetLastError();
}

}  // namespace b200pt
