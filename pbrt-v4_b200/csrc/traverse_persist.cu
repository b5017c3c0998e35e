// traverse_persist.cu - the production closest-hit / any-hit kernels: the 8-wide walk (WIDE, wide.h) and the
// exact binary walk it hands irregular rays to.
//
// Same visit order as BVHAggregate::Intersect / IntersectP (reference
// cpu/aggregates.cpp:529-624) - near child first by dirIsNeg[axis], far child
// stacked - so ids / t stay bit-exact, but organised for the GPU:
//   * persistent warps that pull rays from a global counter and REFILL idle
//     lanes while the others keep traversing (the simple kernel ran at 5.8
//     active threads per warp on incoherent rays: a warp waited for its longest ray);
//   * "while-while": lanes iterate interior nodes until few are left, then the
//     lanes parked at a leaf do their triangle tests together;
//   * one 256-bit ld.global.nc per node (one L1 wavefront per lane instead of two)
//     and a [depth][thread] shared-memory stack (conflict-free: 1 wavefront per
//     warp push/pop; a local-memory stack is up to 32).
// Arithmetic comes from isect.h unchanged.
#include <cuda_runtime.h>

#include <cstdlib>

#include "bvh2.h"
#include "device_scene.h"
#include "instance.h"
#include "isect.h"
#include "kernels.h"
#include "wide.h"
#include "../../include/b200pt.h"

namespace b200pt {

namespace {

constexpr int kBlock = 128;
#ifndef B2_WIDE_BLOCKS
#define B2_WIDE_BLOCKS 5  // resident blocks per SM the wide kernels are compiled for (6 = 80 registers spills: measured slower)
#endif

struct Node8f {
    float v[8];
};

__device__ __forceinline__ Node8f Load256(const void *p) {
    Node8f n;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(n.v[0]), "=f"(n.v[1]), "=f"(n.v[2]), "=f"(n.v[3]), "=f"(n.v[4]), "=f"(n.v[5]),
                   "=f"(n.v[6]), "=f"(n.v[7])
                 : "l"(p));
    return n;
}

enum : int { S_IDLE = 0, S_NODE = 1, S_LEAF = 2, S_DONE = 3 };

}  // namespace

// Traversal stack: the first kSmemDepth entries in shared memory laid out [depth][thread]
// (conflict-free), the rare deeper ones in local memory.  An entry is {child ref, entry distance
// tMin of its box}: the far child's box is tested when its parent is visited, and the only part of
// the reference's test that depends on the ray's current tMax (`tMin < raytMax`,
// util/vecmath.h:1609) is re-evaluated when the entry is popped - i.e. at the moment the
// reference would have tested that node.
constexpr int kSmemDepth = 24;
constexpr int kLocalDepth = 104;

// INTERNAL: closest-hit records for the shading kernels {slot, t, b0, b1, b2, instance + 1} instead of
// the public b200pt_hit (the PrimSlot index saves the shader the (mesh, tri) -> vertices lookup).
// INST: the scene has object instances (TransformedPrimitive, cpu/primitive.cpp:112-131).  A leaf slot
// tagged kInstanceTag switches the lane to the instance's ray (instance.h) and the object's tree; the
// way back is two stack entries under the object's own entries: {slots left in the parent leaf, the
// parent's tMax} and {kReturn | next slot}.  When the object's entries are used up the return entry
// pops: the parent ray is set up again from the ray buffer and the leaf loop resumes.  The parent
// goes on with tMax = the t found in the object, not shifted back by dt, as the reference does.
constexpr int kReturn = int(0x80000000u);
// PATCH: some leaf slots are bilinear patches (kind word of the slot's last 16 bytes); hits report (u, v).
// The exact binary walk: every box test is the reference's own.  Production kernel of B200PT_TRAVERSAL=binary and the
// walk of the rays the wide kernel hands over (`indexList` = that list, `nDev` = its length).
template <bool ANY, bool COUNT, bool INTERNAL, bool INST, bool PATCH>
__global__ void __launch_bounds__(kBlock)
TraverseBinaryKernel(DeviceScene sc, const float *__restrict__ rays, int n, b200pt_hit *__restrict__ hits,
                      unsigned char *__restrict__ occluded, int *__restrict__ rayCounter,
                      unsigned long long *__restrict__ counters, int kRefillMin, int kNodeMin,
                      const int *__restrict__ nDev, const int *__restrict__ indexList) {
    if (nDev) n = min(n, *nDev);  // queue length written by an earlier kernel of the stream
    __shared__ int2 smemStack[kSmemDepth * kBlock];
    int2 *stack = smemStack + threadIdx.x;  // entry d lives at stack[d * kBlock]
    int2 deep[kLocalDepth];
    const char *pairs = reinterpret_cast<const char *>(sc.pairs);
    const char *prims = reinterpret_cast<const char *>(sc.prims);
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;

    int state = S_IDLE, rayIdx = -1;
    bool exhausted = false;
    RayCtx r;
    float tMax = 0, bestB0 = 0, bestB1 = 0, bestB2 = 0;
    int cur = 0, sp = 0, bestSlot = -1;
    int lslot = 0, lrem = 0;                 // leaf cursor: next slot, slots left (0 = read the leaf header)
    int curInst = -1, bestInst = -1;         // instance being traversed / instance of the best hit
    bool instHit = false;                    // a hit was accepted inside the current instance
    unsigned long long cNodes = 0, cLeaves = 0, cPrims = 0;
    auto push = [&](int ref, int tbits) {
        int2 e = make_int2(ref, tbits);
        if (sp < kSmemDepth) stack[sp * kBlock] = e;
        else deep[sp - kSmemDepth] = e;
        ++sp;
    };
    auto pop = [&]() {
        --sp;
        return sp < kSmemDepth ? stack[sp * kBlock] : deep[sp - kSmemDepth];
    };
    auto enter = [&](int ref) {
        cur = ref;
        if (ref < 0) {
            state = S_LEAF;
            lslot = ~ref;
            lrem = 0;
        } else {
            state = S_NODE;
        }
    };
    // next entry whose box the reference would still enter with the current tMax; S_DONE if none
    auto popNext = [&]() {
        state = S_DONE;
        while (sp > 0) {
            int2 e = pop();
            if (INST && e.x < -(1 << 30)) {  // the way back out of an instance
                int2 saved = pop();
                // the parent ray: o, d from the ray buffer, everything derived from them (three 1/d, the shear
                // terms: six divisions) from the four entries stacked under the return entry
                const float *q = rays + 7ll * rayIdx;
                const int2 c3 = pop(), c2 = pop(), c1 = pop(), c0 = pop();
                RayRestore(r, q[0], q[1], q[2], q[3], q[4], q[5], c0, c1, c2, c3);
                if (!instHit) tMax = __int_as_float(saved.y);
                curInst = -1;
                if (saved.x > 0) {
                    state = S_LEAF;
                    lslot = e.x & 0x3fffffff;
                    lrem = saved.x;
                    break;
                }
                continue;
            }
            if (__int_as_float(e.y) < tMax) {
                enter(e.x);
                break;
            }
        }
    };

    while (true) {
        // ---- refill idle lanes from the global ray counter (one atomic per warp)
        unsigned idle = __ballot_sync(FULL, state == S_IDLE);
        if (idle) {
            int nIdle = __popc(idle);
            if (!exhausted && (nIdle >= kRefillMin || idle == FULL)) {
                int base = 0;
                if (lane == 0) base = atomicAdd(rayCounter, nIdle);
                base = __shfl_sync(FULL, base, 0);
                if (base + nIdle >= n) exhausted = true;
                if (state == S_IDLE) {
                    int my = base + __popc(idle & ((1u << lane) - 1));
                    if (my < n) {
                        if (indexList) my = indexList[my];
                        const float *q = rays + 7ll * my;
                        RaySetup(r, q[0], q[1], q[2], q[3], q[4], q[5]);
                        tMax = q[6];
                        rayIdx = my;
                        sp = 0;
                        bestSlot = -1;
                        curInst = -1;
                        bestInst = -1;
                        // the reference tests the root's own box first (cpu/aggregates.cpp:545)
                        if (COUNT) ++cNodes;
                        if (SlabTest(r, sc.rootBox, sc.rootBox + 3, tMax)) enter(sc.rootRef);
                        else state = S_DONE;
                    }
                }
                idle = __ballot_sync(FULL, state == S_IDLE);
            }
            if (idle == FULL && exhausted) break;
        }

        // ---- interior nodes: a warp-uniform loop (every lane evaluates the ballot, lanes that are not
        // at an interior node skip the body), left when too few lanes still have interior work
        unsigned inNode;
        do {
            if (state == S_NODE) {
                {
                if (COUNT) cNodes += 2;
                Node8f a = Load256(pairs + 64ll * cur), b = Load256(pairs + 64ll * cur + 32);
                float t0, t1;
                bool h0 = SlabCore(r, a.v, a.v + 3, t0);
                bool h1 = SlabCore(r, b.v, b.v + 3, t1);
                int ref0 = __float_as_int(a.v[6]), ref1 = __float_as_int(b.v[6]);
                int axis = __float_as_int(a.v[7]);
                bool neg = (axis == 0 ? r.neg[0] : (axis == 1 ? r.neg[1] : r.neg[2])) != 0;
                // near child first: dirIsNeg[axis] ? second : first (cpu/aggregates.cpp:562-568)
                int nearRef = neg ? ref1 : ref0, farRef = neg ? ref0 : ref1;
                float tNear = neg ? t1 : t0, tFar = neg ? t0 : t1;
                // near child: visited now, so the whole test applies now.  far child: stacked on the
                // tMax-independent part only; `tFar < tMax` is decided when it is popped (popNext).
                bool hNear = (neg ? h1 : h0) & (tNear < tMax), hFar = neg ? h0 : h1;
                if (hNear) {
                    if (hFar) push(farRef, __float_as_int(tFar));
                    enter(nearRef);
                } else if (hFar & (tFar < tMax)) {
                    enter(farRef);  // the reference pops it next, with this very tMax: no stack round trip
                } else {
                    popNext();
                }
                }
            }
            inNode = __ballot_sync(FULL, state == S_NODE);
        } while (__popc(inNode) >= kNodeMin);

        // ---- leaves: triangle tests, in slot order, later equal-t hit replaces earlier
        if (state == S_LEAF) {
            if (COUNT && lrem == 0) ++cLeaves;
            bool anyHit = false, entered = false;
            int slot = lslot, left = lrem;
            while (true) {
                Node8f a = Load256(prims + 64ll * slot);
                float4 b = __ldg(reinterpret_cast<const float4 *>(prims + 64ll * slot + 32));
                if (left == 0) left = __float_as_int(b.w);  // leaf header: count in the first slot
                if (INST && __float_as_uint(b.y) == kInstanceTag) {
                    const InstanceDev *in = sc.instances + __float_as_int(b.z);
                    {
                        int2 c0, c1, c2, c3;
                        RaySave(r, c0, c1, c2, c3);
                        push(c0.x, c0.y);
                        push(c1.x, c1.y);
                        push(c2.x, c2.y);
                        push(c3.x, c3.y);
                    }
                    push(left - 1, __float_as_int(tMax));
                    push(kReturn | (slot + 1), 0);
                    const float4 *mi = reinterpret_cast<const float4 *>(in->mInv);
                    float4 r0 = __ldg(mi), r1 = __ldg(mi + 1), r2 = __ldg(mi + 2), r3 = __ldg(mi + 3);
                    float mInv[16] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w,
                                      r2.x, r2.y, r2.z, r2.w, r3.x, r3.y, r3.z, r3.w};
                    float o2[3], d2[3], tMax2;
                    InstanceRay(mInv, r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, tMax, o2, d2, tMax2);
                    RaySetup(r, o2[0], o2[1], o2[2], d2[0], d2[1], d2[2]);
                    tMax = tMax2;
                    curInst = __float_as_int(b.z);
                    instHit = false;
                    int rootRef = __ldg(&in->rootRef);
                    bool go = true;
                    if (rootRef >= 0) {  // a one-primitive object has no BVH and no box test
                        float box[6];
                        for (int k = 0; k < 6; ++k) box[k] = __ldg(&in->rootBox[k]);
                        if (COUNT) ++cNodes;
                        go = SlabTest(r, box, box + 3, tMax);
                    }
                    if (go) enter(rootRef);
                    else popNext();
                    entered = true;
                    break;
                }
                float p0[3] = {a.v[0], a.v[1], a.v[2]}, p1[3] = {a.v[3], a.v[4], a.v[5]},
                      p2[3] = {a.v[6], a.v[7], b.x};
                float t, b0, b1, b2;
                if (COUNT) ++cPrims;
                bool primHit;
                if (PATCH) {
                    float4 c4 = __ldg(reinterpret_cast<const float4 *>(prims + 64ll * slot + 48));
                    const uint32_t kind = __float_as_uint(c4.w);
                    if (kind & 1u) {  // p0..p3 = p00, p10, p01, p11
                        float p3[3] = {c4.x, c4.y, c4.z};
                        b0 = 0.f;
                        primHit = BilinearHit(r, tMax, p0, p1, p2, p3, t, b1, b2);
                    } else {
                        primHit = TriHit(r, tMax, p0, p1, p2, t, b0, b1, b2);
                        if (primHit && (kind & 2u)) primHit = AlphaKeepsHit(r, c4.x);  // alpha cut-out triangle
                    }
                } else
                    primHit = TriHit(r, tMax, p0, p1, p2, t, b0, b1, b2);
                if (primHit) {
                    if (ANY) {
                        anyHit = true;
                        break;
                    }
                    tMax = t;
                    bestSlot = slot;
                    if (INST) {
                        bestInst = curInst;
                        instHit = true;
                    }
                    if (INTERNAL) bestB0 = b0;
                    bestB1 = b1;
                    bestB2 = b2;
                }
                if (--left == 0) break;
                ++slot;
            }
            if (!entered) {
                if (ANY && anyHit) {
                    bestSlot = 0;
                    state = S_DONE;
                } else {
                    popNext();
                }
            }
        }
        __syncwarp();

        // ---- finished rays write their record and go idle
        if (state == S_DONE) {
            if (ANY) {
                occluded[rayIdx] = bestSlot >= 0 ? 1 : 0;
            } else {
                uint2 w0, w1, w2;
                if (bestSlot < 0) {
                    w0 = make_uint2(kMissTag, 0u);
                    w1 = make_uint2(0u, 0u);
                    w2 = make_uint2(0u, 0u);
                } else if (INTERNAL) {
                    w0 = make_uint2((unsigned)bestSlot, __float_as_uint(tMax));
                    w1 = make_uint2(__float_as_uint(bestB0), __float_as_uint(bestB1));
                    w2 = make_uint2(__float_as_uint(bestB2), INST ? unsigned(bestInst + 1) : 0u);
                } else {
                    float4 b = __ldg(reinterpret_cast<const float4 *>(prims + 64ll * bestSlot + 32));
                    // instanceId is the 1-based position in the instance list (scene.cpp:1560)
                    w0 = make_uint2(__float_as_uint(b.y), __float_as_uint(b.z));
                    w1 = make_uint2(INST ? unsigned(bestInst + 1) : 0u, __float_as_uint(tMax));
                    w2 = make_uint2(__float_as_uint(bestB1), __float_as_uint(bestB2));
                }
                uint2 *dst = reinterpret_cast<uint2 *>(hits + rayIdx);
                dst[0] = w0;
                dst[1] = w1;
                dst[2] = w2;
            }
            state = S_IDLE;
            bestSlot = -1;
        }
    }
    if (COUNT) {
        atomicAdd(counters + 0, cNodes);
        atomicAdd(counters + 1, cLeaves);
        atomicAdd(counters + 2, cPrims);
    }
}

// The 8-wide walk (wide.h): interior nodes are Wide8 - eight child boxes per visit under a conservative test, children
// stacked in the reference's visit order by pre-set popcount - and the reference's exact box test is made on each
// candidate leaf's own box when the leaf phase opens it.  A ray that is not regular (isect.h WideRaySetup) - at fetch, or
// as transformed into an instance - is appended to `exactList` and dropped; the launcher runs the binary kernel over
// that list afterwards.  Same loop shape as the binary kernel.  (A phase-per-round variant with a single pop site was
// measured slower, 992 against 1185 Mrays/s closest-hit on config 2: profiles/r02_traversal_variants.txt.)
template <bool ANY, bool COUNT, bool INTERNAL, bool INST, bool PATCH>
__global__ void __launch_bounds__(kBlock, B2_WIDE_BLOCKS)
TraverseWideKernel(DeviceScene sc, const float *__restrict__ rays, int n, b200pt_hit *__restrict__ hits,
                      unsigned char *__restrict__ occluded, int *__restrict__ rayCounter,
                      unsigned long long *__restrict__ counters, int kRefillMin, int kNodeMin,
                      const int *__restrict__ nDev, const int *__restrict__ indexList, int *__restrict__ exactList,
                      int *__restrict__ exactCount) {
    if (nDev) n = min(n, *nDev);  // queue length written by an earlier kernel of the stream
    __shared__ int2 smemStack[kSmemDepth * kBlock];
    int2 *stack = smemStack + threadIdx.x;  // entry d lives at stack[d * kBlock]
    int2 deep[kLocalDepth];
    const char *pairs = reinterpret_cast<const char *>(sc.wide);
    const char *prims = reinterpret_cast<const char *>(sc.prims);
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;

    int state = S_IDLE, rayIdx = -1;
    bool exhausted = false;
    RayCtx r;
    float tMax = 0, bestB0 = 0, bestB1 = 0, bestB2 = 0;
    int cur = 0, sp = 0, bestSlot = -1;
    int lslot = 0, lrem = 0;                 // leaf cursor: next slot, slots left (0 = read the leaf header)
    int curInst = -1, bestInst = -1;         // instance being traversed / instance of the best hit
    bool instHit = false;                    // a hit was accepted inside the current instance
    unsigned long long cNodes = 0, cLeaves = 0, cPrims = 0, cWide = 0;
    unsigned long long cLeafTests = 0;
    bool flip = false;  // dirIsNeg[2] - the stored visit orders are those of dirIsNeg[2] == 0, reversed otherwise
    int cls = 0;        // dirIsNeg[0] | dirIsNeg[1] << 1 of the octant whose order is stored
    WideRay wr;         // per-ray constants of the conservative child test (isect.h)
    wr.cNx = wr.cNy = wr.cNz = wr.cFx = wr.cFy = wr.cFz = 0.f;
    int offNear[3] = {0, 32, 64};  // byte offset of the near plane of each axis inside a node
    // octant-dependent constants of the wide walk for the ray in `r` inside the tree bounded by `treeBox`; false = the
    // ray is not regular there (isect.h WideRaySetup) and has to take the exact binary walk
    auto setWide = [&](const float *treeBox) {
        flip = r.neg[2] != 0;
        cls = (r.neg[0] ^ int(flip)) | ((r.neg[1] ^ int(flip)) << 1);
        offNear[0] = r.neg[0] ? 96 : 0;
        offNear[1] = r.neg[1] ? 128 : 32;
        offNear[2] = r.neg[2] ? 160 : 64;
        return WideRaySetup(r, treeBox, wr);
    };
    // hands the ray over to the exact binary kernel
    auto giveUp = [&]() {
        exactList[atomicAdd(exactCount, 1)] = rayIdx;
        state = S_IDLE;
    };

    auto push = [&](int ref, int tbits) {
        int2 e = make_int2(ref, tbits);
        if (sp < kSmemDepth) stack[sp * kBlock] = e;
        else deep[sp - kSmemDepth] = e;
        ++sp;
    };
    auto pop = [&]() {
        --sp;
        return sp < kSmemDepth ? stack[sp * kBlock] : deep[sp - kSmemDepth];
    };
    auto enter = [&](int ref) {
        cur = ref;
        if (ref < 0) {  // a LeafBox reference: the leaf phase opens it
            state = S_LEAF;
            lrem = 0;
        } else {
            state = S_NODE;
        }
    };
    // next entry whose box the reference would still enter with the current tMax; S_DONE if none
    auto popNext = [&]() {
        state = S_DONE;
        while (sp > 0) {
            int2 e = pop();
            if (INST && e.x < -(1 << 30)) {  // the way back out of an instance
                int2 saved = pop();
                const float *q = rays + 7ll * rayIdx;
                RaySetup(r, q[0], q[1], q[2], q[3], q[4], q[5]);
                (void)setWide(sc.rootBox);  // regular: it was when the ray was fetched
                if (!instHit) tMax = __int_as_float(saved.y);
                curInst = -1;
                if (saved.x > 0) {
                    state = S_LEAF;
                    lslot = e.x & 0x3fffffff;
                    lrem = saved.x;
                    break;
                }
                continue;
            }
            if (__int_as_float(e.y) < tMax) {
                enter(e.x);
                break;
            }
        }
    };

    while (true) {
        // ---- refill idle lanes from the global ray counter (one atomic per warp)
        unsigned idle = __ballot_sync(FULL, state == S_IDLE);
        if (idle) {
            int nIdle = __popc(idle);
            if (!exhausted && (nIdle >= kRefillMin || idle == FULL)) {
                int base = 0;
                if (lane == 0) base = atomicAdd(rayCounter, nIdle);
                base = __shfl_sync(FULL, base, 0);
                if (base + nIdle >= n) exhausted = true;
                if (state == S_IDLE) {
                    int my = base + __popc(idle & ((1u << lane) - 1));
                    if (my < n) {
                        if (indexList) my = indexList[my];
                        const float *q = rays + 7ll * my;
                        RaySetup(r, q[0], q[1], q[2], q[3], q[4], q[5]);
                        tMax = q[6];
                        rayIdx = my;
                        sp = 0;
                        bestSlot = -1;
                        curInst = -1;
                        bestInst = -1;
                        // the reference tests the root's own box first (cpu/aggregates.cpp:545)
                        if (COUNT) ++cNodes;
                        if (!setWide(sc.rootBox)) giveUp();
                        else if (SlabTest(r, sc.rootBox, sc.rootBox + 3, tMax)) enter(sc.rootWide);
                        else state = S_DONE;
                    }
                }
                idle = __ballot_sync(FULL, state == S_IDLE);
            }
            if (idle == FULL && exhausted) break;
        }

        // ---- interior nodes: a warp-uniform loop (every lane evaluates the ballot, lanes that are not
        // at an interior node skip the body), left when too few lanes still have interior work
        unsigned inNode;
        do {
            if (state == S_NODE) {
                {
                    const char *np = pairs + 256ll * cur;
                    if (COUNT) {
                        cNodes += 8;
                        ++cWide;
                    }
                    // children are stacked when the tMax-independent part of the reference's test passes and their
                    // entry distance is below tMax with a margin (tMax can still creep up by rounding, wide.h);
                    // `tMin < tMax` itself is decided when the entry is popped
                    const float tMaxS = __fmaf_rn(fabsf(tMax), 0x1p-10f, tMax);
                    unsigned H = 0;
                    float tm[8];
                    // one 32-byte load per plane of the eight children; the near / far plane of each axis by address
                    const Node8f nx = Load256(np + offNear[0]), ny = Load256(np + offNear[1]), nz = Load256(np + offNear[2]);
                    const Node8f fx = Load256(np + (96 - offNear[0])), fy = Load256(np + (160 - offNear[1])),
                                 fz = Load256(np + (224 - offNear[2]));
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const bool ok = WideChildTest(r, wr, nx.v[i], ny.v[i], nz.v[i], fx.v[i], fy.v[i], fz.v[i], tMaxS, tm[i]);
                        H |= unsigned(ok) << i;
                    }
                    if (H) {
                        Node8f rf = Load256(np + 192);
                        const uint2 pre = __ldg(reinterpret_cast<const uint2 *>(np + 224) + cls);
                        const int nHit = __popc(H);
                        // rank in the reference's visit order = hits among the slots visited before this one; the
                        // first to be visited ends on top of the stack (reversed order when dirIsNeg[2])
                        if (sp + nHit <= kSmemDepth) {
                            int2 *top = stack + (flip ? sp : sp + nHit - 1) * kBlock;
                            const int step = flip ? kBlock : -kBlock;
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const unsigned before = ((i < 4 ? pre.x : pre.y) >> (8 * (i & 3))) & H;
                                int2 *dst = top + __popc(before & 0xffu) * step;
                                if ((H >> i) & 1u) *dst = make_int2(__float_as_int(rf.v[i]), __float_as_int(tm[i]));
                            }
                        } else {
                            const int base = flip ? sp : sp + nHit - 1;
#pragma unroll 1
                            for (int i = 0; i < 8; ++i) {
                                if ((H >> i) & 1u) {
                                    const unsigned before = ((i < 4 ? pre.x : pre.y) >> (8 * (i & 3))) & 0xffu;
                                    const int rk = __popc(H & before);
                                    const int idx = flip ? base + rk : base - rk;
                                    float refBits = i == 0 ? rf.v[0] : i == 1 ? rf.v[1] : i == 2 ? rf.v[2] : i == 3 ? rf.v[3]
                                                  : i == 4 ? rf.v[4] : i == 5 ? rf.v[5] : i == 6 ? rf.v[6] : rf.v[7];
                                    float tBits = i == 0 ? tm[0] : i == 1 ? tm[1] : i == 2 ? tm[2] : i == 3 ? tm[3]
                                                : i == 4 ? tm[4] : i == 5 ? tm[5] : i == 6 ? tm[6] : tm[7];
                                    const int2 e = make_int2(__float_as_int(refBits), __float_as_int(tBits));
                                    if (idx < kSmemDepth) stack[idx * kBlock] = e;
                                    else deep[idx - kSmemDepth] = e;
                                }
                            }
                        }
                        sp += nHit;
                    }
                    popNext();
                }
            }
            inNode = __ballot_sync(FULL, state == S_NODE);
        } while (__popc(inNode) >= kNodeMin);

        // ---- leaves: triangle tests, in slot order, later equal-t hit replaces earlier
        if (state == S_LEAF) {
            bool anyHit = false, entered = false, open = true;
            int slot = lslot, left = lrem;
            if (left == 0) {
                // a fresh leaf: its own box gets the reference's exact test with the tMax of this moment - the one
                // box test of the walk that decides anything (wide.h)
                const Node8f lb = Load256(reinterpret_cast<const char *>(sc.leafBoxes) + 32ll * (~cur));
                const unsigned c = __float_as_uint(lb.v[7]);
                if (!(c & kLeafNoTest)) {
                    if (COUNT) ++cLeafTests;
                    open = SlabTest(r, lb.v, lb.v + 3, tMax);
                }
                slot = __float_as_int(lb.v[6]);
                left = int(c & ~kLeafNoTest);
                if (COUNT && open) ++cLeaves;
            }
            while (open) {
                Node8f a = Load256(prims + 64ll * slot);
                float4 b = __ldg(reinterpret_cast<const float4 *>(prims + 64ll * slot + 32));
                if (INST && __float_as_uint(b.y) == kInstanceTag) {
                    const InstanceDev *in = sc.instances + __float_as_int(b.z);
                    push(left - 1, __float_as_int(tMax));
                    push(kReturn | (slot + 1), 0);
                    const float4 *mi = reinterpret_cast<const float4 *>(in->mInv);
                    float4 r0 = __ldg(mi), r1 = __ldg(mi + 1), r2 = __ldg(mi + 2), r3 = __ldg(mi + 3);
                    float mInv[16] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w,
                                      r2.x, r2.y, r2.z, r2.w, r3.x, r3.y, r3.z, r3.w};
                    float o2[3], d2[3], tMax2;
                    InstanceRay(mInv, r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, tMax, o2, d2, tMax2);
                    RaySetup(r, o2[0], o2[1], o2[2], d2[0], d2[1], d2[2]);
                    tMax = tMax2;
                    curInst = __float_as_int(b.z);
                    instHit = false;
                    float box[6];
                    for (int k = 0; k < 6; ++k) box[k] = __ldg(&in->rootBox[k]);
                    if (!setWide(box)) {  // the transformed ray breaks the wide walk's premise: redo the whole ray exactly
                        giveUp();
                        entered = true;
                        break;
                    }
                    int rootRef = __ldg(&in->rootWide);
                    bool go = true;
                    if (rootRef >= 0) {  // a one-primitive object has no BVH and no box test
                        if (COUNT) ++cNodes;
                        go = SlabTest(r, box, box + 3, tMax);
                    }
                    if (go) enter(rootRef);
                    else popNext();
                    entered = true;
                    break;
                }
                float p0[3] = {a.v[0], a.v[1], a.v[2]}, p1[3] = {a.v[3], a.v[4], a.v[5]},
                      p2[3] = {a.v[6], a.v[7], b.x};
                float t, b0, b1, b2;
                if (COUNT) ++cPrims;
                bool primHit;
                if (PATCH) {
                    float4 c4 = __ldg(reinterpret_cast<const float4 *>(prims + 64ll * slot + 48));
                    const uint32_t kind = __float_as_uint(c4.w);
                    if (kind & 1u) {  // p0..p3 = p00, p10, p01, p11
                        float p3[3] = {c4.x, c4.y, c4.z};
                        b0 = 0.f;
                        primHit = BilinearHit(r, tMax, p0, p1, p2, p3, t, b1, b2);
                    } else {
                        primHit = TriHit(r, tMax, p0, p1, p2, t, b0, b1, b2);
                        if (primHit && (kind & 2u)) primHit = AlphaKeepsHit(r, c4.x);  // alpha cut-out triangle
                    }
                } else
                    primHit = TriHit(r, tMax, p0, p1, p2, t, b0, b1, b2);
                if (primHit) {
                    if (ANY) {
                        anyHit = true;
                        break;
                    }
                    tMax = t;
                    bestSlot = slot;
                    if (INST) {
                        bestInst = curInst;
                        instHit = true;
                    }
                    if (INTERNAL) bestB0 = b0;
                    bestB1 = b1;
                    bestB2 = b2;
                }
                if (--left == 0) break;
                ++slot;
            }
            if (!entered) {
                if (ANY && anyHit) {
                    bestSlot = 0;
                    state = S_DONE;
                } else {
                    popNext();
                }
            }
        }
        __syncwarp();

        // ---- finished rays write their record and go idle
        if (state == S_DONE) {
            if (ANY) {
                occluded[rayIdx] = bestSlot >= 0 ? 1 : 0;
            } else {
                uint2 w0, w1, w2;
                if (bestSlot < 0) {
                    w0 = make_uint2(kMissTag, 0u);
                    w1 = make_uint2(0u, 0u);
                    w2 = make_uint2(0u, 0u);
                } else if (INTERNAL) {
                    w0 = make_uint2((unsigned)bestSlot, __float_as_uint(tMax));
                    w1 = make_uint2(__float_as_uint(bestB0), __float_as_uint(bestB1));
                    w2 = make_uint2(__float_as_uint(bestB2), INST ? unsigned(bestInst + 1) : 0u);
                } else {
                    float4 b = __ldg(reinterpret_cast<const float4 *>(prims + 64ll * bestSlot + 32));
                    // instanceId is the 1-based position in the instance list (scene.cpp:1560)
                    w0 = make_uint2(__float_as_uint(b.y), __float_as_uint(b.z));
                    w1 = make_uint2(INST ? unsigned(bestInst + 1) : 0u, __float_as_uint(tMax));
                    w2 = make_uint2(__float_as_uint(bestB1), __float_as_uint(bestB2));
                }
                uint2 *dst = reinterpret_cast<uint2 *>(hits + rayIdx);
                dst[0] = w0;
                dst[1] = w1;
                dst[2] = w2;
            }
            state = S_IDLE;
            bestSlot = -1;
        }
    }
    if (COUNT) {
        atomicAdd(counters + 0, cNodes);
        atomicAdd(counters + 1, cLeaves);
        atomicAdd(counters + 2, cPrims);
        atomicAdd(counters + 3, cWide);
        atomicAdd(counters + 4, cLeafTests);
    }
}

// One family of kernels - (closest | any | closest with internal records) x (binary | wide) - per translation unit:
// the build compiles this file six times with -DB2_TP_PART=0..5 (kind = part % 3, wide = part / 3) so the 48 variants
// build in parallel.  Part 0 also holds the dispatcher.
struct TraverseLaunch {
    const DeviceScene *sc;
    const float *rays;
    long long n;
    b200pt_hit *hits;
    unsigned char *occluded;
    int *rayCounter;
    unsigned long long *counters;
    int smCount, refillMin, nodeMin;
    cudaStream_t stream;
    const int *nDev;       // queue length on the device, or null
    const int *indexList;  // ray indices to trace instead of 0..n-1, or null
    int *exactScratch;     // wide launches: [0] = length of the hand-over list that follows
};

template <bool ANYV, bool INTV, bool WIDEV>
static cudaError_t LaunchFamily(const TraverseLaunch &a) {
    const DeviceScene &sc = *a.sc;
    auto pick = [&](auto kern) -> cudaError_t {
        int perSM = 0;
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, kern, kBlock, 0);
        if (e != cudaSuccess) return e;
        if (perSM < 1) perSM = 1;
        long long want = (a.n + kBlock - 1) / kBlock;
        long long cap = (long long)a.smCount * perSM;  // persistent: exactly the resident set
        int grid = (int)(want < cap ? want : cap);
        e = cudaMemsetAsync(a.rayCounter, 0, sizeof(int), a.stream);
        if (e != cudaSuccess) return e;
        if (WIDEV) {
            e = cudaMemsetAsync(a.exactScratch, 0, sizeof(int), a.stream);
            if (e != cudaSuccess) return e;
        }
        if constexpr (WIDEV)
            kern<<<grid, kBlock, 0, a.stream>>>(sc, a.rays, (int)a.n, a.hits, a.occluded, a.rayCounter, a.counters, a.refillMin,
                                                a.nodeMin, a.nDev, a.indexList, a.exactScratch + 1, a.exactScratch);
        else
            kern<<<grid, kBlock, 0, a.stream>>>(sc, a.rays, (int)a.n, a.hits, a.occluded, a.rayCounter, a.counters, a.refillMin,
                                                a.nodeMin, a.nDev, a.indexList);
        return cudaGetLastError();
    };
#define B2_KERNEL(COUNTV, INSTV, PATCHV)                                                         \
    [&] {                                                                                        \
        if constexpr (WIDEV) return pick(TraverseWideKernel<ANYV, COUNTV, INTV, INSTV, PATCHV>); \
        else return pick(TraverseBinaryKernel<ANYV, COUNTV, INTV, INSTV, PATCHV>);               \
    }()
#define B2_PICK(INSTV, PATCHV) (a.counters ? B2_KERNEL(true, INSTV, PATCHV) : B2_KERNEL(false, INSTV, PATCHV))
    return sc.nInstances ? (sc.hasPatches ? B2_PICK(true, true) : B2_PICK(true, false))
                         : (sc.hasPatches ? B2_PICK(false, true) : B2_PICK(false, false));
#undef B2_KERNEL
#undef B2_PICK
}

cudaError_t LaunchTP0(const TraverseLaunch &a);
cudaError_t LaunchTP1(const TraverseLaunch &a);
cudaError_t LaunchTP2(const TraverseLaunch &a);
cudaError_t LaunchTP3(const TraverseLaunch &a);
cudaError_t LaunchTP4(const TraverseLaunch &a);
cudaError_t LaunchTP5(const TraverseLaunch &a);

#ifndef B2_TP_PART
#error "compile with -DB2_TP_PART=0..5 (pbrt-v4_b200/_build.py)"
#endif
#if B2_TP_PART == 0
cudaError_t LaunchTP0(const TraverseLaunch &a) { return LaunchFamily<false, false, false>(a); }
#elif B2_TP_PART == 1
cudaError_t LaunchTP1(const TraverseLaunch &a) { return LaunchFamily<true, false, false>(a); }
#elif B2_TP_PART == 2
cudaError_t LaunchTP2(const TraverseLaunch &a) { return LaunchFamily<false, true, false>(a); }
#elif B2_TP_PART == 3
cudaError_t LaunchTP3(const TraverseLaunch &a) { return LaunchFamily<false, false, true>(a); }
#elif B2_TP_PART == 4
cudaError_t LaunchTP4(const TraverseLaunch &a) { return LaunchFamily<true, false, true>(a); }
#elif B2_TP_PART == 5
cudaError_t LaunchTP5(const TraverseLaunch &a) { return LaunchFamily<false, true, true>(a); }
#endif

#if B2_TP_PART == 0
// exactScratch: 1 + n ints of device memory ([0] = length of the list that follows) for the rays the wide kernel
// hands to the exact binary kernel; null = binary kernel only.
cudaError_t LaunchTraversePersist(const DeviceScene &sc, const float *rays, long long n, b200pt_hit *hits,
                                  unsigned char *occluded, bool any, int *rayCounter,
                                  unsigned long long *counters, int smCount, int stackDepth,
                                  int refillMin, int nodeMin, cudaStream_t stream, bool internalHits,
                                  const int *nDev, int *exactScratch) {
    if (n <= 0) return cudaSuccess;
    if (n > 0x7fffffffll - (1ll << 22)) return cudaErrorInvalidValue;  // the fetch counter overshoots n by up to the resident lanes
    if (stackDepth + 4 > kSmemDepth + kLocalDepth) return cudaErrorInvalidValue;  // + the saved parent context of an instance
    const bool wide = exactScratch && sc.wide && int(sc.maxStackWide) + 2 <= kSmemDepth + kLocalDepth;
    TraverseLaunch a{&sc, rays, n, hits, occluded, rayCounter, counters, smCount, refillMin, nodeMin, stream, nDev, nullptr,
                     exactScratch};
    const int kind = any ? 1 : (internalHits ? 2 : 0);
    auto binary = kind == 0 ? LaunchTP0 : kind == 1 ? LaunchTP1 : LaunchTP2;
    if (!wide) return binary(a);
    cudaError_t e = (kind == 0 ? LaunchTP3 : kind == 1 ? LaunchTP4 : LaunchTP5)(a);
    if (e != cudaSuccess) return e;
    // the rays the wide walk declined (not regular, isect.h WideRaySetup), if any: exact binary walk over the list
    a.nDev = exactScratch;
    a.indexList = exactScratch + 1;
    return binary(a);
}
#endif

}  // namespace b200pt
