// tests/native/trace_host.h - TEST-ONLY host statement of the production traversal (PairNode layout,
// {ref, tMin} stack, pop-time re-check, object instances by recursion) shared by hostcheck.cpp and
// rendercheck.cpp.  Same arithmetic headers as the CUDA kernels (isect.h, instance.h).
#pragma once
#include <cmath>
#include <cstdio>
#include <vector>

#include "../../pbrt-v4_b200/csrc/instance.h"
#include "../../pbrt-v4_b200/csrc/isect.h"
#include "../../pbrt-v4_b200/csrc/scene_host.h"

namespace b200pt {

struct Best {
    int slot = -1, inst = -1;  // inst: index into hs.instances, -1 = world geometry
    float b0 = 0, b1 = 0, b2 = 0;
};

// PairNode walk with the {ref, tMin} stack and the pop-time `tMin < tMax` re-check, as
// traverse_persist.cu does it; an instance leaf recurses with the transformed ray (the device kernel
// keeps the way back on its stack instead - same visits, same arithmetic).
inline bool TracePairs(const HostScene &hs, int rootRef, const float *rootBox, const float *o, const float *d,
                       float &tMax, bool any, Best &best, int curInst, bool dbg) {
    RayCtx r;
    RaySetup(r, o[0], o[1], o[2], d[0], d[1], d[2]);
    struct E { int ref; float t; };
    E stack[128];
    int sp = 0, cur = 0;
    bool done = false;
    auto popNext = [&]() {
        done = true;
        while (sp > 0) {
            E e = stack[--sp];
            if (e.t < tMax) { cur = e.ref; done = false; break; }
        }
    };
    // a one-primitive object has no BVH and no box (scene.cpp:1551): rootBox == nullptr
    if (!rootBox || SlabTest(r, rootBox, rootBox + 3, tMax)) cur = rootRef; else done = true;
    while (!done) {
        if (cur >= 0) {
            const PairNode &pn = hs.pairs[cur];
            float t0, t1;
            bool h0 = SlabCore(r, pn.box0, pn.box0 + 3, t0);
            bool h1 = SlabCore(r, pn.box1, pn.box1 + 3, t1);
            if (dbg)
                fprintf(stderr, "pair %d: h0 %d t0 %g ref0 %d | h1 %d t1 %g ref1 %d | tMax %.9g\n", cur, h0, t0, pn.ref0, h1, t1, pn.ref1, tMax);
            bool neg = r.neg[pn.axis] != 0;
            int nearRef = neg ? pn.ref1 : pn.ref0, farRef = neg ? pn.ref0 : pn.ref1;
            float tNear = neg ? t1 : t0, tFar = neg ? t0 : t1;
            bool hNear = (neg ? h1 : h0) && (tNear < tMax), hFar = neg ? h0 : h1;
            if (hFar) stack[sp++] = E{farRef, tFar};
            if (hNear) cur = nearRef;
            else popNext();
        } else {
            int slot = ~cur, cnt = int(hs.prims[slot].flags);
            for (int k = 0; k < cnt; ++k, ++slot) {
                const PrimSlot &s = hs.prims[slot];
                if (s.mesh == kInstanceTag) {
                    const InstanceDev &in = hs.instances[s.tri];
                    float o2[3], d2[3], tMax2;
                    InstanceRay(in.mInv, o[0], o[1], o[2], d[0], d[1], d[2], tMax, o2, d2, tMax2);
                    Best sub;
                    bool h = TracePairs(hs, in.rootRef, in.rootRef < 0 ? nullptr : in.rootBox, o2, d2, tMax2, any, sub,
                                        int(s.tri), dbg);
                    if (any && h) return true;
                    if (!any && sub.slot >= 0) {  // the parent goes on with tMax = the object's t, not shifted back
                        tMax = tMax2;
                        best = sub;
                    }
                    continue;
                }
                float t, b0, b1, b2;
                b0 = 0;
                    if (((s.kind & 1) ? BilinearHit(r, tMax, s.p0, s.p1, s.p2, s.p3, t, b1, b2)
                                      : TriHit(r, tMax, s.p0, s.p1, s.p2, t, b0, b1, b2)) &&
                        (!(s.kind & 2) || AlphaKeepsHit(r, s.p3[0]))) {
                    if (any) return true;
                    tMax = t; best.slot = slot; best.inst = curInst; best.b0 = b0; best.b1 = b1; best.b2 = b2;
                }
            }
            popNext();
        }
    }
    return any ? false : best.slot >= 0;
}

// Wide8 walk as traverse_persist.cu (WIDE) does it: every child box of the wide node gets the CONSERVATIVE test
// (isect.h WideChildTest, against tMax * (1 + 2^-10)), the passing children are stacked by their rank in the
// reference's visit order for the ray's octant (pre-set popcount; reversed when dirIsNeg[2]), stale entries are
// dropped at pop time, and a leaf is entered when its own box (LeafBox) passes the reference's exact test with the
// tMax of that moment.  Rays that are not regular (WideRaySetup) go through TracePairs (`irregular`).
struct WideCounters {
    unsigned long long visits = 0, childHits = 0, leaves = 0, prims = 0, children = 0, leafTests = 0;
};
inline WideCounters &WideStats() {
    static WideCounters c;
    return c;
}

// returns 0 = no hit, 1 = hit, -1 = the ray (or its image inside an instance) is irregular: redo with TracePairs
inline int TraceWide(const HostScene &hs, int rootRef, const float *rootBox, const float *treeBox, const float *o,
                     const float *d, float &tMax, bool any, Best &best, int curInst) {
    RayCtx r;
    RaySetup(r, o[0], o[1], o[2], d[0], d[1], d[2]);
    WideRay wr;
    if (!WideRaySetup(r, treeBox, wr)) return -1;
    struct E { int ref; float t; };
    std::vector<E> stack(hs.maxStackWide + 8);
    int sp = 0, cur = 0, lslot = 0, lcnt = 0;
    bool done = false;
    // cur >= 0: wide node; cur < 0: leaf entered (lslot, lcnt set)
    auto enterLeaf = [&](int ref) {
        const LeafBox &lb = hs.leafBoxes[~ref];
        if (!(lb.count & kLeafNoTest)) {
            WideStats().leafTests++;
            if (!SlabTest(r, lb.box, lb.box + 3, tMax)) return false;
        }
        cur = ref;
        lslot = lb.first;
        lcnt = int(lb.count & ~kLeafNoTest);
        return true;
    };
    auto popNext = [&]() {
        done = true;
        while (sp > 0) {
            E e = stack[--sp];
            if (e.t < tMax) {
                if (e.ref >= 0) { cur = e.ref; done = false; break; }
                if (enterLeaf(e.ref)) { done = false; break; }
            }
        }
    };
    const bool flip = r.neg[2] != 0;
    const int cls = (r.neg[0] ^ int(flip)) | ((r.neg[1] ^ int(flip)) << 1);
    if (!rootBox || SlabTest(r, rootBox, rootBox + 3, tMax)) {
        if (rootRef >= 0) cur = rootRef;
        else if (!enterLeaf(rootRef)) popNext();
    } else
        done = true;
    while (!done) {
        if (cur >= 0) {
            const Wide8 &w = hs.wide[cur];
            const float tMaxS = B2_FMA(fabsf(tMax), 0x1p-10f, tMax);
            unsigned H = 0;
            float tm[8];
            for (int i = 0; i < 8; ++i) {
                // near / far plane of each axis as the kernel picks them by address
                const int nxp = r.neg[0] ? 3 : 0, nyp = r.neg[1] ? 4 : 1, nzp = r.neg[2] ? 5 : 2;
                bool ok = WideChildTest(r, wr, w.plane[nxp][i], w.plane[nyp][i], w.plane[nzp][i], w.plane[3 - nxp][i],
                                        w.plane[5 - nyp][i], w.plane[7 - nzp][i], tMaxS, tm[i]);
                H |= unsigned(ok) << i;
            }
            const int nHit = __builtin_popcount(H);
            WideStats().visits++;
            WideStats().childHits += nHit;
            for (int i = 0; i < 8; ++i) WideStats().children += w.ref[i] != kWideEmpty;
            for (int i = 0; i < 8; ++i)
                if ((H >> i) & 1) {
                    int rk = __builtin_popcount(H & w.pre[cls][i]);
                    int idx = flip ? sp + rk : sp + nHit - 1 - rk;
                    stack[idx] = E{w.ref[i], tm[i]};
                }
            sp += nHit;
            popNext();
        } else {
            int slot = lslot, cnt = lcnt;
            WideStats().leaves++;
            WideStats().prims += cnt;
            for (int k = 0; k < cnt; ++k, ++slot) {
                const PrimSlot &s = hs.prims[slot];
                if (s.mesh == kInstanceTag) {
                    const InstanceDev &in = hs.instances[s.tri];
                    float o2[3], d2[3], tMax2;
                    InstanceRay(in.mInv, o[0], o[1], o[2], d[0], d[1], d[2], tMax, o2, d2, tMax2);
                    Best sub;
                    int h = TraceWide(hs, in.rootWide, in.rootWide < 0 ? nullptr : in.rootBox, in.rootBox, o2, d2, tMax2, any,
                                      sub, int(s.tri));
                    if (h < 0) return -1;
                    if (any && h) return 1;
                    if (!any && sub.slot >= 0) {
                        tMax = tMax2;
                        best = sub;
                    }
                    continue;
                }
                float t, b0, b1, b2;
                b0 = 0;
                if (((s.kind & 1) ? BilinearHit(r, tMax, s.p0, s.p1, s.p2, s.p3, t, b1, b2)
                                  : TriHit(r, tMax, s.p0, s.p1, s.p2, t, b0, b1, b2)) &&
                    (!(s.kind & 2) || AlphaKeepsHit(r, s.p3[0]))) {
                    if (any) return 1;
                    tMax = t; best.slot = slot; best.inst = curInst; best.b0 = b0; best.b1 = b1; best.b2 = b2;
                }
            }
            popNext();
        }
    }
    return any ? 0 : int(best.slot >= 0);
}

}  // namespace b200pt
