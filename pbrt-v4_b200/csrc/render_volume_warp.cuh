// render_volume_warp.cuh - the two volume kernels with one WARP per ray and one LANE per wavelength.
// Delta / ratio tracking updates 31-wide spectra (T_maj, beta, r_u, r_l, T_ray) at every tentative collision while
// every decision (collision distance, event type, roulette) is made on wavelength 0 and on scalar random numbers:
// with a thread per ray the spectra are 31-trip loops and the walks' very different lengths leave ~6 of 32 lanes
// busy (profiles/r01_config5_*); with a lane per wavelength the spectra are one register each, the control flow
// is warp-uniform by construction and the scalar work (DDA, RNG, density lookup) is done once per warp instruction.
// Arithmetic is the per-thread code's (render_volume.h / render_volpath.h) operation for operation - sums over
// wavelengths are taken in index order through shuffles - so both variants produce the same film bit for bit
// (tests/test_render_gpu.py compares them).  Included by render.cu only.
//
// MEASURED (B200, config 5, 16 spp of 1920x1080): thread per ray over medium-sorted lists 1451 ms, this variant
// 2237 ms at 157 registers and 1888 ms capped at 64 registers.  Per tentative collision the scalar work is ~150
// instructions against ~1100 of spectral work, so a warp per ray spends (150 + 1100 / 31) issue slots where a
// diverged thread-per-ray warp (~6 live lanes) spends ~1250 / 6: no gain in issue slots, and one serial dependent
// chain per warp instead of six.  So the thread-per-ray kernels stay the default; B200PT_VOL_KERNEL=warp selects these.
#pragma once

namespace b200pt {

constexpr unsigned kFullMask = 0xFFFFFFFFu;
__device__ __forceinline__ float Lane0(float v) { return __shfl_sync(kFullMask, v, 0); }
__device__ __forceinline__ bool AnyLambda(bool p, int lane) { return __any_sync(kFullMask, p && lane < kNLambda); }
// v[0] + v[1] + ... + v[30], added in index order (SampledSpectrum::Average's order)
__device__ __forceinline__ float SeqSumLambda(float v) {
    float s = __shfl_sync(kFullMask, v, 0);
#pragma unroll
    for (int k = 1; k < kNLambda; ++k) s += __shfl_sync(kFullMask, v, k);
    return s;
}
__device__ __forceinline__ float MaxLambda(float v, int lane) {
    float m = lane < kNLambda ? v : -INFINITY;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(kFullMask, m, o));
    return m;
}
// lambda_k of this lane (the recurrence of LambdaIter, k steps)
__device__ __forceinline__ float LaneLambda(float lambdaU, int k) {
    LambdaIter li(lambdaU);
    for (int i = 0; i < k; ++i) li.Next();
    return li.lam;
}

// SampleT_maj (media.h:734-804), lane k holding sigma_t[k] and T_maj[k]
template <class CB>
__device__ float WarpSampleTMaj(const MediumDev &m, float sigT, float sigT0, V3 ro, V3 rd, float tMax, float u, PCG32 &rng,
                                CB &cb) {
    tMax *= Length(rd);
    rd = Normalize(rd);
    MajorantIter it = MediumSampleRay(m, ro, rd, tMax);
    float T = 1.f;
    bool done = false;
    while (!done) {
        MajorantSeg seg;
        if (!MajorantNext(m, it, seg)) return T;
        const float sigMaj0 = m.type == 0 ? sigT0 : sigT0 * seg.scale;
        const float sigMaj = m.type == 0 ? sigT : sigT * seg.scale;
        if (sigMaj0 == 0) {
            float dt = seg.tMax - seg.tMin;
            if (isinf(dt)) dt = 3.402823466e+38f;
            T *= FastExpRef(-dt * sigMaj);
            continue;
        }
        float tMin = seg.tMin;
        while (true) {
            float t = tMin + (-logf(1 - u) / sigMaj0);
            u = PcgFloat(rng);
            if (t < seg.tMax) {
                const float nd = -(t - tMin);
                T *= FastExpRef(nd * sigMaj);
                V3 p = ro + rd * t;
                if (!cb(p, MediumDensityAt(m, p), m.type == 0 ? 1.f : seg.scale, T)) {
                    done = true;
                    break;
                }
                T = 1.f;
                tMin = t;
            } else {
                float dt = seg.tMax - tMin;
                if (isinf(dt)) dt = 3.402823466e+38f;
                T *= FastExpRef(-dt * sigMaj);
                break;
            }
        }
    }
    return 1.f;
}

// PathMediumCB (render_volume.h) with one wavelength per lane
struct WarpPathCB {
    float beta, ru, rl;
    float sigA, sigS, sigT, sigA0, sigS0, sigT0;
    PCG32 &rng;
    int &depth;
    int maxDepth, lane;
    bool terminated = false, scattered = false;
    V3 pScatter;
    __device__ WarpPathCB(PCG32 &r, int &d, int md, int l) : rng(r), depth(d), maxDepth(md), lane(l) { pScatter = mk(0, 0, 0); }
    __device__ bool operator()(V3 p, float density, float majScale, float T) {
        if (!AnyLambda(beta != 0, lane)) {
            terminated = true;
            return false;
        }
        const float sigMaj0 = sigT0 * majScale;
        float pAbsorb = (sigA0 * density) / sigMaj0;
        float pScat = (sigS0 * density) / sigMaj0;
        float pNull = fmaxf(0.f, 1 - pAbsorb - pScat);
        float um = PcgFloat(rng);
        int mode = SampleDiscrete3(pAbsorb, pScat, pNull, um);
        if (mode == 0) {
            terminated = true;
            return false;
        }
        const float T0 = Lane0(T);
        if (mode == 1) {
            if (depth++ >= maxDepth) {
                terminated = true;
                return false;
            }
            float pdf = T0 * (sigS0 * density);
            float f = T * (sigS * density) / pdf;
            beta *= f;
            ru *= f;
            if (AnyLambda(beta != 0, lane) && AnyLambda(ru != 0, lane)) {
                scattered = true;
                pScatter = p;
            }
            return false;
        }
        float sigMaj = sigT * majScale;
        float sigN = fmaxf(0.f, sigMaj - sigA * density - sigS * density);
        float pdf = T0 * Lane0(sigN);
        beta = pdf == 0 ? 0.f : beta * (T * sigN / pdf);
        ru *= T * sigN / pdf;
        rl *= T * sigMaj / pdf;
        return AnyLambda(beta != 0, lane) && AnyLambda(ru != 0, lane);
    }
};

// ShadowMediumCB (render_volume.h) with one wavelength per lane
struct WarpShadowCB {
    float T_ray, r_l, r_u;
    float sigA, sigS, sigT, sigT0;
    PCG32 &rng;
    int lane;
    __device__ WarpShadowCB(PCG32 &r, int l) : rng(r), lane(l) {}
    __device__ bool operator()(V3, float density, float majScale, float T) {
        float pdf = Lane0(T) * (sigT0 * majScale);
        float sigMaj = sigT * majScale;
        float sigN = fmaxf(0.f, sigMaj - sigA * density - sigS * density);
        T_ray *= T * sigN / pdf;
        r_l *= T * sigMaj / pdf;
        r_u *= T * sigN / pdf;
        float avg = SeqSumLambda(r_l + r_u) / kNLambda;
        float maxTr = MaxLambda(T_ray / avg, lane);
        if (maxTr < 0.05f) {
            const float q = 0.75f;
            if (PcgFloat(rng) < q)
                T_ray = 0.f;
            else
                T_ray /= 1 - q;
        }
        return AnyLambda(T_ray != 0, lane);
    }
};

constexpr int kVolWarpBlock = 128;
#ifndef B2_VOLWARP_MIN_BLOCKS
#define B2_VOLWARP_MIN_BLOCKS 8
#endif

// ---- VolMediumKernel, one warp per ray of the medium work list `kind`
template <int LBVH>
__global__ void __launch_bounds__(kVolWarpBlock, B2_VOLWARP_MIN_BLOCKS)
VolMediumWarpKernel(RenderSceneDev rs, PassBuffers pb, int depth, int nPixels, int width, int kind) {
    const int lane = threadIdx.x & 31, kk = lane < kNLambda ? lane : kNLambda - 1;
    const int warpsPerBlock = kVolWarpBlock / 32;
    const int nWarps = gridDim.x * warpsPerBlock;
    int *counts = pb.counts + kCountStride * depth;
    const int count = counts[6 + kind];
    const int cur = depth & 1, nxt = cur ^ 1;
    const long long n = pb.n;
    for (int qi = blockIdx.x * warpsPerBlock + (threadIdx.x >> 5); qi < count; qi += nWarps) {
        const int t = pb.matQ[kind][qi];
        const int path = pb.rayPath[cur][t];
        const float *r = pb.rays[cur] + 7ll * t;
        PathRegs p;
        LoadHeader(pb, path, p);
        int pix = path % nPixels;
        p.px = rs.film.x0 + pix % width;
        p.py = rs.film.y0 + pix / width;
        p.rayO = mk(r[0], r[1], r[2]);
        p.rayD = mk(r[3], r[4], r[5]);
        p.rayTMax = r[6];
        uint32_t slot = pb.hits[6ll * t];
        float tHit = slot == kMissTag ? INFINITY : __uint_as_float(pb.hits[6ll * t + 1]);
        const MediumDev &m = rs.media[p.medium];
        ZSobol s = SamplerFor(rs, p);
        uint64_t hash0 = Hash1f(ZSobolGet1D(s));
        uint64_t hash1 = Hash1f(ZSobolGet1D(s));
        PCG32 rng = PcgSeed(hash0, hash1);
        float u = ZSobolGet1D(s);
        p.dimension = s.dimension;
        const float lam = LaneLambda(p.lambdaU, kk);
        WarpPathCB cb(rng, p.depth, rs.integrator.maxDepth, lane);
        cb.beta = pb.beta[kk * n + path];
        cb.ru = pb.ru[kk * n + path];
        cb.rl = pb.rl[kk * n + path];
        cb.sigA = MediumSigmaA(rs, m, lam);
        cb.sigS = MediumSigmaS(rs, m, lam);
        cb.sigT = cb.sigA + cb.sigS;
        cb.sigA0 = Lane0(cb.sigA);
        cb.sigS0 = Lane0(cb.sigS);
        cb.sigT0 = Lane0(cb.sigT);
        float T = WarpSampleTMaj(m, cb.sigT, cb.sigT0, p.rayO, p.rayD, tHit, u, rng, cb);
        if (cb.terminated || !AnyLambda(cb.beta != 0, lane) || !AnyLambda(cb.ru != 0, lane)) {
            if (lane == 0) {
                pb.hits[6ll * t] = kConsumedTag;
                pb.ints[3 * n + path] = (int)(p.flags & ~PF_ALIVE);
            }
            continue;
        }
        if (!cb.scattered) {
            float f = T / Lane0(T);
            if (lane < kNLambda) {
                pb.beta[kk * n + path] = cb.beta * f;
                pb.ru[kk * n + path] = cb.ru * f;
                pb.rl[kk * n + path] = cb.rl * f;
            }
            if (lane == 0) StoreHeader(pb, path, p);
            continue;
        }
        // ---- real scattering: light sample, then the phase function (VolMediumStep, render_volpath.h)
        V3 pS = cb.pScatter;
        V3 woN = Normalize(-p.rayD);
        LightCtx ctx;
        ctx.pi = MakeP3i(pS, mk(0, 0, 0));
        ctx.n = ctx.ns = mk(0, 0, 0);
        s = SamplerFor(rs, p);
        float uL = ZSobolGet1D(s);
        float uLight0, uLight1;
        ZSobolGet2D(s, uLight0, uLight1);
        LightPick lp = PickAndSampleLight<LBVH>(rs, ctx, uL, uLight0, uLight1);
        if (lp.ok) {
            float fHat = HenyeyGreenstein(Dot(woN, lp.wi), m.g);
            if (fHat != 0) {
                const LightDev &light = rs.lights[lp.lidx];
                float Le = light.scale * SpectrumAt(rs.spectra[light.spectrum], lam) * lp.atten;
                bool anyL = AnyLambda(Le != 0, lane);
                int sslot = 0;
                if (lane == 0) sslot = atomicAdd(counts + 1, 1);
                sslot = __shfl_sync(kFullMask, sslot, 0);
                if (lane < kNLambda) pb.Ld[kk * n + sslot] = cb.beta * fHat * Le;
                if (lane == 0) {
                    ShadowVolInfo info;
                    info.p_l = lp.pmf * lp.ss.pdf;
                    info.scatterPDF = fHat;
                    info.medium = p.medium;
                    info.lightPi = lp.ss.pi;
                    info.lightN = lp.ss.n;
                    V3 pf = OffsetRayOrigin(ctx.pi, ctx.n, Mid(lp.ss.pi) - pS);
                    V3 pt = OffsetRayOrigin(lp.ss.pi, lp.ss.n, pf - Mid(lp.ss.pi));
                    DevIO io(pb, path, counts + 1, nullptr);
                    io.slot = sslot;
                    io.ShadowEndVol(anyL, pf, pt - pf, 1 - kShadowEpsilon, info);
                }
            }
        }
        float u0, u1;
        ZSobolGet2D(s, u0, u1);
        p.dimension = s.dimension;
        float pdf;
        V3 wi = SampleHenyeyGreenstein(-p.rayD, m.g, u0, u1, &pdf);
        if (lane == 0) pb.hits[6ll * t] = kConsumedTag;
        if (pdf == 0) {
            if (lane == 0) pb.ints[3 * n + path] = (int)(p.flags & ~PF_ALIVE);
            continue;
        }
        if (lane < kNLambda) {
            pb.beta[kk * n + path] = cb.beta * (pdf / pdf);
            pb.ru[kk * n + path] = cb.ru;
            pb.rl[kk * n + path] = cb.ru / pdf;
        }
        if (lane == 0) {
            p.prev = ctx;
            p.rayO = pS;
            p.rayD = wi;
            p.rayTMax = INFINITY;
            p.flags = (p.flags & ~PF_SPECULAR) | PF_ANY_NONSPECULAR;
            int ns = atomicAdd(pb.counts + kCountStride * (depth + 1), 1);
            StoreRay(pb.rays[nxt], ns, p.rayO, p.rayD, p.rayTMax);
            pb.rayPath[nxt][ns] = path;
            StoreHeader(pb, path, p);
        }
    }
}

// ---- ShadowTrKernel, one warp per live light sample of the walk list `kind`
__global__ void __launch_bounds__(kVolWarpBlock, B2_VOLWARP_MIN_BLOCKS)
ShadowTrWarpKernel(RenderSceneDev rs, PassBuffers pb, int depth, int kind) {
    const int lane = threadIdx.x & 31, kk = lane < kNLambda ? lane : kNLambda - 1;
    const int warpsPerBlock = kVolWarpBlock / 32;
    const int nWarps = gridDim.x * warpsPerBlock;
    const int count = pb.counts[kCountStride * depth + 2 + kind];
    const long long n = pb.n;
    for (int qi = blockIdx.x * warpsPerBlock + (threadIdx.x >> 5); qi < count; qi += nWarps) {
        const int t = pb.matQ[kind][qi];
        const int path = pb.shPath[t];
        float *ray = pb.shRays + 7ll * t;
        uint32_t *a = pb.shAux + (long long)kShadowAuxWords * t;
        ShadowRegs sh;
        sh.o = mk(ray[0], ray[1], ray[2]);
        sh.d = mk(ray[3], ray[4], ray[5]);
        const float rayTMax = ray[6];
        sh.medium = (int)a[2];
        sh.lightPi.lo = mk(__uint_as_float(a[8]), __uint_as_float(a[9]), __uint_as_float(a[10]));
        sh.lightPi.hi = mk(__uint_as_float(a[11]), __uint_as_float(a[12]), __uint_as_float(a[13]));
        sh.lightN = mk(__uint_as_float(a[14]), __uint_as_float(a[15]), __uint_as_float(a[16]));
        const float p_l = __uint_as_float(a[0]), scatterPDF = __uint_as_float(a[1]);
        WarpShadowCB cb(sh.rng, lane);
        if (a[3] != 0) {
            sh.rng.state = (uint64_t)a[4] | ((uint64_t)a[5] << 32);
            sh.rng.inc = (uint64_t)a[6] | ((uint64_t)a[7] << 32);
            cb.T_ray = pb.shT[kk * n + t];
            cb.r_l = pb.shRl[kk * n + t];
            cb.r_u = pb.shRu[kk * n + t];
        } else {
            sh.rng = PcgSeed(Hash3f(sh.o.x, sh.o.y, sh.o.z), Hash3f(sh.d.x, sh.d.y, sh.d.z));
            cb.T_ray = cb.r_l = cb.r_u = 1.f;
        }
        uint32_t hit[6];
        for (int i = 0; i < 6; ++i) hit[i] = pb.shHits[6ll * t + i];
        __syncwarp();  // every lane has read the slot before lane 0 rewrites it
        // ---- ShadowTrSegment (render_volpath.h)
        const bool isHit = hit[0] != kMissTag;
        uint32_t meshId = 0;
        SurfHit si;
        int res = -1;
        if (isHit) {
            si = SurfHitFromRecord(rs, sh.d, hit, &meshId);
            if (rs.meshes[meshId].material >= 0) res = SHADOW_BLOCKED;
        }
        if (res < 0 && sh.medium >= 0) {
            const MediumDev &m = rs.media[sh.medium];
            float tMax = isHit ? __uint_as_float(hit[1]) : (1 - kShadowEpsilon);
            float u = PcgFloat(sh.rng);
            const float lam = LaneLambda(pb.scalars[path], kk);
            cb.sigA = MediumSigmaA(rs, m, lam);
            cb.sigS = MediumSigmaS(rs, m, lam);
            cb.sigT = cb.sigA + cb.sigS;
            cb.sigT0 = Lane0(cb.sigT);
            float T = WarpSampleTMaj(m, cb.sigT, cb.sigT0, sh.o, sh.d, tMax, u, sh.rng, cb);
            float f = T / Lane0(T);
            cb.T_ray *= f;
            cb.r_l *= f;
            cb.r_u *= f;
        }
        if (res < 0) {
            if (!AnyLambda(cb.T_ray != 0, lane))
                res = SHADOW_BLOCKED;
            else if (!isHit)
                res = SHADOW_REACHED;
            else {
                V3 pf = OffsetRayOrigin(si.pi, si.n, Mid(sh.lightPi) - Mid(si.pi));
                V3 pt = OffsetRayOrigin(sh.lightPi, sh.lightN, pf - Mid(sh.lightPi));
                sh.o = pf;
                sh.d = pt - pf;
                sh.medium = MediumAfter(rs.meshes[meshId], si.n, sh.d, sh.medium);
                res = (sh.d.x == 0 && sh.d.y == 0 && sh.d.z == 0) ? SHADOW_REACHED : SHADOW_CONTINUE;
            }
        }
        if (res == SHADOW_CONTINUE) {
            if (lane < kNLambda) {
                pb.shT[kk * n + t] = cb.T_ray;
                pb.shRl[kk * n + t] = cb.r_l;
                pb.shRu[kk * n + t] = cb.r_u;
            }
            if (lane == 0) {
                StoreRay(pb.shRays, t, sh.o, sh.d, rayTMax);
                a[2] = (uint32_t)sh.medium;
                a[3] = 1;
                a[4] = (uint32_t)sh.rng.state; a[5] = (uint32_t)(sh.rng.state >> 32);
                a[6] = (uint32_t)sh.rng.inc;   a[7] = (uint32_t)(sh.rng.inc >> 32);
                atomicAdd(pb.shContinue, 1);
            }
            continue;
        }
        if (res == SHADOW_REACHED) {  // ShadowTrFinish
            float rp = pb.ru[kk * n + path];
            float av = cb.r_l * (rp * p_l), bv = cb.r_u * (rp * scatterPDF);
            float avg = SeqSumLambda(av + bv) / kNLambda;
            if (lane < kNLambda) pb.L[kk * n + path] += pb.Ld[kk * n + t] * cb.T_ray / avg;
        }
        if (lane == 0) {
            pb.shPath[t] = -1;
            ray[6] = -1.f;
        }
    }
}

}  // namespace b200pt
