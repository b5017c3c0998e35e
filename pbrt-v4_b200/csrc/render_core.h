// render_core.h - the per-path logic of the render kernels as host+device inline functions:
// sampler, wavelengths, camera, surface interaction, BSDFs, light sampling, one path-vertex step.
// The kernels in render.cu are thin loops over these; tests/native/rendercheck.cpp runs the very
// same functions on the CPU so the pipeline can be debugged against pbrt_ref without a GPU.
//
// Sampler dimensions are consumed in the order of the reference's CPU PathIntegrator
// (cpu/integrators.cpp:228-290 EvaluatePixelSample, :629-762 Li, :764-805 SampleLd), which is the
// image oracle, so with the same sampler the paths track the reference's paths.
#pragma once
#include "device_scene.h"
#include "render_math.h"
#include "render_types.h"

namespace b200pt {

// ------------------------------------------------------------------ ZSobol sampler (samplers.h:224-370)
struct ZSobol {
    int log2Spp, nBase4Digits, seed;
    uint32_t randomize;
    uint64_t mortonIndex;
    int dimension;
};

B2_HD uint64_t LeftShift2(uint64_t x) {  // util/math.h:83-91
    x &= 0xffffffff;
    x = (x ^ (x << 16)) & 0x0000ffff0000ffffull;
    x = (x ^ (x << 8)) & 0x00ff00ff00ff00ffull;
    x = (x ^ (x << 4)) & 0x0f0f0f0f0f0f0f0full;
    x = (x ^ (x << 2)) & 0x3333333333333333ull;
    x = (x ^ (x << 1)) & 0x5555555555555555ull;
    return x;
}
B2_HD uint64_t EncodeMorton2(uint32_t x, uint32_t y) { return (LeftShift2(y) << 1) | LeftShift2(x); }

B2_HD void ZSobolStart(ZSobol &s, const SamplerDev &sd, int px, int py, int index, int dim) {
    s.log2Spp = sd.log2Spp;
    s.nBase4Digits = sd.nBase4Digits;
    s.seed = sd.seed;
    s.randomize = sd.randomize;
    s.dimension = dim;
    s.mortonIndex = (EncodeMorton2(px, py) << sd.log2Spp) | (uint64_t)index;
}

// the 24 permutations of 4 digits in the reference's order (samplers.h:300-326), one byte each
// ({a,b,c,d} -> a | b<<2 | c<<4 | d<<6), packed into three 64-bit words so the lookup is shifts on
// registers (a byte table indexed per thread would live in local memory on the device).
B2_HD int ZSobolPermute(int p, int digit) {
#define B2_PERM(a, b, c, d) uint64_t((a) | (b) << 2 | (c) << 4 | (d) << 6)
#define B2_PERM8(p0, p1, p2, p3, p4, p5, p6, p7) \
    ((p0) | (p1) << 8 | (p2) << 16 | (p3) << 24 | (p4) << 32 | (p5) << 40 | (p6) << 48 | (p7) << 56)
    const uint64_t w0 = B2_PERM8(B2_PERM(0, 1, 2, 3), B2_PERM(0, 1, 3, 2), B2_PERM(0, 2, 1, 3), B2_PERM(0, 2, 3, 1),
                                 B2_PERM(0, 3, 2, 1), B2_PERM(0, 3, 1, 2), B2_PERM(1, 0, 2, 3), B2_PERM(1, 0, 3, 2));
    const uint64_t w1 = B2_PERM8(B2_PERM(1, 2, 0, 3), B2_PERM(1, 2, 3, 0), B2_PERM(1, 3, 2, 0), B2_PERM(1, 3, 0, 2),
                                 B2_PERM(2, 1, 0, 3), B2_PERM(2, 1, 3, 0), B2_PERM(2, 0, 1, 3), B2_PERM(2, 0, 3, 1));
    const uint64_t w2 = B2_PERM8(B2_PERM(2, 3, 0, 1), B2_PERM(2, 3, 1, 0), B2_PERM(3, 1, 2, 0), B2_PERM(3, 1, 0, 2),
                                 B2_PERM(3, 2, 1, 0), B2_PERM(3, 2, 0, 1), B2_PERM(3, 0, 2, 1), B2_PERM(3, 0, 1, 2));
#undef B2_PERM8
#undef B2_PERM
    uint64_t w = p < 8 ? w0 : (p < 16 ? w1 : w2);
    return (int)(w >> (8 * (p & 7) + 2 * digit)) & 3;
}

B2_HD uint64_t ZSobolSampleIndex(const ZSobol &s) {  // samplers.h:298-353
    uint64_t sampleIndex = 0;
    bool pow2Samples = s.log2Spp & 1;
    int lastDigit = pow2Samples ? 1 : 0;
    for (int i = s.nBase4Digits - 1; i >= lastDigit; --i) {
        int digitShift = 2 * i - (pow2Samples ? 1 : 0);
        int digit = (s.mortonIndex >> digitShift) & 3;
        uint64_t higherDigits = s.mortonIndex >> (digitShift + 2);
        int p = (MixBits(higherDigits ^ (0x55555555u * (uint32_t)s.dimension)) >> 24) % 24;
        digit = ZSobolPermute(p, digit);
        sampleIndex |= uint64_t(digit) << digitShift;
    }
    if (pow2Samples) {
        int digit = s.mortonIndex & 1;
        sampleIndex |= digit ^ (MixBits((s.mortonIndex >> 1) ^ (0x55555555u * (uint32_t)s.dimension)) & 1);
    }
    return sampleIndex;
}

// Sobol' generator matrices of dimensions 0 and 1 (the only ones ZSobol uses; SobolMatrices32[0..103],
// util/sobolmatrices.cpp:40-58).  Dimension 0 is the identity (bit j -> 0x80000000 >> j: a bit
// reversal).  Dimension 1 is Pascal's triangle mod 2: column j has bit (31 - i) set iff
// binomial(j, i) is odd, i.e. iff i is a bit-subset of j (Lucas); for j >= 32 only j & 31 matters.
// SobolColumn states that definition (tests compare it with the reference's table); the product with
// an index is then a superset-XOR over the index bits, done in five butterfly steps (SobolDim1).
B2_HD uint32_t SobolColumn(int dim, int j) {
    if (dim == 0) return j < 32 ? (0x80000000u >> j) : 0u;
    uint32_t v = 0;
    for (int i = 0; i < 32; ++i)
        if ((i & ~j) == 0) v |= 0x80000000u >> i;
    return v;
}
B2_HD uint32_t SobolDim1(uint64_t a) {
    uint32_t x = (uint32_t)a ^ (uint32_t)(a >> 32);  // columns repeat with period 32
    x ^= (x >> 1) & 0x55555555u;                     // out[i] = XOR of in[j] over all j that contain i
    x ^= (x >> 2) & 0x33333333u;
    x ^= (x >> 4) & 0x0f0f0f0fu;
    x ^= (x >> 8) & 0x00ff00ffu;
    x ^= (x >> 16) & 0x0000ffffu;
    return ReverseBits32(x);
}
B2_HD uint32_t FastOwen(uint32_t v, uint32_t seed) {  // util/lowdiscrepancy.h:221-237
    v = ReverseBits32(v);
    v ^= v * 0x3d20adea;
    v += seed;
    v *= (seed >> 16) | 1;
    v ^= v * 0x05526c56;
    v ^= v * 0x53a22864;
    return ReverseBits32(v);
}
B2_HD uint32_t OwenScramble(uint32_t v, uint32_t seed) {  // util/lowdiscrepancy.h:240-262
    if (seed & 1) v ^= 1u << 31;
    for (int b = 1; b < 32; ++b) {
        uint32_t mask = (~0u) << (32 - b);
        if ((uint32_t)MixBits((v & mask) ^ seed) & (1u << b)) v ^= 1u << (31 - b);
    }
    return v;
}
B2_HD float SobolSampleZ(uint64_t a, int dim, uint32_t randomize, uint32_t hash) {  // util/lowdiscrepancy.h:168-180
    uint32_t v = dim == 0 ? ReverseBits32((uint32_t)a) : SobolDim1(a);
    if (randomize == 1) v ^= hash;
    else if (randomize == 2) v = FastOwen(v, hash);
    else if (randomize == 3) v = OwenScramble(v, hash);
    return fminf(v * 0x1p-32f, kOneMinusEps);
}
B2_HD_NOINLINE float ZSobolGet1D(ZSobol &s) {  // samplers.h:258-272
    uint64_t idx = ZSobolSampleIndex(s);
    ++s.dimension;
    uint32_t h = (uint32_t)Hash2i(s.dimension, s.seed);
    return SobolSampleZ(idx, 0, s.randomize, h);
}
B2_HD_NOINLINE void ZSobolGet2D(ZSobol &s, float &u0, float &u1) {  // samplers.h:274-293
    uint64_t idx = ZSobolSampleIndex(s);
    s.dimension += 2;
    uint64_t bits = Hash2i(s.dimension, s.seed);
    u0 = SobolSampleZ(idx, 0, s.randomize, (uint32_t)bits);
    u1 = SobolSampleZ(idx, 1, s.randomize, (uint32_t)(bits >> 32));
}

// ------------------------------------------------------------------ the path's sampler
// ZSobol, or the IndependentSampler (samplers.h:442-476): a PCG32 stream per pixel, set to draw number
// sampleIndex * 65536 + dimension; every value is one draw, so the draw count the path state keeps (`dimension`, as
// for ZSobol) is all it takes to pick the stream up again in the next kernel.  The Sobol core is what the out-of-line
// ZSobolGet* see; `independent` stays a local, so a kernel built without the feature (EXT = 0) folds the branches.
struct PathSampler {
    ZSobol z;
    bool independent;
    PCG32 rng;
};
template <int EXT>
B2_HD void PathSamplerStart(PathSampler &s, const SamplerDev &sd, int px, int py, int index, int dim) {
    s.independent = EXT != 0 && sd.type == 1;
    if (s.independent) {  // IndependentSampler::StartPixelSample (samplers.h:457-460)
        unsigned char buf[12];  // Hash(Point2i p, int seed)
        memcpy(buf, &px, 4);
        memcpy(buf + 4, &py, 4);
        memcpy(buf + 8, &sd.seed, 4);
        uint64_t seq = Murmur64A(buf, 12, 0);
        s.rng = PcgSeed(seq, MixBits(seq));
        PcgAdvance(s.rng, uint64_t(index) * 65536ull + uint64_t(dim));
        s.z.dimension = dim;
        return;
    }
    ZSobolStart(s.z, sd, px, py, index, dim);
}
B2_HD float SamplerGet1D(PathSampler &s) {
    if (s.independent) {  // rng.Uniform<Float>() (util/rng.h:120-123)
        ++s.z.dimension;
        return fminf(kOneMinusEps, PcgNext(s.rng) * 0x1p-32f);
    }
    return ZSobolGet1D(s.z);
}
B2_HD void SamplerGet2D(PathSampler &s, float &u0, float &u1) {
    if (s.independent) {
        s.z.dimension += 2;
        u0 = fminf(kOneMinusEps, PcgNext(s.rng) * 0x1p-32f);
        u1 = fminf(kOneMinusEps, PcgNext(s.rng) * 0x1p-32f);
        return;
    }
    ZSobolGet2D(s.z, u0, u1);
}

// ------------------------------------------------------------------ pixel filter
// FindInterval (util/math.h:509-520) over cdf[0..sz-1] with the predicate cdf[i] <= u
B2_HD int FindCdfInterval(const float *cdf, int sz, float u) {
    int size = sz - 2, first = 1;
    while (size > 0) {
        int half = size >> 1, middle = first + half;
        bool pred = cdf[middle] <= u;
        first = pred ? middle + 1 : first;
        size = pred ? size - (half + 1) : half;
    }
    int o = first - 1;
    return o < 0 ? 0 : (o > sz - 2 ? sz - 2 : o);
}
// PiecewiseConstant1D::Sample (util/sampling.h:657-675) over n cells of [lo, hi]; returns x, *offset the cell
B2_HD float SamplePiecewise1D(const float *cdf, int n, float lo, float hi, float u, int *offset) {
    int o = FindCdfInterval(cdf, n + 1, u);
    *offset = o;
    float du = u - cdf[o];
    if (cdf[o + 1] - cdf[o] > 0) du /= cdf[o + 1] - cdf[o];
    return Lerp((o + du) / n, lo, hi);
}
// PiecewiseConstant2D::Sample (util/sampling.h:752-762) over a table f[ny][nx] | conditional cdf[ny][nx + 1] |
// conditional funcInt[ny] | marginal cdf[ny + 1] | marginal funcInt, domain [lox, hix] x [loy, hiy]
B2_HD void SamplePiecewise2D(const float *f, int nx, int ny, float lox, float hix, float loy, float hiy, float u0, float u1,
                             float &x, float &y, float &pdf, int &iu, int &iv) {
    const float *condCdf = f + nx * ny, *condInt = condCdf + ny * (nx + 1);
    const float *margCdf = condInt + ny, margInt = margCdf[ny + 1];
    y = SamplePiecewise1D(margCdf, ny, loy, hiy, u1, &iv);
    float pdf1 = margInt > 0 ? condInt[iv] / margInt : 0.f;
    x = SamplePiecewise1D(condCdf + iv * (nx + 1), nx, lox, hix, u0, &iu);
    float pdf0 = condInt[iv] > 0 ? fabsf(f[iv * nx + iu]) / condInt[iv] : 0.f;
    pdf = pdf0 * pdf1;
}
// PiecewiseConstant2D::PDF (util/sampling.h:764-771) for p already in [0, 1]^2 of the domain
B2_HD float Piecewise2DPDF(const float *f, int nx, int ny, float px, float py) {
    int iu = (int)(px * nx), iv = (int)(py * ny);
    iu = iu < 0 ? 0 : (iu > nx - 1 ? nx - 1 : iu);
    iv = iv < 0 ? 0 : (iv > ny - 1 ? ny - 1 : iv);
    const float margInt = f[nx * ny + ny * (nx + 1) + ny + ny + 1];
    return fabsf(f[iv * nx + iu]) / margInt;
}

// Filter::Sample(u): BoxFilter (filters.h:60-64), or FilterSampler::Sample (filters.h:27-33): the point from the
// tabulated distribution (PiecewiseConstant2D::Sample, util/sampling.h:752-762), weight f[cell] / pdf
// SampleTent (util/sampling.h:196-201): SampleDiscrete({0.5, 0.5}, u, nullptr, &u) picks the half, SampleLinear the point
B2_HD float SampleTent(float u, float r) {
    float up = u;  // u * sumWeights, sumWeights = 1
    if (up == 1.f) up = NextDown(up);
    const bool upper = 0.5f <= up;
    float ur = fminf((up - (upper ? 0.5f : 0.f)) / 0.5f, kOneMinusEps);
    return upper ? r * SampleLinear(ur, 1, 0) : -r + r * SampleLinear(ur, 0, 1);
}
B2_HD void SampleFilter(const FilmDev &film, float u0, float u1, float &fx, float &fy, float &weight) {
    if (film.filterType == 0) {
        fx = Lerp(u0, -film.filterRadiusX, film.filterRadiusX);
        fy = Lerp(u1, -film.filterRadiusY, film.filterRadiusY);
        weight = 1.f;
        return;
    }
    if (film.filterType == 4) {  // TriangleFilter::Sample (filters.h:224-227): tent-distributed, weight 1
        fx = SampleTent(u0, film.filterRadiusX);
        fy = SampleTent(u1, film.filterRadiusY);
        weight = 1.f;
        return;
    }
    const int nx = film.filterNx, ny = film.filterNy;
    int iv, iu;
    float pdf;
    SamplePiecewise2D(film.filterTable, nx, ny, -film.filterRadiusX, film.filterRadiusX, -film.filterRadiusY,
                      film.filterRadiusY, u0, u1, fx, fy, pdf, iu, iv);
    weight = film.filterTable[iv * nx + iu] / pdf;
}

// ------------------------------------------------------------------ wavelengths and spectra
// SampledWavelengths::SampleUniform (util/spectrum.h:318-336): all 31 wavelengths follow from u.
B2_HD void SampleWavelengths(float u, float lambda[kNLambda]) {
    lambda[0] = Lerp(u, kLambdaMin, kLambdaMax);
    float delta = (kLambdaMax - kLambdaMin) / kNLambda;
    for (int i = 1; i < kNLambda; ++i) {
        lambda[i] = lambda[i - 1] + delta;
        if (lambda[i] > kLambdaMax) lambda[i] = kLambdaMin + (lambda[i] - kLambdaMax);
    }
}
B2_HD float SpectrumAt(const SpectrumTable &t, float lambda) {
    if (t.interp == 0) {  // DenselySampledSpectrum::Sample (util/spectrum.h:420-430)
        // std::lround(lambda); for 0 < lambda < 2^22, lambda + 0.5f is exact or a tie that floors alike
        int offset = (int)floorf(lambda + 0.5f) - (int)kLambdaMin;
        return (offset < 0 || offset >= kSpecTable) ? 0.f : t.values[offset];
    }
    float x = lambda - kLambdaMin;
    int i = (int)floorf(x);
    if (i < 0) return t.values[0];
    if (i >= kSpecTable - 1) return t.values[kSpecTable - 1];
    return Lerp(x - i, t.values[i], t.values[i + 1]);
}

// lambda_k of SampledWavelengths::SampleUniform (util/spectrum.h:318-336), one at a time
struct LambdaIter {
    float lam;
    B2_HD explicit LambdaIter(float u) : lam(Lerp(u, kLambdaMin, kLambdaMax)) {}
    B2_HD void Next() {
        lam = lam + (kLambdaMax - kLambdaMin) / kNLambda;
        if (lam > kLambdaMax) lam = kLambdaMin + (lam - kLambdaMax);
    }
};

// ------------------------------------------------------------------ camera (cameras.cpp:433-456)
B2_HD V3 XformPoint(const float *m, V3 p) {  // Transform::operator()(Point3) util/transform.h
    float x = m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3];
    float y = m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7];
    float z = m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11];
    float w = m[12] * p.x + m[13] * p.y + m[14] * p.z + m[15];
    if (w == 1) return mk(x, y, z);
    return mk(x, y, z) / w;
}
B2_HD V3 XformVector(const float *m, V3 v) {
    return mk(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z,
              m[8] * v.x + m[9] * v.y + m[10] * v.z);
}
B2_HD void GeneratePerspectiveRay(const CameraDev &c, float pFilmX, float pFilmY, float uLens0, float uLens1,
                                  V3 &o, V3 &d) {
    V3 pCamera = XformPoint(c.cameraFromRaster, mk(pFilmX, pFilmY, 0));
    V3 ro = mk(0, 0, 0), rd = Normalize(pCamera);
    if (c.lensRadius > 0) {
        float lx, ly;
        SampleUniformDiskConcentric(uLens0, uLens1, lx, ly);
        lx *= c.lensRadius;
        ly *= c.lensRadius;
        float ft = c.focalDistance / rd.z;
        V3 pFocus = ro + rd * ft;
        ro = mk(lx, ly, 0);
        rd = Normalize(pFocus - ro);
    }
    // RenderFromCamera(ray) = Transform::operator()(Ray) (util/transform.h:336-348): with an exact origin the
    // error box is gamma(3) * sum |m_ij o_j| + |m_i3|; a pinhole at the render-space origin has none.
    o = XformPoint(c.renderFromCamera, ro);
    d = XformVector(c.renderFromCamera, rd);
    const float *m = c.renderFromCamera;
    if (ro.x != 0 || ro.y != 0 || ro.z != 0 || m[3] != 0 || m[7] != 0 || m[11] != 0) {
        const float g3 = GammaF(3);
        V3 e = mk(g3 * (fabsf(m[0] * ro.x) + fabsf(m[1] * ro.y) + fabsf(m[2] * ro.z) + fabsf(m[3])),
                  g3 * (fabsf(m[4] * ro.x) + fabsf(m[5] * ro.y) + fabsf(m[6] * ro.z) + fabsf(m[7])),
                  g3 * (fabsf(m[8] * ro.x) + fabsf(m[9] * ro.y) + fabsf(m[10] * ro.z) + fabsf(m[11])));
        P3i pi = MakeP3i(o, e);
        float lengthSquared = LengthSquared(d);
        if (lengthSquared > 0) {
            float dt = Dot(Abs(d), Err(pi)) / lengthSquared;
            V3 s = d * dt;
            pi.lo = mk(NextDown(pi.lo.x + s.x), NextDown(pi.lo.y + s.y), NextDown(pi.lo.z + s.z));
            pi.hi = mk(NextUp(pi.hi.x + s.x), NextUp(pi.hi.y + s.y), NextUp(pi.hi.z + s.z));
        }
        o = Mid(pi);
    }
}

}  // namespace b200pt

#include "render_shade.h"
