// render_types.h - flat POD tables of the render path (camera .. film), device side.
// Everything the shading kernels read is in these structs; no host pointers, no virtual
// dispatch (the reference's TaggedPointer dispatch becomes a switch on a small tag).
#pragma once
#include <cstdint>

namespace b200pt {

constexpr float kCIE_Y_integral = 106.856895f;  // util/spectrum.h:31
constexpr int kNLambda = 31;          // NSpectrumSamples of the ISET fork (util/spectrum.h:36)
constexpr float kLambdaMin = 395.f;   // Lambda_min / Lambda_max (util/spectrum.h:34-35)
constexpr float kLambdaMax = 705.f;
constexpr int kSpecTable = 311;       // 395..705 nm at 1 nm (DenselySampledSpectrum default range)

struct Spec {
    float v[kNLambda];
};

// One spectrum as a 1 nm table.  interp 0: nearest (DenselySampledSpectrum::Sample,
// util/spectrum.h:420-430: lround(lambda) - lambda_min); interp 1: linear (stands in for the
// smooth RGBAlbedoSpectrum / piecewise-linear spectra the host evaluated for us).
struct SpectrumTable {
    uint32_t interp;
    float values[kSpecTable];
};

enum MaterialType : uint32_t { MAT_DIFFUSE = 0, MAT_CONDUCTOR = 1, MAT_DIELECTRIC = 2, MAT_COATED_DIFFUSE = 3, MAT_NONE = 0xFFFFFFFFu };

struct MaterialDev {
    uint32_t type;
    int32_t spectrum[4];  // diffuse: [0] reflectance
    float params[12];     // layout per material: include/b200pt.h b200pt_material
};

// DiffuseAreaLight on one triangle (lights.h:410-470) + its alias-table bin (util/sampling.h:804).
struct LightDev {
    uint32_t type;       // 0 = diffuse area light on a triangle, 1 = uniform infinite light; 2 point, 3 spot, 4 distant:
                         //   the delta lights, whose placement is in RenderSceneDev::lightGeom (kLightGeomFloats per light)
    uint32_t slot;       // PrimSlot of its triangle
    int32_t spectrum;    // Lemit
    float scale;
    uint32_t twoSided;
    float cosFalloffEnd, tanFalloffEnd, normalizeFalloffEnd;  // ISET spread angle (lights.cpp:716-718)
    float aliasQ, aliasP;
    int32_t alias;
    uint32_t flipNormal; // mesh reverseOrientation ^ transformSwapsHandedness
};

enum : uint32_t { LIGHT_AREA = 0, LIGHT_UNIFORM_INFINITE = 1, LIGHT_POINT = 2, LIGHT_SPOT = 3, LIGHT_DISTANT = 4,
                  LIGHT_IMAGE_INFINITE = 5 };
// lightGeom record of light i: [0..2] point / spot: renderFromLight(0,0,0); distant: Normalize(renderFromLight(0,0,1));
// [3] distant: sceneRadius (DistantLight::Preprocess, lights.h:277-279); [4..12] spot: the upper 3 x 3 of
// renderFromLight's inverse, row-major; [13] cosFalloffStart, [14] cosFalloffEnd (lights.cpp:1376-1385).
// Image infinite light: [0..8] the upper 3 x 3 of renderFromLight, [9..17] of its inverse, [18] sceneRadius.
constexpr int kLightGeomFloats = 24;

struct MeshAttrDev {
    int32_t material;
    int32_t lightFirst;  // light index of triangle 0 of this mesh, -1 if not emissive
    uint32_t flipNormal;
    uint32_t nOff;  // per-corner normals of triangle t at cornerAttr[nOff + 9 t], kNoCornerAttr = none
    int32_t mediumInside, mediumOutside;  // MediumInterface; equal = no transition (the ray keeps its medium)
    uint32_t uvOff;  // per-corner uv of triangle t at cornerAttr[uvOff + 6 t], kNoCornerAttr = none
    uint32_t sOff;   // per-corner shading tangents of triangle t at cornerAttr[sOff + 9 t], kNoCornerAttr = none
};
constexpr uint32_t kNoCornerAttr = 0xFFFFFFFFu;

// One participating medium (media.h:217-345): homogeneous, or a density grid with its majorant grid.
struct MediumDev {
    uint32_t type;              // 0 homogeneous, 1 grid
    int32_t sigmaA, sigmaS;     // spectrum tables (nearest lookup, already scaled)
    float g;                    // Henyey-Greenstein asymmetry
    float bounds[6];            // grid: medium-space bounds
    float mediumFromRender[16];
    float renderFromMedium[16];
    int32_t nx, ny, nz;
    int32_t mres[3];
    const float *density;
    const float *majorants;
    // emission (media.h:217-262, 288-318): Le_spec (spectrum table; a homogeneous medium's is already scaled) times, for a
    // grid medium, the LeScale grid at the point (its own resolution: 1 x 1 x 1 when the scene gave one number)
    uint32_t emissive;
    int32_t le;
    int32_t lnx, lny, lnz;
    const float *leScale;  // null: homogeneous
    // a grid medium's temperature grid (the density's resolution): blackbody emission instead of Le_spec where
    // (T - temperatureOffset) * temperatureScale > 100 K (media.h:306-314); null = none
    const float *temperature;
    float temperatureScale, temperatureOffset;
};

struct CameraDev {
    uint32_t type;               // 0 perspective (cameras.cpp:433-456), 1 realistic (cameras.cpp:820-1019), 2 omni (:2823-3373)
    float cameraFromRaster[16];  // row major
    float renderFromCamera[16];
    float lensRadius, focalDistance;
    // realistic: lens elements {curvatureRadius, thickness, eta, apertureRadius} front to rear (metres),
    // exit-pupil bounds {xmin, ymin, xmax, ymax} per film-radius bin, film extent and diagonal
    // omni: the interfaces' {transform[16], inverse[16]} follow the table at lens[4 nLens + 32 i]
    const float *lens;
    const float *pupilBounds;
    int32_t nLens, nPupilBounds;
    float physicalExtent[4];     // Bounds2f {pMin.x, pMin.y, pMax.x, pMax.y}
    float filmDiagonal;
    int32_t medium;              // medium camera rays start in, -1 = none
    // omni with a microlens array: its table, transforms and per-lens offsets follow the main lens's in `lens`:
    // micro = lens + 36 nLens: [4 nMicro][32 nMicro][2 microDimX microDimY]
    int32_t nMicro, microDimX, microDimY, microRadius;
    float microOffsetFromSensor;
};

struct FilmDev {
    uint32_t type;               // 0 = RGBFilm (film.h:232-310), 1 = SpectralFilm (film.h:430-560), 2 = GBufferFilm (film.h:322-408)
    int32_t nBuckets;            // spectral film
    int32_t x0, y0, x1, y1;      // pixelBounds
    int32_t fullResX, fullResY;
    uint32_t filterType;         // 0 box; otherwise sampled through filterTable (FilterSampler, filters.cpp:92-118)
    float filterRadiusX, filterRadiusY;
    // FilterSampler: the tabulated filter function and its PiecewiseConstant2D over [-radius, radius]^2, one array:
    // f[ny][nx] | conditional cdf[ny][nx + 1] | conditional funcInt[ny] | marginal cdf[ny + 1] | marginal funcInt
    int32_t filterNx, filterNy;
    const float *filterTable;
    float imagingRatio;
    float maxComponentValue;
    float outputRGBFromSensorRGB[9];
    int32_t sensorSpectrum[3];   // r_bar, g_bar, b_bar tables
    // GBufferFilm: render space -> the AOVs' output space (points through outM; normals through the transpose of outMInv)
    float outM[16], outMInv[16];
    // GBufferFilm's Albedo.* (film.cpp:640-644): computed when gbufferAlbedo != 0.  albedo * illuminant -> XYZ through the
    // CIE matching functions (spectra tables) -> rgbFromXYZ (the film colour space's)
    uint32_t gbufferAlbedo;
    int32_t xyzSpectrum[3], albedoIlluminant;
    float rgbFromXYZ[9];
};

struct SamplerDev {
    uint32_t type;               // 0 = ZSobol (samplers.h:224-370), 1 = independent
    int32_t samplesPerPixel;
    int32_t seed;
    uint32_t randomize;          // 0 none, 1 permutedigits, 2 fastowen, 3 owen
    int32_t log2Spp;
    int32_t nBase4Digits;
};

struct IntegratorDev {
    int32_t maxDepth;
    uint32_t regularize;
    uint32_t lightSampler;       // 0 uniform, 1 power, 2 bvh
    uint32_t type;               // 0 path (cpu/integrators.cpp:629-805), 1 volpath (:953-1405)
};

struct RenderSceneDev {
    CameraDev camera;
    FilmDev film;
    SamplerDev sampler;
    IntegratorDev integrator;
    const SpectrumTable *spectra;
    const MaterialDev *materials;
    const LightDev *lights;
    const MeshAttrDev *meshes;
    const void *prims;           // PrimSlot[] of the geometry scene
    const void *instances;       // InstanceDev[] of the geometry scene (null if none)
    uint32_t nSpectra, nMaterials, nLights, nMeshes;
    // BVHLightSampler (render_lightbvh.h) and the infinite lights (indices into lights)
    const void *lightBvhNodes;
    const uint32_t *lightBitTrails;
    const int32_t *infiniteLights;
    uint32_t nLightBvhNodes, nInfiniteLights;
    float lightBvhBounds[6];
    const MediumDev *media;
    uint32_t nMedia;
    uint32_t geomExt;         // some mesh is a bilinear patch mesh or has per-corner attributes, or the scene uses a material
                              // variant (thin dielectric, diffuse transmission), the independent sampler or a delta light:
                              // the EXT = 1 kernels run; else the EXT = 0 ones, which carry none of those branches
    const float *cornerAttr;  // per-corner normals / uv of the meshes that have them (MeshAttrDev::nOff, uvOff)
    // ImageInfiniteLight (lights.h:557-643), at most one per scene: per pixel of its equal-area square image the
    // RGBIlluminantSpectrum's {c0, c1, c2, scale} (util/spectrum.h RGBSigmoidPolynomial), and the compensated sampling
    // distribution in the layout of FilmDev::filterTable: f[res][res] | conditional cdf | conditional funcInt | marginal
    // cdf | marginal funcInt
    const float *envImage;
    const float *envDist;
    int32_t envRes, envLight;  // resolution; index of the light (-1: none)
    const float *lightGeom;   // delta lights' placement (kLightGeomFloats per light, all lights), or null when there are none
    const int32_t *materialIds;  // GBufferFilm: the host's Material::materialId per material (materials.cpp:738), or null
};

// Per-path state, structure of arrays over `n` paths of a pass.  Spectral arrays are [k][path].
struct PathStateDev {
    int n;
    float *beta;       // 31 x n
    float *L;          // 31 x n
    float *Ld;         // 31 x n  pending direct-light contribution (already x beta), valid if shadowPending
    float *rays;       // n x 7   next closest-hit ray
    float *shadowRays; // n x 7
    uint32_t *hits;    // n x 6   internal hit record {slot, t, b0, b1, b2, 0}
    uint8_t *occluded; // n
    int *pixel;        // pixel index in the film (y * width + x)
    int *sampleIndex;
    float *lambdaU;
    int *dimension;    // sampler dimension consumed so far
    int *depth;
    uint32_t *flags;   // bit0 alive, bit1 specularBounce, bit2 anyNonSpecularBounces, bit3 shadowPending
    float *etaScale;
    float *pB;         // p_b of the last BSDF sample
    float *prevCtx;    // n x 12: p, pError, n, ns of the previous vertex (LightSampleContext)
    float *filterWeight;
};

}  // namespace b200pt
