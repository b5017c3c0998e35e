/* b200pt.h - C ABI of the B200-native wavefront path-tracing backend for
 * pbrt-v4 (ISET fork).  Plain pointers and sizes only; no C++/torch types.
 *
 * This is the drop-in boundary.  Each entry point names the
 * reference interface it replaces (paths relative to the reference's src/pbrt):
 *
 *   b200pt_scene_create        CPUAggregate::CPUAggregate -> BasicScene::CreateAggregate
 *                              (wavefront/aggregate.cpp:20-31, scene.cpp:1430-1596) and the
 *                              BVHAggregate ctor / SAH build / flattenBVH
 *                              (cpu/aggregates.cpp:140-190, 192-387, 505-522)
 *   b200pt_scene_bounds        WavefrontAggregate::Bounds (wavefront/integrator.h:36,
 *                              wavefront/aggregate.h:33)
 *   b200pt_intersect_closest   WavefrontAggregate::IntersectClosest (wavefront/integrator.h:38-43,
 *                              wavefront/aggregate.cpp:34-59) = BVHAggregate::Intersect
 *                              (cpu/aggregates.cpp:529-579) over a ray batch
 *   b200pt_intersect_any       WavefrontAggregate::IntersectShadow (wavefront/integrator.h:45-46,
 *                              wavefront/aggregate.cpp:61-70) = BVHAggregate::IntersectP
 *                              (cpu/aggregates.cpp:581-624) over a ray batch
 *   b200pt_bvh2_build          BVHAggregate::buildRecursive + flattenBVH, host only
 *   b200pt_render / film_*     RenderWavefront (wavefront/wavefront.cpp:14-70):
 *                              WavefrontPathIntegrator::Render (wavefront/integrator.cpp:290-493)
 *
 * Conventions: every function returns 0 (B200PT_OK) or a B200PT_ERR_* code and
 * never aborts or throws (the reference LOG_FATALs; this boundary is stricter).
 * The caller owns every host buffer it passes; the library owns all device
 * memory.  One context = one CUDA device + one stream; handles are not
 * thread-safe, distinct handles are independent.  There is NO CPU fallback:
 * without a usable sm_100 device, context creation fails with
 * B200PT_ERR_NO_DEVICE and nothing else can be called.
 */
#ifndef B200PT_H
#define B200PT_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200PT_VERSION 100

enum {
    B200PT_OK = 0,
    B200PT_ERR_INVALID = 1,   /* bad argument / malformed description */
    B200PT_ERR_CUDA = 2,      /* a CUDA call failed; see b200pt_last_error */
    B200PT_ERR_NO_DEVICE = 3, /* no CUDA device: there is no CPU path */
    B200PT_ERR_NOMEM = 4
};

/* where a ray / hit buffer lives */
enum { B200PT_MEM_HOST = 0, B200PT_MEM_DEVICE = 1 };

typedef struct b200pt_ctx b200pt_ctx;
typedef struct b200pt_scene b200pt_scene;

/* One triangle mesh in render space (util/mesh.h:19-44: p[], vertexIndices[]).
 * Its position in the meshes[] array is the reference's Triangle::meshIndex
 * (shapes.cpp:280).  object < 0: part of the world; otherwise the id of the
 * object definition (instanced geometry, scene.cpp:1527-1555) it belongs to. */
typedef struct b200pt_mesh {
    int32_t object;
    uint32_t n_vertices;
    uint32_t n_triangles;   /* number of primitives: triangles, or patches for a bilinear mesh */
    const float *p;         /* 3 * n_vertices */
    const int32_t *indices; /* 3 per triangle; 4 per bilinear patch: p00 p10 p01 p11 */
    /* 0 = triangle mesh; 1 = bilinear patch mesh (BilinearPatchMesh, util/mesh.h:47-77; BilinearPatch::Intersect
     * / IntersectP, shapes.cpp:1131-1156 over IntersectBilinearPatch, shapes.h:1279-1347).  Hits on a patch report
     * prim = the patch index and (b1, b2) = the patch's (u, v).  Both the traversal entry points and the render
     * path take both kinds.
     * | B200PT_MESH_ALPHA (triangle meshes only): every primitive of the mesh carries the constant alpha texture
     * `alpha` (GeometricPrimitive::Intersect / IntersectP, cpu/primitive.cpp:50-109): a hit counts when alpha >= 1, or
     * alpha > 0 and HashFloat(ray.o, ray.d) <= alpha (the ray as the primitive receives it: the instance-space ray
     * inside an object instance); otherwise the primitive is passed over.  `alpha` is read only under that flag. */
    uint32_t kind;
    float alpha;
} b200pt_mesh;
enum { B200PT_MESH_TRIANGLES = 0, B200PT_MESH_BILINEAR = 1, B200PT_MESH_ALPHA = 2 };

/* One object instance (TransformedPrimitive, cpu/primitive.cpp:112-131).
 * instanceId reported in hits = 1 + index in instances[] (scene.cpp:1560-1573). */
typedef struct b200pt_instance {
    uint32_t object;
    float render_from_instance[16]; /* row major 4x4: Transform::m    (util/transform.h) */
    float instance_from_render[16]; /* its inverse as the host holds it: Transform::mInv */
} b200pt_instance;

/* A binary BVH handed over by the host in the reference's own flat layout
 * (LinearBVHNode, cpu/aggregates.cpp:129-137: 32 B = bounds 24 B, i32
 * primitivesOffset | secondChildOffset, u16 nPrimitives, u8 axis, pad).
 * prims: 3 u32 per ordered primitive {kind, a, b}: kind 0 = triangle
 * (a = mesh, b = tri); kind 1 = instance (a = instanceId, 1-based); kind 2 = bilinear patch (a = mesh, b = patch). */
typedef struct b200pt_bvh2 {
    int32_t object; /* -1 = top level */
    uint32_t n_nodes;
    uint32_t n_prims;
    const void *nodes;
    const uint32_t *prims;
} b200pt_bvh2;

typedef struct b200pt_scene_desc {
    uint32_t n_meshes;
    const b200pt_mesh *meshes;
    uint32_t n_instances;
    const b200pt_instance *instances;
    /* n_bvhs == 0: the library builds the trees itself with the reference's
     * SAH decisions (top level maxPrimsInNode 4, object level 1).  Otherwise
     * one entry per BVH (top level + every object), e.g. read out of the
     * host's BVHAggregate objects. */
    uint32_t n_bvhs;
    const b200pt_bvh2 *bvhs;
    uint32_t flags; /* reserved, 0 */
} b200pt_scene_desc;

/* Closest-hit record.  miss: mesh == 0xFFFFFFFF.  (mesh, prim) =
 * Triangle{meshIndex, triIndex} (shapes.h:1188); instance = instanceId or 0;
 * t, b1, b2 = TriangleIntersection{t, b1, b2} (shapes.cpp:247-272), bit-exact. */
typedef struct b200pt_hit {
    uint32_t mesh;
    uint32_t prim;
    uint32_t instance;
    float t;
    float b1;
    float b2;
} b200pt_hit;

typedef struct b200pt_scene_stats {
    uint64_t n_triangles;   /* unique triangles resident */
    uint64_t n_instances;
    uint64_t n_bvh2_nodes;  /* all trees */
    uint64_t n_wide_nodes;  /* device wide-BVH nodes, all trees */
    uint64_t device_bytes;  /* HBM held by the scene */
    double build_seconds;   /* host SAH build + flatten */
    double upload_seconds;
} b200pt_scene_stats;

/* Per-call trace counters (the reference's "BVH/Nodes visited" cpu/aggregates.cpp:27,577
 * and "Ray-Triangle intersection tests" shapes.cpp:148), filled only when
 * B200PT_TRACE_COUNTERS is passed; the counting kernel variant is slower. */
typedef struct b200pt_trace_stats {
    uint64_t rays;
    uint64_t nodes_visited;
    uint64_t leaves_visited;
    uint64_t prims_tested;
    float kernel_ms; /* CUDA-event time of the traversal kernel(s) of the last call (B200PT_MEM_DEVICE); for a
                      * host-buffer call: first launch to last launch of the chunk pipeline, which includes waits on
                      * the copies in */
    /* 8-wide nodes fetched (256 B each; nodes_visited counts their 8 box tests each) and rays that went through
     * the exact binary kernel instead (non-finite 1/d or origin, at the top level or inside an instance) */
    uint64_t wide_nodes_visited;
    uint64_t exact_rays;
    /* leaves whose own box got the reference's exact test in the 8-wide walk (32 B each; the wide nodes' own eight
     * tests are a conservative filter, csrc/wide.h) */
    uint64_t leaf_boxes_tested;
} b200pt_trace_stats;

enum { B200PT_TRACE_COUNTERS = 1u << 8 };

int b200pt_version(void);
int b200pt_device_count(int *count);
int b200pt_context_create(int device, b200pt_ctx **out);
/* run on the caller's CUDA stream (cudaStream_t passed as void*; NULL = the legacy default
 * stream) instead of the context's own non-blocking stream */
int b200pt_context_set_stream(b200pt_ctx *ctx, void *cuda_stream);
int b200pt_context_sync(b200pt_ctx *ctx);
int b200pt_context_destroy(b200pt_ctx *ctx);
/* message of the last failure on this context (or of the last failed create when ctx == NULL) */
const char *b200pt_last_error(const b200pt_ctx *ctx);

int b200pt_scene_create(b200pt_ctx *ctx, const b200pt_scene_desc *desc, b200pt_scene **out);
int b200pt_scene_destroy(b200pt_scene *scene);
int b200pt_scene_bounds(const b200pt_scene *scene, float bounds_min_max[6]);
int b200pt_scene_get_stats(const b200pt_scene *scene, b200pt_scene_stats *out);

/* rays: n x 7 floats {o.xyz, d.xyz, tMax}; hits: n records; occluded: n bytes (0/1); n <= 2^31 - 2^22.
 * mem: B200PT_MEM_HOST (copied straight from / to the caller's memory in 2^21-ray chunks over three streams - the
 * copies overlap the traversal only when that memory is pinned; the call returns after the results are in `hits`)
 * or B200PT_MEM_DEVICE (asynchronous on the context's stream), optionally or-ed with B200PT_TRACE_COUNTERS. */
int b200pt_intersect_closest(b200pt_scene *scene, const float *rays, int64_t n,
                             b200pt_hit *hits, uint32_t mem);
int b200pt_intersect_any(b200pt_scene *scene, const float *rays, int64_t n,
                         uint8_t *occluded, uint32_t mem);
int b200pt_get_trace_stats(const b200pt_scene *scene, b200pt_trace_stats *out);

/* b200pt_intersect_closest for callers that may not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same kernel and same results as b200pt_intersect_closest with B200PT_MEM_DEVICE (rays_dev / hits_dev are device
 * buffers of the caller; asynchronous on the context's stream), but the kernel's scratch (ray-fetch counter + the list
 * of rays handed to the exact walk) is a static __device__ array of 4 * (65 + B200PT_NOALLOC_MAX_RAYS) bytes (~4 MB)
 * that every GPU loading the library holds until exit, and the call makes no cudaMalloc / cudaFree.
 * n > B200PT_NOALLOC_MAX_RAYS -> B200PT_ERR_INVALID (message in b200pt_last_error).  All calls on one device share the
 * array: calls that overlap on the GPU (other streams or contexts of the same device) race; calls queued on one
 * stream do not.  Trace counters / kernel_ms are not recorded for these calls. */
#define B200PT_NOALLOC_MAX_RAYS (1 << 20)
int b200pt_intersect_closest_noalloc(b200pt_scene *scene, const float *rays_dev, int64_t n, b200pt_hit *hits_dev);

/* ------------------------------------------------------------------ render path
 * b200pt_render_create replaces the object construction of RenderWavefront / the
 * WavefrontPathIntegrator constructor (wavefront/wavefront.cpp:14-70, wavefront/integrator.cpp:89-199):
 * the host (pbrt's parser + scene objects) hands over flat tables; b200pt_render_run replaces
 * WavefrontPathIntegrator::Render (wavefront/integrator.cpp:290-493) for a range of sample indices;
 * b200pt_film_* replace the film read-out before Film::WriteImage (film.cpp:542-574).
 * Image parity is judged against the CPU PathIntegrator (cpu/integrators.cpp:629-805), whose
 * sampler-dimension order the kernels follow. */

/* spectrum = 311 values at 395..705 nm, 1 nm apart.  interp 0: nearest (DenselySampledSpectrum,
 * util/spectrum.h:420-430); 1: linear between samples. */
#define B200PT_SPECTRUM_SAMPLES 311
typedef struct b200pt_spectrum {
    uint32_t interp;
    float values[B200PT_SPECTRUM_SAMPLES];
} b200pt_spectrum;

enum { B200PT_MAT_DIFFUSE = 0, B200PT_MAT_CONDUCTOR = 1, B200PT_MAT_DIELECTRIC = 2, B200PT_MAT_COATED_DIFFUSE = 3 };
typedef struct b200pt_material {
    uint32_t type;
    int32_t spectrum[4]; /* diffuse: [0] reflectance (materials.h:430-476); with params[4] != 0 the material is a
                          * DiffuseTransmissionMaterial (materials.h:869-927, DiffuseTransmissionBxDF bxdfs.h:218-297):
                          * [0] reflectance, [1] transmittance, params[5] = scale */
    float params[12];    /* dielectric: {uRoughness, vRoughness, remapRoughness, eta is a ConstantSpectrum, THIN};
                          * THIN != 0 = ThinDielectricMaterial (materials.h:215-262, ThinDielectricBxDF bxdfs.h:342-410):
                          * specular reflection + straight-through transmission, the roughness entries are ignored.
                          * coated diffuse: {interface uRoughness, vRoughness, remapRoughness, eta is a ConstantSpectrum,
                          * thickness, g, maxdepth, nsamples, COATED_CONDUCTOR, conductor uRoughness, vRoughness};
                          * spectrum = {reflectance, albedo (black), interface eta}.  COATED_CONDUCTOR != 0 =
                          * CoatedConductorMaterial (materials.h:695-790, materials.cpp:392-438; LayeredBxDF over a
                          * ConductorBxDF): spectrum = {conductor eta or -1, k, interface eta, reflectance when eta is -1} */
} b200pt_material;

/* DiffuseAreaLight on one triangle or one bilinear patch (lights.h:407-500; `tri` is then the patch index of a
 * B200PT_MESH_BILINEAR mesh, sampled as BilinearPatch::Sample, shapes.cpp:1158-1372) or a UniformInfiniteLight (lights.h:512-554:
 * spectrum + scale only; mesh/tri unused) + its PowerLightSampler alias-table bin
 * (lightsamplers.cpp:76-96, util/sampling.cpp:562-617).  Order = the host's `lights` vector. */
/* B200PT_LIGHT_POINT / SPOT / DISTANT: PointLight (lights.h:189-239), SpotLight (lights.h:747-801), DistantLight
 * (lights.h:242-297) - `spectrum` is I / Lemit, `scale` the light's scale; mesh / tri and the falloff fields are
 * unused, the placement comes from b200pt_render_desc::light_geometry. */
/* B200PT_LIGHT_IMAGE_INFINITE: ImageInfiniteLight (lights.h:557-643) - `spectrum` is the image colour space's illuminant,
 * `scale` the light's scale; image and sampling distribution come through b200pt_render_desc::env_*; one per scene. */
enum { B200PT_LIGHT_AREA_TRIANGLE = 0, B200PT_LIGHT_UNIFORM_INFINITE = 1, B200PT_LIGHT_POINT = 2, B200PT_LIGHT_SPOT = 3,
       B200PT_LIGHT_DISTANT = 4, B200PT_LIGHT_IMAGE_INFINITE = 5 };
typedef struct b200pt_light {
    uint32_t mesh, tri;
    int32_t spectrum; /* Lemit */
    float scale;
    uint32_t two_sided;
    float cos_falloff_end, tan_falloff_end, normalize_falloff_end;
    float alias_q, alias_p;
    int32_t alias;
    uint32_t type;
} b200pt_light;

typedef struct b200pt_mesh_attr {
    int32_t material;    /* index into materials, -1 = interface (no BSDF) */
    int32_t light_first; /* index of the light on triangle 0, -1 if not emissive */
    uint32_t flip_normal; /* reverseOrientation ^ transformSwapsHandedness (util/mesh.h:42) */
    /* TriangleMesh::n / ::uv (util/mesh.h:44-47) of a triangle mesh, one entry per triangle CORNER in index order
     * (3 x 3 floats, 3 x 2 floats per triangle), or NULL: shading normals and the (u,v) parameterisation of
     * Triangle::InteractionFromIntersection (shapes.h:884-1010).  A bilinear patch mesh takes corner_n with FOUR
     * entries per patch (4 x 3 floats, index order p00 p10 p01 p11: BilinearPatchMesh::n, shapes.h:1474-1494) and
     * corner_uv likewise (4 x 2 floats: BilinearPatchMesh::uv, the (s, t) that dpdu / dpdv follow, shapes.h:1412-1440). */
    const float *corner_n;
    const float *corner_uv;
    /* TriangleMesh::s, the per-vertex shading tangents (shapes.h:950-959), per corner like corner_n (3 x 3 floats per
     * triangle), or NULL: the interpolated tangent replaces dpdu before it is made orthogonal to the shading normal.
     * Triangle meshes only. */
    const float *corner_s;
} b200pt_mesh_attr;

/* One participating medium with a Henyey-Greenstein phase function and no emission.
 * type 0 = HomogeneousMedium (media.h:217-262): sigma_a / sigma_s spectra (already times the medium's scale,
 * as its DenselySampledSpectrum members hold them: interp 0 tables).
 * type 1 = GridMedium (media.h:265-345, media.cpp GridMedium ctor): density on an nx * ny * nz grid over
 * `bounds` in medium space (SampledGrid<Float>, x fastest), the ray is taken to medium space with
 * medium_from_render (= renderFromMedium's inverse as the host holds it), majorants = the host's MajorantGrid
 * (media.h:105-133): mres[3] voxels of max density, x fastest. */
enum { B200PT_MEDIUM_HOMOGENEOUS = 0, B200PT_MEDIUM_GRID = 1 };
typedef struct b200pt_medium {
    uint32_t type;
    int32_t sigma_a, sigma_s; /* spectrum indices */
    float g;
    float bounds[6];
    float medium_from_render[16];
    float render_from_medium[16];
    int32_t nx, ny, nz;
    const float *density;
    int32_t mres[3];
    const float *majorants;
    /* emission (HomogeneousMedium / GridMedium::SamplePoint, media.h:217-262, 288-318): emissive != 0 adds, at every
     * tentative collision, sigma_a * Le (cpu/integrators.cpp:990-1001).  le_spectrum = Le_spec (a homogeneous medium's
     * already times its Lescale / photometric normalisation, as the object holds it).  Grid media: le_scale = the LeScale
     * grid, le_nx x le_ny x le_nz values (1 x 1 x 1 when the scene gave no "Lescale" array; trilinear with zeros outside,
     * like the density).  temperature != NULL: a grid of the density's resolution; where (T - temperature_offset) *
     * temperature_scale > 100 the emission is le_scale times the normalised blackbody spectrum of that temperature
     * instead of Le_spec (media.h:306-314; le_spectrum is then unused but must be a valid index). */
    uint32_t emissive;
    int32_t le_spectrum;
    int32_t le_nx, le_ny, le_nz;
    const float *le_scale;
    const float *temperature;
    float temperature_scale, temperature_offset;
} b200pt_medium;

typedef struct b200pt_render_desc {
    /* camera: 0 = perspective (cameras.cpp:433-456), 1 = realistic (fields at the end of this struct) */
    uint32_t camera_type;
    float camera_from_raster[16];
    float render_from_camera[16];
    float lens_radius, focal_distance;
    /* film: 0 = RGBFilm (film.h:232-310), 1 = SpectralFilm (film.h:430-560; n_buckets below), 2 = GBufferFilm, the
     * ISET spectral film (film.h:322-408, film.cpp:597-666: 31 radiance buckets over 395..705 nm + AOVs; fields at
     * the end of this struct); box filter */
    uint32_t film_type;
    int32_t pixel_bounds[4]; /* x0 y0 x1 y1 */
    int32_t full_resolution[2];
    float filter_radius[2];
    float imaging_ratio, max_component_value;
    float output_rgb_from_sensor_rgb[9];
    int32_t sensor_spectrum[3]; /* r_bar g_bar b_bar */
    /* sampler: B200PT_SAMPLER_ZSOBOL (samplers.h:224-370) or B200PT_SAMPLER_INDEPENDENT (samplers.h:442-476: one PCG32
     * stream per pixel; `randomize` is unused) */
    uint32_t sampler_type;
    int32_t samples_per_pixel, seed;
    uint32_t randomize; /* 0 none 1 permutedigits 2 fastowen 3 owen */
    /* integrator: path (cpu/integrators.cpp:629-805) */
    int32_t max_depth;
    uint32_t regularize;
    uint32_t light_sampler; /* 0 uniform, 1 power, 2 bvh */
    uint32_t n_spectra, n_materials, n_lights, n_meshes;
    const b200pt_spectrum *spectra;
    const b200pt_material *materials;
    const b200pt_light *lights;
    const b200pt_mesh_attr *meshes; /* one per mesh of the geometry scene */
    /* BVHLightSampler (lightsamplers.h:266-375), built by the host (lightsamplers.cpp:109-232).
     * light_bvh_nodes: n_light_bvh_nodes x 32 B = eight u32 {w.x | w.y << 16 (OctahedralVector), phi,
     * qCosTheta_o | qCosTheta_e << 15 | twoSided << 30, qb[0][0] | qb[0][1] << 16, qb[0][2] | qb[1][0] << 16,
     * qb[1][1] | qb[1][2] << 16, childOrLightIndex | isLeaf << 31, 0} (LightBVHNode, lightsamplers.h:203-264);
     * light_bvh_bounds = allLightBounds; light_bit_trails[n_lights] = lightToBitTrail, 0xFFFFFFFF for a
     * light that is not in the tree.  Needed only when light_sampler == 2. */
    uint32_t n_light_bvh_nodes;
    float light_bvh_bounds[6];
    const void *light_bvh_nodes;
    const uint32_t *light_bit_trails;
    /* SpectralFilm (film_type 1): n_buckets equal bands over [lambda_min, lambda_max]; the wavelengths of a
     * sample are drawn uniformly over the same range (film.h:443-445), which has to be the build's 395..705 nm
     * (the ISET 31-band setting); 1 <= n_buckets <= 64. */
    uint32_t n_buckets;
    float lambda_min, lambda_max;
    /* camera_type 1 = RealisticCamera (cameras.cpp:820-884 TraceLensesFromFilm, :968-1019 SampleExitPupil /
     * GenerateRay).  The host keeps the constructor's work (cameras.cpp:770-817: lens file, FocusThickLens,
     * BoundExitPupil) and hands over: lens_elements = n x {curvatureRadius, thickness, eta, apertureRadius} in
     * metres, front element first (RealisticCamera::elementInterfaces; curvatureRadius 0 = the aperture stop);
     * exit_pupil_bounds = n x {pMin.x, pMin.y, pMax.x, pMax.y} (exitPupilBounds); physical_extent likewise;
     * film_diagonal = Film::Diagonal() in metres.  Aperture images are not supported (circular stops).
     * camera_from_raster, lens_radius and focal_distance are ignored for this camera. */
    uint32_t n_lens_elements;
    const float *lens_elements;
    uint32_t n_exit_pupil_bounds;
    const float *exit_pupil_bounds;
    float physical_extent[4];
    float film_diagonal;
    /* integrator_type 1 = VolPathIntegrator (cpu/integrators.cpp:953-1405; wavefront/media.cpp:22-354): participating
     * media.  media[] are the scene's media; mesh_media[2 * m + {0, 1}] = {inside, outside} medium of mesh m's
     * MediumInterface (cpu/primitive.cpp:50-109, interaction.h GetMedium), -1 = vacuum; a mesh whose two entries
     * are equal is no medium transition (the ray keeps its medium).  NULL = no mesh has an interface.
     * camera_medium = the medium camera rays start in (CameraBase::medium), -1 = none.
     * integrator_type 0 = PathIntegrator ignores all of this. */
    uint32_t integrator_type;
    uint32_t n_media;
    const struct b200pt_medium *media;
    const int32_t *mesh_media;
    int32_t camera_medium;
    /* camera_type 2 = the ISET OmniCamera without a microlens array (cameras.cpp:2823-2989 TraceElement /
     * TraceLensesFromFilm, :3118-3138 SampleExitPupil, :3329-3373 GenerateRay): the camera_type 1 tables (from
     * OmniCamera's own constructor) plus lens_element_transforms = n x {transform[16], inverse[16]} row major, the
     * LensElementInterface::transform of each interface (cameras.h:856-878; decentred / tilted elements), NULL =
     * all identity. */
    const float *lens_element_transforms;
    /* camera_type 2 with a microlens array in front of the sensor (OmniCamera::microlens, cameras.h:940-960;
     * TraceFullLensSystemFromFilm cameras.cpp:3259-3327, SampleMicrolensPupil :3167-3187,
     * MicrolensElementFromIndex :3221-3244): n_microlens_elements > 0 switches it on.  microlens_elements /
     * microlens_element_transforms as for the main lens; microlens_dims = lenses across the film in x and y;
     * microlens_offsets = dims[0] * dims[1] x {dx, dy} (metres, row y * dims[0] + x); microlens_offset_from_sensor
     * and microlens_simulation_radius as the OmniCamera object holds them (microlens.simulationRadius). */
    uint32_t n_microlens_elements;
    const float *microlens_elements;
    const float *microlens_element_transforms;
    int32_t microlens_dims[2];
    const float *microlens_offsets;
    float microlens_offset_from_sensor;
    int32_t microlens_simulation_radius;
    /* film_type 2 = GBufferFilm.  Radiance buckets: n_buckets = 31 over [395, 705] nm is hard-coded in the reference
     * (film.h:383-389) and required here.  gbuffer_output_from_render / _inv: the transform (m, mInv, row major) that
     * takes the first visible surface from render space to the AOVs' output space - for the default "camera"
     * coordinate system that is the INVERSE of GBufferFilm::outputFromRender (applyInverse, film.cpp:1088-1090), i.e.
     * CameraFromRender.  material_ids[n_materials]: Material::materialId of each material (materials.cpp:738), NULL =
     * 1 + index.  Not built: Albedo.*, dzdx / dzdy (they need bsdf.rho and camera-ray differentials) and the SVD basis. */
    float gbuffer_output_from_render[16];
    float gbuffer_output_from_render_inv[16];
    const int32_t *material_ids;
    /* the film's pixel filter (filters.h).  B200PT_FILTER_BOX (BoxFilter, filters.h:37-77, radius = filter_radius):
     * uniform sample offsets, weight 1.  The filters the reference samples through a FilterSampler (Gaussian, Mitchell,
     * LanczosSinc; filters.h:80-230, filters.cpp:92-118) need filter_table: the host's tabulation of Filter::Evaluate
     * at the centres of filter_table_size[0] x filter_table_size[1] cells over [-radius, radius]^2, row-major with x
     * fastest (FilterSampler::f: int(32 * radius) cells per axis) - the sampling distribution is built from it as
     * the reference builds its PiecewiseConstant2D.  B200PT_FILTER_TRIANGLE (TriangleFilter, filters.h:204-239) samples its
     * tent analytically (SampleTent), weight 1, and needs no table. */
    uint32_t filter_type;
    /* placement of the delta lights, B200PT_LIGHT_GEOMETRY_FLOATS floats per light for ALL n_lights lights (rows of
     * other light types are ignored); NULL when the scene has none.  Row: [0..2] point / spot: renderFromLight(0,0,0)
     * (lights.h:224, 787); distant: Normalize(renderFromLight(0,0,1)) (lights.h:285); [3] distant: sceneRadius
     * (DistantLight::Preprocess, lights.h:277-279); [4..12] spot: upper 3 x 3 of renderFromLight's inverse, row major
     * (lights.h:790); [13] spot cosFalloffStart, [14] spot cosFalloffEnd (lights.cpp:1376-1385).
     * Image infinite light: [0..8] upper 3 x 3 of renderFromLight, [9..17] of its inverse (both row major), [18]
     * sceneRadius (ImageInfiniteLight::Preprocess). */
    const float *light_geometry;
    const float *filter_table; /* see filter_type; NULL for the box filter */
    int32_t filter_table_size[2];
    /* the B200PT_LIGHT_IMAGE_INFINITE light's equal-area square image, env_resolution pixels a side, row-major:
     * env_image = per pixel the RGBIlluminantSpectrum of its (clamped) RGB - {c0, c1, c2} of the RGBSigmoidPolynomial
     * and the spectrum's scale (util/spectrum.cpp RGBIlluminantSpectrum, util/color.h) - so the library needs no
     * RGB-to-spectrum table; env_distribution = the function of the light's compensatedDistribution
     * (lights.cpp:1064-1070: max(luminance - average, 0), or 1 everywhere), from which the library builds the
     * PiecewiseConstant2D as the reference does.  NULL / 0 without such a light. */
    const float *env_image;
    const float *env_distribution;
    int32_t env_resolution;
    /* GBufferFilm's Albedo.R G B (film.cpp:640-644, 846-858): with gbuffer_albedo != 0 the first visible surface's
     * albedo is estimated as the reference does (bsdf.rho over its 16 fixed sample points, cpu/integrators.cpp:690-711)
     * and summed as (albedo * illuminant).ToRGB(lambda, colorSpace): gbuffer_xyz_spectrum = Spectra::X / Y / Z,
     * gbuffer_illuminant = the film colour space's illuminant (both indices into `spectra`), gbuffer_rgb_from_xyz =
     * its RGBFromXYZ, row major.  Costs 16 BSDF samples per camera ray; leave 0 when the film does not save albedo. */
    uint32_t gbuffer_albedo;
    int32_t gbuffer_xyz_spectrum[3];
    int32_t gbuffer_illuminant;
    float gbuffer_rgb_from_xyz[9];
} b200pt_render_desc;
#define B200PT_LIGHT_GEOMETRY_FLOATS 24
enum { B200PT_SAMPLER_ZSOBOL = 0, B200PT_SAMPLER_INDEPENDENT = 1 };
enum { B200PT_FILTER_BOX = 0, B200PT_FILTER_GAUSSIAN = 1, B200PT_FILTER_MITCHELL = 2, B200PT_FILTER_SINC = 3, B200PT_FILTER_TRIANGLE = 4 };

typedef struct b200pt_render b200pt_render;

typedef struct b200pt_render_stats {
    uint64_t samples;        /* pixel samples completed by the last run */
    uint64_t closest_rays, shadow_rays;
    uint32_t kernel_launches;
    float gpu_ms;            /* CUDA-event time of the last run */
    /* volpath only: paths still alive when a pass stopped after its 96 wavefront iterations, and light-sample walks
     * still crossing interfaces after 64 segments - both are dropped there (the reference's loops are bounded by
     * maxdepth and roulette only); non-zero counts mean the image is darker than the reference's */
    uint64_t vol_paths_cut, vol_walks_cut;
} b200pt_render_stats;

int b200pt_render_create(b200pt_scene *scene, const b200pt_render_desc *desc, b200pt_render **out);
int b200pt_render_destroy(b200pt_render *r);
/* Renders sample indices first, first+stride, ... (n_samples of them) of every pixel and adds them to
 * the device film; asynchronous on the context's stream.  GPU g of G calls (g, spp/G, G): the union
 * over GPUs is exactly the single-GPU sample set. */
int b200pt_render_run(b200pt_render *r, int32_t first_sample, int32_t n_samples, int32_t sample_stride);
int b200pt_film_clear(b200pt_render *r);
/* film accumulators {rgbSum[3], weightSum} as 4 doubles per pixel, row major over pixel_bounds
 * (RGBFilm::Pixel, film.h:305-310).  A spectral film appends, after those 4 * n_pixels doubles, n_buckets
 * planes of bucketSums and n_buckets planes of weightSums, n_pixels doubles each (SpectralFilm::Pixel,
 * film.h:547-556): n_doubles = n_pixels * (4 + 2 * n_buckets).  device_ptr: for an in-place NCCL reduction
 * by the host; film_read copies all n_doubles. */
/* (GBufferFilm: n_pixels x {4 RGB doubles}, then planes of n_pixels doubles: 31 radiance buckets, 17 AOV sums - a
 * reduction that only needs R G B + Radiance can stop after the first 4 + 31 planes) */
int b200pt_film_device_ptr(b200pt_render *r, void **ptr, uint64_t *n_doubles);
int b200pt_film_read(b200pt_render *r, double *rgb_weight);
/* Multi-GPU hand-back without the host touching the buffer: ncclReduce (double, sum, in place) of the first
 * n_doubles doubles of this rank's film (0 = all of it) to rank `root` of `nccl_comm` - an ncclComm_t the host made,
 * one rank per render object - on the context's stream; afterwards root's film holds the frame (every rank rendered
 * its stripe of sample indices, b200pt_render_run).  The NCCL the process has already loaded is used (dlopen; the
 * library does not link it).  The reference has no counterpart: its film lives in one address space. */
int b200pt_film_reduce(b200pt_render *r, void *nccl_comm, int32_t root, uint64_t n_doubles);
/* spectral film only: the S0.* channels of SpectralFilm::GetImage (film.cpp:1221-1273), n_buckets floats per
 * pixel (bucketSums / weightSums, 0 where no weight). */
int b200pt_film_spectral_image(b200pt_render *r, float *buckets);
/* final image, 3 floats per pixel: outputRGBFromSensorRGB * (rgbSum / weightSum) (film.h:261-277) */
int b200pt_film_image(b200pt_render *r, float *rgb);
/* GBufferFilm::GetImage (film.cpp:743-1044) for film_type 2: R, G, B and, in the reference's channel order, the groups
 * selected in `channels` (the film's save* switches, film.cpp:1062-1073): Radiance.C01-31 (bucket sum / sum of all sample
 * weights: the CPU build's observed behaviour), Albedo.R-B (rgbAlbedoSum / rgbWeightSum), P.X-Z, MaterialId, InstanceId, Ns.X-Z, N.X-Z,
 * Variance.R-B, RelativeVariance.R-B.  out: height x width x (*n_channels) floats, channels last; out == NULL only
 * reports *n_channels.  Under multi-GPU rendering the sums reduce with the film buffer; the two id channels are this
 * rank's (the reference keeps the last sample's). */
enum {
    B200PT_GBUFFER_RADIANCE = 1, B200PT_GBUFFER_POSITION = 2, B200PT_GBUFFER_MATERIAL = 4, B200PT_GBUFFER_INSTANCE = 8,
    B200PT_GBUFFER_NS = 16, B200PT_GBUFFER_NORMAL = 32, B200PT_GBUFFER_VARIANCE = 64, B200PT_GBUFFER_RELATIVE_VARIANCE = 128,
    B200PT_GBUFFER_ALBEDO = 256 /* Albedo.R G B, between Radiance and P in the channel order; needs gbuffer_albedo */
};
int b200pt_film_gbuffer_image(b200pt_render *r, uint32_t channels, float *out, int32_t *n_channels);
/* Film::WriteImage for the .exr case (film.cpp:542-574 -> Image::WriteEXR, util/image.cpp:1008+): the film's image
 * finalised and packed ON THE DEVICE into the chunks of an uncompressed scanline OpenEXR file, then written to `path`.
 * RGBFilm: R G B; GBufferFilm: R G B + the groups of `channels` under the reference's channel names ("Radiance.C01" ...,
 * "P.X", "MaterialId", "Ns.X", ..., film.cpp:752-766); SpectralFilm: R G B + "S0.<lambda>nm" as the reference names them.
 * fp16 != 0 writes HALF channels (values clamped to 65504 as the reference does), else FLOAT.  dataWindow =
 * pixel_bounds, displayWindow = full_resolution.  No compression, no metadata beyond the required attributes. */
int b200pt_film_write_exr(b200pt_render *r, uint32_t channels, const char *path, int32_t fp16);
int b200pt_render_get_stats(const b200pt_render *r, b200pt_render_stats *out);

/* Host-only: build the binary SAH BVH for one primitive list exactly as the
 * reference does (cpu/aggregates.cpp:192-387, 505-522) and return it in the
 * LinearBVHNode layout.  bounds: n x 6 floats (pMin, pMax).  nodes_out must
 * hold 2n-1 nodes of 32 B; order_out n indices into the input list (the
 * reference's ordered `primitives`, leaves in depth-first order).  Needs no GPU. */
int b200pt_bvh2_build(const float *bounds, uint32_t n, int max_prims_in_node,
                      void *nodes_out, uint32_t *n_nodes_out, uint32_t *order_out);

/* Host-only: a PLY mesh file as the reference's `Shape "plymesh"` reads it (TriQuadMesh::ReadPLY,
 * util/mesh.cpp:322-423): vertex x y z, optional nx ny nz, optional (u, v) named u v / s t / texture_u texture_v /
 * texture_s texture_t, faces of three or four vertices from the list property `vertex_indices` (other lengths are
 * passed over), optional face_indices; ascii, binary_little_endian, binary_big_endian (gzip refused).  A quad comes
 * out in bilinear-patch order p00 p10 p01 p11 (mesh.cpp:299-305), ready for a B200PT_MESH_BILINEAR mesh;
 * whether quads are split into triangles instead (TriQuadMesh::ConvertToOnlyTriangles, mesh.cpp:425-442) is the host's choice.
 * The arrays are malloc'ed; b200pt_ply_free releases them.  On failure: B200PT_ERR_INVALID and, if err != NULL, the
 * reason in err (the reference's message where it has one).  Needs no GPU. */
typedef struct b200pt_ply_mesh {
    uint64_t n_vertices, n_triangles, n_quads, n_face_indices;
    float *p;              /* 3 per vertex */
    float *n;              /* 3 per vertex, or NULL */
    float *uv;             /* 2 per vertex, or NULL */
    int32_t *tri_indices;  /* 3 per triangle, or NULL */
    int32_t *quad_indices; /* 4 per quad, or NULL */
    int32_t *face_indices; /* n_face_indices entries, or NULL */
} b200pt_ply_mesh;
int b200pt_ply_read(const char *path, b200pt_ply_mesh *out, char *err, uint64_t err_len);
void b200pt_ply_free(b200pt_ply_mesh *m);

#ifdef __cplusplus
}
#endif
#endif /* B200PT_H */
