// oracle/pbrt_b200.cpp - the reference-side binding, compiled: pbrt's own parser / scene objects / film writer with
// libb200pt.so behind them.  TEST INFRASTRUCTURE (it links the reference, oracle/_ref/libpbrt_ref.a) - this is the
// code a pbrt maintainer would add next to wavefront/wavefront.cpp and wavefront/aggregate.cpp.
//
//   pbrt_b200 [--spp n] [--nthreads n] [--outfile f] scene.pbrt
//       Level 2, the whole frame: RenderB200(BasicScene&) stands where RenderWavefront stands
//       (wavefront/wavefront.cpp:14-70, selected at cmd/pbrt.cpp:296-299): builds the host objects as
//       WavefrontPathIntegrator's constructor does (wavefront/integrator.cpp:89-199; the SAH BVHAggregate build
//       stays on the host), flattens them (flatten_ref.h), hands the tables over through the C ABI
//       (b200pt_scene_create, b200pt_render_create, b200pt_render_run), copies the device film accumulators back
//       into the reference's own Film object (b200pt_film_read -> RGBFilm::Pixel / SpectralFilm::Pixel) and lets
//       the reference write the image (Film::WriteImage, film.cpp:542-574; wavefront/wavefront.cpp:64-69).
//   pbrt_b200 --wavefront [...] scene.pbrt
//       Level 1, the intersector plug: the reference's own WavefrontPathIntegrator (its CPU kernels for camera rays,
//       shading, lights, film) with B200Aggregate : WavefrontAggregate (wavefront/integrator.h:32-54) set in place
//       of the CPUAggregate it constructs at wavefront/integrator.cpp:157-169.  IntersectClosest / IntersectShadow
//       run on the GPU through b200pt_intersect_closest / b200pt_intersect_any; the SurfaceInteraction of a hit is
//       rebuilt by the reference's own Primitive::Intersect on the ONE primitive the GPU names.
#include "flatten_ref.h"

#define private public
#define protected public
#include <pbrt/wavefront/aggregate.h>
#include <pbrt/wavefront/integrator.h>
#include <pbrt/wavefront/intersect.h>
#undef private
#undef protected
#include <pbrt/util/args.h>
#include <pbrt/util/parallel.h>

#include <chrono>
#include <unordered_map>

#include "../include/b200pt.h"

// ------------------------------------------------------------------ named tables in memory (the .b2rs layout)
struct Table {
    uint32_t dtype = 0;  // 0 f32, 1 i32, 2 u32
    std::vector<uint32_t> dims;
    std::vector<uint32_t> words;
    size_t rows() const { return dims.empty() ? 0 : dims[0]; }
    size_t cols() const { return dims.size() < 2 ? 1 : dims[1]; }
    const float *f() const { return reinterpret_cast<const float *>(words.data()); }
    const int32_t *i() const { return reinterpret_cast<const int32_t *>(words.data()); }
    const uint32_t *u() const { return words.data(); }
};
struct MemTables {
    std::map<std::string, Table> t;
    void put(const std::string &name, uint32_t dtype, std::vector<uint32_t> dims, const void *data) {
        Table &tb = t[name];
        tb.dtype = dtype;
        tb.dims = dims;
        size_t n = 1;
        for (uint32_t d : dims) n *= d;
        tb.words.resize(n);
        if (n) memcpy(tb.words.data(), data, n * 4);
    }
    void f32(const std::string &name, std::vector<uint32_t> dims, const std::vector<float> &v) { put(name, 0, dims, v.data()); }
    void i32(const std::string &name, std::vector<uint32_t> dims, const std::vector<int32_t> &v) { put(name, 1, dims, v.data()); }
    void u32(const std::string &name, std::vector<uint32_t> dims, const std::vector<uint32_t> &v) { put(name, 2, dims, v.data()); }
    bool has(const std::string &n) const { return t.count(n) != 0; }
    const Table &at(const std::string &n) const {
        auto it = t.find(n);
        if (it == t.end()) die("table missing: " + n);
        return it->second;
    }
};

static void check(int rc, b200pt_ctx *ctx, const char *what) {
    if (rc != B200PT_OK) die(std::string(what) + ": " + b200pt_last_error(ctx));
}

// ------------------------------------------------------------------ tables -> include/b200pt.h descriptions
struct Marshalled {
    std::vector<b200pt_mesh> meshes;
    std::vector<b200pt_instance> instances;
    std::vector<b200pt_bvh2> bvhs;
    b200pt_scene_desc scene{};
    std::vector<b200pt_spectrum> spectra;
    std::vector<b200pt_material> materials;
    std::vector<b200pt_light> lights;
    std::vector<b200pt_mesh_attr> attrs;
    std::vector<b200pt_medium> media;
    b200pt_render_desc render{};
};

static void MarshalGeometry(const MemTables &T, Marshalled &m, bool hostBvh) {
    const Table &nv = T.at("mesh_nverts"), &nt = T.at("mesh_ntris");
    const size_t nMesh = nv.rows();
    const uint32_t *kind = T.has("mesh_kind") ? T.at("mesh_kind").u() : nullptr;
    const int32_t *obj = T.at("mesh_object").i();
    const float *alpha = T.has("mesh_alpha") ? T.at("mesh_alpha").f() : nullptr;
    const float *P = T.at("P").f();
    const int32_t *idx = T.at("indices").i();
    const int32_t *pidx = T.has("patch_indices") ? T.at("patch_indices").i() : nullptr;
    size_t vo = 0, to = 0, po = 0;
    for (size_t i = 0; i < nMesh; ++i) {
        b200pt_mesh me{};
        me.object = obj[i];
        me.n_vertices = nv.u()[i];
        me.n_triangles = nt.u()[i];
        me.p = P + 3 * vo;
        const bool patch = kind && kind[i] == 1;
        me.indices = patch ? pidx + 4 * po : idx + 3 * to;
        me.kind = patch ? B200PT_MESH_BILINEAR : B200PT_MESH_TRIANGLES;
        me.alpha = 1.f;
        if (alpha && alpha[i] >= 0) {
            me.kind |= B200PT_MESH_ALPHA;
            me.alpha = alpha[i];
        }
        vo += me.n_vertices;
        (patch ? po : to) += me.n_triangles;
        m.meshes.push_back(me);
    }
    if (T.has("instance_object")) {
        const Table &io = T.at("instance_object");
        const float *im = T.at("instance_m").f(), *imi = T.at("instance_minv").f();
        for (size_t k = 0; k < io.rows(); ++k) {
            b200pt_instance in{};
            in.object = uint32_t(io.i()[k]);
            memcpy(in.render_from_instance, im + 16 * k, 64);
            memcpy(in.instance_from_render, imi + 16 * k, 64);
            m.instances.push_back(in);
        }
    }
    if (hostBvh && T.has("bvh_nodes")) {  // the host's own BVHAggregate trees (LinearBVHNode layout)
        const Table &bn = T.at("bvh_nodes"), &bp = T.at("bvh_prims");
        m.bvhs.push_back(b200pt_bvh2{-1, uint32_t(bn.rows()), uint32_t(bp.rows()), bn.u(), bp.u()});
        for (int k = 0; T.has("obj_bvh_nodes_" + std::to_string(k)); ++k) {
            const Table &on = T.at("obj_bvh_nodes_" + std::to_string(k)), &op = T.at("obj_bvh_prims_" + std::to_string(k));
            m.bvhs.push_back(b200pt_bvh2{k, uint32_t(on.rows()), uint32_t(op.rows()), on.u(), op.u()});
        }
    }
    m.scene.n_meshes = uint32_t(m.meshes.size());
    m.scene.meshes = m.meshes.data();
    m.scene.n_instances = uint32_t(m.instances.size());
    m.scene.instances = m.instances.data();
    m.scene.n_bvhs = uint32_t(m.bvhs.size());
    m.scene.bvhs = m.bvhs.data();
}

static void MarshalRender(const MemTables &T, Marshalled &m) {
    b200pt_render_desc &r = m.render;
    r.camera_type = T.at("camera_type").u()[0];
    memcpy(r.render_from_camera, T.at("render_from_camera").f(), 64);
    if (r.camera_type == 0) {
        memcpy(r.camera_from_raster, T.at("camera_from_raster").f(), 64);
        r.lens_radius = T.at("camera_params").f()[0];
        r.focal_distance = T.at("camera_params").f()[1];
    } else {
        const Table &lens = T.at("camera_lens"), &pupil = T.at("camera_exit_pupil");
        r.n_lens_elements = uint32_t(lens.rows());
        r.lens_elements = lens.f();
        r.n_exit_pupil_bounds = uint32_t(pupil.rows());
        r.exit_pupil_bounds = pupil.f();
        memcpy(r.physical_extent, T.at("camera_extent").f(), 16);
        r.film_diagonal = T.at("camera_extent").f()[4];
        if (r.camera_type == 2 && T.has("camera_lens_transforms")) r.lens_element_transforms = T.at("camera_lens_transforms").f();
        if (r.camera_type == 2 && T.has("microlens_lens")) {
            const int32_t *mi = T.at("microlens_ints").i();
            r.n_microlens_elements = uint32_t(T.at("microlens_lens").rows());
            r.microlens_elements = T.at("microlens_lens").f();
            r.microlens_element_transforms = T.at("microlens_transforms").f();
            r.microlens_offsets = T.at("microlens_offsets").f();
            r.microlens_dims[0] = mi[0];
            r.microlens_dims[1] = mi[1];
            r.microlens_simulation_radius = mi[2];
            r.microlens_offset_from_sensor = T.at("microlens_params").f()[0];
        }
    }
    r.film_type = T.at("film_type").u()[0];
    if (r.film_type == 2) {  // GBufferFilm: the AOVs' output space and the host's material ids
        memcpy(r.gbuffer_output_from_render, T.at("gbuffer_from_render").f(), 64);
        memcpy(r.gbuffer_output_from_render_inv, T.at("gbuffer_from_render_inv").f(), 64);
        r.material_ids = T.at("material_ids").i();
        if (T.has("gbuffer_albedo")) {
            const int32_t *ga = T.at("gbuffer_albedo").i();
            r.gbuffer_albedo = 1;
            memcpy(r.gbuffer_xyz_spectrum, ga, 12);
            r.gbuffer_illuminant = ga[3];
            memcpy(r.gbuffer_rgb_from_xyz, T.at("gbuffer_rgb_from_xyz").f(), 36);
        }
    }
    if (r.film_type == 1 || r.film_type == 2) {
        const float *fs = T.at("film_spectral").f();
        r.n_buckets = uint32_t(fs[0]);
        r.lambda_min = fs[1];
        r.lambda_max = fs[2];
    }
    const int32_t *fi = T.at("film_ints").i();
    memcpy(r.pixel_bounds, fi, 16);
    memcpy(r.full_resolution, fi + 4, 8);
    const float *fp = T.at("film_params").f();
    r.filter_type = T.has("filter_type") ? T.at("filter_type").u()[0] : 0;
    if (T.has("filter_table")) {
        const Table &ft = T.at("filter_table");
        r.filter_table = ft.f();
        r.filter_table_size[0] = int32_t(ft.cols());
        r.filter_table_size[1] = int32_t(ft.rows());
    }
    r.filter_radius[0] = fp[0];
    r.filter_radius[1] = fp[1];
    r.imaging_ratio = fp[2];
    r.max_component_value = fp[3];
    memcpy(r.output_rgb_from_sensor_rgb, T.at("film_matrix").f(), 36);
    memcpy(r.sensor_spectrum, T.at("film_sensor").i(), 12);
    const int32_t *s = T.at("sampler").i();
    r.sampler_type = uint32_t(s[0]);
    r.samples_per_pixel = s[1];
    r.seed = s[2];
    r.randomize = uint32_t(s[3]);
    const int32_t *ig = T.at("integrator").i();
    r.max_depth = ig[0];
    r.regularize = uint32_t(ig[1]);
    r.light_sampler = uint32_t(ig[2]);
    r.integrator_type = T.at("integrator_type").u()[0];
    const Table &sv = T.at("spectra_values");
    m.spectra.resize(std::max<size_t>(1, sv.rows()));
    for (size_t i = 0; i < sv.rows(); ++i) {
        m.spectra[i].interp = T.at("spectra_interp").u()[i];
        memcpy(m.spectra[i].values, sv.f() + 311 * i, 311 * 4);
    }
    const Table &mi = T.at("materials_i"), &mf = T.at("materials_f");
    m.materials.resize(std::max<size_t>(1, mi.rows()));
    for (size_t i = 0; i < mi.rows(); ++i) {
        const int32_t *rec = mi.i() + 5 * i;
        m.materials[i].type = rec[0] >= 0 ? uint32_t(rec[0]) : 0u;
        memcpy(m.materials[i].spectrum, rec + 1, 16);
        memset(m.materials[i].params, 0, sizeof(m.materials[i].params));
        memcpy(m.materials[i].params, mf.f() + mf.cols() * i, 4 * std::min<size_t>(mf.cols(), 12));
    }
    const Table &li = T.at("lights_i"), &lf = T.at("lights_f");
    m.lights.resize(std::max<size_t>(1, li.rows()));
    for (size_t i = 0; i < li.rows(); ++i) {
        const int32_t *a = li.i() + 7 * i;
        const float *b = lf.f() + 6 * i;
        b200pt_light &L = m.lights[i];
        L.mesh = uint32_t(std::max(a[0], 0));
        L.tri = uint32_t(std::max(a[1], 0));
        L.spectrum = a[2];
        L.scale = b[0];
        L.two_sided = uint32_t(a[3]);
        L.cos_falloff_end = b[1];
        L.tan_falloff_end = b[2];
        L.normalize_falloff_end = b[3];
        L.alias_q = b[4];
        L.alias_p = b[5];
        L.alias = a[4];
        L.type = uint32_t(a[6]);
    }
    if (T.has("lights_geom")) r.light_geometry = T.at("lights_geom").f();  // B200PT_LIGHT_GEOMETRY_FLOATS columns
    if (T.has("env_image")) {
        r.env_image = T.at("env_image").f();
        r.env_distribution = T.at("env_distribution").f();
        r.env_resolution = int32_t(T.at("env_distribution").rows());
    }
    const Table &ma = T.at("mesh_attr");
    const size_t nMesh = ma.rows();
    m.attrs.resize(std::max<size_t>(1, nMesh));
    const uint32_t *hasN = T.has("mesh_has_n") ? T.at("mesh_has_n").u() : nullptr;
    const uint32_t *hasUV = T.has("mesh_has_uv") ? T.at("mesh_has_uv").u() : nullptr;
    const uint32_t *hasS = T.has("mesh_has_s") ? T.at("mesh_has_s").u() : nullptr;
    const uint32_t *kind = T.has("mesh_kind") ? T.at("mesh_kind").u() : nullptr;
    const uint32_t *ntris = T.at("mesh_ntris").u();
    size_t nAt = 0, uvAt = 0, sAt = 0;
    for (size_t i = 0; i < nMesh; ++i) {
        const int32_t *a = ma.i() + 3 * i;
        b200pt_mesh_attr &A = m.attrs[i];
        A.material = a[0];
        A.light_first = a[1];
        A.flip_normal = uint32_t(std::max(a[2], 0));
        A.corner_n = A.corner_uv = A.corner_s = nullptr;
        if (hasN && hasN[i]) {
            const size_t per = (kind && kind[i] == 1) ? 4 : 3;
            A.corner_n = T.at("corner_n").f() + 3 * nAt;
            nAt += per * ntris[i];
        }
        if (hasUV && hasUV[i]) {
            A.corner_uv = T.at("corner_uv").f() + 2 * uvAt;
            uvAt += ((kind && kind[i] == 1) ? 4 : 3) * size_t(ntris[i]);
        }
        if (hasS && hasS[i]) {
            A.corner_s = T.at("corner_s").f() + 3 * sAt;
            sAt += 3 * size_t(ntris[i]);
        }
    }
    r.n_spectra = uint32_t(sv.rows());
    r.n_materials = uint32_t(mi.rows());
    r.n_lights = uint32_t(li.rows());
    r.n_meshes = uint32_t(nMesh);
    r.spectra = m.spectra.data();
    r.materials = m.materials.data();
    r.lights = m.lights.data();
    r.meshes = m.attrs.data();
    r.camera_medium = -1;
    if (r.integrator_type == 1) {
        const Table &mdi = T.at("media_i"), &mdf = T.at("media_f");
        m.media.resize(std::max<size_t>(1, mdi.rows()));
        for (size_t k = 0; k < mdi.rows(); ++k) {
            const int32_t *a = mdi.i() + 9 * k;
            const float *b = mdf.f() + 39 * k;
            b200pt_medium &M = m.media[k];
            memset(&M, 0, sizeof M);
            M.type = uint32_t(a[0]);
            M.sigma_a = a[1];
            M.sigma_s = a[2];
            M.nx = a[3];
            M.ny = a[4];
            M.nz = a[5];
            memcpy(M.mres, a + 6, 12);
            M.g = b[0];
            memcpy(M.bounds, b + 1, 24);
            memcpy(M.medium_from_render, b + 7, 64);
            memcpy(M.render_from_medium, b + 23, 64);
            if (M.type == 1) {
                M.density = T.at("medium_density_" + std::to_string(k)).f();
                M.majorants = T.at("medium_majorants_" + std::to_string(k)).f();
            }
            if (T.has("media_le") && T.at("media_le").i()[4 * k] >= 0) {
                const int32_t *le = T.at("media_le").i() + 4 * k;
                M.emissive = 1;
                M.le_spectrum = le[0];
                if (M.type == 1) {
                    M.le_nx = le[1];
                    M.le_ny = le[2];
                    M.le_nz = le[3];
                    M.le_scale = T.at("medium_lescale_" + std::to_string(k)).f();
                    if (T.has("medium_temperature_" + std::to_string(k))) {
                        M.temperature = T.at("medium_temperature_" + std::to_string(k)).f();
                        M.temperature_scale = T.at("medium_temperature_params_" + std::to_string(k)).f()[0];
                        M.temperature_offset = T.at("medium_temperature_params_" + std::to_string(k)).f()[1];
                    }
                }
            }
        }
        r.n_media = uint32_t(mdi.rows());
        r.media = m.media.data();
        r.mesh_media = T.at("mesh_media").i();
        r.camera_medium = T.at("camera_medium").i()[0];
    }
    if (r.light_sampler == 2) {
        r.n_light_bvh_nodes = uint32_t(T.at("lightbvh_nodes").rows());
        memcpy(r.light_bvh_bounds, T.at("lightbvh_bounds").f(), 24);
        r.light_bvh_nodes = T.at("lightbvh_nodes").u();
        r.light_bit_trails = T.at("lightbvh_trails").u();
    }
}

// ------------------------------------------------------------------ Level 2: the frame
static void RenderB200(BasicScene &scene, int device, bool hostBvh) {
    using clk = std::chrono::steady_clock;
    RefObjects ro;
    CreateRefObjects(scene, ro);
    MemTables T;
    FlattenScene(ro, T, hostBvh);
    Marshalled m;
    MarshalGeometry(T, m, hostBvh);
    MarshalRender(T, m);

    b200pt_ctx *ctx = nullptr;
    if (b200pt_context_create(device, &ctx) != B200PT_OK) die(std::string("b200pt_context_create: ") + b200pt_last_error(nullptr));
    b200pt_scene *sc = nullptr;
    check(b200pt_scene_create(ctx, &m.scene, &sc), ctx, "b200pt_scene_create");
    b200pt_render *rd = nullptr;
    check(b200pt_render_create(sc, &m.render, &rd), ctx, "b200pt_render_create");
    auto t0 = clk::now();
    check(b200pt_render_run(rd, 0, m.render.samples_per_pixel, 1), ctx, "b200pt_render_run");
    check(b200pt_context_sync(ctx), ctx, "b200pt_context_sync");
    const double seconds = std::chrono::duration<double>(clk::now() - t0).count();

    if (GBufferFilm *gf = ro.film.CastOrNullptr<GBufferFilm>()) {
        // GBufferFilm writes EXR only (film.cpp:1081-1083) and the oracle build of the reference has no EXR writer
        // (oracle/ref.mk): the library's own - image finalised and packed on the device - writes the file, with the
        // channel groups the film's save* switches ask for.  (Albedo, dzdx / dzdy and the SVD basis are not built.)
        if (gf->writeDz || gf->writeBasis) die("pbrt_b200: saveDz / saveRadianceAsBasis are not built");
        uint32_t ch = (gf->writeRadiance ? B200PT_GBUFFER_RADIANCE : 0) | (gf->writePosition ? B200PT_GBUFFER_POSITION : 0) |
                      (gf->writeAlbedo ? B200PT_GBUFFER_ALBEDO : 0) |
                      (gf->writeMaterial ? B200PT_GBUFFER_MATERIAL : 0) | (gf->writeInstance ? B200PT_GBUFFER_INSTANCE : 0) |
                      (gf->writeNs ? B200PT_GBUFFER_NS : 0) | (gf->writeNormal ? B200PT_GBUFFER_NORMAL : 0) |
                      (gf->writeVariance ? B200PT_GBUFFER_VARIANCE : 0) |
                      (gf->writeRelativeVariance ? B200PT_GBUFFER_RELATIVE_VARIANCE : 0);
        check(b200pt_film_write_exr(rd, ch, gf->GetFilename().c_str(), gf->writeFP16 ? 1 : 0), ctx, "b200pt_film_write_exr");
        b200pt_render_stats st{};
        b200pt_render_get_stats(rd, &st);
        fprintf(stderr, "pbrt_b200: %llu samples in %.3f s (gpu %.1f ms, %u launches, %llu + %llu rays) -> %s\n",
                (unsigned long long)st.samples, seconds, st.gpu_ms, st.kernel_launches, (unsigned long long)st.closest_rays,
                (unsigned long long)st.shadow_rays, gf->GetFilename().c_str());
        b200pt_render_destroy(rd);
        b200pt_scene_destroy(sc);
        b200pt_context_destroy(ctx);
        return;
    }
    // the device film accumulators, back into the reference's Film object (row major over pixelBounds, Array2D order)
    Bounds2i pb = ro.film.PixelBounds();
    const size_t nPix = size_t(pb.Area());
    const size_t nB = m.render.film_type == 1 ? m.render.n_buckets : 0;
    std::vector<double> acc(nPix * (4 + 2 * nB));
    check(b200pt_film_read(rd, acc.data()), ctx, "b200pt_film_read");
    const int W = pb.pMax.x - pb.pMin.x;
    if (RGBFilm *rf = ro.film.CastOrNullptr<RGBFilm>()) {
        for (Point2i p : pb) {
            const size_t k = size_t(p.y - pb.pMin.y) * W + (p.x - pb.pMin.x);
            auto &px = rf->pixels[p];
            for (int c = 0; c < 3; ++c) px.rgbSum[c] = acc[4 * k + c];
            px.weightSum = acc[4 * k + 3];
        }
    } else if (SpectralFilm *sf = ro.film.CastOrNullptr<SpectralFilm>()) {
        const double *bs = acc.data() + 4 * nPix, *ws = bs + nB * nPix;
        for (Point2i p : pb) {
            const size_t k = size_t(p.y - pb.pMin.y) * W + (p.x - pb.pMin.x);
            auto &px = sf->pixels[p];
            for (int c = 0; c < 3; ++c) px.rgbSum[c] = acc[4 * k + c];
            px.rgbWeightSum = acc[4 * k + 3];
            for (size_t b = 0; b < nB; ++b) {
                px.bucketSums[b] = bs[b * nPix + k];
                px.weightSums[b] = ws[b * nPix + k];
            }
        }
    } else
        die("RenderB200: film type not handled");
    b200pt_render_stats st{};
    b200pt_render_get_stats(rd, &st);
    fprintf(stderr, "pbrt_b200: %llu samples in %.3f s (gpu %.1f ms, %u launches, %llu + %llu rays)\n",
            (unsigned long long)st.samples, seconds, st.gpu_ms, st.kernel_launches, (unsigned long long)st.closest_rays,
            (unsigned long long)st.shadow_rays);
    b200pt_render_destroy(rd);
    b200pt_scene_destroy(sc);
    b200pt_context_destroy(ctx);

    ImageMetadata metadata;  // wavefront/wavefront.cpp:64-69
    ro.camera.InitMetadata(&metadata);
    metadata.renderTimeSeconds = seconds;
    metadata.samplesPerPixel = ro.sampler.SamplesPerPixel();
    ro.film.WriteImage(metadata);
}

// ------------------------------------------------------------------ Level 1: the intersector plug
class B200Aggregate : public WavefrontAggregate {
  public:
    // `accel`: the aggregate BasicScene::CreateAggregate built (the host keeps parsing and the SAH build);
    // its trees are handed over as they are, so every ray visits what the reference's BVHAggregate visits.
    B200Aggregate(Primitive accel, int device) : aggregate(accel) {
        if (!accel.Is<BVHAggregate>()) die("B200Aggregate: top-level aggregate is not a BVH");
        const BVHAggregate *bvh = accel.Cast<BVHAggregate>();
        const auto &triMeshes = *Triangle::allMeshes;
        nTriMeshes = triMeshes.size();
        const size_t nPatchMeshes = BilinearPatch::allMeshes ? BilinearPatch::allMeshes->size() : 0;
        std::vector<int32_t> meshObject(nTriMeshes + nPatchMeshes, -2);  // -2 = not referenced by the aggregate
        std::vector<float> meshAlpha(nTriMeshes + nPatchMeshes, -1.f);
        auto visit = [&](const Primitive &p, int object, std::vector<uint32_t> &out) {
            Shape shape;
            if (const SimplePrimitive *sp = p.CastOrNullptr<SimplePrimitive>()) shape = sp->shape;
            else if (const GeometricPrimitive *gp = p.CastOrNullptr<GeometricPrimitive>()) {
                shape = gp->shape;
                if (gp->alpha) {
                    const FloatConstantTexture *ct = gp->alpha.CastOrNullptr<FloatConstantTexture>();
                    if (!ct || !shape.Is<Triangle>()) die("B200Aggregate: only constant alpha textures on triangle meshes");
                    meshAlpha[shape.Cast<Triangle>()->meshIndex] = ct->Evaluate(TextureEvalContext());
                }
            } else
                die("B200Aggregate: only triangle / bilinear patch primitives and object instances");
            uint32_t m, k, kind;
            if (const Triangle *t = shape.CastOrNullptr<Triangle>()) {
                m = uint32_t(t->meshIndex);
                k = uint32_t(t->triIndex);
                kind = 0;
            } else if (const BilinearPatch *bp = shape.CastOrNullptr<BilinearPatch>()) {
                m = uint32_t(nTriMeshes + bp->meshIndex);
                k = uint32_t(bp->blpIndex);
                kind = 2;
            } else
                die("B200Aggregate: only triangles and bilinear patches");
            meshObject[m] = object;
            out.insert(out.end(), {kind, m, k});
            primOf[uint64_t(m) << 32 | k] = p;
        };
        std::vector<uint32_t> topPrims;
        std::vector<std::vector<uint32_t>> objPrims;
        std::vector<std::vector<uint32_t>> objNodes;
        std::map<const void *, int> objectId;
        std::map<int, std::pair<int, const TransformedPrimitive *>> byId;
        for (const Primitive &p : bvh->primitives) {
            const TransformedPrimitive *tp = p.CastOrNullptr<TransformedPrimitive>();
            if (!tp) {
                visit(p, -1, topPrims);
                continue;
            }
            const void *key = tp->primitive.ptr();
            if (!objectId.count(key)) {
                const int id = int(objectId.size());
                objectId[key] = id;
                objPrims.emplace_back();
                objNodes.emplace_back();
                if (const BVHAggregate *ob = tp->primitive.CastOrNullptr<BVHAggregate>()) {
                    for (const Primitive &q : ob->primitives) visit(q, id, objPrims[id]);
                    const int nn = countNodes(ob->nodes);
                    objNodes[id].resize(size_t(nn) * 8);
                    memcpy(objNodes[id].data(), ob->nodes, size_t(nn) * 32);
                } else {  // a one-primitive object has no BVH (scene.cpp:1551): a one-leaf tree of its bounds
                    visit(tp->primitive, id, objPrims[id]);
                    LinearBVHNode leaf;
                    memset(&leaf, 0, sizeof leaf);
                    leaf.bounds = tp->primitive.Bounds();
                    leaf.nPrimitives = 1;
                    objNodes[id].resize(8);
                    memcpy(objNodes[id].data(), &leaf, 32);
                }
            }
            topPrims.insert(topPrims.end(), {1u, uint32_t(tp->instanceId), 0u});
            byId[tp->instanceId] = {objectId[key], tp};
        }
        // meshes the aggregate refers to, renumbered densely; refMesh maps back to the reference's numbering
        std::vector<b200pt_mesh> used;
        std::vector<uint32_t> remap(meshObject.size(), 0xFFFFFFFFu);
        for (size_t mi = 0; mi < meshObject.size(); ++mi) {
            if (meshObject[mi] == -2) continue;
            b200pt_mesh me;
            memset(&me, 0, sizeof me);
            me.object = meshObject[mi];
            me.alpha = 1.f;
            if (mi < nTriMeshes) {
                const TriangleMesh *tm = triMeshes[mi];
                me.n_vertices = uint32_t(tm->nVertices);
                me.n_triangles = uint32_t(tm->nTriangles);
                me.p = reinterpret_cast<const float *>(tm->p);  // Point3f = 3 floats
                me.indices = tm->vertexIndices;
                if (meshAlpha[mi] >= 0) {
                    me.kind = B200PT_MESH_ALPHA;
                    me.alpha = meshAlpha[mi];
                }
            } else {
                const BilinearPatchMesh *bm = (*BilinearPatch::allMeshes)[mi - nTriMeshes];
                me.n_vertices = uint32_t(bm->nVertices);
                me.n_triangles = uint32_t(bm->nPatches);
                me.p = reinterpret_cast<const float *>(bm->p);
                me.indices = bm->vertexIndices;
                me.kind = B200PT_MESH_BILINEAR;
            }
            remap[mi] = uint32_t(used.size());
            refMesh.push_back(uint32_t(mi));
            used.push_back(me);
        }
        auto renumber = [&](std::vector<uint32_t> &pr) {
            for (size_t i = 0; i < pr.size(); i += 3)
                if (pr[i] != 1) pr[i + 1] = remap[pr[i + 1]];
        };
        renumber(topPrims);
        for (auto &op : objPrims) renumber(op);
        std::vector<b200pt_instance> insts;
        for (size_t k = 0; k < byId.size(); ++k) {
            auto it = byId.find(int(k) + 1);
            if (it == byId.end()) die("B200Aggregate: instance ids are not 1..n");
            b200pt_instance in{};
            in.object = uint32_t(it->second.first);
            std::vector<float> a = Mat(it->second.second->renderFromPrimitive->GetMatrix()),
                               b = Mat(it->second.second->renderFromPrimitive->GetInverseMatrix());
            memcpy(in.render_from_instance, a.data(), 64);
            memcpy(in.instance_from_render, b.data(), 64);
            insts.push_back(in);
            instances.push_back(it->second.second);
        }
        const int nTop = countNodes(bvh->nodes);
        std::vector<b200pt_bvh2> bvhs;
        bvhs.push_back(b200pt_bvh2{-1, uint32_t(nTop), uint32_t(topPrims.size() / 3), bvh->nodes, topPrims.data()});
        for (size_t k = 0; k < objNodes.size(); ++k)
            bvhs.push_back(b200pt_bvh2{int32_t(k), uint32_t(objNodes[k].size() / 8), uint32_t(objPrims[k].size() / 3),
                                       objNodes[k].data(), objPrims[k].data()});
        b200pt_scene_desc desc{};
        desc.n_meshes = uint32_t(used.size());
        desc.meshes = used.data();
        desc.n_instances = uint32_t(insts.size());
        desc.instances = insts.data();
        desc.n_bvhs = uint32_t(bvhs.size());
        desc.bvhs = bvhs.data();
        if (b200pt_context_create(device, &ctx) != B200PT_OK)
            die(std::string("b200pt_context_create: ") + b200pt_last_error(nullptr));
        check(b200pt_scene_create(ctx, &desc, &scene), ctx, "b200pt_scene_create");
    }
    ~B200Aggregate() override {
        fprintf(stderr, "B200Aggregate: %llu closest + %llu shadow rays on the GPU, %llu single-ray calls, %llu t mismatches\n",
                (unsigned long long)nClosest.load(), (unsigned long long)nShadow.load(), (unsigned long long)nSingle.load(),
                (unsigned long long)nMismatch.load());
        b200pt_scene_destroy(scene);
        b200pt_context_destroy(ctx);
    }

    Bounds3f Bounds() const override {
        float b[6];
        b200pt_scene_bounds(scene, b);
        return Bounds3f(Point3f(b[0], b[1], b[2]), Point3f(b[3], b[4], b[5]));
    }

    // The SurfaceInteraction of a GPU hit: the reference's own Primitive::Intersect on the one primitive the hit
    // names (GeometricPrimitive::Intersect sets material / area light / medium interface, cpu/primitive.cpp:50-95;
    // an instance goes through TransformedPrimitive::Intersect's steps, cpu/primitive.cpp:112-127).
    pstd::optional<ShapeIntersection> Rebuild(const Ray &ray, Float tMax, const b200pt_hit &h) const {
        auto it = primOf.find(uint64_t(refMesh[h.mesh]) << 32 | h.prim);
        if (it == primOf.end()) die("B200Aggregate: hit on an unknown primitive");
        pstd::optional<ShapeIntersection> si;
        if (h.instance == 0) {
            si = it->second.Intersect(ray, tMax);
        } else {
            const TransformedPrimitive *tp = instances[h.instance - 1];
            Float tm = tMax;
            Ray r = tp->renderFromPrimitive->ApplyInverse(ray, &tm);
            si = it->second.Intersect(r, tm);
            if (si) {
                si->intr = (*tp->renderFromPrimitive)(si->intr);
                si->intr.instanceId = tp->instanceId;
            }
        }
        if (!si || si->tHit != h.t) ++nMismatch;
        return si;
    }

    static void Pack(const Ray &r, Float tMax, float *q) {
        q[0] = r.o.x; q[1] = r.o.y; q[2] = r.o.z;
        q[3] = r.d.x; q[4] = r.d.y; q[5] = r.d.z;
        q[6] = tMax;
    }

    void IntersectClosest(int maxRays, const RayQueue *rayQ, EscapedRayQueue *escapedRayQ,
                          HitAreaLightQueue *hitAreaLightQ, MaterialEvalQueue *basicMtlQ,
                          MaterialEvalQueue *universalMtlQ, MediumSampleQueue *mediumSampleQ,
                          RayQueue *nextRayQ) const override {
        const int n = rayQ->Size();
        if (n == 0) return;
        std::vector<float> rays(size_t(n) * 7);
        ParallelFor(0, n, [&](int64_t i) { Pack((*rayQ)[i].ray, Infinity, &rays[size_t(i) * 7]); });
        std::vector<b200pt_hit> hits(n);
        check(b200pt_intersect_closest(scene, rays.data(), n, hits.data(), B200PT_MEM_HOST), ctx, "b200pt_intersect_closest");
        nClosest += n;
        ParallelFor(0, n, [&](int64_t i) {  // wavefront/aggregate.cpp:42-57, with the search done on the GPU
            const RayWorkItem r = (*rayQ)[i];
            pstd::optional<ShapeIntersection> si;
            if (hits[i].mesh != 0xFFFFFFFFu) si = Rebuild(r.ray, Infinity, hits[i]);
            if (!si)
                EnqueueWorkAfterMiss(r, mediumSampleQ, escapedRayQ);
            else
                EnqueueWorkAfterIntersection(r, r.ray.medium, si->tHit, si->intr, mediumSampleQ, nextRayQ, hitAreaLightQ,
                                             basicMtlQ, universalMtlQ);
        });
    }

    void IntersectShadow(int maxRays, ShadowRayQueue *q, SOA<PixelSampleState> *pixelSampleState) const override {
        const int n = q->Size();
        if (n == 0) return;
        std::vector<float> rays(size_t(n) * 7);
        ParallelFor(0, n, [&](int64_t i) {
            const ShadowRayWorkItem w = (*q)[i];
            Pack(w.ray, w.tMax, &rays[size_t(i) * 7]);
        });
        std::vector<uint8_t> occ(n);
        check(b200pt_intersect_any(scene, rays.data(), n, occ.data(), B200PT_MEM_HOST), ctx, "b200pt_intersect_any");
        nShadow += n;
        ParallelFor(0, n, [&](int64_t i) { RecordShadowRayResult((*q)[i], pixelSampleState, occ[i] != 0); });
    }

    // one ray through the GPU (the walks below are sequential per ray: segment k + 1 starts where segment k ended)
    pstd::optional<ShapeIntersection> IntersectOne(const Ray &ray, Float tMax) const {
        float q[7];
        Pack(ray, tMax, q);
        b200pt_hit h;
        {
            std::lock_guard<std::mutex> lock(mu);  // one context = one stream: calls are serialised
            check(b200pt_intersect_closest(scene, q, 1, &h, B200PT_MEM_HOST), ctx, "b200pt_intersect_closest");
        }
        ++nSingle;
        if (h.mesh == 0xFFFFFFFFu) return {};
        return Rebuild(ray, tMax, h);
    }

    // wavefront/aggregate.cpp:72-90: TraceTransmittance (wavefront/intersect.h:164-274) with the reference's own
    // ratio-tracking code and this aggregate as the intersector.  (The product's transmittance walks run inside
    // b200pt_render_run; this adaptor is the test vehicle for the intersector plug.)
    void IntersectShadowTr(int maxRays, ShadowRayQueue *q, SOA<PixelSampleState> *pixelSampleState) const override {
        for (int index = 0; index < q->Size(); ++index) {
            const ShadowRayWorkItem w = (*q)[index];
            pstd::optional<ShapeIntersection> si;
            TraceTransmittance(
                w, pixelSampleState,
                [&](Ray ray, Float tMax) -> TransmittanceTraceResult {
                    si = IntersectOne(ray, tMax);
                    if (!si) return TransmittanceTraceResult{false, Point3f(), pbrt::Material()};
                    return TransmittanceTraceResult{true, si->intr.p(), si->intr.material};
                },
                [&](Point3f p) -> Ray { return si->intr.SpawnRayTo(p); });
        }
    }

    // wavefront/aggregate.cpp:92-116
    void IntersectOneRandom(int maxRays, SubsurfaceScatterQueue *subsurfaceScatterQueue) const override {
        for (int index = 0; index < subsurfaceScatterQueue->Size(); ++index) {
            const SubsurfaceScatterWorkItem &w = (*subsurfaceScatterQueue)[index];
            uint64_t seed = Hash(w.p0, w.p1);
            WeightedReservoirSampler<SubsurfaceInteraction> wrs(seed);
            Interaction base(w.p0, 0.f, Medium());
            while (true) {
                Ray r = base.SpawnRayTo(w.p1);
                if (r.d == Vector3f(0, 0, 0)) break;
                pstd::optional<ShapeIntersection> si = IntersectOne(r, 1);
                if (!si) break;
                base = si->intr;
                if (si->intr.material == w.material) wrs.Add(SubsurfaceInteraction(si->intr), 1.f);
            }
            if (wrs.HasSample()) {
                subsurfaceScatterQueue->reservoirPDF[index] = wrs.SampleProbability();
                subsurfaceScatterQueue->ssi[index] = wrs.GetSample();
            } else
                subsurfaceScatterQueue->reservoirPDF[index] = 0;
        }
    }

  private:
    Primitive aggregate;
    b200pt_ctx *ctx = nullptr;
    b200pt_scene *scene = nullptr;
    size_t nTriMeshes = 0;
    std::vector<uint32_t> refMesh;  // library mesh id -> reference mesh id (patch meshes after the triangle meshes)
    std::unordered_map<uint64_t, Primitive> primOf;
    std::vector<const TransformedPrimitive *> instances;
    mutable std::mutex mu;
    mutable std::atomic<uint64_t> nClosest{0}, nShadow{0}, nSingle{0}, nMismatch{0};
};

// RenderWavefront (wavefront/wavefront.cpp:14-70) with the aggregate the integrator constructed
// (wavefront/integrator.cpp:157-169) replaced by B200Aggregate over the same host-built BVH.
static void RenderWavefrontB200(BasicScene &scene, int device) {
    WavefrontPathIntegrator *integrator = new WavefrontPathIntegrator(pstd::pmr::get_default_resource(), scene);
    CPUAggregate *cpu = dynamic_cast<CPUAggregate *>(integrator->aggregate);
    if (!cpu) die("the integrator's aggregate is not the CPUAggregate");
    integrator->aggregate = new B200Aggregate(cpu->aggregate, device);
    Float seconds = integrator->Render();
    delete integrator->aggregate;
    ImageMetadata metadata;
    integrator->camera.InitMetadata(&metadata);
    metadata.renderTimeSeconds = seconds;
    metadata.samplesPerPixel = integrator->sampler.SamplesPerPixel();
    integrator->film.WriteImage(metadata);
}

int main(int argc, char *argv[]) {
    PBRTOptions options;
    options.quiet = true;
    std::vector<std::string> filenames;
    bool wavefront = false, ownBvh = false;
    int device = 0;
    auto onError = [](const std::string &err) {
        fprintf(stderr, "pbrt_b200: %s\n", err.c_str());
        exit(1);
    };
    std::vector<std::string> args = GetCommandLineArguments(argv);
    for (auto iter = args.begin(); iter != args.end(); ++iter) {
        if ((*iter)[0] != '-') {
            filenames.push_back(*iter);
            continue;
        }
        if (ParseArg(&iter, args.end(), "nthreads", &options.nThreads, onError) ||
            ParseArg(&iter, args.end(), "spp", &options.pixelSamples, onError) ||
            ParseArg(&iter, args.end(), "seed", &options.seed, onError) ||
            ParseArg(&iter, args.end(), "outfile", &options.imageFile, onError) ||
            ParseArg(&iter, args.end(), "wavefront", &wavefront, onError) ||
            ParseArg(&iter, args.end(), "own-bvh", &ownBvh, onError) ||  // let the library build the trees itself
            ParseArg(&iter, args.end(), "device", &device, onError))
            continue;
        onError("unknown option " + *iter);
    }
    if (filenames.empty()) onError("usage: pbrt_b200 [--wavefront] [--spp n] [--nthreads n] [--outfile f] scene.pbrt");
    options.wavefront = wavefront;
    InitPBRT(options);
    {
        BasicScene scene;
        BasicSceneBuilder builder(&scene);
        ParseFiles(&builder, filenames);
        if (wavefront) RenderWavefrontB200(scene, device);  // cmd/pbrt.cpp:296-299
        else RenderB200(scene, device, !ownBvh);
    }
    fflush(stdout);
    fflush(stderr);
    _Exit(0);
}
