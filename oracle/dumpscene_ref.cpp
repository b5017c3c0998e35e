// oracle/dumpscene_ref.cpp - parses a .pbrt file with the REFERENCE's own parser and scene
// construction (BasicSceneBuilder / BasicScene, the host side the north_star keeps) and writes the
// flat tables a host adaptor would hand to b200pt_render_create as a `.b2rs`
// container of named arrays.  TEST INFRASTRUCTURE ONLY (see oracle/raycast_ref.cpp header).
//
//   dumpscene_ref scene.pbrt out.b2rs
//
// Follows RenderCPU's construction order (cpu/render.cpp:25-55); reads private members of the
// created objects (camera matrices, film sensor, alias table, BVH) in this one TU.
#include "flatten_ref.h"

// ---- .b2rs container: 'B2RS', u32 count, then {u32 nameLen, name, u32 dtype (0 f32, 1 i32, 2 u32),
// u32 ndim, u32 dims[], data}
struct Writer {
    FILE *f;
    uint32_t count = 0;
    explicit Writer(const char *fn) {
        f = fopen(fn, "wb");
        if (!f) die("cannot open output");
        uint32_t hdr[2] = {0x53523242u, 0};
        fwrite(hdr, 4, 2, f);
    }
    void put(const std::string &name, uint32_t dtype, std::vector<uint32_t> dims, const void *data) {
        uint32_t nl = name.size(), nd = dims.size();
        size_t n = 1;
        for (uint32_t d : dims) n *= d;
        fwrite(&nl, 4, 1, f);
        fwrite(name.data(), 1, nl, f);
        fwrite(&dtype, 4, 1, f);
        fwrite(&nd, 4, 1, f);
        fwrite(dims.data(), 4, nd, f);
        if (n) fwrite(data, 4, n, f);
        ++count;
    }
    void f32(const std::string &name, std::vector<uint32_t> dims, const std::vector<float> &v) { put(name, 0, dims, v.data()); }
    void i32(const std::string &name, std::vector<uint32_t> dims, const std::vector<int32_t> &v) { put(name, 1, dims, v.data()); }
    void u32(const std::string &name, std::vector<uint32_t> dims, const std::vector<uint32_t> &v) { put(name, 2, dims, v.data()); }
    void close() {
        fseek(f, 4, SEEK_SET);
        fwrite(&count, 4, 1, f);
        fclose(f);
    }
};

int main(int argc, char **argv) {
    if (argc < 3) die("usage: dumpscene_ref scene.pbrt out.b2rs");
    PBRTOptions opt;
    opt.quiet = true;
    InitPBRT(opt);
    BasicScene scene;
    BasicSceneBuilder builder(&scene);
    std::vector<std::string> files{argv[1]};
    ParseFiles(&builder, files);

    // ---- construction order of RenderCPU (cpu/render.cpp:25-55)
    RefObjects ro;
    CreateRefObjects(scene, ro);
    Primitive &accel = ro.accel;
    Camera &camera = ro.camera;
    Film &film = ro.film;
    Sampler &sampler = ro.sampler;
    std::vector<Light> &lights = ro.lights;

    // ---- `dumpscene_ref scene.pbrt --render out.b2rs`: the reference's own render loop body
    // (ImageTileIntegrator::Render, cpu/integrators.cpp:60-140: EvaluatePixelSample per pixel sample, samples of
    // a pixel in index order) followed by Film::GetImage, every channel written as float32.  For films whose
    // file format the oracle build cannot write (spectral film -> EXR).
    if (std::string(argv[2]) == "--render") {
        if (argc < 4) die("usage: dumpscene_ref scene.pbrt --render out.b2rs");
        std::unique_ptr<Integrator> integ(scene.CreateIntegrator(camera, sampler, accel, lights));
        ImageTileIntegrator *ti = dynamic_cast<ImageTileIntegrator *>(integ.get());
        if (!ti) die("--render needs an image-tile integrator");
        Bounds2i pb = film.PixelBounds();
        int spp = sampler.SamplesPerPixel();
        ParallelFor2D(pb, [&](Bounds2i tile) {
            ScratchBuffer scratch(65536);
            Sampler s = sampler.Clone(Allocator());
            for (Point2i p : tile)
                for (int si = 0; si < spp; ++si) {
                    s.StartPixelSample(p, si);
                    ti->EvaluatePixelSample(p, si, s, scratch);
                    scratch.Reset();
                }
        });
        ImageMetadata md;
        Image im = film.GetImage(&md);
        int W = im.Resolution().x, H = im.Resolution().y, C = im.NChannels();
        std::vector<float> px(size_t(W) * H * C);
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x)
                for (int c = 0; c < C; ++c) px[(size_t(y) * W + x) * C + c] = im.GetChannel(Point2i(x, y), c);
        Writer w(argv[3]);
        w.f32("image", {uint32_t(H), uint32_t(W), uint32_t(C)}, px);
        w.close();
        std::string names;
        for (const std::string &n : im.ChannelNames()) names += (names.empty() ? "\"" : ", \"") + n + "\"";
        printf("{\"width\": %d, \"height\": %d, \"channels\": [%s]}\n", W, H, names.c_str());
        fflush(stdout);
        _Exit(0);
    }

    // ---- debug mode: `dumpscene_ref scene.pbrt --path x y sampleIndex` prints, for one pixel sample, the
    // radiance the reference's PathIntegrator::Li returns when maxDepth is cut at 0, 1, 2, ... (the
    // increments are the per-vertex contributions), to localise a divergence of our kernels.
    if (std::string(argv[2]) == "--path") {
        if (argc < 6) die("usage: dumpscene_ref scene.pbrt --path x y sampleIndex");
        int px = atoi(argv[3]), py = atoi(argv[4]), si = atoi(argv[5]);
        std::unique_ptr<Integrator> integ(scene.CreateIntegrator(camera, sampler, accel, lights));
        PathIntegrator *pi = dynamic_cast<PathIntegrator *>(integ.get());
        if (!pi) die("--path needs the path integrator");
        int fullDepth = pi->maxDepth;
        ScratchBuffer scratch(65536);
        for (int d = 0; d <= fullDepth; ++d) {
            pi->maxDepth = d;
            Sampler s = sampler.Clone(Allocator());
            s.StartPixelSample(Point2i(px, py), si, 0);
            Float lu = s.Get1D();
            SampledWavelengths lambda = film.SampleWavelengths(lu);
            CameraSample cs = GetCameraSample(s, Point2i(px, py), film.GetFilter());
            pstd::optional<CameraRayDifferential> cr = camera.GenerateRayDifferential(cs, lambda);
            SampledSpectrum L = pi->Li(cr->ray, lambda, s, scratch, nullptr);
            printf("maxdepth %d  lambda0 %.9g  o %.9g %.9g %.9g  d %.9g %.9g %.9g  L0 %.9g L1 %.9g L15 %.9g L30 %.9g  dim %d\n", d,
                   lambda[0], cr->ray.o.x, cr->ray.o.y, cr->ray.o.z, cr->ray.d.x, cr->ray.d.y, cr->ray.d.z, L[0], L[1],
                   L[15], L[30], s.Is<ZSobolSampler>() ? s.Cast<ZSobolSampler>()->dimension : -1);
            scratch.Reset();
        }
        // the same path once more, vertex by vertex, through the reference's own Intersect / GetBSDF /
        // Sample_f (a driver loop over the reference API, printing the state Li keeps to itself)
        {
            Sampler s = sampler.Clone(Allocator());
            s.StartPixelSample(Point2i(px, py), si, 0);
            Float lu = s.Get1D();
            SampledWavelengths lambda = film.SampleWavelengths(lu);
            CameraSample cs = GetCameraSample(s, Point2i(px, py), film.GetFilter());
            RayDifferential ray = camera.GenerateRayDifferential(cs, lambda)->ray;
            SampledSpectrum beta(1.f);
            Float etaScale = 1;
            for (int depth = 0; depth < fullDepth; ++depth) {
                pstd::optional<ShapeIntersection> isr = accel.Intersect(ray, Infinity);
                if (!isr) break;
                SurfaceInteraction &isect = isr->intr;
                BSDF bsdf = isect.GetBSDF(ray, lambda, camera, scratch, s);
                if (!bsdf) break;
                int dim0 = s.Cast<ZSobolSampler>()->dimension;
                if (IsNonSpecular(bsdf.Flags())) { s.Get1D(); s.Get2D(); }
                Vector3f wo = -ray.d;
                Float u = s.Get1D();
                pstd::optional<BSDFSample> bs = bsdf.Sample_f(wo, u, s.Get2D());
                if (!bs) break;
                beta *= bs->f * AbsDot(bs->wi, isect.shading.n) / bs->pdf;
                if (bs->IsTransmission()) etaScale *= Sqr(bs->eta);
                SampledSpectrum rrBeta = beta * etaScale;
                printf("   ref vertex %d: t %.9g dim %d f0 %.9g absdot %.9g pdf %.9g eta %.9g -> beta0 %.9g etaScale %.9g rr %.9g wi %.9g %.9g %.9g\n",
                       depth, isr->tHit, dim0, bs->f[0], AbsDot(bs->wi, isect.shading.n), bs->pdf, bs->eta, beta[0], etaScale,
                       rrBeta.MaxComponentValue(), bs->wi.x, bs->wi.y, bs->wi.z);
                ray = isect.SpawnRay(ray, bsdf, bs->wi, bs->flags, bs->eta);
                if (rrBeta.MaxComponentValue() < 1 && depth + 1 > 1) {
                    Float q = std::max<Float>(0, 1 - rrBeta.MaxComponentValue());
                    if (s.Get1D() < q) break;
                    beta /= 1 - q;
                }
            }
        }
        fflush(stdout);
        _Exit(0);
    }

    Writer w(argv[2]);
    // `--no-bvh`: leave the reference's BVHs out (large scenes; our host builds the same trees itself)
    const bool dumpBvh = !(argc > 3 && std::string(argv[3]) == "--no-bvh");
    size_t nMesh = 0, nTri = 0, nMat = 0, nSpec = 0;
    FlattenScene(ro, w, dumpBvh, &nMesh, &nTri, &nMat, &nSpec);
    w.close();
    printf("{\"meshes\": %zu, \"triangles\": %zu, \"lights\": %zu, \"materials\": %zu, \"spectra\": %zu}\n", nMesh,
           nTri, lights.size(), nMat, nSpec);
    fflush(stdout);
    _Exit(0);
}
