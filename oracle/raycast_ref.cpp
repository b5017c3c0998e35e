// oracle/raycast_ref.cpp - ray-batch driver around the REFERENCE's own
// BVHAggregate (compiled in place from /root/reference by oracle/ref.mk).
// TEST INFRASTRUCTURE ONLY: nothing under oracle/ is linked into, imported by
// or executed from the product (pbrt-v4_b200/, libb200pt.so).  Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// run this binary, and only as the checker / the timed CPU baseline.
//
// What it calls (reference file:line):
//   TriangleMesh ctor                 util/mesh.h:24, util/mesh.cpp:38
//   Triangle::CreateTriangles         shapes.cpp:276
//   SimplePrimitive                   cpu/primitive.cpp:88
//   TransformedPrimitive              cpu/primitive.cpp:112 (instanceId: scene.cpp:1560-1573)
//   BVHAggregate ctor (SAH)           cpu/aggregates.cpp:140 (top level maxPrimsInNode 4
//                                     as BVHAggregate::Create :744; instance level 1 as scene.cpp:1542)
//   BVHAggregate::Intersect/IntersectP  cpu/aggregates.cpp:529 / :581
//
// Usage: raycast_ref <scene.b2sc> <command> ...
//   dump-bvh  <out.bin>                              flat nodes + ordered (mesh,tri) ids, per BVH
//   closest   <rays.f32> <hits.bin> [nthreads] [reps]
//   any       <rays.f32> <occl.u8>  [nthreads] [reps]
// Scene file (little endian), written by pbrt-v4_b200/scenes.py:
//   u32 magic 'B2SC', u32 version=2, u32 nMeshes, u32 nInstances
//   per mesh: i32 object (-1 = world), u32 nVerts, u32 nPrims, u32 kind (0 triangles, 1 bilinear patches),
//             f32 P[3*nVerts], i32 idx[(3 or 4)*nPrims]
//   (a bilinear mesh becomes a BilinearPatchMesh + BilinearPatch::CreatePatches, util/mesh.h:47-77, shapes.cpp:1005)
//   per instance: u32 object, f32 renderFromInstance[16], f32 instanceFromRender[16] (row major)
// Rays: n x 7 f32 (o.xyz, d.xyz, tMax).  Hits: n x 12 words
//   {u32 mesh, u32 tri, u32 instanceId, f32 t, f32 u, f32 v, f32 p[3], f32 n[3]}, miss: mesh = 0xFFFFFFFF.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
// The driver reads BVHAggregate::{nodes,primitives}, SimplePrimitive::shape and
// Triangle::{meshIndex,triIndex}; layouts are unchanged, so this one TU links
// against the normally compiled library.
#define private public
#include <pbrt/pbrt.h>
#include <pbrt/cpu/aggregates.h>
#include <pbrt/cpu/primitive.h>
#include <pbrt/materials.h>
#include <pbrt/options.h>
#include <pbrt/shapes.h>
#include <pbrt/textures.h>
#include <pbrt/util/mesh.h>
#include <pbrt/util/parallel.h>
#include <pbrt/util/transform.h>
#undef private

namespace pbrt {
// Same layout as the definition private to cpu/aggregates.cpp:129-137.
struct alignas(32) LinearBVHNode {
    Bounds3f bounds;
    union {
        int primitivesOffset;
        int secondChildOffset;
    };
    uint16_t nPrimitives;
    uint8_t axis;
};
}  // namespace pbrt
using namespace pbrt;

static void die(const char *msg) {
    fprintf(stderr, "raycast_ref: %s\n", msg);
    exit(2);
}
template <typename T>
static std::vector<T> readVec(FILE *f, size_t n) {
    std::vector<T> v(n);
    if (n && fread(v.data(), sizeof(T), n, f) != n)
        die("short read");
    return v;
}

// One fake Material per mesh: SimplePrimitive only stores and forwards it
// (cpu/primitive.cpp:105), so pointer identity gives us the mesh of a hit.
static char *g_matBase;
static Material meshMaterial(uint32_t mesh) {
    return Material(reinterpret_cast<DiffuseMaterial *>(g_matBase + 64 * size_t(mesh + 1)));
}
static uint32_t meshOfMaterial(Material m) {
    return uint32_t(((char *)m.ptr() - g_matBase) / 64 - 1);
}

struct Scene {
    Primitive aggregate;
    std::vector<BVHAggregate *> bvhs;  // [0] = top level, then one per object
    std::vector<int> bvhObject;
};

static Scene loadScene(const char *fn) {
    FILE *f = fopen(fn, "rb");
    if (!f)
        die("cannot open scene");
    uint32_t hdr[4];
    if (fread(hdr, 4, 4, f) != 4 || hdr[0] != 0x43533242u || hdr[1] != 2)
        die("bad scene header");
    uint32_t nMeshes = hdr[2], nInst = hdr[3];
    g_matBase = (char *)aligned_alloc(64, 64 * size_t(nMeshes + 2));
    Allocator alloc;
    std::vector<Primitive> world;
    std::map<int, std::vector<Primitive>> objPrims;
    static Transform identity;
    for (uint32_t m = 0; m < nMeshes; ++m) {
        int32_t object;
        uint32_t nv, nt, kind;
        if (fread(&object, 4, 1, f) != 1 || fread(&nv, 4, 1, f) != 1 || fread(&nt, 4, 1, f) != 1 ||
            fread(&kind, 4, 1, f) != 1)
            die("short read");
        // kind | 2: the mesh's primitives are GeometricPrimitives with a constant alpha texture (cpu/primitive.cpp:50-109)
        float alpha = 1.f;
        const bool hasAlpha = (kind & 2) != 0;
        if (hasAlpha && fread(&alpha, 4, 1, f) != 1) die("short read");
        kind &= 1;
        std::vector<float> P = readVec<float>(f, 3 * size_t(nv));
        std::vector<int> idx = readVec<int>(f, (kind == 1 ? 4 : 3) * size_t(nt));
        std::vector<Point3f> p(nv);
        for (uint32_t i = 0; i < nv; ++i)
            p[i] = Point3f(P[3 * i], P[3 * i + 1], P[3 * i + 2]);
        std::vector<int> face(nt);
        for (uint32_t i = 0; i < nt; ++i)
            face[i] = int(i);  // faceIndex := triIndex, read back from the hit
        std::vector<Primitive> &dst = object < 0 ? world : objPrims[object];
        if (kind == 1) {
            BilinearPatchMesh *bm = alloc.new_object<BilinearPatchMesh>(
                identity, false, std::move(idx), std::move(p), std::vector<Normal3f>(), std::vector<Point2f>(),
                std::move(face), nullptr, alloc);
            for (Shape s : BilinearPatch::CreatePatches(bm, alloc)) dst.push_back(new SimplePrimitive(s, meshMaterial(m)));
            continue;
        }
        TriangleMesh *mesh = alloc.new_object<TriangleMesh>(
            identity, false, std::move(idx), std::move(p), std::vector<Vector3f>(),
            std::vector<Normal3f>(), std::vector<Point2f>(), std::move(face), alloc);
        pstd::vector<Shape> tris = Triangle::CreateTriangles(mesh, alloc);
        for (Shape s : tris) {
            if (hasAlpha)
                dst.push_back(new GeometricPrimitive(s, meshMaterial(m), nullptr, MediumInterface(),
                                                     FloatTexture(alloc.new_object<FloatConstantTexture>(alpha))));
            else
                dst.push_back(new SimplePrimitive(s, meshMaterial(m)));
        }
    }
    Scene sc;
    sc.bvhs.push_back(nullptr);
    sc.bvhObject.push_back(-1);
    std::map<int, Primitive> objAgg;
    for (auto &kv : objPrims) {
        if (kv.second.size() < 2)
            die("object definitions need >= 2 triangles");
        BVHAggregate *b = new BVHAggregate(std::move(kv.second));  // scene.cpp:1542
        objAgg[kv.first] = b;
        sc.bvhs.push_back(b);
        sc.bvhObject.push_back(kv.first);
    }
    for (uint32_t i = 0; i < nInst; ++i) {
        uint32_t object;
        float m[16], mi[16];
        if (fread(&object, 4, 1, f) != 1 || fread(m, 4, 16, f) != 16 || fread(mi, 4, 16, f) != 16)
            die("short read");
        if (!objAgg.count(object))
            die("instance of undefined object");
        SquareMatrix<4> M(m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], m[8], m[9], m[10],
                          m[11], m[12], m[13], m[14], m[15]);
        SquareMatrix<4> MI(mi[0], mi[1], mi[2], mi[3], mi[4], mi[5], mi[6], mi[7], mi[8], mi[9],
                           mi[10], mi[11], mi[12], mi[13], mi[14], mi[15]);
        Transform *t = new Transform(M, MI);
        world.push_back(new TransformedPrimitive(objAgg[object], t, i + 1));  // scene.cpp:1560
    }
    fclose(f);
    if (world.empty())
        die("empty world");
    sc.bvhs[0] = new BVHAggregate(std::move(world), 4, BVHAggregate::SplitMethod::SAH);
    sc.aggregate = sc.bvhs[0];
    return sc;
}

static int countNodes(const LinearBVHNode *nodes) {
    // DFS layout (cpu/aggregates.cpp:505): walk to find the extent.
    int maxIdx = 0;
    std::vector<int> stack{0};
    while (!stack.empty()) {
        int i = stack.back();
        stack.pop_back();
        maxIdx = std::max(maxIdx, i);
        if (nodes[i].nPrimitives == 0) {
            stack.push_back(i + 1);
            stack.push_back(nodes[i].secondChildOffset);
        }
    }
    return maxIdx + 1;
}

static void dumpBVH(const Scene &sc, const char *out) {
    FILE *f = fopen(out, "wb");
    if (!f)
        die("cannot open output");
    uint32_t nb = sc.bvhs.size();
    fwrite(&nb, 4, 1, f);
    for (size_t b = 0; b < sc.bvhs.size(); ++b) {
        const BVHAggregate *bvh = sc.bvhs[b];
        int32_t object = sc.bvhObject[b];
        uint32_t nNodes = countNodes(bvh->nodes), nPrims = bvh->primitives.size();
        fwrite(&object, 4, 1, f);
        fwrite(&nNodes, 4, 1, f);
        fwrite(&nPrims, 4, 1, f);
        fwrite(bvh->nodes, sizeof(LinearBVHNode), nNodes, f);
        // per ordered primitive: {kind 0 = triangle, mesh, tri} or {kind 1 = instance, instanceId, 0}
        for (const Primitive &p : bvh->primitives) {
            uint32_t rec[3] = {0, 0, 0};
            if (const SimplePrimitive *sp = p.CastOrNullptr<SimplePrimitive>()) {
                // the mesh id of the scene file is carried by the per-mesh material (triangle and bilinear
                // meshes have separate index spaces inside the reference)
                rec[1] = meshOfMaterial(sp->material);
                if (const BilinearPatch *bp = sp->shape.CastOrNullptr<BilinearPatch>()) {
                    rec[0] = 2;
                    rec[2] = bp->blpIndex;
                } else
                    rec[2] = sp->shape.Cast<Triangle>()->triIndex;
            } else if (const GeometricPrimitive *gp = p.CastOrNullptr<GeometricPrimitive>()) {
                rec[1] = meshOfMaterial(gp->material);
                rec[2] = gp->shape.Cast<Triangle>()->triIndex;
            } else if (const TransformedPrimitive *tp = p.CastOrNullptr<TransformedPrimitive>()) {
                rec[0] = 1;
                rec[1] = tp->instanceId;
            } else
                die("unexpected primitive kind");
            fwrite(rec, 4, 3, f);
        }
    }
    fclose(f);
}

int main(int argc, char **argv) {
    if (argc < 4)
        die("usage: raycast_ref scene.b2sc {dump-bvh out | closest rays hits [nthreads] [reps] | any rays occl [nthreads] [reps]}");
    std::string cmd = argv[2];
    int nThreads = argc > 5 ? atoi(argv[5]) : 0;
    int reps = argc > 6 ? atoi(argv[6]) : 1;
    PBRTOptions opt;
    opt.quiet = true;
    opt.nThreads = nThreads;
    InitPBRT(opt);
    auto t0 = std::chrono::steady_clock::now();
    Scene sc = loadScene(argv[1]);
    double buildS = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    if (cmd == "dump-bvh") {
        dumpBVH(sc, argv[3]);
        printf("{\"build_s\": %.4f, \"bvhs\": %zu}\n", buildS, sc.bvhs.size());
        return 0;
    }
    if (argc < 5)
        die("need rays and output files");
    FILE *rf = fopen(argv[3], "rb");
    if (!rf)
        die("cannot open rays");
    fseek(rf, 0, SEEK_END);
    size_t n = ftell(rf) / 28;
    fseek(rf, 0, SEEK_SET);
    std::vector<float> rays = readVec<float>(rf, n * 7);
    fclose(rf);

    Primitive agg = sc.aggregate;
    double bestS = 1e30;
    if (cmd == "closest") {
        std::vector<uint32_t> hits(n * 12);
        for (int r = 0; r < reps; ++r) {
            auto s = std::chrono::steady_clock::now();
            ParallelFor(0, (int64_t)n, [&](int64_t i) {
                const float *q = &rays[7 * i];
                Ray ray(Point3f(q[0], q[1], q[2]), Vector3f(q[3], q[4], q[5]));
                pstd::optional<ShapeIntersection> si = agg.Intersect(ray, q[6]);
                uint32_t *h = &hits[12 * i];
                if (!si) {
                    memset(h, 0, 48);
                    h[0] = 0xFFFFFFFFu;
                    return;
                }
                const SurfaceInteraction &in = si->intr;
                h[0] = meshOfMaterial(in.material);
                h[1] = in.faceIndex;
                h[2] = in.instanceId;
                float v[9] = {si->tHit, in.uv.x, in.uv.y, in.p().x, in.p().y, in.p().z,
                              in.n.x,   in.n.y,  in.n.z};
                memcpy(h + 3, v, 36);
            });
            bestS = std::min(bestS, std::chrono::duration<double>(std::chrono::steady_clock::now() - s).count());
        }
        FILE *of = fopen(argv[4], "wb");
        fwrite(hits.data(), 4, hits.size(), of);
        fclose(of);
    } else if (cmd == "any") {
        std::vector<uint8_t> occ(n);
        for (int r = 0; r < reps; ++r) {
            auto s = std::chrono::steady_clock::now();
            ParallelFor(0, (int64_t)n, [&](int64_t i) {
                const float *q = &rays[7 * i];
                Ray ray(Point3f(q[0], q[1], q[2]), Vector3f(q[3], q[4], q[5]));
                occ[i] = agg.IntersectP(ray, q[6]) ? 1 : 0;
            });
            bestS = std::min(bestS, std::chrono::duration<double>(std::chrono::steady_clock::now() - s).count());
        }
        FILE *of = fopen(argv[4], "wb");
        fwrite(occ.data(), 1, occ.size(), of);
        fclose(of);
    } else
        die("unknown command");
    printf("{\"cmd\": \"%s\", \"rays\": %zu, \"seconds\": %.6f, \"mrays_per_s\": %.4f, \"threads\": %d, "
           "\"build_s\": %.4f}\n",
           cmd.c_str(), n, bestS, n / bestS * 1e-6, RunningThreads(), buildS);
    fflush(stdout);
    _Exit(0);  // skip pbrt's static teardown
}
