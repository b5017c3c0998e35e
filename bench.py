#!/usr/bin/env python
"""bench.py - BASELINE.json's metric on its config 2: Mrays/s of BVH traversal
(closest-hit + shadow any-hit) over the procedural 1M-triangle scene.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One "step" = one pass of the hot path over one batch: `--rays` primary rays through
b200pt_intersect_closest plus as many shadow rays through b200pt_intersect_any.
`value` times the kernels with the batches resident in HBM; `e2e` times the same
calls through the C ABI with pinned HOST buffers (H2D + kernels + D2H inside the timed
region).  Under torchrun every rank traces its own full batch against a replicated scene
(weak scaling, no data-path collective); time = max over ranks.
`--impl reference` times the reference's own CPU BVHAggregate (oracle/_ref/raycast_ref)
on a bounded sample of the same batches, on all host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "traversal_mrays_per_s"
UNIT = "Mrays/s"


HEADLINE_NCU = os.path.join(ROOT, "profiles", "r02_headline_ncu.json")  # scripts/ncu_extract.py of the capture below


def traversal_bytes(ts, any_hit=False):
    """Algorithmic bytes per ray of a counted launch: the ray in, the record out, every node / leaf
    box / primitive record the walk fetched.  wide walk: 256 B per Wide8 node, 32 B per leaf box, 48 B of each 64 B
    primitive slot; binary walk: 32 B per box test (two per 64 B pair node), 48 B per primitive."""
    r = max(ts["rays"], 1)
    if ts["wide_nodes_visited"]:
        node_bytes = 256 * ts["wide_nodes_visited"] + 32 * ts["leaf_boxes_tested"]
    else:
        node_bytes = 32 * ts["nodes_visited"]
    return 28 + (1 if any_hit else 24) + (node_bytes + 48 * ts["prims_tested"]) / r


def headline_ncu(kernel_substring):
    """Counters of the headline kernel from the committed ncu capture of `scripts/r2_headline_prof.sh` (same scene, same
    kernel build): DRAM bytes per ray, issue-slot utilisation, active lanes.  None when there is no capture."""
    try:
        for l in json.load(open(HEADLINE_NCU))["launches"]:
            if kernel_substring in l["kernel"] and l.get("rays"):
                return l
    except Exception:
        pass
    return None


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tris", type=int, default=1_000_000)
    ap.add_argument("--rays", type=int, default=1 << 24, help="primary rays per step (as many shadow rays)")
    ap.add_argument("--cpu-sample", type=int, default=1 << 21, help="rays of each kind in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-render", action="store_true", help="skip the config-1 end-to-end render section")
    ap.add_argument("--render-cpu-spp", type=int, default=64, help="spp of the bounded pbrt_ref sample")
    ap.add_argument("--no-config3", action="store_true", help="skip the config-3 (10M instanced triangles) section")
    ap.add_argument("--config3-spp", type=int, default=32, help="spp of config 3's frame rendered per step")
    ap.add_argument("--no-config4", action="store_true", help="skip the config-4 (realistic lens + spectral film) section")
    ap.add_argument("--config4-spp", type=int, default=64, help="spp of config 4's frame rendered per step")
    ap.add_argument("--config4-cpu-spp", type=int, default=4, help="spp of the bounded reference sample of config 4")
    ap.add_argument("--no-config5", action="store_true", help="skip the config-5 (volpath, 256^3 grid medium) section")
    ap.add_argument("--config5-spp", type=int, default=16, help="spp of config 5's frame rendered per step")
    ap.add_argument("--config5-grid", type=int, default=256, help="density grid resolution of config 5")
    ap.add_argument("--config5-cpu-res", type=int, nargs=2, default=[960, 540], help="frame of the bounded reference sample")
    ap.add_argument("--config5-cpu-spp", type=int, default=4, help="spp of the bounded reference sample of config 5")
    ap.add_argument("--config3-cpu-spp", type=int, default=2, help="spp of the bounded pbrt_ref sample of config 3")
    ap.add_argument("--config3-grid", type=int, default=224, help="tessellation of config 3's object definitions (2 * grid^2 triangles each)")
    ap.add_argument("--parity-sample", type=int, default=1 << 20, help="rays of each kind compared with the reference inside the bench")
    return ap.parse_args()


def workload_config(args):
    return {"workload": f"config2: procedural {args.tris}-triangle random-mesh scene (seed 12111), "
                        f"{args.rays} primary closest-hit + {args.rays} shadow any-hit rays per step",
            "triangles": args.tris, "rays_per_step": 2 * args.rays,
            "l2_policy": "ray/hit buffers (28+24 B/ray x %d rays) exceed the 126 MB L2; no flush needed" % args.rays,
            "parallelism": "rays sharded by rank, scene replicated"}


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if not self.proc:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
            except ValueError:
                continue
            for nm, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.path)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                   "samples": len(sm)}
        return out


def make_batches(P, agg, n, seed):
    prim = P.scenes.primary_rays(n, seed=seed)
    hits = np.empty(n, dtype=P.capi.HIT_DTYPE)
    chunk = 1 << 22
    for a in range(0, n, chunk):
        hits[a:a + chunk] = agg.IntersectClosest(prim[a:a + chunk])
    shad = P.scenes.shadow_rays(prim, hits, seed=seed + 1000)
    return prim, shad


def _bin(img, k):
    h, w = img.shape[:2]
    h2, w2 = h // k * k, w // k * k
    return np.asarray(img, np.float64)[:h2, :w2].reshape(h2 // k, k, w2 // k, k, -1).mean(axis=(1, 3))


def image_parity(img, ref, ref2=None):
    """SURVEY 8(d)'s image gate at the size the reference sample was rendered at.  With a second reference image (the
    next sampler seed): rRMSE(ours, ref) <= 1.25 x rRMSE(ref2, ref), per pixel and on 8 x 8 pixel bins (where the
    Monte-Carlo noise is 8 x smaller), and |mean difference| <= 1 %.  Without: the path-for-path gate rRMSE <= 1e-3,
    |mean difference| <= 1e-4 (the kernels consume the sampler dimensions in the CPU integrator's order)."""
    def rr(a, b):
        return float(np.sqrt(np.mean((np.asarray(a, np.float64) - b) ** 2)) / np.mean(b))
    out = {"rrmse": rr(img, ref), "rrmse_binned8": rr(_bin(img, 8), _bin(ref, 8)),
           "mean_rel_diff": float(abs(np.mean(img, dtype=np.float64) - np.mean(ref, dtype=np.float64)) / np.mean(ref, dtype=np.float64))}
    if ref2 is not None:
        out["noise_floor"] = rr(ref2, ref)
        out["noise_floor_binned8"] = rr(_bin(ref2, 8), _bin(ref, 8))
        out["gate"] = "rrmse <= 1.25 x noise_floor per pixel and on 8x8 bins, |mean diff| <= 1 %"
        out["pass"] = bool(out["rrmse"] <= 1.25 * out["noise_floor"] and
                           out["rrmse_binned8"] <= 1.25 * out["noise_floor_binned8"] and out["mean_rel_diff"] <= 0.01)
    else:
        out["gate"] = "path for path: rrmse <= 1e-3, |mean diff| <= 1e-4"
        out["pass"] = bool(out["rrmse"] <= 1e-3 and out["mean_rel_diff"] <= 1e-4)
    if not out["pass"]:
        print("bench.py: IMAGE PARITY GATE FAILED: " + json.dumps(out), file=sys.stderr, flush=True)
    return out


def gpu_image_of_tables(P, tables, local, stream, spectral=False, gbuffer=False):
    """One frame of the scene tables a host adaptor handed over (here: the reference's own dump of the very file the
    reference rendered), all of its samples, on this rank's GPU."""
    rsc = P.RenderScene(tables)
    r = P.B200Renderer(rsc, device=local, stream=stream.cuda_stream)
    try:
        r.render()
        if gbuffer:
            return r.gbuffer_image(P.scenes.GBUFFER_ALL)
        if spectral:
            return np.concatenate([r.image(), r.spectral_image()], axis=-1)
        return r.image()
    finally:
        r.close()


RENDER_CASE = "cornell_r512_s64"  # BASELINE.json configs[0]; tables in tests/golden (make_golden.py TABLE_CASES)
RENDER_KW = dict(res=512, spp=64, maxdepth=5, lightsampler="power", seed=0)


class _DevArray:
    """A device pointer as __cuda_array_interface__, so torch can alias the library's film buffer."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def render_cpu_reference(args, P, local, stream):
    """pbrt_ref (the reference's own main, CPU PathIntegrator) on a bounded sample of config 1:
    the same scene and resolution at --render-cpu-spp samples per pixel, all host cores."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refrun
    if not refrun.render_available():
        return {"value": None, "unit": "Msamples/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    kw = dict(RENDER_KW, spp=args.render_cpu_spp)
    with tempfile.TemporaryDirectory() as tmp:
        pbrt, pfm = os.path.join(tmp, "c.pbrt"), os.path.join(tmp, "c.pfm")
        P.scenes.cornell_pbrt(pbrt, out=pfm, **kw)
        ref, secs = refrun.render(pbrt, pfm)
        tables = P.scenes.load_b2rs(refrun.dump_scene(pbrt, os.path.join(tmp, "c.b2rs"))) if local is not None else None
    parity = None  # (--impl reference: the reference alone, no GPU in that process)
    if local is not None:
        parity = image_parity(gpu_image_of_tables(P, tables, local, stream), ref)
        parity["sample"] = f"{kw['res']}x{kw['res']}, {kw['spp']} spp, same sampler seed"
    n = kw["res"] * kw["res"] * kw["spp"]
    return {"value": n / secs * 1e-6, "parity": parity, "unit": "Msamples/s", "cores": os.cpu_count(), "kind": "reference",
            "seconds": secs,
            "sample": f"Cornell {kw['res']}x{kw['res']} at {kw['spp']} spp (config 1 is 64 spp), wall clock of the "
                      f"whole pbrt_ref process incl. parse + BVH build (36 triangles)"}


def render_section(args, P, torch, dist, rank, world, local, stream, barrier):
    """Config 1 end to end: Msamples/s of the whole frame (sample stripes per rank, NCCL film reduce,
    film read-back on rank 0)."""
    rsc = P.RenderScene.load(os.path.join(ROOT, "tests", "golden", RENDER_CASE + ".b2rs"))
    spp = rsc.spp
    r = P.B200Renderer(rsc, device=local, stream=stream.cuda_stream)
    ptr, nd = r.film_device_ptr()
    film_t = torch.as_tensor(_DevArray(ptr, nd, "<f8"), device="cuda")

    def frame(read):
        r.clear()
        r.render(*P.parallel.sample_stripe(spp, rank, world))
        P.parallel.reduce_film(film_t, dist)
        return r.image() if (read and rank == 0) else None

    for _ in range(args.warmup):
        frame(False)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        frame(False)
    e1.record(stream)
    barrier()
    dev_ms = e0.elapsed_time(e1)
    st = r.stats()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        img = frame(True)
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([dev_ms, e2e_s], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_s = (float(x) for x in t.cpu())
    w, h = rsc.resolution
    n = w * h * spp * args.steps
    out = None
    if rank == 0:
        out = {"metric": "render_msamples_per_s", "unit": "Msamples/s", "value": n / (dev_ms * 1e-3) * 1e-6,
               "ms_per_frame": dev_ms / args.steps, "scaling": "strong",
               "workload": f"config1: Cornell box {w}x{h}, {spp} spp, path maxdepth 5, zsobol, power light sampler; "
                           f"sample stripes over {world} rank(s), film reduced by NCCL",
               "e2e": {"value": n / e2e_s * 1e-6, "unit": "Msamples/s", "h2d_bytes_per_step": 0,
                       "d2h_bytes_per_step": w * h * 12,
                       "note": "scene tables resident; per frame: clear + render + reduce + final image to host"},
               "rays_per_frame": (st["closest_rays"] + st["shadow_rays"]) * world,
               "launches_per_frame": st["kernel_launches"], "image_mean": float(img.mean())}
        if not args.no_cpu_baseline:
            out["cpu_baseline"] = render_cpu_reference(args, P, local, stream)
    r.close()
    return out


CONFIG3_CASE = "config3_full"  # BASELINE.json configs[2]; tables in tests/golden (make_golden.py CONFIG3_FULL)


def config3_rays(P, rsc, n, seed):
    """Primary rays of the scene's own perspective camera (jittered over the whole frame, n of them)."""
    d = rsc.d
    rng = np.random.default_rng(seed)
    w, h = rsc.resolution
    cfr = np.asarray(d["camera_from_raster"], np.float64).reshape(4, 4)
    rfc = np.asarray(d["render_from_camera"], np.float64).reshape(4, 4)
    # generation order of a wavefront pass (GenerateKernel; wavefront/camera.cpp:32-80): sample after sample, each
    # in raster order, one jittered position per pixel
    pix = np.arange(n, dtype=np.int64) % (w * h)
    pf = np.stack([pix % w + rng.uniform(0, 1, n), pix // w + rng.uniform(0, 1, n), np.zeros(n), np.ones(n)], axis=1)
    pc = pf @ cfr.T
    pc = pc[:, :3] / pc[:, 3:4]
    dirs = pc / np.linalg.norm(pc, axis=1, keepdims=True)
    rays = np.empty((n, 7), np.float32)
    rays[:, 0:3] = rfc[:3, 3]
    rays[:, 3:6] = dirs @ rfc[:3, :3].T
    rays[:, 6] = np.inf
    return rays


def config3_shadow_rays(P, rsc, prim, hits, seed):
    """From each primary hit towards a random emitter of the scene (unnormalised d, tMax = 1 - 1e-4, as
    pbrt's SpawnRayTo); rays that escaped start at a random point above the ground instead."""
    d = rsc.d
    rng = np.random.default_rng(seed)
    n = len(prim)
    li = d["lights_i"]
    li = li[li[:, 0] >= 0]
    first = np.concatenate([[0], np.cumsum(d["mesh_nverts"].astype(np.int64))[:-1]])
    tfirst = np.concatenate([[0], np.cumsum(d["mesh_ntris"].astype(np.int64))[:-1]])
    tri = d["indices"][tfirst[li[:, 0]] + li[:, 1]] + first[li[:, 0]][:, None]
    cent = d["P"][tri].mean(axis=1)  # emitters are world-level meshes: render-space vertices
    target = cent[rng.integers(0, len(cent), n)]
    miss = hits["mesh"] == 0xFFFFFFFF
    t = np.where(miss, 0.0, hits["t"]).astype(np.float32)
    o = prim[:, 0:3] + prim[:, 3:6] * t[:, None]
    # (render space is camera-world: the world origin sits at -eye = (0, -14, 48))
    o[miss] = np.stack([rng.uniform(-30, 30, miss.sum()), rng.uniform(0.2, 10, miss.sum()) - 14.0,
                        rng.uniform(-30, 30, miss.sum()) + 48.0], axis=1)
    dirn = target - o
    o = o + dirn * 1e-4  # step off the surface
    rays = np.empty((n, 7), np.float32)
    rays[:, 0:3] = o
    rays[:, 3:6] = target - o
    rays[:, 6] = 1 - 1e-4
    return rays


def config3_section(args, P, torch, dist, rank, world, local, stream, barrier):
    """BASELINE configs[2], the north_star's headline scene at the size SURVEY 8(d) fixes: 100 object definitions x
    100 352 triangles (10 M unique) placed 1000 times (100 M instanced), glass / metal / coated / diffuse, 20 000 emitters
    through the light BVH.  The committed tables hold the scene at a coarse tessellation; the definitions are
    re-tessellated here by the scene's own generator (scenes.config3_refine).
    (a) traversal of camera + shadow rays, weak scaling like the main line; (b) the frame at
    --config3-spp of its 1024 spp per step, sample stripes over the ranks."""
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_golden
    full = make_golden.CONFIG3_FULL
    rsc = P.RenderScene(P.scenes.config3_refine(P.scenes.load_b2rs(os.path.join(ROOT, "tests", "golden", CONFIG3_CASE + ".b2rs")),
                                                full["n_objects"], full["n_instances"], args.config3_grid))
    geo = rsc.geometry
    agg = P.B200Aggregate(geo, device=local, stream=stream.cuda_stream)
    st = agg.stats()
    n = args.rays
    prim = config3_rays(P, rsc, n, seed=11 + rank)
    hits = np.empty(n, dtype=P.capi.HIT_DTYPE)
    for a in range(0, n, 1 << 22):
        hits[a:a + (1 << 22)] = agg.IntersectClosest(prim[a:a + (1 << 22)])
    shad = config3_shadow_rays(P, rsc, prim, hits, seed=1011 + rank)
    sample = slice(0, min(n, 1 << 22))
    agg.IntersectClosest(prim[sample], counters=True)
    tc = agg.trace_stats()
    agg.IntersectShadow(shad[sample], counters=True)
    ta = agg.trace_stats()
    d_prim, d_shad = torch.from_numpy(prim).cuda(), torch.from_numpy(shad).cuda()
    d_hits = torch.empty(n * 24, dtype=torch.uint8, device="cuda")
    d_occ = torch.empty(n, dtype=torch.uint8, device="cuda")
    for _ in range(args.warmup):
        agg.intersect_closest_ptr(d_prim.data_ptr(), n, d_hits.data_ptr(), P.capi.MEM_DEVICE)
        agg.intersect_any_ptr(d_shad.data_ptr(), n, d_occ.data_ptr(), P.capi.MEM_DEVICE)
    barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * args.steps + 1)]
    ev[0].record(stream)
    for k in range(args.steps):
        agg.intersect_closest_ptr(d_prim.data_ptr(), n, d_hits.data_ptr(), P.capi.MEM_DEVICE)
        ev[2 * k + 1].record(stream)
        agg.intersect_any_ptr(d_shad.data_ptr(), n, d_occ.data_ptr(), P.capi.MEM_DEVICE)
        ev[2 * k + 2].record(stream)
    barrier()
    ms_closest = float(np.mean([ev[2 * k].elapsed_time(ev[2 * k + 1]) for k in range(args.steps)]))
    ms_any = float(np.mean([ev[2 * k + 1].elapsed_time(ev[2 * k + 2]) for k in range(args.steps)]))
    hit_frac = float((np.frombuffer(d_hits.cpu().numpy(), dtype=P.capi.HIT_DTYPE)["mesh"] != 0xFFFFFFFF).mean())
    occ_frac = float(d_occ.float().mean().item())
    del d_prim, d_shad, d_hits, d_occ
    agg.close()

    # ---- the frame
    spp = min(args.config3_spp, rsc.spp)
    r = P.B200Renderer(rsc, device=local, stream=stream.cuda_stream)
    ptr, nd = r.film_device_ptr()
    film_t = torch.as_tensor(_DevArray(ptr, nd, "<f8"), device="cuda")

    def frame(read):
        r.clear()
        r.render(*P.parallel.sample_stripe(spp, rank, world))
        P.parallel.reduce_film(film_t, dist)
        return r.image() if (read and rank == 0) else None

    for _ in range(min(args.warmup, 3)):
        frame(False)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        frame(False)
    e1.record(stream)
    barrier()
    dev_ms = e0.elapsed_time(e1)
    rst = r.stats()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        img = frame(True)
    barrier()
    e2e_s = time.perf_counter() - t0
    r.close()
    t = torch.tensor([ms_closest, ms_any, dev_ms, e2e_s], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_closest, ms_any, dev_ms, e2e_s = (float(x) for x in t.cpu())
    if rank != 0:
        return None
    w, h = rsc.resolution
    ns = w * h * spp * args.steps
    peak, _ = peaks()
    bytes_closest = traversal_bytes(tc)
    per_object = {}
    for m in geo.meshes:
        if m.object >= 0:
            per_object[m.object] = per_object.get(m.object, 0) + len(m.indices)
    ntri = int(sum(len(m.indices) for m in geo.meshes if m.object < 0) + sum(per_object[i.object] for i in geo.instances))
    nunique = int(sum(len(m.indices) for m in geo.meshes))
    out = {"workload": f"config3: {nunique} unique / {ntri} instanced triangles ({len(geo.instances)} placements of "
                       f"{len(set(i.object for i in geo.instances))} object definitions), {len(rsc.d['lights_i'])} lights, "
                       f"{w}x{h}; traversal: {n} camera rays (jittered, in the order a pass generates them: raster order, sample after sample) + {n} shadow rays from their hits towards random emitters per step per rank",
           "triangles": ntri, "unique_triangles": nunique,
           "traversal": {"metric": METRIC, "unit": UNIT,
                         "value": 2 * n * world / ((ms_closest + ms_any) * 1e-3) * 1e-6,
                         "closest_mrays_per_s": n * world / (ms_closest * 1e-3) * 1e-6,
                         "any_mrays_per_s": n * world / (ms_any * 1e-3) * 1e-6,
                         "box_tests_per_ray": tc["nodes_visited"] / tc["rays"], "prims_per_ray": tc["prims_tested"] / tc["rays"],
                         "wide_nodes_per_ray": tc["wide_nodes_visited"] / tc["rays"],
                         "leaf_boxes_per_ray": tc["leaf_boxes_tested"] / tc["rays"],
                         "box_tests_per_ray_any": ta["nodes_visited"] / ta["rays"],
                         "bytes_per_ray": bytes_closest,
                         "roofline_frac": bytes_closest * n / (ms_closest * 1e-3) * 1e-9 / peak,
                         "hit_fraction": hit_frac, "occluded_fraction": occ_frac,
                         "device_bytes": st["device_bytes"], "build_s": st["build_seconds"]},
           "render": {"metric": "render_msamples_per_s", "unit": "Msamples/s", "value": ns / (dev_ms * 1e-3) * 1e-6,
                      "ms_per_frame": dev_ms / args.steps, "scaling": "strong",
                      "sample": f"{spp} of the frame's {rsc.spp} spp per step (sample indices 0..{spp - 1}), path maxdepth 5, "
                                f"bvh light sampler; sample stripes over {world} rank(s), film reduced by NCCL",
                      "e2e": {"value": ns / e2e_s * 1e-6, "unit": "Msamples/s", "h2d_bytes_per_step": 0,
                              "d2h_bytes_per_step": w * h * 12},
                      "rays_per_frame": (rst["closest_rays"] + rst["shadow_rays"]) * world,
                      "launches_per_frame": rst["kernel_launches"], "image_mean": float(img.mean())}}
    if not args.no_cpu_baseline:
        out["render"]["cpu_baseline"] = config3_cpu_reference(args, P, local, stream)
        out["traversal"]["parity"] = config3_traversal_parity(args, P, geo, prim, shad, local, stream)
    return out


def config3_traversal_parity(args, P, geo, prim, shad, local, stream):
    """Hit records of a sample of the timed batches against the reference's BVHAggregate (raycast_ref builds its own
    two-level BVH over the 10 M unique triangles): 0 mismatches in (mesh, prim, instance, t bits, b1, b2)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refrun
    from conftest import hit_mismatches
    if not refrun.available():
        return {"mismatches": None, "sample": "oracle/_ref not built"}
    m = min(args.parity_sample, len(prim))
    rs = refrun.RefScene(geo)
    agg = P.B200Aggregate(geo, device=local, stream=stream.cuda_stream)
    ref, _ = rs.closest(prim[:m])
    bad = int(hit_mismatches(agg.IntersectClosest(prim[:m]), ref).sum())
    occ, _ = rs.any(shad[:m])
    bad_any = int((agg.IntersectShadow(shad[:m]) != occ).sum())
    agg.close()
    if bad or bad_any:
        print(f"bench.py: CONFIG 3 TRAVERSAL PARITY FAILED: {bad} closest / {bad_any} any-hit mismatches", file=sys.stderr, flush=True)
    return {"mismatches": bad + bad_any, "rays": 2 * m,
            "sample": f"first {m} camera + {m} shadow rays of the timed batches vs oracle/_ref/raycast_ref, bit-exact compare"}


def config3_cpu_reference(args, P, local, stream):
    """pbrt_ref on a bounded sample of config 3: the whole full-size scene (object definitions as binary PLY) and frame at
    --config3-cpu-spp samples per pixel, sampler seeds 0 and 1; the GPU renders the reference's own dump of the seed-0
    file and is held to SURVEY 8(d)'s noise-floor gate."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refrun
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    if not refrun.render_available():
        return {"value": None, "unit": "Msamples/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    import make_golden
    kw = dict(make_golden.CONFIG3_FULL, grid=args.config3_grid, spp=args.config3_cpu_spp)
    with tempfile.TemporaryDirectory() as tmp:
        pbrt, pfm = os.path.join(tmp, "c3.pbrt"), os.path.join(tmp, "c3.pfm")
        os.makedirs(os.path.join(tmp, "ply"))
        P.scenes.config3_pbrt(pbrt, out=pfm, ply_dir=os.path.join(tmp, "ply"), **kw)
        ref, secs = refrun.render(pbrt, pfm)
        tables = P.scenes.load_b2rs(refrun.dump_scene(pbrt, os.path.join(tmp, "c3.b2rs"), bvh=False))
        P.scenes.config3_pbrt(pbrt, out=pfm, ply_dir=os.path.join(tmp, "ply"), **dict(kw, seed=1))
        ref2, _ = refrun.render(pbrt, pfm)
    parity = image_parity(gpu_image_of_tables(P, tables, local, stream), ref, ref2)
    parity["sample"] = f"{kw['res'][0]}x{kw['res'][1]}, {kw['spp']} spp, seeds 0 / 1 for the floor"
    n = kw["res"][0] * kw["res"][1] * kw["spp"]
    return {"value": n / secs * 1e-6, "unit": "Msamples/s", "cores": os.cpu_count(), "kind": "reference",
            "seconds": secs, "parity": parity,
            "sample": f"the full-size config-3 scene at {kw['res'][0]}x{kw['res'][1]}, {kw['spp']} spp (config 3 is 1024 spp), wall "
                      f"clock of the whole pbrt_ref process incl. parse + BVH builds over 10 M triangles"}


CONFIG4_CASE = "config4_full"  # BASELINE.json configs[3]; tables in tests/golden (make_golden.py CONFIG4_FULL)


def config4_section(args, P, torch, dist, rank, world, local, stream, barrier):
    """Config 4 end to end: RealisticCamera rays through the lens table, path tracing, the ISET GBufferFilm (RGB + 31
    radiance channels); --config4-spp of its 256 spp per step, sample stripes over the ranks, the film's RGB
    sums + 31 radiance planes (fp64) reduced by NCCL, the 34-channel image read back into pinned memory on rank 0."""
    rsc = P.RenderScene.load(os.path.join(ROOT, "tests", "golden", CONFIG4_CASE + ".b2rs"))
    spp = min(args.config4_spp, rsc.spp)
    r = P.B200Renderer(rsc, device=local, stream=stream.cuda_stream)
    ptr, nd = r.film_device_ptr()
    film_t = torch.as_tensor(_DevArray(ptr, nd, "<f8"), device="cuda")
    w, h = rsc.resolution
    gbuffer = int(rsc.d["film_type"][0]) == 2
    # the image this frame delivers needs the RGB sums and the radiance planes; the AOV planes behind them stay local
    film_used = film_t[:w * h * (4 + rsc.n_buckets)] if gbuffer else film_t
    pinned = torch.empty((h, w, 3 + rsc.n_buckets if gbuffer else rsc.n_buckets), dtype=torch.float32).pin_memory()

    def frame(read):
        r.clear()
        r.render(*P.parallel.sample_stripe(spp, rank, world))
        P.parallel.reduce_film(film_used, dist)
        if not (read and rank == 0):
            return None
        return r.gbuffer_image(out=pinned.numpy()) if gbuffer else r.spectral_image()

    for _ in range(min(args.warmup, 3)):
        frame(False)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        frame(False)
    e1.record(stream)
    barrier()
    dev_ms = e0.elapsed_time(e1)
    st = r.stats()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        img = frame(True)
    barrier()
    e2e_s = time.perf_counter() - t0
    r.close()
    t = torch.tensor([dev_ms, e2e_s], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_s = (float(x) for x in t.cpu())
    if rank != 0:
        return None
    ns = w * h * spp * args.steps
    out = {"metric": "render_msamples_per_s", "unit": "Msamples/s", "value": ns / (dev_ms * 1e-3) * 1e-6,
           "ms_per_frame": dev_ms / args.steps, "scaling": "strong",
           "workload": f"config4: Cornell room + 24-patch spectral chart, realistic camera ({len(rsc.d['camera_lens'])} lens "
                       f"interfaces, double-Gauss 50 mm), {'gbuffer' if gbuffer else 'spectral'} film {rsc.n_buckets} bands, {w}x{h}; {spp} of the frame's "
                       f"{rsc.spp} spp per step, path maxdepth 5; sample stripes over {world} rank(s), film reduced by NCCL",
           "e2e": {"value": ns / e2e_s * 1e-6, "unit": "Msamples/s", "h2d_bytes_per_step": 0,
                   "d2h_bytes_per_step": int(pinned.numel()) * 4},
           "film_bytes_reduced": int(film_used.numel()) * 8,
           "rays_per_frame": (st["closest_rays"] + st["shadow_rays"]) * world,
           "launches_per_frame": st["kernel_launches"], "image_mean": float(img.mean())}
    if not args.no_cpu_baseline:
        out["cpu_baseline"] = config4_cpu_reference(args, P, local, stream)
    return out


def config4_cpu_reference(args, P, local, stream):
    """The reference's render loop + Film::GetImage (dumpscene_ref --render: the oracle build has no EXR writer,
    which pbrt_ref's spectral film would need) on a bounded sample of config 4: the whole frame at
    --config4-cpu-spp samples per pixel, all host cores."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refrun
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    if not refrun.render_available():
        return {"value": None, "unit": "Msamples/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    import make_golden
    kw = dict(make_golden.CONFIG4_FULL, spp=args.config4_cpu_spp)
    with tempfile.TemporaryDirectory() as tmp:
        pbrt = os.path.join(tmp, "c4.pbrt")
        P.scenes.config4_pbrt(pbrt, out=os.path.join(tmp, "c4.exr"), film="gbuffer", **kw)
        ref, _, secs = refrun.render_channels(pbrt)
        tables = P.scenes.load_b2rs(refrun.dump_scene(pbrt, os.path.join(tmp, "c4.b2rs")))
    # path for path (no smooth dielectric in this scene): R G B + Radiance.C01-31 of GBufferFilm::GetImage, and its AOVs
    img = gpu_image_of_tables(P, tables, local, stream, gbuffer=True)
    parity = image_parity(img[..., :34], ref[..., :34])
    aov = {"position": slice(34, 37), "ns": slice(39, 42), "n": slice(42, 45), "variance": slice(45, 48)}
    parity["aov_rrmse"] = {k: float(np.sqrt(np.mean((img[..., s].astype(np.float64) - ref[..., s]) ** 2)) /
                                    max(float(np.mean(np.abs(ref[..., s]))), 1e-12)) for k, s in aov.items()}
    parity["id_mismatches"] = int((img[..., 37:39] != ref[..., 37:39]).sum())
    # (the variance is a second moment of 4 samples: it doubles the image's own rRMSE; positions and normals are exact)
    parity["pass"] = bool(parity["pass"] and max(v for k, v in parity["aov_rrmse"].items() if k != "variance") <= 1e-3 and
                          parity["aov_rrmse"]["variance"] <= 5e-3)
    parity["sample"] = f"{kw['res']}x{kw['res']}, {kw['spp']} spp, R G B + 31 radiance channels; AOV groups apart"
    n = kw["res"] * kw["res"] * kw["spp"]
    return {"value": n / secs * 1e-6, "parity": parity, "unit": "Msamples/s", "cores": os.cpu_count(), "kind": "reference",
            "seconds": secs,
            "sample": f"the config-4 scene at {kw['res']}x{kw['res']}, {kw['spp']} spp (config 4 is 256 spp), wall clock of "
                      f"the whole process incl. parse and the camera's exit-pupil table (64 x 2^20 lens traces)"}


CONFIG5_CASE = "config5_full"  # BASELINE.json configs[4]; frame tables in tests/golden, the 256^3 grid is made here


def config5_section(args, P, torch, dist, rank, world, local, stream, barrier):
    """Config 5 end to end: VolPathIntegrator through a homogeneous fog and a --config5-grid^3 density grid inside
    an interface shell, 1920x1080; --config5-spp of its 512 spp per step, sample stripes over the ranks."""
    d = P.scenes.load_b2rs(os.path.join(ROOT, "tests", "golden", CONFIG5_CASE + ".b2rs"))
    rsc = P.RenderScene(P.scenes.with_density_grid(d, P.scenes.procedural_density(args.config5_grid)))
    spp = min(args.config5_spp, rsc.spp)
    r = P.B200Renderer(rsc, device=local, stream=stream.cuda_stream)
    ptr, nd = r.film_device_ptr()
    film_t = torch.as_tensor(_DevArray(ptr, nd, "<f8"), device="cuda")

    def frame(read):
        r.clear()
        r.render(*P.parallel.sample_stripe(spp, rank, world))
        P.parallel.reduce_film(film_t, dist)
        return r.image() if (read and rank == 0) else None

    for _ in range(min(args.warmup, 3)):
        frame(False)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        frame(False)
    e1.record(stream)
    barrier()
    dev_ms = e0.elapsed_time(e1)
    st = r.stats()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        img = frame(True)
    barrier()
    e2e_s = time.perf_counter() - t0
    r.close()
    t = torch.tensor([dev_ms, e2e_s], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_s = (float(x) for x in t.cpu())
    if rank != 0:
        return None
    w, h = rsc.resolution
    ns = w * h * spp * args.steps
    g = args.config5_grid
    out = {"metric": "render_msamples_per_s", "unit": "Msamples/s", "value": ns / (dev_ms * 1e-3) * 1e-6,
           "ms_per_frame": dev_ms / args.steps, "scaling": "strong",
           "workload": f"config5: volpath maxdepth 5 in the Cornell room, homogeneous fog + {g}^3 procedural density grid "
                       f"(16^3 majorants) inside an interface shell, area light, {w}x{h}; {spp} of the frame's {rsc.spp} spp "
                       f"per step; sample stripes over {world} rank(s), film reduced by NCCL",
           "e2e": {"value": ns / e2e_s * 1e-6, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": w * h * 12},
           "rays_per_frame": (st["closest_rays"] + st["shadow_rays"]) * world,
           "launches_per_frame": st["kernel_launches"], "image_mean": float(img.mean())}
    if not args.no_cpu_baseline:
        out["cpu_baseline"] = config5_cpu_reference(args, P, local, stream)
    return out


def config5_cpu_reference(args, P, local, stream):
    """pbrt_ref (CPU VolPathIntegrator) on a bounded sample of config 5: the same scene and grid, a smaller frame at
    --config5-cpu-spp; the parse of the inline grid (16.7 M values) is timed apart with an 8x8, 1 spp run and
    subtracted."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refrun
    if not refrun.render_available():
        return {"value": None, "unit": "Msamples/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    w, h = args.config5_cpu_res
    dens = P.scenes.procedural_density(args.config5_grid)
    with tempfile.TemporaryDirectory() as tmp:
        pbrt, pfm = os.path.join(tmp, "c5.pbrt"), os.path.join(tmp, "c5.pfm")
        P.scenes.config5_pbrt(pbrt, out=pfm, res=(8, 8), spp=1, grid=args.config5_grid, density=dens)
        _, setup = refrun.render(pbrt, pfm)
        P.scenes.config5_pbrt(pbrt, out=pfm, res=(w, h), spp=args.config5_cpu_spp, grid=args.config5_grid, density=dens)
        ref, secs = refrun.render(pbrt, pfm)
        # tables of the same frame from the reference's parser (small grid: the dump keeps the frame, the full grid is
        # swapped in as in the timed section), and the next seed for the noise floor
        P.scenes.config5_pbrt(pbrt, out=pfm, res=(w, h), spp=args.config5_cpu_spp, grid=16)
        tables = P.scenes.with_density_grid(P.scenes.load_b2rs(refrun.dump_scene(pbrt, os.path.join(tmp, "c5.b2rs"))), dens)
        P.scenes.config5_pbrt(pbrt, out=pfm, res=(w, h), spp=args.config5_cpu_spp, grid=args.config5_grid, density=dens, seed=1)
        ref2, _ = refrun.render(pbrt, pfm)
    parity = image_parity(gpu_image_of_tables(P, tables, local, stream), ref, ref2)
    parity["sample"] = f"{w}x{h}, {args.config5_cpu_spp} spp, seeds 0 / 1 for the floor"
    n = w * h * args.config5_cpu_spp
    render = max(secs - setup, 1e-3)
    return {"value": n / render * 1e-6, "unit": "Msamples/s", "cores": os.cpu_count(), "kind": "reference",
            "seconds": render, "setup_seconds": setup, "parity": parity,
            "sample": f"the config-5 scene ({args.config5_grid}^3 grid) at {w}x{h}, {args.config5_cpu_spp} spp (config 5 is "
                      f"1920x1080, 512 spp); wall clock of pbrt_ref minus that of an 8x8 1-spp run (parse + majorant grid)"}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU BVHAggregate on all host cores, bounded sample per step."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refrun
    import pbrt_v4_b200 as P
    n = min(args.rays, args.cpu_sample)
    sc = P.scenes.random_soup(args.tris, seed=12111)
    prim = P.scenes.primary_rays(n, seed=1)
    rs = refrun.RefScene(sc)
    hits, info0 = rs.closest(prim)  # also the first warm-up; gives shadow origins
    gold = np.zeros(n, dtype=P.capi.HIT_DTYPE)
    gold["mesh"], gold["t"] = hits["mesh"], hits["t"]
    shad = P.scenes.shadow_rays(prim, gold, seed=1001)
    # raycast_ref repeats the batch `reps` times inside one process and reports the best pass;
    # we want K timed steps after W warm-ups: run reps = W + K and use per-pass best as the step time
    # (the tree build, ~1 s, is outside the timer as for the GPU arm).
    reps = args.warmup + args.steps
    _, ic = rs.closest(prim, nthreads=0, reps=reps)
    _, ia = rs.any(shad, nthreads=0, reps=reps)
    sec = ic["seconds"] + ia["seconds"]
    val = 2 * n / sec * 1e-6
    cores = ic["threads"]
    line = {"metric": METRIC, "value": val, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args),
            "closest_mrays_per_s": ic["mrays_per_s"], "any_mrays_per_s": ia["mrays_per_s"],
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "reference",
                             "sample": f"{n} primary + {n} shadow rays of the same workload, best of "
                                       f"{reps} passes, BVH build excluded ({ic['build_s']:.2f} s)"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if not args.no_render:
        rc = render_cpu_reference(args, P, None, None)
        line["render"] = {"metric": "render_msamples_per_s", "unit": "Msamples/s", "value": rc["value"],
                          "cpu_baseline": rc}
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    import torch
    import pbrt_v4_b200 as P
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the b200pt path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- scene (replicated) and this rank's batches
    sc = P.scenes.random_soup(args.tris, seed=12111)
    stream = torch.cuda.Stream()  # the library launches on this stream; torch events are recorded on it
    torch.cuda.set_stream(stream)
    agg = P.B200Aggregate(sc, device=local, stream=stream.cuda_stream)
    st = agg.stats()
    n = args.rays
    prim, shad = make_batches(P, agg, n, seed=1 + rank)

    # algorithmic bytes per ray from the counting variant (outside the timed region)
    sample = slice(0, min(n, 1 << 22))
    agg.IntersectClosest(prim[sample], counters=True)
    tc = agg.trace_stats()
    agg.IntersectShadow(shad[sample], counters=True)
    ta = agg.trace_stats()
    bytes_closest, bytes_any = traversal_bytes(tc), traversal_bytes(ta, any_hit=True)

    # ---- device-resident timing (value)
    d_prim = torch.from_numpy(prim).cuda()
    d_shad = torch.from_numpy(shad).cuda()
    d_hits = torch.empty(n * 24, dtype=torch.uint8, device="cuda")
    d_occ = torch.empty(n, dtype=torch.uint8, device="cuda")

    def step_device():
        agg.intersect_closest_ptr(d_prim.data_ptr(), n, d_hits.data_ptr(), P.capi.MEM_DEVICE)
        agg.intersect_any_ptr(d_shad.data_ptr(), n, d_occ.data_ptr(), P.capi.MEM_DEVICE)

    for _ in range(args.warmup):
        step_device()
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3 * args.steps + 1)]
    ev[0].record(stream)
    for k in range(args.steps):
        agg.intersect_closest_ptr(d_prim.data_ptr(), n, d_hits.data_ptr(), P.capi.MEM_DEVICE)
        ev[3 * k + 1].record(stream)
        agg.intersect_any_ptr(d_shad.data_ptr(), n, d_occ.data_ptr(), P.capi.MEM_DEVICE)
        ev[3 * k + 2].record(stream)
        ev[3 * k + 3].record(stream)
    barrier()
    total_ms = ev[0].elapsed_time(ev[3 * args.steps])
    ms_closest = float(np.mean([ev[3 * k].elapsed_time(ev[3 * k + 1]) for k in range(args.steps)]))
    ms_any = float(np.mean([ev[3 * k + 1].elapsed_time(ev[3 * k + 2]) for k in range(args.steps)]))
    clk = clocks.stop() if rank == 0 else None

    # ---- end to end through the C ABI with pinned host buffers (e2e)
    h_prim = torch.from_numpy(prim).pin_memory()
    h_shad = torch.from_numpy(shad).pin_memory()
    h_hits = torch.empty(n * 24, dtype=torch.uint8).pin_memory()
    h_occ = torch.empty(n, dtype=torch.uint8).pin_memory()

    def step_host():
        agg.intersect_closest_ptr(h_prim.data_ptr(), n, h_hits.data_ptr(), P.capi.MEM_HOST)
        agg.intersect_any_ptr(h_shad.data_ptr(), n, h_occ.data_ptr(), P.capi.MEM_HOST)

    for _ in range(max(1, args.warmup // 2)):
        step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_host()
    barrier()
    e2e_s = time.perf_counter() - t0
    # the e2e result really is the device result
    if rank == 0:
        assert np.array_equal(h_hits.numpy(), d_hits.cpu().numpy()) and np.array_equal(h_occ.numpy(), d_occ.cpu().numpy())

    cpu_base = None
    if rank == 0 and not args.no_cpu_baseline:
        cpu_base = cpu_baseline(args, P, sc, prim, shad, agg)

    render = None
    if not args.no_render:
        render = render_section(args, P, torch, dist, rank, world, local, stream, barrier)

    config3 = None
    if not args.no_config3:
        agg.close()
        agg = None
        del d_prim, d_shad, d_hits, d_occ
        config3 = config3_section(args, P, torch, dist, rank, world, local, stream, barrier)

    config4 = None
    if not args.no_config4:
        if agg is not None:
            agg.close()
            agg = None
        config4 = config4_section(args, P, torch, dist, rank, world, local, stream, barrier)

    config5 = None
    if not args.no_config5:
        if agg is not None:
            agg.close()
            agg = None
        config5 = config5_section(args, P, torch, dist, rank, world, local, stream, barrier)

    # ---- max over ranks
    t = torch.tensor([total_ms, e2e_s, ms_closest, ms_any], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_s, ms_closest, ms_any = (float(x) for x in t.cpu())

    if rank == 0:
        peak, peak_src = peaks()
        rays_total = 2.0 * n * world * args.steps
        value = rays_total / (total_ms * 1e-3) * 1e-6
        ach = bytes_closest * n / (ms_closest * 1e-3) * 1e-9
        wide = tc["wide_nodes_visited"] > 0
        cap = headline_ncu("TraverseWideKernel<0" if wide else "TraverseBinaryKernel<0") if args.tris == 1_000_000 else None
        roofline = {
            # what bounds the kernel according to ncu (profiles/): instruction issue at ~45 % lane occupancy plus the L1
            # tag stage (one wavefront per lane per 32-byte node fetch); the tree is L2-resident, HBM is almost idle
            "bound": "issue", "kernel": "closest-hit traversal (%s walk)" % ("8-wide" if wide else "binary"),
            # contract figure: ALGORITHMIC bytes per launch / CUDA-event time, against the measured HBM copy peak.  These
            # bytes are served by L2, so this is an L2-level rate; `frac_dram` below is the HBM-level one.
            "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "peak_source": peak_src,
            "bytes_per_ray": bytes_closest, "bytes_per_ray_any": bytes_any, "compulsory_bytes_per_ray": 52,
            "box_tests_per_ray": tc["nodes_visited"] / tc["rays"],
            "nodes_per_ray": (tc["wide_nodes_visited"] if wide else tc["nodes_visited"] / 2) / tc["rays"],
            "node_bytes": 256 if wide else 64,
            "leaf_boxes_per_ray": tc["leaf_boxes_tested"] / tc["rays"],
            "prims_per_ray": tc["prims_tested"] / tc["rays"],
            "traffic": None, "frac_dram": None,
        }
        if cap:
            dram = cap["dram_bytes_per_ray"]
            roofline.update({
                # dram__bytes_read.sum + dram__bytes_write.sum of this kernel per ray (ncu --set full capture of the same
                # kernel on the same scene, file below) x this launch's rays
                "traffic": dram * n, "traffic_bytes_per_ray": dram, "traffic_source": os.path.relpath(HEADLINE_NCU, ROOT),
                "frac_dram": dram * n / (ms_closest * 1e-3) * 1e-9 / peak,
                "issue_slots_busy": cap.get("issue_active_pct", 0.0) / 100.0,
                "active_threads_per_warp": cap.get("active_threads_per_warp"),
                "issue_x_lanes": cap.get("issue_active_pct", 0.0) / 100.0 * (cap.get("active_threads_per_warp") or 0.0) / 32.0,
                "l1tex_throughput_pct": cap.get("l1tex_throughput_pct"), "l2_throughput_pct": cap.get("l2_throughput_pct"),
                "warp_instructions_per_ray": cap.get("warp_instructions_per_ray")})
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args),
            "closest_mrays_per_s": n * world / (ms_closest * 1e-3) * 1e-6,
            "any_mrays_per_s": n * world / (ms_any * 1e-3) * 1e-6,
            "roofline": roofline,
            "e2e": {"value": rays_total / e2e_s * 1e-6, "unit": UNIT,
                    "h2d_bytes_per_step": 2 * n * 28, "d2h_bytes_per_step": n * 24 + n},
            "gpu_launches": 2 * args.steps,
            "render": render,
            "config3": config3,
            "config4": config4,
            "config5": config5,
            "clocks": clk,
            "scene": {"bvh2_nodes": st["n_bvh2_nodes"], "wide_nodes": st["n_wide_nodes"],
                      "device_bytes": st["device_bytes"], "build_s": st["build_seconds"]},
        }
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
            line["parity"] = cpu_base.pop("parity")
        print(json.dumps(line), flush=True)
    if agg is not None:
        agg.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def cpu_baseline(args, P, sc, prim, shad, agg):
    """The reference's own BVHAggregate (oracle/_ref/raycast_ref) on a bounded sample, all host cores; its answers are
    also the bit-exactness check of the GPU's records for the same rays (`parity`)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refrun
    if not refrun.available():
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "oracle/_ref not built",
                "parity": {"mismatches": None}}
    from conftest import hit_mismatches
    m = min(args.cpu_sample, prim.shape[0])
    rs = refrun.RefScene(sc)
    ref, ic = rs.closest(prim[:m], nthreads=0, reps=2)
    occ, ia = rs.any(shad[:m], nthreads=0, reps=2)
    bad = int(hit_mismatches(agg.IntersectClosest(prim[:m]), ref).sum()) + int((agg.IntersectShadow(shad[:m]) != occ).sum())
    adv = P.scenes.adversarial_rays(sc, min(args.parity_sample, 1 << 20), seed=300)
    ref_adv, _ = rs.closest(adv)
    bad += int(hit_mismatches(agg.IntersectClosest(adv), ref_adv).sum())
    if bad:
        print(f"bench.py: TRAVERSAL PARITY FAILED: {bad} mismatches against the reference", file=sys.stderr, flush=True)
    sec = ic["seconds"] + ia["seconds"]
    return {"value": 2 * m / sec * 1e-6, "unit": UNIT, "cores": ic["threads"], "kind": "reference",
            "closest_mrays_per_s": ic["mrays_per_s"], "any_mrays_per_s": ia["mrays_per_s"],
            "sample": f"first {m} primary + {m} shadow rays of the timed batches, best of 2 passes, "
                      f"BVH build excluded ({ic['build_s']:.2f} s)",
            "parity": {"mismatches": bad, "rays": 2 * m + len(adv),
                       "sample": f"the CPU sample's {m} primary + {m} shadow rays and {len(adv)} adversarial rays (through vertices / "
                                 f"edges, axis-aligned, origins in bounding planes): (mesh, prim, instance, t bits, b1, b2) and "
                                 f"occlusion bits against BVHAggregate::Intersect / IntersectP"}}


if __name__ == "__main__":
    main()
