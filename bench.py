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


DRAM_BYTES_PER_RAY_CLOSEST = (692695040 + 109081856) / (1 << 22)  # profiles/r01_headline_traffic_ncu.csv


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


RENDER_CASE = "cornell_r512_s64"  # BASELINE.json configs[0]; tables in tests/golden (make_golden.py TABLE_CASES)
RENDER_KW = dict(res=512, spp=64, maxdepth=5, lightsampler="power", seed=0)


class _DevArray:
    """A device pointer as __cuda_array_interface__, so torch can alias the library's film buffer."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def render_cpu_reference(args, P):
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
        _, secs = refrun.render(pbrt, pfm)
    n = kw["res"] * kw["res"] * kw["spp"]
    return {"value": n / secs * 1e-6, "unit": "Msamples/s", "cores": os.cpu_count(), "kind": "reference",
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
            out["cpu_baseline"] = render_cpu_reference(args, P)
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
    """BASELINE configs[2], the north_star's headline scene: ~10 M instanced triangles (32 definitions x 2048
    triangles x 4883 placements), glass / metal / coated / diffuse, 2048 emitters through the light BVH.
    (a) traversal of camera + shadow rays, weak scaling like the main line; (b) the frame at
    --config3-spp of its 1024 spp per step, sample stripes over the ranks."""
    rsc = P.RenderScene.load(os.path.join(ROOT, "tests", "golden", CONFIG3_CASE + ".b2rs"))
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
    NODE_B, PRIM_B = 32, 48
    bytes_closest = 28 + 24 + (NODE_B * tc["nodes_visited"] + PRIM_B * tc["prims_tested"]) / tc["rays"]
    ntri = int(sum(len(m.indices) for m in geo.meshes if m.object < 0) +
               sum(sum(len(m.indices) for m in geo.meshes if m.object == i.object) for i in geo.instances))
    out = {"workload": f"config3: {ntri} instanced triangles ({len(geo.instances)} placements of "
                       f"{len(set(i.object for i in geo.instances))} object definitions), {len(rsc.d['lights_i'])} lights, "
                       f"{w}x{h}; traversal: {n} camera rays (jittered, in the order a pass generates them: raster order, sample after sample) + {n} shadow rays from their hits towards random emitters per step per rank",
           "triangles": ntri,
           "traversal": {"metric": METRIC, "unit": UNIT,
                         "value": 2 * n * world / ((ms_closest + ms_any) * 1e-3) * 1e-6,
                         "closest_mrays_per_s": n * world / (ms_closest * 1e-3) * 1e-6,
                         "any_mrays_per_s": n * world / (ms_any * 1e-3) * 1e-6,
                         "nodes_per_ray": tc["nodes_visited"] / tc["rays"], "prims_per_ray": tc["prims_tested"] / tc["rays"],
                         "nodes_per_ray_any": ta["nodes_visited"] / ta["rays"],
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
        out["render"]["cpu_baseline"] = config3_cpu_reference(args, P)
    return out


def config3_cpu_reference(args, P):
    """pbrt_ref on a bounded sample of config 3: the whole scene and frame at --config3-cpu-spp samples per pixel."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refrun
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    if not refrun.render_available():
        return {"value": None, "unit": "Msamples/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    import make_golden
    kw = dict(make_golden.CONFIG3_FULL, spp=args.config3_cpu_spp)
    with tempfile.TemporaryDirectory() as tmp:
        pbrt, pfm = os.path.join(tmp, "c3.pbrt"), os.path.join(tmp, "c3.pfm")
        P.scenes.config3_pbrt(pbrt, out=pfm, **kw)
        _, secs = refrun.render(pbrt, pfm)
    n = kw["res"][0] * kw["res"][1] * kw["spp"]
    return {"value": n / secs * 1e-6, "unit": "Msamples/s", "cores": os.cpu_count(), "kind": "reference",
            "seconds": secs,
            "sample": f"the config-3 scene at {kw['res'][0]}x{kw['res'][1]}, {kw['spp']} spp (config 3 is 1024 spp), wall "
                      f"clock of the whole pbrt_ref process incl. parse + BVH builds"}


CONFIG4_CASE = "config4_full"  # BASELINE.json configs[3]; tables in tests/golden (make_golden.py CONFIG4_FULL)


def config4_section(args, P, torch, dist, rank, world, local, stream, barrier):
    """Config 4 end to end: RealisticCamera rays through the lens table, path tracing, 31-band SpectralFilm;
    --config4-spp of its 256 spp per step, sample stripes over the ranks, the whole film (RGB sums + 2 x 31
    bucket planes, fp64) reduced by NCCL, the spectral image read back on rank 0."""
    rsc = P.RenderScene.load(os.path.join(ROOT, "tests", "golden", CONFIG4_CASE + ".b2rs"))
    spp = min(args.config4_spp, rsc.spp)
    r = P.B200Renderer(rsc, device=local, stream=stream.cuda_stream)
    ptr, nd = r.film_device_ptr()
    film_t = torch.as_tensor(_DevArray(ptr, nd, "<f8"), device="cuda")

    def frame(read):
        r.clear()
        r.render(*P.parallel.sample_stripe(spp, rank, world))
        P.parallel.reduce_film(film_t, dist)
        return r.spectral_image() if (read and rank == 0) else None

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
    out = {"metric": "render_msamples_per_s", "unit": "Msamples/s", "value": ns / (dev_ms * 1e-3) * 1e-6,
           "ms_per_frame": dev_ms / args.steps, "scaling": "strong",
           "workload": f"config4: Cornell room + 24-patch spectral chart, realistic camera ({len(rsc.d['camera_lens'])} lens "
                       f"interfaces, double-Gauss 50 mm), spectral film {rsc.n_buckets} bands, {w}x{h}; {spp} of the frame's "
                       f"{rsc.spp} spp per step, path maxdepth 5; sample stripes over {world} rank(s), film reduced by NCCL",
           "e2e": {"value": ns / e2e_s * 1e-6, "unit": "Msamples/s", "h2d_bytes_per_step": 0,
                   "d2h_bytes_per_step": w * h * 4 * rsc.n_buckets},
           "film_bytes_reduced": nd * 8,
           "rays_per_frame": (st["closest_rays"] + st["shadow_rays"]) * world,
           "launches_per_frame": st["kernel_launches"], "image_mean": float(img.mean())}
    if not args.no_cpu_baseline:
        out["cpu_baseline"] = config4_cpu_reference(args, P)
    return out


def config4_cpu_reference(args, P):
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
        P.scenes.config4_pbrt(pbrt, out=os.path.join(tmp, "c4.exr"), **kw)
        _, _, secs = refrun.render_channels(pbrt)
    n = kw["res"] * kw["res"] * kw["spp"]
    return {"value": n / secs * 1e-6, "unit": "Msamples/s", "cores": os.cpu_count(), "kind": "reference",
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
        out["cpu_baseline"] = config5_cpu_reference(args, P)
    return out


def config5_cpu_reference(args, P):
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
        _, secs = refrun.render(pbrt, pfm)
    n = w * h * args.config5_cpu_spp
    render = max(secs - setup, 1e-3)
    return {"value": n / render * 1e-6, "unit": "Msamples/s", "cores": os.cpu_count(), "kind": "reference",
            "seconds": render, "setup_seconds": setup,
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
        rc = render_cpu_reference(args, P)
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
    NODE_B, PRIM_B = 32, 48
    bytes_closest = 28 + 24 + (NODE_B * tc["nodes_visited"] + PRIM_B * tc["prims_tested"]) / tc["rays"]
    bytes_any = 28 + 1 + (NODE_B * ta["nodes_visited"] + PRIM_B * ta["prims_tested"]) / ta["rays"]

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
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args),
            "closest_mrays_per_s": n * world / (ms_closest * 1e-3) * 1e-6,
            "any_mrays_per_s": n * world / (ms_any * 1e-3) * 1e-6,
            "roofline": {"bound": "hbm", "kernel": "closest-hit traversal", "achieved": ach, "peak": peak,
                         "unit": "GB/s", "frac": ach / peak,
                         # DRAM bytes per launch: dram__bytes_read.sum + dram__bytes_write.sum of this kernel from ncu
                         # (profiles/r01_headline_traffic_ncu.csv: 801.8 MB for a 2^22-ray launch on this scene =
                         # 191.2 B/ray, against 52 B/ray compulsory and ~3.7 KB/ray algorithmic, which the 126 MB
                         # L2 serves: lts__t_bytes = 3.97 KB/ray), scaled to this launch's ray count
                         "traffic": (DRAM_BYTES_PER_RAY_CLOSEST * n if args.tris == 1_000_000 else None),
                         "traffic_note": "ncu dram bytes per ray (2^22-ray launch) x rays per launch; the algorithmic "
                                         "bytes are served by L2 (lts__t_bytes 3.97 KB/ray), so `achieved` is an "
                                         "L2-level rate reported against the HBM peak as the contract asks",
                         "peak_source": peak_src,
                         "bytes_per_ray": bytes_closest, "bytes_per_ray_any": bytes_any,
                         "nodes_per_ray": tc["nodes_visited"] / tc["rays"],
                         "prims_per_ray": tc["prims_tested"] / tc["rays"],
                         "compulsory_bytes_per_ray": 52},
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
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args, P, sc, prim, shad)
        print(json.dumps(line), flush=True)
    if agg is not None:
        agg.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def cpu_baseline(args, P, sc, prim, shad):
    """The reference's own BVHAggregate (oracle/_ref/raycast_ref) on a bounded sample, all host cores."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refrun
    if not refrun.available():
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    m = min(args.cpu_sample, prim.shape[0])
    rs = refrun.RefScene(sc)
    _, ic = rs.closest(prim[:m], nthreads=0, reps=2)
    _, ia = rs.any(shad[:m], nthreads=0, reps=2)
    sec = ic["seconds"] + ia["seconds"]
    return {"value": 2 * m / sec * 1e-6, "unit": UNIT, "cores": ic["threads"], "kind": "reference",
            "closest_mrays_per_s": ic["mrays_per_s"], "any_mrays_per_s": ia["mrays_per_s"],
            "sample": f"first {m} primary + {m} shadow rays of the timed batches, best of 2 passes, "
                      f"BVH build excluded ({ic['build_s']:.2f} s)"}


if __name__ == "__main__":
    main()
