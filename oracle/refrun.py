"""oracle/refrun.py - runs the compiled reference (oracle/_ref/raycast_ref, pbrt_ref).
TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs; never by the product package."""
import json
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
RAYCAST = os.path.join(REF_DIR, "raycast_ref")
PBRT = os.path.join(REF_DIR, "pbrt_ref")

REF_HIT_DTYPE = np.dtype([("mesh", "<u4"), ("prim", "<u4"), ("instance", "<u4"), ("t", "<f4"),
                          ("uv", "<f4", 2), ("p", "<f4", 3), ("n", "<f4", 3)])
NODE32_DTYPE = np.dtype([("bmin", "<f4", 3), ("bmax", "<f4", 3), ("offset", "<i4"),
                         ("nprims", "<u2"), ("axis", "u1"), ("pad", "u1")])


def available():
    return os.access(RAYCAST, os.X_OK)


def _run(args, timeout=3600):
    r = subprocess.run(args, capture_output=True, text=True, timeout=timeout)
    if r.returncode != 0:
        raise RuntimeError(f"{' '.join(args)} failed ({r.returncode}): {r.stderr[-2000:]}")
    lines = [l for l in r.stdout.strip().splitlines() if l.startswith("{")]
    return json.loads(lines[-1]) if lines else {}


class RefScene:
    """A scene file on disk + the reference driver."""

    def __init__(self, scene, workdir=None):
        self._tmp = None
        if workdir is None:
            self._tmp = tempfile.TemporaryDirectory(prefix="b2ref_")
            workdir = self._tmp.name
        self.dir = workdir
        self.path = os.path.join(workdir, "scene.b2sc")
        scene.write(self.path)

    def closest(self, rays, nthreads=0, reps=1):
        rp, hp = os.path.join(self.dir, "rays.f32"), os.path.join(self.dir, "hits.bin")
        np.ascontiguousarray(rays, dtype="<f4").tofile(rp)
        info = _run([RAYCAST, self.path, "closest", rp, hp, str(nthreads), str(reps)])
        return np.fromfile(hp, dtype=REF_HIT_DTYPE), info

    def any(self, rays, nthreads=0, reps=1):
        rp, hp = os.path.join(self.dir, "rays.f32"), os.path.join(self.dir, "occ.u8")
        np.ascontiguousarray(rays, dtype="<f4").tofile(rp)
        info = _run([RAYCAST, self.path, "any", rp, hp, str(nthreads), str(reps)])
        return np.fromfile(hp, dtype=np.uint8), info

    def dump_bvh(self):
        """-> list of (object, nodes[NODE32_DTYPE], prims (n,3) u4) as the reference built them."""
        out = os.path.join(self.dir, "bvh.bin")
        _run([RAYCAST, self.path, "dump-bvh", out])
        trees = []
        with open(out, "rb") as f:
            nb = int(np.fromfile(f, dtype="<u4", count=1)[0])
            for _ in range(nb):
                obj = int(np.fromfile(f, dtype="<i4", count=1)[0])
                nn, npr = (int(x) for x in np.fromfile(f, dtype="<u4", count=2))
                nodes = np.fromfile(f, dtype=NODE32_DTYPE, count=nn)
                prims = np.fromfile(f, dtype="<u4", count=3 * npr).reshape(-1, 3)
                trees.append((obj, nodes, prims))
        return trees


# ------------------------------------------------------------------ whole-image reference (pbrt_ref)
DUMPSCENE = os.path.join(REF_DIR, "dumpscene_ref")


def render_available():
    return os.access(PBRT, os.X_OK) and os.access(DUMPSCENE, os.X_OK)


def read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = map(int, f.readline().split())
        scale = float(f.readline())
        a = np.fromfile(f, "<f4" if scale < 0 else ">f4").reshape(h, w, 3)
    return np.ascontiguousarray(a[::-1])  # PFM rows go bottom-up


def render(pbrt_file, out_pfm, nthreads=None, timeout=3600):
    """Runs the reference CPU renderer (its own main: cmd/pbrt.cpp) on a .pbrt file whose Film
    writes `out_pfm`; returns (image, seconds)."""
    import time
    args = [PBRT, "--quiet"] + (["--nthreads", str(nthreads)] if nthreads else []) + [pbrt_file]
    t = time.time()
    r = subprocess.run(args, capture_output=True, text=True, timeout=timeout, cwd=os.path.dirname(pbrt_file))
    dt = time.time() - t
    if r.returncode != 0:
        raise RuntimeError(f"pbrt_ref failed ({r.returncode}): {r.stderr[-2000:]}")
    return read_pfm(out_pfm), dt


PBRT_B200 = os.path.join(REF_DIR, "pbrt_b200")


def b200_available():
    return os.access(PBRT_B200, os.X_OK)


def render_b200(pbrt_file, out_pfm, wavefront=False, extra=(), timeout=3600):
    """The compiled reference-side binding (oracle/pbrt_b200.cpp): the reference's parser, scene objects and film
    writer with libb200pt.so behind them.  wavefront=True: the reference's own WavefrontPathIntegrator over
    B200Aggregate.  Returns (image, seconds, stderr)."""
    import time
    args = [PBRT_B200] + (["--wavefront"] if wavefront else []) + list(extra) + [pbrt_file]
    t = time.time()
    r = subprocess.run(args, capture_output=True, text=True, timeout=timeout, cwd=os.path.dirname(pbrt_file))
    dt = time.time() - t
    if r.returncode != 0:
        raise RuntimeError(f"pbrt_b200 failed ({r.returncode}): {r.stderr[-2000:]}")
    return (read_pfm(out_pfm) if out_pfm else None), dt, r.stderr


def render_wavefront(pbrt_file, out_pfm, nthreads=None, timeout=3600):
    """pbrt_ref --wavefront: the reference's CPU wavefront path (CPUAggregate)."""
    args = [PBRT, "--quiet", "--wavefront"] + (["--nthreads", str(nthreads)] if nthreads else []) + [pbrt_file]
    r = subprocess.run(args, capture_output=True, text=True, timeout=timeout, cwd=os.path.dirname(pbrt_file))
    if r.returncode != 0:
        raise RuntimeError(f"pbrt_ref --wavefront failed ({r.returncode}): {r.stderr[-2000:]}")
    return read_pfm(out_pfm)


def dump_scene(pbrt_file, out_b2rs, bvh=True):
    """Scene tables as the reference's own parser + object creation produced them (dumpscene_ref.cpp).
    bvh=False leaves the reference's BVH node arrays out of the file."""
    r = subprocess.run([DUMPSCENE, pbrt_file, out_b2rs] + ([] if bvh else ["--no-bvh"]), capture_output=True, text=True,
                       cwd=os.path.dirname(pbrt_file))
    if r.returncode != 0:
        raise RuntimeError(f"dumpscene_ref failed ({r.returncode}): {r.stderr[-2000:]}")
    return out_b2rs


def render_channels(pbrt_file):
    """The reference's render loop + Film::GetImage through dumpscene_ref --render: every channel of the
    film's image as (H, W, C) float32, and the channel names.  For films the oracle build has no file
    writer for (the spectral film's EXR)."""
    import time
    out = pbrt_file + ".img.b2rs"
    t = time.time()
    r = subprocess.run([DUMPSCENE, pbrt_file, "--render", out], capture_output=True, text=True,
                       cwd=os.path.dirname(pbrt_file))
    dt = time.time() - t
    if r.returncode != 0:
        raise RuntimeError(f"dumpscene_ref --render failed ({r.returncode}): {r.stderr[-2000:]}")
    info = json.loads([l for l in r.stdout.strip().splitlines() if l.startswith("{")][-1])
    import importlib.util
    spec = importlib.util.spec_from_file_location("b2_scenes", os.path.join(os.path.dirname(HERE), "pbrt-v4_b200", "scenes.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    img = mod.load_b2rs(out)["image"]
    os.unlink(out)
    return np.ascontiguousarray(img, np.float32), info["channels"], dt
