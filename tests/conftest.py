import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def P():
    import pbrt_v4_b200 as pkg
    if not os.path.exists(pkg.capi.LIB_PATH):
        pkg.build()
    return pkg


@pytest.fixture(scope="session")
def refrun():
    import refrun as r
    if not r.available():
        pytest.skip("oracle/_ref/raycast_ref not built (needs /root/reference: make -f oracle/ref.mk)")
    return r


@pytest.fixture(scope="session")
def hostcheck_bin():
    """TEST-ONLY host build of the device arithmetic (tests/native/hostcheck.cpp)."""
    out = os.path.join(ROOT, "tests", "native", "_build", "hostcheck")
    src = [os.path.join(ROOT, "tests", "native", "hostcheck.cpp"),
           os.path.join(ROOT, "pbrt-v4_b200", "csrc", "scene_host.cpp"),
           os.path.join(ROOT, "pbrt-v4_b200", "csrc", "bvh2_build.cpp"),
           os.path.join(ROOT, "pbrt-v4_b200", "csrc", "wide_build.cpp")]
    deps = src + [os.path.join(ROOT, "pbrt-v4_b200", "csrc", h) for h in ("isect.h", "instance.h", "scene_host.h", "bvh2.h", "wide.h")] \
        + [os.path.join(ROOT, "tests", "native", "trace_host.h")]
    if not os.path.exists(out) or any(os.path.getmtime(s) > os.path.getmtime(out) for s in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-march=x86-64-v3", "-o", out] + src)
    return out


def hit_mismatches(mine, ref, scene=None):
    """Bit-exact comparison of product hits (capi.HIT_DTYPE) with reference hits
    (refrun.REF_HIT_DTYPE).  ids and t by bits; barycentrics through the reference's uv
    (default uv (0,0),(1,0),(1,1): u = b1 + b2, v = b2; `+ 0.0` folds -0, which the
    reference's b0*0 + ... sum also does, shapes.h:884-1010).  Hits on bilinear patches (meshes of
    `scene` with kind 1) carry the patch's own (u, v), compared as they are (shapes.h:1396-1472)."""
    hit = ref["mesh"] != 0xFFFFFFFF
    z = np.float32(0.0)
    bad = mine["mesh"] != ref["mesh"]
    bad |= hit & (mine["prim"] != ref["prim"])
    bad |= hit & (mine["instance"] != ref["instance"])
    bad |= hit & (mine["t"].view("u4") != ref["t"].view("u4"))
    patch = np.zeros(len(ref), bool)
    if scene is not None:
        kinds = np.array([getattr(m, "kind", 0) for m in scene.meshes] + [0], np.int64)
        patch = hit & (kinds[np.minimum(ref["mesh"], len(kinds) - 1)] == 1)
    tri = hit & ~patch
    bad |= tri & ((mine["b2"] + z).view("u4") != (ref["uv"][:, 1] + z).view("u4"))
    bad |= tri & (((mine["b1"] + mine["b2"]).astype("f4") + z).view("u4") != (ref["uv"][:, 0] + z).view("u4"))
    bad |= patch & (mine["b1"].view("u4") != ref["uv"][:, 0].view("u4"))
    bad |= patch & (mine["b2"].view("u4") != ref["uv"][:, 1].view("u4"))
    return bad


@pytest.fixture(scope="session")
def rendercheck(P):
    """TEST-ONLY host build of the render headers (tests/native/rendercheck.cpp): the same per-path
    functions the CUDA kernels call, looped on the CPU.  Returns run(rscene, first, n, stride) -> film
    accumulators [H, W, 4] float64."""
    import ctypes as C
    csrc = os.path.join(ROOT, "pbrt-v4_b200", "csrc")
    out = os.path.join(ROOT, "tests", "native", "_build", "librendercheck.so")
    src = [os.path.join(ROOT, "tests", "native", "rendercheck.cpp")] + \
          [os.path.join(csrc, f) for f in ("scene_host.cpp", "bvh2_build.cpp", "wide_build.cpp", "render_host.cpp")]
    deps = src + [os.path.join(csrc, h) for h in os.listdir(csrc) if h.endswith(".h")] \
        + [os.path.join(ROOT, "tests", "native", "trace_host.h")]
    if not os.path.exists(out) or any(os.path.getmtime(s) > os.path.getmtime(out) for s in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-march=x86-64-v3", "-shared",
                               "-fPIC", "-pthread", "-o", out] + src)
    L = C.CDLL(out)

    def run(rscene, first=0, n=None, stride=1, threads=8, use_host_bvh=False, full=False):
        keep = []
        sd = P.render.scene_desc(rscene.geometry, rscene.bvh if use_host_bvh else None, keep)
        rd = rscene.desc(keep)
        w, h = rscene.resolution
        nb = rscene.n_buckets
        film = np.zeros(h * w * (4 + 2 * nb), np.float64)
        msg = C.create_string_buffer(512)
        rc = L.rendercheck_run(C.byref(sd), C.byref(rd), int(first), int(rscene.spp if n is None else n),
                               int(stride), film.ctypes.data_as(C.POINTER(C.c_double)), msg, 512, threads)
        assert rc == 0, msg.value.decode()
        rgbw = film[:4 * h * w].reshape(h, w, 4)
        if not full:
            return rgbw
        planes = film[4 * h * w:].reshape(2, nb, h, w)  # spectral film: bucketSums, weightSums
        return rgbw, planes[0], planes[1]
    return run


def rrmse(img, ref):
    """relative RMSE the way SURVEY.md 8c states the image gate: sqrt(mean((a-b)^2)) / mean(ref)."""
    return float(np.sqrt(np.mean((np.asarray(img, np.float64) - ref) ** 2)) / np.mean(ref))


def read_exr(path):
    """Minimal reader of an uncompressed scanline OpenEXR file (what b200pt_film_write_exr writes; the layout is the
    OpenEXR file-format document's): -> (dict name -> (H, W) float32 array, dataWindow, displayWindow)."""
    import struct
    b = open(path, "rb").read()
    assert struct.unpack_from("<i", b, 0)[0] == 20000630 and struct.unpack_from("<i", b, 4)[0] == 2
    p, attrs = 8, {}
    while b[p] != 0:
        e = b.index(b"\0", p)
        name = b[p:e].decode()
        p = e + 1
        e = b.index(b"\0", p)
        typ = b[p:e].decode()
        p = e + 1
        size = struct.unpack_from("<i", b, p)[0]
        p += 4
        attrs[name] = (typ, b[p:p + size])
        p += size
    p += 1
    chans, q, v = [], 0, attrs["channels"][1]
    while v[q] != 0:
        e = v.index(b"\0", q)
        nm = v[q:e].decode()
        ptype, _, xs, ys = struct.unpack_from("<i4sii", v, e + 1)
        assert xs == 1 and ys == 1
        chans.append((nm, ptype))
        q = e + 1 + 16
    assert [c[0] for c in chans] == sorted(c[0] for c in chans), "EXR channel lists are sorted by name"
    assert attrs["compression"][1] == b"\0" and attrs["lineOrder"][1] == b"\0"
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    disp = struct.unpack("<4i", attrs["displayWindow"][1])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    offs = np.frombuffer(b, "<u8", h, p)
    out = {nm: np.empty((h, w), np.float32) for nm, _ in chans}
    for k in range(h):
        o = int(offs[k])
        y, nbytes = struct.unpack_from("<ii", b, o)
        assert y == y0 + k
        o += 8
        for nm, ptype in chans:
            dt = "<f2" if ptype == 1 else "<f4"
            out[nm][k] = np.frombuffer(b, dt, w, o).astype(np.float32)
            o += w * (2 if ptype == 1 else 4)
        assert o - int(offs[k]) - 8 == nbytes
    return out, (x0, y0, x1, y1), disp
