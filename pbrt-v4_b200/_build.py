"""Builds libb200pt.so (the C-ABI library of include/b200pt.h) in-tree.

nvcc cross-compiles sm_100a without a GPU.  Host translation units go through
g++ with -ffp-contract=off (the SAH builder must reproduce the reference's float
decisions); traversal .cu files use --fmad=false on top of explicit rn-intrinsics.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libb200pt.so")

NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
CUDA_HOME = os.path.dirname(os.path.dirname(NVCC))
GENCODE = ["-gencode", "arch=compute_100a,code=sm_100a"]

CPP_SOURCES = ["bvh2_build.cpp", "wide_build.cpp", "scene_host.cpp", "render_host.cpp", "api.cpp", "render_api.cpp", "ply_read.cpp"]
# (file, extra nvcc flags)
CU_SOURCES = [
    ("traverse_bvh2.cu", ["--fmad=false"]),
    # one family of kernels per part (see the end of traverse_persist.cu), built in parallel
    ("traverse_persist.cu", ["--fmad=false", "-DB2_TP_PART=0"], "traverse_persist_0"),
    ("traverse_persist.cu", ["--fmad=false", "-DB2_TP_PART=1"], "traverse_persist_1"),
    ("traverse_persist.cu", ["--fmad=false", "-DB2_TP_PART=2"], "traverse_persist_2"),
    ("traverse_persist.cu", ["--fmad=false", "-DB2_TP_PART=3"], "traverse_persist_3"),
    ("traverse_persist.cu", ["--fmad=false", "-DB2_TP_PART=4"], "traverse_persist_4"),
    ("traverse_persist.cu", ["--fmad=false", "-DB2_TP_PART=5"], "traverse_persist_5"),
    # one family of render kernels per part (see render.cu), built in parallel
    ("render.cu", ["--fmad=false", "-DB2_RENDER_PART=0"], "render_0"),
    ("render.cu", ["--fmad=false", "-DB2_RENDER_PART=1"], "render_1"),
    ("render.cu", ["--fmad=false", "-DB2_RENDER_PART=2"], "render_2"),
    ("render.cu", ["--fmad=false", "-DB2_RENDER_PART=3"], "render_3"),
    ("render.cu", ["--fmad=false", "-DB2_RENDER_PART=4", "-DB2_SPECTRAL_UNROLL=4", "-DB2_VOL_MIN_BLOCKS=4"], "render_4"),
    ("noalloc.cu", []),
    ("exr_pack.cu", []),
]


def _newer(src, dst, extra_deps):
    if not os.path.exists(dst):
        return True
    t = os.path.getmtime(dst)
    return any(os.path.getmtime(p) > t for p in [src] + extra_deps)


def _run(cmd, verbose):
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("build failed: " + " ".join(cmd))
    if verbose and (r.stdout or r.stderr):
        print(r.stdout + r.stderr, flush=True)


def build(verbose=False, force=False):
    from concurrent.futures import ThreadPoolExecutor
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".h") or f.endswith(".cuh")]
    headers.append(os.path.join(os.path.dirname(HERE), "include", "b200pt.h"))
    objs, jobs = [], []
    for f in CPP_SOURCES:
        src, obj = os.path.join(CSRC, f), os.path.join(OBJ, f + ".o")
        objs.append(obj)
        if force or _newer(src, obj, headers):
            jobs.append(["g++", "-std=c++17", "-O2", "-fPIC", "-ffp-contract=off", "-fno-fast-math",
                         "-march=x86-64-v3", "-Wall", "-I" + os.path.join(CUDA_HOME, "include"),
                         "-c", src, "-o", obj])
    for entry in CU_SOURCES:
        f, extra = entry[0], entry[1]
        src, obj = os.path.join(CSRC, f), os.path.join(OBJ, (entry[2] if len(entry) > 2 else f) + ".o")
        objs.append(obj)
        only = os.environ.get("B200PT_BUILD_ONLY")  # rebuild just these objects (comma separated), e.g. render_4
        if (only and os.path.basename(obj)[:-2] in only.split(",")) or (not only and (force or _newer(src, obj, headers))):
            jobs.append([NVCC, "-std=c++17", "-O3", "-lineinfo", "-Xptxas", "-v", "-Xcompiler", "-fPIC",
                         "-Xcompiler", "-ffp-contract=off"] + GENCODE + extra + os.environ.get("B200PT_NVCC_EXTRA", "").split() + ["-c", src, "-o", obj])
    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            list(ex.map(lambda c: _run(c, verbose), jobs))
    if force or jobs or any(_newer(o, LIB, []) for o in objs):
        _run([NVCC, "-shared", "-o", LIB] + objs + GENCODE + ["-cudart", "static", "-ldl"], verbose)
    return LIB


if __name__ == "__main__":
    print(build(verbose=True, force="--force" in sys.argv))
