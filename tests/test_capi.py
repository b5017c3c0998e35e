"""The C-ABI library loads and exports every symbol include/b200pt.h declares (no GPU needed)."""
import ctypes
import os
import re

from conftest import ROOT


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "b200pt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200pt_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(P):
    names = declared_symbols()
    assert len(names) >= 15
    L = ctypes.CDLL(P.capi.LIB_PATH)
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/b200pt.h but not exported"
    assert set(names) == set(P.capi.SYMBOLS), "capi.SYMBOLS out of sync with the header"


def test_version_and_struct_sizes(P):
    assert P.capi.lib().b200pt_version() == 100
    assert ctypes.sizeof(P.capi.Mesh) == 40  # + kind, padded to the pointers
    assert ctypes.sizeof(P.capi.Instance) == 4 + 128
    assert ctypes.sizeof(P.capi.SceneDesc) == 56
    assert P.capi.HIT_DTYPE.itemsize == 24


def test_no_cpu_fallback_without_device(P):
    """Without a GPU the context cannot be created; nothing computes on the CPU."""
    L = P.capi.lib()
    n = ctypes.c_int(-1)
    assert L.b200pt_device_count(ctypes.byref(n)) == 0
    if n.value == 0:
        ctx = ctypes.c_void_p()
        rc = L.b200pt_context_create(0, ctypes.byref(ctx))
        assert rc == P.capi.ERR_NO_DEVICE and not ctx.value
        assert b"no CPU path" in L.b200pt_last_error(None)


def test_invalid_arguments_are_reported_not_fatal(P):
    L = P.capi.lib()
    assert L.b200pt_device_count(None) == P.capi.ERR_INVALID
    assert L.b200pt_context_create(0, None) == P.capi.ERR_INVALID
    assert L.b200pt_scene_bounds(None, None) == P.capi.ERR_INVALID
    assert L.b200pt_intersect_closest(None, None, 0, None, 0) == P.capi.ERR_INVALID
    assert L.b200pt_bvh2_build(None, 0, 4, None, None, None) == P.capi.ERR_INVALID
