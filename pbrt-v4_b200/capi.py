"""ctypes binding of include/b200pt.h (the same stub a maintainer would use from
Python; the C++ reference binds the identical symbols directly)."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb200pt.so")

OK, ERR_INVALID, ERR_CUDA, ERR_NO_DEVICE, ERR_NOMEM = 0, 1, 2, 3, 4
MEM_HOST, MEM_DEVICE = 0, 1
TRACE_COUNTERS = 1 << 8
NOALLOC_MAX_RAYS = 1 << 20  # B200PT_NOALLOC_MAX_RAYS
LIGHT_GEOMETRY_FLOATS = 24  # B200PT_LIGHT_GEOMETRY_FLOATS
MISS = 0xFFFFFFFF

HIT_DTYPE = np.dtype([("mesh", "<u4"), ("prim", "<u4"), ("instance", "<u4"),
                      ("t", "<f4"), ("b1", "<f4"), ("b2", "<f4")])
NODE32_DTYPE = np.dtype([("bmin", "<f4", 3), ("bmax", "<f4", 3), ("offset", "<i4"),
                         ("nprims", "<u2"), ("axis", "u1"), ("pad", "u1")])
assert HIT_DTYPE.itemsize == 24 and NODE32_DTYPE.itemsize == 32


class Mesh(C.Structure):
    _fields_ = [("object", C.c_int32), ("n_vertices", C.c_uint32), ("n_triangles", C.c_uint32),
                ("p", C.POINTER(C.c_float)), ("indices", C.POINTER(C.c_int32)), ("kind", C.c_uint32),
                ("alpha", C.c_float)]  # read under kind & MESH_ALPHA only


MESH_TRIANGLES, MESH_BILINEAR, MESH_ALPHA = 0, 1, 2


def mesh_kind_alpha(m):
    """(kind word, alpha) of a scenes.Mesh for the C struct."""
    a = getattr(m, "alpha", None)
    return int(getattr(m, "kind", 0)) | (MESH_ALPHA if a is not None else 0), float(a if a is not None else 1.0)


class Instance(C.Structure):
    _fields_ = [("object", C.c_uint32), ("render_from_instance", C.c_float * 16),
                ("instance_from_render", C.c_float * 16)]


class BVH2(C.Structure):
    _fields_ = [("object", C.c_int32), ("n_nodes", C.c_uint32), ("n_prims", C.c_uint32),
                ("nodes", C.c_void_p), ("prims", C.POINTER(C.c_uint32))]


class SceneDesc(C.Structure):
    _fields_ = [("n_meshes", C.c_uint32), ("meshes", C.POINTER(Mesh)),
                ("n_instances", C.c_uint32), ("instances", C.POINTER(Instance)),
                ("n_bvhs", C.c_uint32), ("bvhs", C.POINTER(BVH2)), ("flags", C.c_uint32)]


class SceneStats(C.Structure):
    _fields_ = [("n_triangles", C.c_uint64), ("n_instances", C.c_uint64), ("n_bvh2_nodes", C.c_uint64),
                ("n_wide_nodes", C.c_uint64), ("device_bytes", C.c_uint64),
                ("build_seconds", C.c_double), ("upload_seconds", C.c_double)]


class TraceStats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("nodes_visited", C.c_uint64), ("leaves_visited", C.c_uint64),
                ("prims_tested", C.c_uint64), ("kernel_ms", C.c_float), ("wide_nodes_visited", C.c_uint64),
                ("exact_rays", C.c_uint64), ("leaf_boxes_tested", C.c_uint64)]


# every symbol include/b200pt.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "b200pt_version": (C.c_int, []),
    "b200pt_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "b200pt_context_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "b200pt_context_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200pt_context_sync": (C.c_int, [C.c_void_p]),
    "b200pt_context_destroy": (C.c_int, [C.c_void_p]),
    "b200pt_last_error": (C.c_char_p, [C.c_void_p]),
    "b200pt_scene_create": (C.c_int, [C.c_void_p, C.POINTER(SceneDesc), C.POINTER(C.c_void_p)]),
    "b200pt_scene_destroy": (C.c_int, [C.c_void_p]),
    "b200pt_scene_bounds": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "b200pt_scene_get_stats": (C.c_int, [C.c_void_p, C.POINTER(SceneStats)]),
    "b200pt_intersect_closest": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_uint32]),
    "b200pt_intersect_any": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_uint32]),
    "b200pt_intersect_closest_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "b200pt_get_trace_stats": (C.c_int, [C.c_void_p, C.POINTER(TraceStats)]),
    "b200pt_bvh2_build": (C.c_int, [C.POINTER(C.c_float), C.c_uint32, C.c_int, C.c_void_p,
                                    C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "b200pt_ply_read": (C.c_int, [C.c_char_p, C.c_void_p, C.c_char_p, C.c_uint64]),
    "b200pt_ply_free": (None, [C.c_void_p]),
}


class PlyMesh(C.Structure):  # b200pt_ply_mesh
    _fields_ = [("n_vertices", C.c_uint64), ("n_triangles", C.c_uint64), ("n_quads", C.c_uint64), ("n_face_indices", C.c_uint64),
                ("p", C.POINTER(C.c_float)), ("n", C.POINTER(C.c_float)), ("uv", C.POINTER(C.c_float)),
                ("tri_indices", C.POINTER(C.c_int32)), ("quad_indices", C.POINTER(C.c_int32)),
                ("face_indices", C.POINTER(C.c_int32))]


def read_ply(path):
    """b200pt_ply_read -> dict of numpy arrays {p [nv,3], n | None, uv | None, tri [nt,3], quad [nq,4], face_indices | None}
    (copies; the library's arrays are freed before returning).  Raises B200PTError with the library's message."""
    import numpy as np
    m, err = PlyMesh(), C.create_string_buffer(512)
    rc = lib().b200pt_ply_read(str(path).encode(), C.byref(m), err, 512)
    if rc != 0:
        raise B200PTError(rc, err.value.decode())
    try:
        def arr(ptr, n, cols, dt):
            if not ptr or n == 0:
                return None
            return np.ctypeslib.as_array(ptr, shape=(int(n) * cols,)).astype(dt).reshape(int(n), cols).copy()
        out = {"p": arr(m.p, m.n_vertices, 3, np.float32), "n": arr(m.n, m.n_vertices, 3, np.float32),
               "uv": arr(m.uv, m.n_vertices, 2, np.float32), "tri": arr(m.tri_indices, m.n_triangles, 3, np.int32),
               "quad": arr(m.quad_indices, m.n_quads, 4, np.int32)}
        fi = arr(m.face_indices, m.n_face_indices, 1, np.int32)
        out["face_indices"] = None if fi is None else fi.reshape(-1)
        return out
    finally:
        lib().b200pt_ply_free(C.byref(m))

_lib = None


class B200PTError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"b200pt error {code}: {msg}")
        self.code = code


def lib():
    """Loads libb200pt.so; fails loudly (no fallback) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200PTError(-1, f"{LIB_PATH} is missing - run `python -c 'import __graft_entry__ as g; g.build()'`; "
                                  "there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(code, ctx=None):
    if code != OK:
        msg = lib().b200pt_last_error(ctx)
        raise B200PTError(code, (msg or b"").decode())


def fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def bvh2_build(boxes, max_prims_in_node=4):
    """Host-only SAH build (no GPU): boxes (n,6) float32 -> (nodes[NODE32_DTYPE], order[u4])."""
    boxes = np.ascontiguousarray(boxes, dtype=np.float32).reshape(-1, 6)
    n = boxes.shape[0]
    nodes = np.zeros(max(2 * n - 1, 1), dtype=NODE32_DTYPE)
    order = np.zeros(n, dtype=np.uint32)
    nn = C.c_uint32(0)
    check(lib().b200pt_bvh2_build(fptr(boxes), n, int(max_prims_in_node), nodes.ctypes.data,
                                  C.byref(nn), order.ctypes.data_as(C.POINTER(C.c_uint32))))
    return nodes[:nn.value].copy(), order


# ------------------------------------------------------------------ render path structs
class Spectrum(C.Structure):
    _fields_ = [("interp", C.c_uint32), ("values", C.c_float * 311)]


class Material(C.Structure):
    _fields_ = [("type", C.c_uint32), ("spectrum", C.c_int32 * 4), ("params", C.c_float * 12)]


class Light(C.Structure):
    _fields_ = [("mesh", C.c_uint32), ("tri", C.c_uint32), ("spectrum", C.c_int32), ("scale", C.c_float),
                ("two_sided", C.c_uint32), ("cos_falloff_end", C.c_float), ("tan_falloff_end", C.c_float),
                ("normalize_falloff_end", C.c_float), ("alias_q", C.c_float), ("alias_p", C.c_float),
                ("alias", C.c_int32), ("type", C.c_uint32)]


class MeshAttr(C.Structure):
    _fields_ = [("material", C.c_int32), ("light_first", C.c_int32), ("flip_normal", C.c_uint32),
                ("corner_n", C.POINTER(C.c_float)), ("corner_uv", C.POINTER(C.c_float)),
                ("corner_s", C.POINTER(C.c_float))]


class Medium(C.Structure):
    _fields_ = [("type", C.c_uint32), ("sigma_a", C.c_int32), ("sigma_s", C.c_int32), ("g", C.c_float),
                ("bounds", C.c_float * 6), ("medium_from_render", C.c_float * 16),
                ("render_from_medium", C.c_float * 16), ("nx", C.c_int32), ("ny", C.c_int32), ("nz", C.c_int32),
                ("density", C.POINTER(C.c_float)), ("mres", C.c_int32 * 3), ("majorants", C.POINTER(C.c_float)),
                ("emissive", C.c_uint32), ("le_spectrum", C.c_int32), ("le_nx", C.c_int32), ("le_ny", C.c_int32),
                ("le_nz", C.c_int32), ("le_scale", C.POINTER(C.c_float)), ("temperature", C.POINTER(C.c_float)),
                ("temperature_scale", C.c_float), ("temperature_offset", C.c_float)]


class RenderDesc(C.Structure):
    _fields_ = [("camera_type", C.c_uint32), ("camera_from_raster", C.c_float * 16),
                ("render_from_camera", C.c_float * 16), ("lens_radius", C.c_float), ("focal_distance", C.c_float),
                ("film_type", C.c_uint32), ("pixel_bounds", C.c_int32 * 4), ("full_resolution", C.c_int32 * 2),
                ("filter_radius", C.c_float * 2), ("imaging_ratio", C.c_float), ("max_component_value", C.c_float),
                ("output_rgb_from_sensor_rgb", C.c_float * 9), ("sensor_spectrum", C.c_int32 * 3),
                ("sampler_type", C.c_uint32), ("samples_per_pixel", C.c_int32), ("seed", C.c_int32),
                ("randomize", C.c_uint32), ("max_depth", C.c_int32), ("regularize", C.c_uint32),
                ("light_sampler", C.c_uint32), ("n_spectra", C.c_uint32), ("n_materials", C.c_uint32),
                ("n_lights", C.c_uint32), ("n_meshes", C.c_uint32), ("spectra", C.POINTER(Spectrum)),
                ("materials", C.POINTER(Material)), ("lights", C.POINTER(Light)), ("meshes", C.POINTER(MeshAttr)),
                ("n_light_bvh_nodes", C.c_uint32), ("light_bvh_bounds", C.c_float * 6),
                ("light_bvh_nodes", C.c_void_p), ("light_bit_trails", C.POINTER(C.c_uint32)),
                ("n_buckets", C.c_uint32), ("lambda_min", C.c_float), ("lambda_max", C.c_float),
                ("n_lens_elements", C.c_uint32), ("lens_elements", C.POINTER(C.c_float)),
                ("n_exit_pupil_bounds", C.c_uint32), ("exit_pupil_bounds", C.POINTER(C.c_float)),
                ("physical_extent", C.c_float * 4), ("film_diagonal", C.c_float),
                ("integrator_type", C.c_uint32), ("n_media", C.c_uint32), ("media", C.POINTER(Medium)),
                ("mesh_media", C.POINTER(C.c_int32)), ("camera_medium", C.c_int32),
                ("lens_element_transforms", C.POINTER(C.c_float)),
                ("n_microlens_elements", C.c_uint32), ("microlens_elements", C.POINTER(C.c_float)),
                ("microlens_element_transforms", C.POINTER(C.c_float)), ("microlens_dims", C.c_int32 * 2),
                ("microlens_offsets", C.POINTER(C.c_float)), ("microlens_offset_from_sensor", C.c_float),
                ("microlens_simulation_radius", C.c_int32),
                ("gbuffer_output_from_render", C.c_float * 16), ("gbuffer_output_from_render_inv", C.c_float * 16),
                ("material_ids", C.POINTER(C.c_int32)), ("filter_type", C.c_uint32),
                ("light_geometry", C.POINTER(C.c_float)),
                ("filter_table", C.POINTER(C.c_float)), ("filter_table_size", C.c_int32 * 2),
                ("env_image", C.POINTER(C.c_float)), ("env_distribution", C.POINTER(C.c_float)), ("env_resolution", C.c_int32),
                ("gbuffer_albedo", C.c_uint32), ("gbuffer_xyz_spectrum", C.c_int32 * 3), ("gbuffer_illuminant", C.c_int32),
                ("gbuffer_rgb_from_xyz", C.c_float * 9)]


GBUFFER_RADIANCE, GBUFFER_POSITION, GBUFFER_MATERIAL, GBUFFER_INSTANCE = 1, 2, 4, 8
GBUFFER_NS, GBUFFER_NORMAL, GBUFFER_VARIANCE, GBUFFER_RELATIVE_VARIANCE = 16, 32, 64, 128
GBUFFER_ALBEDO = 256


class RenderStats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("kernel_launches", C.c_uint32), ("gpu_ms", C.c_float),
                ("vol_paths_cut", C.c_uint64), ("vol_walks_cut", C.c_uint64)]


SYMBOLS.update({
    "b200pt_render_create": (C.c_int, [C.c_void_p, C.POINTER(RenderDesc), C.POINTER(C.c_void_p)]),
    "b200pt_render_destroy": (C.c_int, [C.c_void_p]),
    "b200pt_render_run": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32]),
    "b200pt_film_clear": (C.c_int, [C.c_void_p]),
    "b200pt_film_device_ptr": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]),
    "b200pt_film_reduce": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_uint64]),
    "b200pt_film_read": (C.c_int, [C.c_void_p, C.POINTER(C.c_double)]),
    "b200pt_film_image": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "b200pt_film_write_exr": (C.c_int, [C.c_void_p, C.c_uint32, C.c_char_p, C.c_int32]),
    "b200pt_film_gbuffer_image": (C.c_int, [C.c_void_p, C.c_uint32, C.POINTER(C.c_float), C.POINTER(C.c_int32)]),
    "b200pt_film_spectral_image": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "b200pt_render_get_stats": (C.c_int, [C.c_void_p, C.POINTER(RenderStats)]),
})
