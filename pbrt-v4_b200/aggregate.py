"""Host-side mirror of the reference's intersector plug.

`B200Aggregate` follows `WavefrontAggregate` (reference wavefront/integrator.h:32-54;
CPU backend wavefront/aggregate.{h,cpp}) method for method - Bounds, IntersectClosest,
IntersectShadow - over ray batches instead of work queues: the queue plumbing
(EnqueueWorkAfterMiss/Intersection, wavefront/intersect.h) stays with the caller.
All work happens in libb200pt.so on the GPU; there is no CPU fallback.
"""
import ctypes as C

import numpy as np

from . import capi


class B200Aggregate:
    def __init__(self, scene, device=0, bvhs=None, stream=None):
        """scene: scenes.Scene.  bvhs: optional list of (object, nodes[NODE32_DTYPE], prims (n,3) u4)
        handed over by the host (e.g. read out of the reference's BVHAggregate); by default the
        library builds the SAH trees itself.  stream: CUDA stream handle (int) to run on."""
        L = capi.lib()
        self._L = L
        self._ctx = C.c_void_p()
        self._scene = C.c_void_p()
        capi.check(L.b200pt_context_create(int(device), C.byref(self._ctx)))
        if stream is not None:
            capi.check(L.b200pt_context_set_stream(self._ctx, C.c_void_p(int(stream))), self._ctx)
        self._keep = []
        meshes = (capi.Mesh * max(1, len(scene.meshes)))()
        for k, m in enumerate(scene.meshes):
            p = np.ascontiguousarray(m.p, dtype=np.float32)
            idx = np.ascontiguousarray(m.indices, dtype=np.int32)
            self._keep += [p, idx]
            meshes[k] = capi.Mesh(int(m.object), p.shape[0], idx.shape[0], capi.fptr(p),
                                  idx.ctypes.data_as(C.POINTER(C.c_int32)), *capi.mesh_kind_alpha(m))
        insts = (capi.Instance * max(1, len(scene.instances)))()
        for k, i in enumerate(scene.instances):
            insts[k].object = int(i.object)
            insts[k].render_from_instance[:] = [float(x) for x in i.render_from_instance.reshape(-1)]
            insts[k].instance_from_render[:] = [float(x) for x in i.instance_from_render.reshape(-1)]
        nb = 0
        barr = (capi.BVH2 * max(1, len(bvhs) if bvhs else 1))()
        if bvhs:
            for k, (obj, nodes, prims) in enumerate(bvhs):
                nodes = np.ascontiguousarray(nodes)
                prims = np.ascontiguousarray(prims, dtype=np.uint32).reshape(-1, 3)
                self._keep += [nodes, prims]
                barr[k] = capi.BVH2(int(obj), nodes.shape[0], prims.shape[0], nodes.ctypes.data,
                                    prims.ctypes.data_as(C.POINTER(C.c_uint32)))
            nb = len(bvhs)
        desc = capi.SceneDesc(len(scene.meshes), meshes, len(scene.instances), insts, nb, barr, 0)
        capi.check(L.b200pt_scene_create(self._ctx, C.byref(desc), C.byref(self._scene)), self._ctx)
        self._keep = []  # the library copied everything

    # ---- WavefrontAggregate::Bounds (wavefront/integrator.h:36)
    def Bounds(self):
        b = (C.c_float * 6)()
        capi.check(self._L.b200pt_scene_bounds(self._scene, b), self._ctx)
        return np.array(b[:], dtype=np.float32).reshape(2, 3)

    def stats(self):
        s = capi.SceneStats()
        capi.check(self._L.b200pt_scene_get_stats(self._scene, C.byref(s)), self._ctx)
        return {k: getattr(s, k) for k, _ in capi.SceneStats._fields_}

    def trace_stats(self):
        s = capi.TraceStats()
        capi.check(self._L.b200pt_get_trace_stats(self._scene, C.byref(s)), self._ctx)
        return {k: getattr(s, k) for k, _ in capi.TraceStats._fields_}

    # ---- WavefrontAggregate::IntersectClosest (wavefront/aggregate.cpp:34-59) over a host batch
    def IntersectClosest(self, rays, counters=False):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 7)
        hits = np.empty(rays.shape[0], dtype=capi.HIT_DTYPE)
        flags = capi.MEM_HOST | (capi.TRACE_COUNTERS if counters else 0)
        capi.check(self._L.b200pt_intersect_closest(self._scene, rays.ctypes.data, rays.shape[0],
                                                    hits.ctypes.data, flags), self._ctx)
        return hits

    # ---- WavefrontAggregate::IntersectShadow (wavefront/aggregate.cpp:61-70) over a host batch
    def IntersectShadow(self, rays, counters=False):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 7)
        occ = np.empty(rays.shape[0], dtype=np.uint8)
        flags = capi.MEM_HOST | (capi.TRACE_COUNTERS if counters else 0)
        capi.check(self._L.b200pt_intersect_any(self._scene, rays.ctypes.data, rays.shape[0],
                                                occ.ctypes.data, flags), self._ctx)
        return occ

    # ---- raw-pointer variants (device or pinned-host buffers owned by the caller, e.g. torch tensors)
    def intersect_closest_ptr(self, rays_ptr, n, hits_ptr, mem=capi.MEM_DEVICE):
        capi.check(self._L.b200pt_intersect_closest(self._scene, C.c_void_p(rays_ptr), int(n),
                                                    C.c_void_p(hits_ptr), int(mem)), self._ctx)

    def intersect_any_ptr(self, rays_ptr, n, occ_ptr, mem=capi.MEM_DEVICE):
        capi.check(self._L.b200pt_intersect_any(self._scene, C.c_void_p(rays_ptr), int(n),
                                                C.c_void_p(occ_ptr), int(mem)), self._ctx)

    def sync(self):
        capi.check(self._L.b200pt_context_sync(self._ctx), self._ctx)

    def close(self):
        if self._scene:
            self._L.b200pt_scene_destroy(self._scene)
            self._scene = C.c_void_p()
        if self._ctx:
            self._L.b200pt_context_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
