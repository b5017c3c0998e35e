"""Host side of the render path: a RenderScene (the flat tables a pbrt host adaptor hands over, e.g.
a `.b2rs` dump) marshalled into include/b200pt.h structs, and B200Renderer, the Python mirror of
RenderWavefront (reference wavefront/wavefront.cpp:14-70): create -> run sample ranges -> read film."""
import ctypes as C

import numpy as np

from . import capi, scenes
from .aggregate import B200Aggregate


class RenderScene:
    """Geometry + render tables.  `d` is a dict of arrays with the .b2rs names (scenes.load_b2rs)."""

    def __init__(self, d):
        self.d = d
        nv, nt = d["mesh_nverts"].astype(np.int64), d["mesh_ntris"].astype(np.int64)
        # mesh_kind 1: a bilinear patch mesh, its patches' four indices each in patch_indices
        kind = d["mesh_kind"].astype(np.int64) if "mesh_kind" in d else np.zeros(len(nv), np.int64)
        vo = np.concatenate([[0], np.cumsum(nv)])
        to = np.concatenate([[0], np.cumsum(nt * (kind == 0))])
        po = np.concatenate([[0], np.cumsum(nt * (kind == 1))])
        obj = d["mesh_object"] if "mesh_object" in d else np.full(len(nv), -1, np.int32)
        meshes = [scenes.Mesh(d["P"][vo[i]:vo[i + 1]],
                              d["patch_indices"][po[i]:po[i + 1]] if kind[i] else d["indices"][to[i]:to[i + 1]],
                              object=int(obj[i]), kind=int(kind[i]),
                              alpha=(float(d["mesh_alpha"][i]) if "mesh_alpha" in d and d["mesh_alpha"][i] >= 0 else None))
                  for i in range(len(nv))]
        insts = []
        if "instance_object" in d:
            insts = [scenes.Instance(int(o), m, mi)
                     for o, m, mi in zip(d["instance_object"], d["instance_m"], d["instance_minv"])]
        self.geometry = scenes.Scene(meshes, insts)
        self.bvh = None
        if "bvh_nodes" in d:
            def nodes(a):
                return np.ascontiguousarray(a).view(capi.NODE32_DTYPE).reshape(-1)
            self.bvh = [(-1, nodes(d["bvh_nodes"]), d["bvh_prims"])]
            k = 0
            while f"obj_bvh_nodes_{k}" in d:
                self.bvh.append((k, nodes(d[f"obj_bvh_nodes_{k}"]), d[f"obj_bvh_prims_{k}"]))
                k += 1

    @classmethod
    def load(cls, path):
        return cls(scenes.load_b2rs(path))

    @property
    def resolution(self):
        b = self.d["film_ints"]
        return int(b[2] - b[0]), int(b[3] - b[1])

    @property
    def n_buckets(self):
        """Bands of a spectral film, 0 for an RGB film."""
        d = self.d
        return int(d["film_spectral"][0]) if "film_type" in d and int(d["film_type"][0]) in (1, 2) else 0

    @property
    def spp(self):
        return int(self.d["sampler"][1])

    def desc(self, keep):
        """-> capi.RenderDesc (arrays it points to are appended to `keep`)."""
        d = self.d
        r = capi.RenderDesc()
        r.camera_type = int(d["camera_type"][0])
        r.render_from_camera[:] = [float(x) for x in d["render_from_camera"]]
        if r.camera_type in (1, 2):  # RealisticCamera / OmniCamera: lens table + exit-pupil bounds from the host
            lens = np.ascontiguousarray(d["camera_lens"], np.float32).reshape(-1, 4)
            pupil = np.ascontiguousarray(d["camera_exit_pupil"], np.float32).reshape(-1, 4)
            r.n_lens_elements, r.lens_elements = len(lens), capi.fptr(lens)
            r.n_exit_pupil_bounds, r.exit_pupil_bounds = len(pupil), capi.fptr(pupil)
            r.physical_extent[:] = [float(x) for x in d["camera_extent"][:4]]
            r.film_diagonal = float(d["camera_extent"][4])
            keep += [lens, pupil]
            if r.camera_type == 2 and "camera_lens_transforms" in d:
                xf = np.ascontiguousarray(d["camera_lens_transforms"], np.float32).reshape(-1, 32)
                assert len(xf) == len(lens)
                r.lens_element_transforms = capi.fptr(xf)
                keep.append(xf)
            if r.camera_type == 2 and "microlens_lens" in d:
                ml = np.ascontiguousarray(d["microlens_lens"], np.float32).reshape(-1, 4)
                mxf = np.ascontiguousarray(d["microlens_transforms"], np.float32).reshape(-1, 32)
                off = np.ascontiguousarray(d["microlens_offsets"], np.float32).reshape(-1, 2)
                mi = [int(x) for x in d["microlens_ints"]]
                assert len(mxf) == len(ml) and len(off) == mi[0] * mi[1]
                r.n_microlens_elements, r.microlens_elements = len(ml), capi.fptr(ml)
                r.microlens_element_transforms, r.microlens_offsets = capi.fptr(mxf), capi.fptr(off)
                r.microlens_dims[:] = mi[:2]
                r.microlens_simulation_radius = mi[2]
                r.microlens_offset_from_sensor = float(d["microlens_params"][0])
                keep += [ml, mxf, off]
        else:
            r.camera_from_raster[:] = [float(x) for x in d["camera_from_raster"]]
            r.lens_radius, r.focal_distance = float(d["camera_params"][0]), float(d["camera_params"][1])
        r.film_type = int(d["film_type"][0]) if "film_type" in d else 0
        if r.film_type == 2:  # GBufferFilm: output transform of the AOVs, the host's material ids
            r.gbuffer_output_from_render[:] = [float(x) for x in d["gbuffer_from_render"].reshape(-1)]
            r.gbuffer_output_from_render_inv[:] = [float(x) for x in d["gbuffer_from_render_inv"].reshape(-1)]
            ids = np.ascontiguousarray(d["material_ids"], np.int32)
            keep.append(ids)
            r.material_ids = ids.ctypes.data_as(C.POINTER(C.c_int32))
            if "gbuffer_albedo" in d:  # saveAlbedo: {X, Y, Z, illuminant} spectrum indices, RGBFromXYZ
                r.gbuffer_albedo = 1
                r.gbuffer_xyz_spectrum[:] = [int(x) for x in d["gbuffer_albedo"][:3]]
                r.gbuffer_illuminant = int(d["gbuffer_albedo"][3])
                r.gbuffer_rgb_from_xyz[:] = [float(x) for x in d["gbuffer_rgb_from_xyz"].reshape(-1)]
        if r.film_type in (1, 2):  # SpectralFilm / GBufferFilm buckets
            r.n_buckets = int(d["film_spectral"][0])
            r.lambda_min, r.lambda_max = float(d["film_spectral"][1]), float(d["film_spectral"][2])
        r.pixel_bounds[:] = [int(x) for x in d["film_ints"][:4]]
        r.full_resolution[:] = [int(x) for x in d["film_ints"][4:6]]
        r.filter_radius[:] = [float(x) for x in d["film_params"][:2]]
        r.filter_type = int(d["filter_type"][0]) if "filter_type" in d else 0  # 0 = box (older dumps: box only)
        if "filter_table" in d:  # FilterSampler::f of the gaussian / mitchell / sinc filters, [ny, nx]
            ft = np.ascontiguousarray(d["filter_table"], np.float32)
            keep.append(ft)
            r.filter_table = ft.ctypes.data_as(C.POINTER(C.c_float))
            r.filter_table_size[:] = [ft.shape[1], ft.shape[0]]
        r.imaging_ratio, r.max_component_value = float(d["film_params"][2]), float(d["film_params"][3])
        r.output_rgb_from_sensor_rgb[:] = [float(x) for x in d["film_matrix"]]
        r.sensor_spectrum[:] = [int(x) for x in d["film_sensor"]]
        s = d["sampler"]
        r.sampler_type, r.samples_per_pixel, r.seed, r.randomize = int(s[0]), int(s[1]), int(s[2]), int(s[3])
        r.max_depth, r.regularize, r.light_sampler = (int(x) for x in d["integrator"])
        ns, nm, nl, nme = len(d["spectra_interp"]), len(d["materials_i"]), len(d["lights_i"]), len(d["mesh_attr"])
        spectra = (capi.Spectrum * max(ns, 1))()
        for i in range(ns):
            spectra[i].interp = int(d["spectra_interp"][i])
            C.memmove(spectra[i].values, np.ascontiguousarray(d["spectra_values"][i], np.float32).ctypes.data, 311 * 4)
        mats = (capi.Material * max(nm, 1))()
        for i in range(nm):
            mi = d["materials_i"][i]
            mats[i].type = int(mi[0]) if mi[0] >= 0 else 0
            mats[i].spectrum[:] = [int(x) for x in mi[1:5]]
            mf = [float(x) for x in d["materials_f"][i]]  # 8 columns in dumps older than the coated conductor, 12 since
            mats[i].params[:] = mf + [0.0] * (12 - len(mf))
        lights = (capi.Light * max(nl, 1))()
        for i in range(nl):
            li, lf = d["lights_i"][i], d["lights_f"][i]
            ltype = int(li[6]) if len(li) > 6 else 0  # 1: uniform infinite light (no triangle)
            lights[i] = capi.Light(max(int(li[0]), 0), max(int(li[1]), 0), int(li[2]), float(lf[0]), int(li[3]),
                                   float(lf[1]), float(lf[2]), float(lf[3]), float(lf[4]), float(lf[5]), int(li[4]),
                                   ltype)
        if "env_image" in d:  # image infinite light: per-pixel spectrum coefficients [res, res, 4], distribution [res, res]
            ei, ed = np.ascontiguousarray(d["env_image"], np.float32), np.ascontiguousarray(d["env_distribution"], np.float32)
            keep += [ei, ed]
            r.env_image, r.env_distribution = (a.ctypes.data_as(C.POINTER(C.c_float)) for a in (ei, ed))
            r.env_resolution = ed.shape[0]
        if "lights_geom" in d:  # point / spot / distant lights: their placement (b200pt_render_desc::light_geometry)
            geom = np.ascontiguousarray(d["lights_geom"], np.float32).reshape(nl, -1)
            if geom.shape[1] < capi.LIGHT_GEOMETRY_FLOATS:  # 16 columns in dumps older than the image infinite light
                geom = np.ascontiguousarray(np.pad(geom, ((0, 0), (0, capi.LIGHT_GEOMETRY_FLOATS - geom.shape[1]))))
            keep.append(geom)
            r.light_geometry = geom.ctypes.data_as(C.POINTER(C.c_float))
        meshes = (capi.MeshAttr * max(nme, 1))()
        # per-corner normals / uv: the meshes that have them take their 3 * ntris rows from corner_n / corner_uv in order
        has_n = d["mesh_has_n"] if "mesh_has_n" in d else np.zeros(nme, np.uint32)
        has_uv = d["mesh_has_uv"] if "mesh_has_uv" in d else np.zeros(nme, np.uint32)
        has_s = d["mesh_has_s"] if "mesh_has_s" in d else np.zeros(nme, np.uint32)
        ntris = d["mesh_ntris"].astype(np.int64)
        n_at = uv_at = s_at = 0
        for i in range(nme):
            a = d["mesh_attr"][i]
            meshes[i] = capi.MeshAttr(int(a[0]), int(a[1]), int(max(a[2], 0)))
            if has_n[i]:
                per = 4 if "mesh_kind" in d and d["mesh_kind"][i] == 1 else 3  # corners per primitive
                cn = np.ascontiguousarray(d["corner_n"][n_at:n_at + per * ntris[i]], np.float32)
                n_at += per * int(ntris[i])
                meshes[i].corner_n = capi.fptr(cn)
                keep.append(cn)
            if has_uv[i]:
                per = 4 if "mesh_kind" in d and d["mesh_kind"][i] == 1 else 3
                cuv = np.ascontiguousarray(d["corner_uv"][uv_at:uv_at + per * ntris[i]], np.float32)
                uv_at += per * int(ntris[i])
                meshes[i].corner_uv = capi.fptr(cuv)
                keep.append(cuv)
            if has_s[i]:  # per-vertex shading tangents
                cs = np.ascontiguousarray(d["corner_s"][s_at:s_at + 3 * ntris[i]], np.float32)
                s_at += 3 * int(ntris[i])
                meshes[i].corner_s = capi.fptr(cs)
                keep.append(cs)
        r.n_spectra, r.n_materials, r.n_lights, r.n_meshes = ns, nm, nl, nme
        r.spectra, r.materials, r.lights, r.meshes = spectra, mats, lights, meshes
        keep += [spectra, mats, lights, meshes]
        r.integrator_type = int(d["integrator_type"][0]) if "integrator_type" in d else 0
        r.camera_medium = -1
        if r.integrator_type == 1:  # VolPathIntegrator: media, medium interfaces, camera medium
            mi, mf = d["media_i"].reshape(-1, 9), d["media_f"].reshape(-1, 39)
            media = (capi.Medium * max(len(mi), 1))()
            for k in range(len(mi)):
                m = media[k]
                m.type, m.sigma_a, m.sigma_s = int(mi[k][0]), int(mi[k][1]), int(mi[k][2])
                m.nx, m.ny, m.nz = (int(x) for x in mi[k][3:6])
                m.mres[:] = [int(x) for x in mi[k][6:9]]
                m.g = float(mf[k][0])
                m.bounds[:] = [float(x) for x in mf[k][1:7]]
                m.medium_from_render[:] = [float(x) for x in mf[k][7:23]]
                m.render_from_medium[:] = [float(x) for x in mf[k][23:39]]
                if m.type == 1:
                    dens = np.ascontiguousarray(d[f"medium_density_{k}"], np.float32).reshape(-1)
                    maj = np.ascontiguousarray(d[f"medium_majorants_{k}"], np.float32).reshape(-1)
                    m.density, m.majorants = capi.fptr(dens), capi.fptr(maj)
                    keep += [dens, maj]
                if "media_le" in d and int(d["media_le"].reshape(-1, 4)[k][0]) >= 0:  # emissive: Le_spec, LeScale grid
                    le = d["media_le"].reshape(-1, 4)[k]
                    m.emissive, m.le_spectrum = 1, int(le[0])
                    if m.type == 1:
                        lg = np.ascontiguousarray(d[f"medium_lescale_{k}"], np.float32).reshape(-1)
                        m.le_nx, m.le_ny, m.le_nz = int(le[1]), int(le[2]), int(le[3])
                        m.le_scale = capi.fptr(lg)
                        keep.append(lg)
                        if f"medium_temperature_{k}" in d:  # blackbody emission from a temperature grid
                            tg = np.ascontiguousarray(d[f"medium_temperature_{k}"], np.float32).reshape(-1)
                            m.temperature = capi.fptr(tg)
                            m.temperature_scale, m.temperature_offset = (float(x) for x in d[f"medium_temperature_params_{k}"])
                            keep.append(tg)
            mm = np.ascontiguousarray(d["mesh_media"], np.int32).reshape(-1)
            r.n_media, r.media = len(mi), media
            r.mesh_media = mm.ctypes.data_as(C.POINTER(C.c_int32))
            r.camera_medium = int(d["camera_medium"][0])
            keep += [media, mm]
        if r.light_sampler == 2:  # BVHLightSampler tables
            nodes = np.ascontiguousarray(d["lightbvh_nodes"], np.uint32)
            trails = np.ascontiguousarray(d["lightbvh_trails"], np.uint32)
            r.n_light_bvh_nodes = len(nodes)
            r.light_bvh_bounds[:] = [float(x) for x in d["lightbvh_bounds"]]
            r.light_bvh_nodes = nodes.ctypes.data
            r.light_bit_trails = trails.ctypes.data_as(C.POINTER(C.c_uint32))
            keep += [nodes, trails]
        return r


def scene_desc(geometry, bvh, keep):
    """scenes.Scene (+ optional host BVH) -> capi.SceneDesc."""
    meshes = (capi.Mesh * max(1, len(geometry.meshes)))()
    for k, m in enumerate(geometry.meshes):
        p, idx = np.ascontiguousarray(m.p, np.float32), np.ascontiguousarray(m.indices, np.int32)
        keep += [p, idx]
        meshes[k] = capi.Mesh(int(m.object), p.shape[0], idx.shape[0], capi.fptr(p),
                              idx.ctypes.data_as(C.POINTER(C.c_int32)), *capi.mesh_kind_alpha(m))
    barr = (capi.BVH2 * max(1, len(bvh) if bvh else 1))()
    nb = 0
    if bvh:
        for k, (obj, nodes, prims) in enumerate(bvh):
            nodes, prims = np.ascontiguousarray(nodes), np.ascontiguousarray(prims, np.uint32).reshape(-1, 3)
            keep += [nodes, prims]
            barr[k] = capi.BVH2(int(obj), nodes.shape[0], prims.shape[0], nodes.ctypes.data,
                                prims.ctypes.data_as(C.POINTER(C.c_uint32)))
        nb = len(bvh)
    insts = (capi.Instance * max(1, len(geometry.instances)))()
    for k, i in enumerate(geometry.instances):
        insts[k].object = int(i.object)
        insts[k].render_from_instance[:] = [float(x) for x in i.render_from_instance.reshape(-1)]
        insts[k].instance_from_render[:] = [float(x) for x in i.instance_from_render.reshape(-1)]
    keep += [meshes, barr, insts]
    return capi.SceneDesc(len(geometry.meshes), meshes, len(geometry.instances), insts, nb, barr, 0)


class B200Renderer:
    """create -> render(sample range) -> image, all on the GPU through libb200pt.so."""

    def __init__(self, rscene, device=0, stream=None, use_host_bvh=False):
        self.rscene = rscene
        self.agg = B200Aggregate(rscene.geometry, device=device, stream=stream,
                                 bvhs=rscene.bvh if use_host_bvh else None)
        self._L = capi.lib()
        self._r = C.c_void_p()
        keep = []
        desc = rscene.desc(keep)
        capi.check(self._L.b200pt_render_create(self.agg._scene, C.byref(desc), C.byref(self._r)), self.agg._ctx)
        self.width, self.height = rscene.resolution

    def render(self, first_sample=0, n_samples=None, stride=1):
        n = self.rscene.spp if n_samples is None else n_samples
        capi.check(self._L.b200pt_render_run(self._r, int(first_sample), int(n), int(stride)), self.agg._ctx)

    def clear(self):
        capi.check(self._L.b200pt_film_clear(self._r), self.agg._ctx)

    def film(self):
        """{rgbSum[3], weightSum} per pixel as (H, W, 4) doubles."""
        return self.film_all()[0]

    def film_all(self):
        """-> (rgbw (H, W, 4), bucketSums (nB, H, W), weightSums (nB, H, W)); the last two None for an RGB film."""
        nb, n = self.rscene.n_buckets, self.height * self.width
        d = self.rscene.d
        gb = "film_type" in d and int(d["film_type"][0]) == 2  # GBufferFilm: 31 radiance planes + 17 AOV planes (kGBufferPlanes)
        buf = np.empty(n * (4 + nb + 17 if gb else 4 + 2 * nb), dtype=np.float64)
        capi.check(self._L.b200pt_film_read(self._r, buf.ctypes.data_as(C.POINTER(C.c_double))), self.agg._ctx)
        rgbw = buf[:4 * n].reshape(self.height, self.width, 4)
        if not nb:
            return rgbw, None, None
        if gb:
            return rgbw, buf[4 * n:(4 + nb) * n].reshape(nb, self.height, self.width), None
        planes = buf[4 * n:(4 + 2 * nb) * n].reshape(2, nb, self.height, self.width)
        return rgbw, planes[0], planes[1]

    def spectral_image(self):
        """The S0.* channels of SpectralFilm::GetImage as (H, W, nB) float32."""
        out = np.empty((self.height, self.width, self.rscene.n_buckets), dtype=np.float32)
        capi.check(self._L.b200pt_film_spectral_image(self._r, out.ctypes.data_as(C.POINTER(C.c_float))), self.agg._ctx)
        return out

    def gbuffer_image(self, channels=capi.GBUFFER_RADIANCE, out=None):
        """GBufferFilm::GetImage: (H, W, C) float32, R G B then the groups of `channels` in the reference's order.
        out: optional float32 array of that shape to fill (e.g. pinned host memory: the copy then runs at PCIe speed)."""
        nc = C.c_int32(0)
        capi.check(self._L.b200pt_film_gbuffer_image(self._r, int(channels), None, C.byref(nc)), self.agg._ctx)
        if out is None:
            out = np.empty((self.height, self.width, nc.value), np.float32)
        assert out.dtype == np.float32 and out.size == self.height * self.width * nc.value and out.flags.c_contiguous
        capi.check(self._L.b200pt_film_gbuffer_image(self._r, int(channels), out.ctypes.data_as(C.POINTER(C.c_float)),
                                                     C.byref(nc)), self.agg._ctx)
        return out

    def write_exr(self, path, channels=capi.GBUFFER_RADIANCE, fp16=True):
        """Film::WriteImage: the film's image as an (uncompressed scanline) OpenEXR file, packed on the device."""
        capi.check(self._L.b200pt_film_write_exr(self._r, int(channels), str(path).encode(), 1 if fp16 else 0), self.agg._ctx)

    def film_device_ptr(self):
        p, n = C.c_void_p(), C.c_uint64()
        capi.check(self._L.b200pt_film_device_ptr(self._r, C.byref(p), C.byref(n)), self.agg._ctx)
        return p.value, n.value

    def image(self):
        out = np.empty((self.height, self.width, 3), dtype=np.float32)
        capi.check(self._L.b200pt_film_image(self._r, out.ctypes.data_as(C.POINTER(C.c_float))), self.agg._ctx)
        return out

    def stats(self):
        s = capi.RenderStats()
        capi.check(self._L.b200pt_render_get_stats(self._r, C.byref(s)), self.agg._ctx)
        return {k: getattr(s, k) for k, _ in capi.RenderStats._fields_}

    def sync(self):
        self.agg.sync()

    def close(self):
        if self._r:
            self._L.b200pt_render_destroy(self._r)
            self._r = C.c_void_p()
        self.agg.close()


def film_to_image(film_rgbw, matrix):
    """RGBFilm::GetPixelRGB (film.h:261-277): rgbSum / weightSum, then outputRGBFromSensorRGB."""
    w = film_rgbw[..., 3:4]
    rgb = np.where(w != 0, film_rgbw[..., :3] / np.where(w != 0, w, 1), film_rgbw[..., :3])
    return (rgb.astype(np.float32) @ np.asarray(matrix, np.float32).reshape(3, 3).T).astype(np.float32)
