"""Procedural scenes and ray batches for BASELINE.json's configs, plus
the little-endian scene file (`.b2sc`) that carries the same geometry to both
the product (through capi / B200Aggregate) and the reference driver
(oracle/raycast_ref.cpp).  numpy only; nothing here computes intersections.

File layout:  u32 magic 'B2SC', u32 version=2, u32 nMeshes, u32 nInstances
  per mesh:     i32 object (-1 world), u32 nVerts, u32 nPrims, u32 kind (0 triangles, 1 bilinear patches; | 2: an
                f32 constant alpha follows),
                f32 P[3nV], i32 idx[(3 or 4) nPrims]
  per instance: u32 object, f32 renderFromInstance[16], f32 instanceFromRender[16]
"""
from dataclasses import dataclass, field
from typing import List

import os

import numpy as np

MAGIC = 0x43533242


@dataclass
class Mesh:
    p: np.ndarray            # (nv, 3) float32, render space
    indices: np.ndarray      # (nt, 3) int32; bilinear patches: (np, 4) = p00 p10 p01 p11
    object: int = -1         # -1: world; >= 0: object definition id
    kind: int = 0            # 0 triangles, 1 bilinear patches
    alpha: float = None      # constant alpha cut-out of a triangle mesh (None: opaque, no alpha texture)

    def __post_init__(self):
        self.p = np.ascontiguousarray(self.p, dtype=np.float32).reshape(-1, 3)
        self.indices = np.ascontiguousarray(self.indices, dtype=np.int32).reshape(-1, 4 if self.kind == 1 else 3)


@dataclass
class Instance:
    object: int
    render_from_instance: np.ndarray   # (4,4) float32
    instance_from_render: np.ndarray   # (4,4) float32

    def __post_init__(self):
        self.render_from_instance = np.ascontiguousarray(self.render_from_instance, dtype=np.float32).reshape(4, 4)
        self.instance_from_render = np.ascontiguousarray(self.instance_from_render, dtype=np.float32).reshape(4, 4)


@dataclass
class Scene:
    meshes: List[Mesh] = field(default_factory=list)
    instances: List[Instance] = field(default_factory=list)

    @property
    def n_triangles(self):
        return sum(m.indices.shape[0] for m in self.meshes)

    def write(self, path):
        with open(path, "wb") as f:
            np.array([MAGIC, 2, len(self.meshes), len(self.instances)], dtype="<u4").tofile(f)
            for m in self.meshes:
                np.array([m.object], dtype="<i4").tofile(f)
                np.array([m.p.shape[0], m.indices.shape[0], m.kind | (2 if m.alpha is not None else 0)], dtype="<u4").tofile(f)
                if m.alpha is not None:
                    np.array([m.alpha], dtype="<f4").tofile(f)
                m.p.astype("<f4").tofile(f)
                m.indices.astype("<i4").tofile(f)
            for i in self.instances:
                np.array([i.object], dtype="<u4").tofile(f)
                i.render_from_instance.astype("<f4").tofile(f)
                i.instance_from_render.astype("<f4").tofile(f)

    def triangle_vertices(self, mesh):
        m = self.meshes[mesh]
        return m.p[m.indices]  # (nt, 3, 3)

    def triangle_boxes(self, meshes=None):
        """(n,6) float32 boxes in reference primitive order (mesh order, then triangle order)."""
        out = []
        for k, m in enumerate(self.meshes):
            if meshes is not None and k not in meshes:
                continue
            v = m.p[m.indices]
            out.append(np.concatenate([v.min(axis=1), v.max(axis=1)], axis=1))
        return np.concatenate(out, axis=0).astype(np.float32)


# ------------------------------------------------------------------ scenes

def random_soup(n_tris=1_000_000, seed=12111, s=0.01, n_meshes=1):
    """BASELINE config 2: unit cube filled with random triangles: centre ~U[0,1]^3,
    vertices = centre + U[-s,s]^3 (the construction of the reference's own
    Triangle.Watertight test, shapes_test.cpp:33-60, scaled up)."""
    rng = np.random.default_rng(seed)
    meshes = []
    per = [n_tris // n_meshes + (1 if i < n_tris % n_meshes else 0) for i in range(n_meshes)]
    for cnt in per:
        c = rng.random((cnt, 1, 3), dtype=np.float32)
        v = (c + (rng.random((cnt, 3, 3), dtype=np.float32) * 2 - 1) * np.float32(s)).astype(np.float32)
        meshes.append(Mesh(v.reshape(-1, 3), np.arange(3 * cnt, dtype=np.int32).reshape(-1, 3)))
    return Scene(meshes)


def instanced_scene(n_objects=4, tris_per_object=2000, n_instances=40, world_tris=3000, seed=77, s=0.05):
    """BASELINE config 3 in miniature: object definitions (random soups in [-1/2, 1/2]^3) placed many
    times by rotation x uniform scale x translation inside the unit cube, over a world-level soup.
    Both matrices of each instance are stored in float32 (the host hands over pbrt's m and mInv)."""
    rng = np.random.default_rng(seed)
    meshes = []
    if world_tris:
        meshes += random_soup(world_tris, seed=seed + 1, s=0.02).meshes
    for o in range(n_objects):
        cnt = tris_per_object if o else 2  # object 0: the smallest BVH an object can have
        c = rng.random((cnt, 1, 3)) - 0.5
        v = (c + (rng.random((cnt, 3, 3)) * 2 - 1) * s).astype(np.float32)
        meshes.append(Mesh(v.reshape(-1, 3), np.arange(3 * cnt, dtype=np.int32).reshape(-1, 3), object=o))
    insts = []
    for i in range(n_instances):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        w, x, y, z = q
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                      [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                      [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
        if i % 7 == 3:
            R = np.eye(3)  # some axis-aligned placements: exact zeros in the matrices
        sc = rng.uniform(0.08, 0.3)
        t = rng.uniform(0.15, 0.85, 3)
        M = np.eye(4)
        M[:3, :3] = R * sc
        M[:3, 3] = t
        insts.append(Instance(i % n_objects, M.astype(np.float32), np.linalg.inv(M).astype(np.float32)))
    return Scene(meshes, insts)


def flattened(scene):
    """All geometry of `scene` as world-level meshes in render space (instances applied in float64);
    only for aiming test rays at instanced geometry."""
    out = [m for m in scene.meshes if m.object < 0]
    for inst in scene.instances:
        M = inst.render_from_instance.astype(np.float64)
        for m in scene.meshes:
            if m.object == inst.object:
                out.append(Mesh((m.p.astype(np.float64) @ M[:3, :3].T + M[:3, 3]).astype(np.float32), m.indices))
    return Scene(out)


def patch_scene(n_patches=20000, n_tris=10000, seed=91, s=0.03):
    """Bilinear patches (BilinearPatchMesh: 4 indices per patch, p00 p10 p01 p11) in the unit cube, one mesh of
    twisted (non-planar) patches, one of planar parallelograms (the quadratic's a = 0 branch) and axis-aligned
    squares, next to a triangle mesh, so leaves mix both kinds."""
    rng = np.random.default_rng(seed)
    def quads(n, twist, axis_aligned=False):
        c = rng.uniform(0.05, 0.95, (n, 3))
        e0 = rng.uniform(-s, s, (n, 3))
        e1 = rng.uniform(-s, s, (n, 3))
        if axis_aligned:
            ax = rng.integers(0, 3, n)
            e0[:] = 0
            e1[:] = 0
            e0[np.arange(n), (ax + 1) % 3] = s
            e1[np.arange(n), (ax + 2) % 3] = s
        p = np.stack([c, c + e0, c + e1, c + e0 + e1], axis=1)
        p[:, 3] += twist * rng.uniform(-s, s, (n, 3))
        return Mesh(p.reshape(-1, 3), np.arange(4 * n).reshape(n, 4), kind=1)
    n1 = n_patches // 2
    n2 = (n_patches - n1) // 2
    meshes = [quads(n1, 1.0), quads(n2, 0.0), quads(n_patches - n1 - n2, 0.0, axis_aligned=True)]
    if n_tris:
        meshes.append(random_soup(n_tris, seed=seed + 1, s=s).meshes[0])
    return Scene(meshes)


def alpha_scene(seed=43):
    """Alpha cut-outs at every level: an instanced scene whose world mesh and one object mesh carry a
    constant alpha in (0, 1) (stochastic cut-out by HashFloat(ray.o, ray.d), the instance-space ray inside objects),
    plus world soups with alpha 0 (never hit), alpha 1 (opaque through the GeometricPrimitive path) and none."""
    sc = instanced_scene(n_objects=3, tris_per_object=500, n_instances=20, world_tris=1200, seed=seed)
    world = [m for m in sc.meshes if m.object < 0]
    objs = [m for m in sc.meshes if m.object >= 0]
    world[0].alpha = 0.6
    objs[0].alpha = 0.35
    extra = random_soup(1500, seed=seed + 1, s=0.04, n_meshes=3).meshes
    extra[0].alpha, extra[1].alpha = 0.0, 1.0
    return Scene(sc.meshes + extra, sc.instances)


def _quad(p0, p1, p2, p3):
    return Mesh(np.array([p0, p1, p2, p3], dtype=np.float32), np.array([[0, 1, 2], [0, 2, 3]], dtype=np.int32))


def _box(lo, hi, rot_y=0.0, centre=None):
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    c = np.array([[lo[0], lo[1], lo[2]], [hi[0], lo[1], lo[2]], [hi[0], hi[1], lo[2]], [lo[0], hi[1], lo[2]],
                  [lo[0], lo[1], hi[2]], [hi[0], lo[1], hi[2]], [hi[0], hi[1], hi[2]], [lo[0], hi[1], hi[2]]])
    if rot_y:
        ctr = (lo + hi) / 2 if centre is None else np.asarray(centre, np.float64)
        ca, sa = np.cos(rot_y), np.sin(rot_y)
        R = np.array([[ca, 0, sa], [0, 1, 0], [-sa, 0, ca]])
        c = (c - ctr) @ R.T + ctr
    quads = [[0, 3, 2, 1], [4, 5, 6, 7], [0, 1, 5, 4], [2, 3, 7, 6], [1, 2, 6, 5], [0, 4, 7, 3]]
    idx = []
    for q in quads:
        idx += [[q[0], q[1], q[2]], [q[0], q[2], q[3]]]
    return Mesh(c.astype(np.float32), np.array(idx, dtype=np.int32))


def cornell_box():
    """BASELINE config 1 geometry: 5 wall quads, a ceiling light quad and 2 boxes
    (axis-aligned walls make exact t ties along shared diagonals - the tie-breaking
    case of SURVEY F7)."""
    m = [
        _quad([0, 0, 0], [1, 0, 0], [1, 0, 1], [0, 0, 1]),      # floor
        _quad([0, 1, 0], [0, 1, 1], [1, 1, 1], [1, 1, 0]),      # ceiling
        _quad([0, 0, 1], [1, 0, 1], [1, 1, 1], [0, 1, 1]),      # back
        _quad([0, 0, 0], [0, 0, 1], [0, 1, 1], [0, 1, 0]),      # left (red)
        _quad([1, 0, 0], [1, 1, 0], [1, 1, 1], [1, 0, 1]),      # right (green)
        _quad([0.35, 0.999, 0.35], [0.65, 0.999, 0.35], [0.65, 0.999, 0.65], [0.35, 0.999, 0.65]),  # light
        _box([0.15, 0, 0.5], [0.45, 0.6, 0.8], rot_y=0.3),      # tall box
        _box([0.55, 0, 0.2], [0.85, 0.3, 0.5], rot_y=-0.3),     # short box
    ]
    return Scene(m)


def grid_sheet(nx=64, ny=64, z=0.5, jitter=0.0, seed=1):
    """Indexed mesh with shared vertices/edges (watertightness + tie cases)."""
    rng = np.random.default_rng(seed)
    xs, ys = np.meshgrid(np.linspace(0, 1, nx + 1), np.linspace(0, 1, ny + 1), indexing="ij")
    zs = np.full_like(xs, z) + (rng.random(xs.shape) * 2 - 1) * jitter
    p = np.stack([xs, ys, zs], axis=-1).reshape(-1, 3).astype(np.float32)
    idx = []
    for i in range(nx):
        for j in range(ny):
            a, b, c, d = i * (ny + 1) + j, (i + 1) * (ny + 1) + j, (i + 1) * (ny + 1) + j + 1, i * (ny + 1) + j + 1
            idx += [[a, b, c], [a, c, d]]
    return Mesh(p, np.array(idx, dtype=np.int32))


# ------------------------------------------------------------------ ray batches
# rays: (n,7) float32 rows {o.xyz, d.xyz, tMax}

def _unit_sphere(rng, n):
    z = 1 - 2 * rng.random(n)
    r = np.sqrt(np.maximum(0, 1 - z * z))
    phi = 2 * np.pi * rng.random(n)
    return np.stack([r * np.cos(phi), r * np.sin(phi), z], axis=1)


def primary_rays(n, seed=1, centre=(0.5, 0.5, 0.5), radius=2.0, lo=0.1, hi=0.9):
    """Origins on a sphere about the cube, aimed at uniform points in [lo,hi]^3, tMax = inf
    (SURVEY 8d config 2)."""
    rng = np.random.default_rng(seed)
    o = np.asarray(centre) + radius * _unit_sphere(rng, n)
    tgt = lo + (hi - lo) * rng.random((n, 3))
    d = tgt - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.empty((n, 7), dtype=np.float32)
    rays[:, 0:3], rays[:, 3:6], rays[:, 6] = o, d, np.inf
    return rays


def shadow_rays(rays, hits, seed=2, light_lo=(0, 1.0, 0), light_hi=(1, 1.0, 1)):
    """From each primary hit point (offset 1e-3 back along the ray) to a uniform point on the
    cube's top face: d unnormalised, tMax = 1 - ShadowEpsilon (cpu/integrators.cpp:297-314,
    ray.h:95-109).  Misses shoot from the origin of the primary ray."""
    rng = np.random.default_rng(seed)
    n = rays.shape[0]
    t = np.where(hits["mesh"] != 0xFFFFFFFF, hits["t"], 0).astype(np.float64)
    o = rays[:, 0:3].astype(np.float64) + rays[:, 3:6].astype(np.float64) * (t - 1e-3)[:, None]
    l0, l1 = np.asarray(light_lo, np.float64), np.asarray(light_hi, np.float64)
    tgt = l0 + (l1 - l0) * rng.random((n, 3))
    out = np.empty((n, 7), dtype=np.float32)
    out[:, 0:3] = o
    out[:, 3:6] = tgt - o
    out[:, 6] = np.float32(1 - 1e-4)
    return out


def adversarial_rays(scene, n, seed=3):
    """Rays that stress tie-breaking and NaN/inf paths: through mesh vertices, through edge
    midpoints, along +-axes (zero direction components -> invDir = +-inf), and from origins
    lying exactly in bounding planes."""
    rng = np.random.default_rng(seed)
    verts = np.concatenate([m.p[m.indices].reshape(-1, 3) for m in scene.meshes if m.object < 0], axis=0)
    # (a bilinear patch contributes its first three corners: edges p00-p10, p10-p01 (a diagonal), p01-p00)
    tris = np.concatenate([m.p[m.indices[:, :3]] for m in scene.meshes if m.object < 0], axis=0)
    k = n // 4
    rays = np.empty((4 * k, 7), dtype=np.float32)
    # 1) through vertices from random outside origins
    o = 0.5 + 2.0 * _unit_sphere(rng, k)
    v = verts[rng.integers(0, verts.shape[0], k)]
    rays[0:k, 0:3], rays[0:k, 3:6] = o, v - o
    # 2) through edge midpoints
    t = tris[rng.integers(0, tris.shape[0], k)]
    e = rng.integers(0, 3, k)
    mid = 0.5 * (t[np.arange(k), e] + t[np.arange(k), (e + 1) % 3])
    o = 0.5 + 2.0 * _unit_sphere(rng, k)
    rays[k:2 * k, 0:3], rays[k:2 * k, 3:6] = o, mid - o
    # 3) axis-aligned rays through vertices (two zero components in d)
    ax = rng.integers(0, 3, k)
    sgn = rng.choice([-1.0, 1.0], k)
    v = verts[rng.integers(0, verts.shape[0], k)].astype(np.float64)
    d = np.zeros((k, 3))
    d[np.arange(k), ax] = sgn
    o = v - 3.0 * d
    rays[2 * k:3 * k, 0:3], rays[2 * k:3 * k, 3:6] = o, d
    # 4) one zero component, origin sharing a coordinate with a vertex (in-plane rays)
    v = verts[rng.integers(0, verts.shape[0], k)].astype(np.float64)
    ax = rng.integers(0, 3, k)
    d = _unit_sphere(rng, k)
    d[np.arange(k), ax] = 0.0
    d[np.arange(k), (ax + 1) % 3] += 1e-3
    o = v - 2.5 * d
    rays[3 * k:, 0:3], rays[3 * k:, 3:6] = o, d
    rays[:, 6] = np.inf
    return rays


def camera_rays(width, height, eye, look, up=(0, 1, 0), fov_deg=40.0, spp=1, seed=0):
    """Simple pinhole batch (pixel centres or jittered) for config-1 style parity batches."""
    rng = np.random.default_rng(seed)
    eye, look, up = (np.asarray(a, np.float64) for a in (eye, look, up))
    w = look - eye
    w /= np.linalg.norm(w)
    u = np.cross(up, w)
    u /= np.linalg.norm(u)
    v = np.cross(w, u)
    ys, xs = np.meshgrid(np.arange(height), np.arange(width), indexing="ij")
    xs = np.repeat(xs.reshape(-1), spp).astype(np.float64)
    ys = np.repeat(ys.reshape(-1), spp).astype(np.float64)
    jx, jy = (rng.random(xs.shape), rng.random(xs.shape)) if spp > 1 else (0.5, 0.5)
    s = np.tan(np.radians(fov_deg) / 2)
    px = (2 * (xs + jx) / width - 1) * s * (width / height)
    py = (1 - 2 * (ys + jy) / height) * s
    d = px[:, None] * u + py[:, None] * v + w
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.empty((xs.shape[0], 7), dtype=np.float32)
    rays[:, 0:3], rays[:, 3:6], rays[:, 6] = eye, d, np.inf
    return rays


# ------------------------------------------------------------------ .pbrt text for the reference (config 1)

def _mesh_pbrt(m):
    idx = " ".join(str(int(i)) for i in m.indices.reshape(-1))
    pts = " ".join(repr(float(x)) for x in m.p.reshape(-1))
    return f'Shape "trianglemesh" "integer indices" [{idx}] "point3 P" [{pts}]'


# Material lines for cornell_pbrt(boxes=...): the BxDF families of BASELINE.json configs[2]
MATERIALS = {
    "conductor_rough": 'Material "conductor" "spectrum eta" "metal-Cu-eta" "spectrum k" "metal-Cu-k" "float roughness" 0.05',
    "conductor_smooth": 'Material "conductor" "spectrum eta" "metal-Au-eta" "spectrum k" "metal-Au-k" "float roughness" 0',
    "conductor_reflectance": 'Material "conductor" "rgb reflectance" [0.8 0.6 0.3] "float uroughness" 0.2 "float vroughness" 0.05',
    "dielectric_smooth": 'Material "dielectric" "float eta" 1.5',
    "dielectric_rough": 'Material "dielectric" "float eta" 1.5 "float roughness" 0.1',
    "dielectric_glass": 'Material "dielectric" "spectrum eta" "glass-BK7"',
    "thin_dielectric": 'Material "thindielectric" "float eta" 1.5',
    "thin_dielectric_glass": 'Material "thindielectric" "spectrum eta" "glass-BK7"',
    "diffuse_transmission": 'Material "diffusetransmission" "rgb reflectance" [0.3 0.5 0.2] "rgb transmittance" [0.5 0.4 0.2]',
    "diffuse_transmission_scaled": 'Material "diffusetransmission" "rgb reflectance" [0.1 0.1 0.1] "rgb transmittance" [0.3 0.45 0.6] "float scale" 1.8',
    # CoatedConductorMaterial (materials.h:695-790): a clear coat over copper (rough base under a smooth coat), and a
    # rough coat over a mirror-like base given by its reflectance
    "coated_conductor": 'Material "coatedconductor" "spectrum conductor.eta" "metal-Cu-eta" "spectrum conductor.k" "metal-Cu-k" '
                        '"float conductor.roughness" 0.15 "float interface.roughness" 0',
    "coated_conductor_rough": 'Material "coatedconductor" "rgb reflectance" [0.9 0.7 0.4] "float conductor.roughness" 0 '
                              '"float interface.roughness" 0.1 "float thickness" 0.02',
    "coated_conductor_both": 'Material "coatedconductor" "spectrum conductor.eta" "metal-Au-eta" "spectrum conductor.k" "metal-Au-k" '
                             '"float conductor.uroughness" 0.2 "float conductor.vroughness" 0.08 "float interface.roughness" 0.05 '
                             '"spectrum interface.eta" "glass-BK7"',
    "coated_diffuse": 'Material "coateddiffuse" "rgb reflectance" [0.2 0.3 0.7] "float roughness" 0',
    "coated_diffuse_rough": 'Material "coateddiffuse" "rgb reflectance" [0.7 0.3 0.2] "float roughness" 0.1 "float thickness" 0.02',
}


# LightSource lines for cornell_pbrt / config5_pbrt (lights=...): the delta lights (lights.h: PointLight 189-239,
# SpotLight 747-801, DistantLight 242-297).  The distant light shines in through the room's open front.
LIGHTS = {
    "point": 'LightSource "point" "point3 from" [0.3 0.72 0.4] "rgb I" [0.5 0.4 0.25]',
    "spot": 'LightSource "spot" "point3 from" [0.85 0.9 0.15] "point3 to" [0.4 0 0.6] "float coneangle" 35 '
            '"float conedeltaangle" 12 "rgb I" [2 2.2 3]',
    "distant": 'LightSource "distant" "point3 from" [0.2 1.3 -1.5] "point3 to" [0.6 0.2 0.5] "rgb L" [0.9 0.8 0.6]',
}


# GBufferFilm, the ISET spectral film (film.cpp:1053-1106): 31 radiance channels + the AOV groups built here; the
# channel order of its image is R G B, Radiance.C01-31, P.X-Z, MaterialId, InstanceId, Ns.X-Z, N.X-Z, Variance.R-B,
# RelativeVariance.R-B (film.cpp:766-831)
GBUFFER_ALL = 1 | 2 | 4 | 8 | 16 | 32 | 64 | 128


def _gbuffer_film(out, xres, yres, albedo=False):
    return (f'Film "gbuffer" "string filename" "{os.path.splitext(out)[0]}.exr" "integer xresolution" {xres} '
            f'"integer yresolution" {yres} "bool savefp16" false "bool saveRadiance" true "bool saveDepth" true '
            '"bool saveMaterial" true "bool saveInstance" true "bool saveNormal" true "bool saveNs" true '
            '"bool saveVariance" true "bool saveRelativeVariance" true' + (' "bool saveAlbedo" true' if albedo else ''))


def _envmap_lines(path, res):
    """ImageInfiniteLight (lights.h:557-643): an equal-area square image written next to the scene as a PFM - a dim
    bluish sky with two bright warm patches - turned so that the image's pole does not sit on a scene axis."""
    g = (np.arange(res) + 0.5) / res
    u, v = np.meshgrid(g, g)
    img = np.stack([0.25 + 0.2 * v, 0.3 + 0.25 * u, 0.45 + 0.3 * u * v], axis=-1)
    img += 40.0 * np.exp(-((u - 0.62) ** 2 + (v - 0.35) ** 2) / 0.004)[..., None] * np.array([1.0, 0.8, 0.5])
    img += 12.0 * np.exp(-((u - 0.2) ** 2 + (v - 0.8) ** 2) / 0.01)[..., None] * np.array([0.9, 0.9, 1.0])
    pfm = os.path.splitext(path)[0] + "_env.pfm"
    with open(pfm, "wb") as f:  # PFM: rows bottom to top, little endian
        f.write(f"PF\n{res} {res}\n-1.0\n".encode())
        f.write(np.ascontiguousarray(img[::-1], "<f4").tobytes())
    return ['AttributeBegin', '  Rotate 200 0 1 0', '  Rotate 65 1 0 0',
            f'  LightSource "infinite" "string filename" "{pfm}" "float scale" 0.6', 'AttributeEnd']


def cornell_pbrt(path, res=512, spp=64, maxdepth=5, lightsampler="power", out="cornell.pfm", seed=0,
                 boxes=(None, None), regularize=False, instanced=False, many_lights=0, infinite=None,
                 film="rgb", nbuckets=31, patches=False, smooth=False, alpha=False, lights=(), sampler="zsobol",
                 pixelfilter="box", envmap=0):
    """BASELINE config 1 as a .pbrt file over cornell_box()'s geometry: diffuse walls (white / red / green),
    ceiling quad emitter `rgb L [17 12 4]`, perspective camera fov 39.3, rgb film, box filter, zsobol."""
    sc = cornell_box()
    white, red, green = (0.73, 0.73, 0.73), (0.65, 0.05, 0.05), (0.12, 0.45, 0.15)
    colours = [white, white, white, red, green, None, white, white]
    override = {6: boxes[0], 7: boxes[1]}  # the short and the tall box
    lines = [
        'LookAt 0.5 0.5 -1.4  0.5 0.5 0.5  0 1 0',
        'Camera "perspective" "float fov" 39.3',
        f'Sampler "{sampler}" "integer pixelsamples" {spp} "integer seed" {seed}',
        f'Integrator "path" "integer maxdepth" {maxdepth} "string lightsampler" "{lightsampler}"'
        + (' "bool regularize" true' if regularize else ''),
        'PixelFilter ' + (pixelfilter if pixelfilter.startswith('"') else f'"{pixelfilter}"'),  # a name, or name + parameters
        (f'Film "rgb" "string filename" "{out}" "integer xresolution" {res} "integer yresolution" {res} "bool savefp16" false'
         if film == "rgb" else
         _gbuffer_film(out, res, res, albedo=film == "gbuffer_albedo") if film in ("gbuffer", "gbuffer_albedo") else
         # SpectralFilm (film.h:430-560) only accepts an .exr name; the oracle reads its Image out directly
         f'Film "spectral" "string filename" "{os.path.splitext(out)[0]}.exr" "integer nbuckets" {nbuckets} '
         f'"integer xresolution" {res} "integer yresolution" {res} "bool savefp16" false'),
        'WorldBegin',
    ]
    if instanced:
        # the two boxes become ONE object definition (a unit cube) placed twice (config 3's mechanism):
        # ObjectBegin/ObjectInstance -> TransformedPrimitive over a per-object BVH (scene.cpp:1523-1575)
        unit = _box((-0.5, -0.5, -0.5), (0.5, 0.5, 0.5))
        lines += ['AttributeBegin', '  ' + MATERIALS.get(boxes[0], boxes[0] or f'Material "diffuse" "rgb reflectance" [{white[0]} {white[1]} {white[2]}]'),
                  '  ObjectBegin "cube"', '    ' + _mesh_pbrt(unit), '  ObjectEnd', 'AttributeEnd']
        for (tx, ty, tz, sx, sy, sz, rot) in [(0.3, 0.3, 0.65, 0.3, 0.6, 0.3, 17.0), (0.7, 0.15, 0.35, 0.3, 0.3, 0.3, -18.0)]:
            lines += ['AttributeBegin', f'  Translate {tx} {ty} {tz}', f'  Rotate {rot} 0 1 0', f'  Scale {sx} {sy} {sz}',
                      '  ObjectInstance "cube"', 'AttributeEnd']
    for k, (m, c) in enumerate(zip(sc.meshes, colours)):
        if instanced and k in (6, 7):
            continue
        lines.append('AttributeBegin')
        if override.get(k):
            lines.append('  ' + MATERIALS.get(override[k], override[k]))
        elif c is None:
            lines.append('  AreaLightSource "diffuse" "rgb L" [17 12 4]')
            lines.append('  Material "diffuse" "rgb reflectance" [0 0 0]')
            if patches == "emissive":
                # the ceiling emitter as ONE rectangular bilinear patch (spherical-rectangle sampling, shapes.cpp:1281-1330)
                # and a twisted one-sided emissive patch in front of the back wall (area sampling, :1266-1280)
                q = m.p[[0, 1, 3, 2]]
                lines.append('  Shape "bilinearmesh" "integer indices" [0 1 2 3] "point3 P" ['
                             + " ".join(repr(float(x)) for x in q.reshape(-1)) + ']')
                lines += ['AttributeEnd', 'AttributeBegin', '  AreaLightSource "diffuse" "rgb L" [3 6 9]',
                          '  Material "diffuse" "rgb reflectance" [0 0 0]', '  ReverseOrientation',
                          '  Shape "bilinearmesh" "integer indices" [0 1 2 3] "point3 P" '
                          '[0.3 0.55 0.93  0.5 0.5 0.97  0.32 0.8 0.97  0.52 0.75 0.9]', 'AttributeEnd']
                continue
        else:
            lines.append(f'  Material "diffuse" "rgb reflectance" [{c[0]} {c[1]} {c[2]}]')
        # alpha: the tall box is a stochastic cut-out (constant alpha 0.5), cpu/primitive.cpp:57-70
        lines.append('  ' + _mesh_pbrt(m) + (' "float alpha" 0.5' if alpha and k == 6 else ''))
        lines.append('AttributeEnd')
    for name in lights:  # LIGHTS names or LightSource lines
        lines += ['AttributeBegin', '  ' + LIGHTS.get(name, name), 'AttributeEnd']
    if many_lights:
        # config 3's "many-light" mechanism at test size: small one-sided emissive quads of varied power,
        # position and orientation inside the box (what the light BVH's importance has to tell apart)
        rng = np.random.default_rng(1234 + many_lights)
        for _ in range(many_lights):
            c = rng.uniform(0.08, 0.92, 3)
            n = rng.normal(size=3)
            n /= np.linalg.norm(n)
            a = np.cross(n, [0.0, 1.0, 0.0] if abs(n[1]) < 0.9 else [1.0, 0.0, 0.0])
            a /= np.linalg.norm(a)
            b = np.cross(n, a)
            h = rng.uniform(0.01, 0.03)
            q = [c - h * a - h * b, c + h * a - h * b, c + h * a + h * b, c - h * a + h * b]
            L = rng.uniform(0.5, 40.0) * rng.uniform(0.3, 1.0, 3)
            pts = " ".join(f"{float(v):.7g}" for pt in q for v in pt)
            lines += ['AttributeBegin', f'  AreaLightSource "diffuse" "rgb L" [{L[0]:.5g} {L[1]:.5g} {L[2]:.5g}]',
                      '  Material "diffuse" "rgb reflectance" [0 0 0]',
                      f'  Shape "trianglemesh" "point3 P" [{pts}] "integer indices" [0 1 2 0 2 3]', 'AttributeEnd']
    if patches:
        # two `bilinearmesh` shapes (shapes.cpp:1117-1200) of non-planar patches: a diffuse wavy sheet above the short
        # box and a rough copper saddle by the left wall, the second under ReverseOrientation
        def sheet(x0, x1, z0, z1, n, height):
            g = np.linspace(0.0, 1.0, n + 1)
            pts = [(x0 + (x1 - x0) * u, height(u, v), z0 + (z1 - z0) * v) for v in g for u in g]
            idx = [k for j in range(n) for i in range(n)
                   for k in (j * (n + 1) + i, j * (n + 1) + i + 1, (j + 1) * (n + 1) + i, (j + 1) * (n + 1) + i + 1)]
            txt = (f'Shape "bilinearmesh" "integer indices" [{" ".join(map(str, idx))}] "point3 P" '
                   f'[{" ".join(f"{float(c):.7g}" for pt in pts for c in pt)}]')
            if patches == "normals":  # vertex normals from the height field's finite differences (BilinearPatchMesh::n)
                e = 1e-3
                nrm = []
                for v in g:
                    for u in g:
                        dx = (height(u + e, v) - height(u - e, v)) / (2 * e) / (x1 - x0)
                        dz = (height(u, v + e) - height(u, v - e)) / (2 * e) / (z1 - z0)
                        nv = np.array([-dx, 1.0, -dz])
                        nrm.append(nv / np.linalg.norm(nv))
                txt += f' "normal N" [{" ".join(f"{float(c):.7g}" for nv in nrm for c in nv)}]'
            if patches == "uv":  # a skewed, slightly non-affine (s, t) of its own (BilinearPatchMesh::uv): dpdu / dpdv follow it
                st = [(0.7 * u + 0.5 * v + 0.2 * u * v, -0.3 * u + 0.9 * v) for v in g for u in g]
                txt += f' "point2 uv" [{" ".join(f"{float(c):.7g}" for q in st for c in q)}]'
            return txt
        lines += ['AttributeBegin', ('  ' + MATERIALS["conductor_reflectance"]) if patches == "uv" else
                  '  Material "diffuse" "rgb reflectance" [0.25 0.35 0.75]',
                  '  ' + sheet(0.5, 0.92, 0.12, 0.55, 4, lambda u, v: 0.42 + 0.07 * np.sin(5.0 * u) * np.cos(4.0 * v) + 0.1 * u * v),
                  'AttributeEnd',
                  'AttributeBegin', '  ' + MATERIALS["conductor_reflectance" if patches == "uv" else "conductor_rough"], '  ReverseOrientation',
                  '  ' + sheet(0.05, 0.3, 0.1, 0.45, 1, lambda u, v: 0.75 + 0.18 * (2 * u - 1) * (2 * v - 1)),
                  'AttributeEnd']
        if patches == "normals":
            # a twisted one-sided emitter whose vertex normals (towards the room) decide its emitting side
            lines += ['AttributeBegin', '  AreaLightSource "diffuse" "rgb L" [3 6 9]',
                      '  Material "diffuse" "rgb reflectance" [0 0 0]',
                      '  Shape "bilinearmesh" "integer indices" [0 1 2 3] "point3 P" '
                      '[0.3 0.55 0.93  0.5 0.5 0.97  0.32 0.8 0.97  0.52 0.75 0.9] '
                      '"normal N" [0 0 -1  0.1 0 -1  0 0.1 -1  0.1 0.1 -1]', 'AttributeEnd']
    if alpha:
        # a 30 % screen in front of the camera and a fully cut-out (alpha 0) blocker under the light
        lines += ['AttributeBegin', '  Material "diffuse" "rgb reflectance" [0.2 0.7 0.3]',
                  '  Shape "trianglemesh" "point3 P" [0.2 0.2 0.05  0.8 0.2 0.05  0.8 0.8 0.1  0.2 0.8 0.1] '
                  '"integer indices" [0 1 2 0 2 3] "float alpha" 0.3', 'AttributeEnd',
                  'AttributeBegin', '  Material "diffuse" "rgb reflectance" [0.9 0.1 0.1]',
                  '  Shape "trianglemesh" "point3 P" [0.3 0.9 0.3  0.7 0.9 0.3  0.7 0.9 0.7  0.3 0.9 0.7] '
                  '"integer indices" [0 1 2 0 2 3] "float alpha" 0', 'AttributeEnd']
    if smooth:
        # triangle meshes with per-vertex normals and uv (Triangle::InteractionFromIntersection's shading geometry,
        # shapes.h:934-1005): an anisotropic metal ball, a diffuse ball with normals only, and a small emissive
        # dome whose normals decide its emitting side (Triangle::Sample's FaceForward(n, ns), shapes.h:1024-1029)
        def ball(c, r, nu, nv, vmax=np.pi, with_uv=True, inward=False, with_n=True, tangents=None):
            th = np.linspace(0.0, vmax, nv + 1)
            ph = np.linspace(0.0, 2 * np.pi, nu + 1)
            nrm = np.array([[np.sin(t) * np.cos(f), np.cos(t), np.sin(t) * np.sin(f)] for t in th for f in ph])
            pts = np.asarray(c) + r * nrm
            uv = np.array([[f / (2 * np.pi), t / np.pi] for t in th for f in ph])
            idx = []
            for j in range(nv):
                for i in range(nu):
                    a, b = j * (nu + 1) + i, (j + 1) * (nu + 1) + i
                    if j > 0:
                        idx += [a, a + 1, b]
                    if j < nv - 1 or vmax < np.pi:
                        idx += [a + 1, b + 1, b]
            if inward:
                nrm = -nrm
            txt = (f'Shape "trianglemesh" "integer indices" [{" ".join(map(str, idx))}] "point3 P" '
                   f'[{" ".join(f"{float(x):.7g}" for x in pts.reshape(-1))}]')
            if with_n:
                txt += f' "normal N" [{" ".join(f"{float(x):.7g}" for x in nrm.reshape(-1))}]'
            if tangents is not None:
                # per-vertex shading tangents (TriangleMesh::s, shapes.h:950-959): a swirl about `tangents`, an axis that
                # is not the sphere's own; zero at the two vertices where the axis pierces it (falls back to dpdu)
                sv = np.cross(nrm, np.asarray(tangents, float) / np.linalg.norm(tangents))
                txt += f' "vector3 S" [{" ".join(f"{float(x):.7g}" for x in sv.reshape(-1))}]'
            if with_uv:
                txt += f' "point2 uv" [{" ".join(f"{float(x):.7g}" for x in uv.reshape(-1))}]'
            return txt
        if smooth == "instanced":
            # the metal ball as an object instance under a rotation and a non-uniform scale: the interaction's
            # shading normal and tangent go through Transform::operator()(SurfaceInteraction) (util/transform.cpp:227-258)
            lines += ['AttributeBegin', '  ' + MATERIALS["conductor_reflectance"], '  ObjectBegin "ball"',
                      '    ' + ball((0.0, 0.0, 0.0), 1.0, 12, 8), '  ObjectEnd', 'AttributeEnd',
                      'AttributeBegin', '  Translate 0.68 0.5 0.3', '  Rotate 35 0.3 1 0.2', '  Scale 0.17 0.12 0.15',
                      '  ObjectInstance "ball"', 'AttributeEnd',
                      'AttributeBegin', '  Translate 0.3 0.25 0.2', '  Rotate -20 1 0 0.4', '  Scale 0.08 0.1 0.08',
                      '  ObjectInstance "ball"', 'AttributeEnd']
        elif smooth == "tangents":
            # the anisotropic metal ball with normals, uv AND tangents; a second, flat-shaded one with tangents only
            lines += ['AttributeBegin', '  ' + MATERIALS["conductor_reflectance"],
                      '  ' + ball((0.68, 0.5, 0.3), 0.16, 12, 8, tangents=(0.3, 0.2, 1.0)), 'AttributeEnd',
                      'AttributeBegin', '  ' + MATERIALS["conductor_reflectance"],
                      '  ' + ball((0.3, 0.3, 0.25), 0.12, 10, 6, with_uv=False, with_n=False, tangents=(1.0, 0.4, 0.1)),
                      'AttributeEnd']
        else:
            lines += ['AttributeBegin', '  ' + MATERIALS["conductor_reflectance"], '  ' + ball((0.68, 0.5, 0.3), 0.16, 12, 8),
                      'AttributeEnd']
        lines += [
                  'AttributeBegin', '  Material "diffuse" "rgb reflectance" [0.3 0.6 0.7]',
                  '  ' + ball((0.25, 0.78, 0.35), 0.13, 10, 6, with_uv=False), 'AttributeEnd',
                  'AttributeBegin', '  AreaLightSource "diffuse" "rgb L" [4 5 9]',
                  '  Material "diffuse" "rgb reflectance" [0 0 0]',
                  '  ' + ball((0.15, 0.05, 0.2), 0.1, 8, 3, vmax=0.5 * np.pi, with_uv=False), 'AttributeEnd']
    if infinite is not None:
        lines.append(f'LightSource "infinite" "rgb L" [{infinite[0]} {infinite[1]} {infinite[2]}]')
    if envmap:  # an image infinite light of envmap x envmap pixels; reaches the room through its open front
        lines += _envmap_lines(path, int(envmap))
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return sc


def _bumpy_torus(rng, nu=32, nv=32):
    """Closed surface without poles, 2 * nu * nv triangles inside [-1/2, 1/2]^3: a torus whose tube radius
    is modulated by a few random harmonics."""
    u = np.arange(nu) * (2 * np.pi / nu)
    v = np.arange(nv) * (2 * np.pi / nv)
    U, V = np.meshgrid(u, v, indexing="ij")
    r = np.full_like(U, 0.13)
    for _ in range(4):
        a, b = rng.integers(1, 5, 2)
        r += rng.uniform(-0.025, 0.025) * np.cos(a * U + b * V + rng.uniform(0, 2 * np.pi))
    R = 0.3
    x = (R + r * np.cos(V)) * np.cos(U)
    z = (R + r * np.cos(V)) * np.sin(U)
    y = r * np.sin(V) * rng.uniform(0.6, 1.4)
    p = np.stack([x, y, z], axis=-1).reshape(-1, 3).astype(np.float32)
    i, j = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    a = (i * nv + j).reshape(-1)
    b = (((i + 1) % nu) * nv + j).reshape(-1)
    c = (((i + 1) % nu) * nv + (j + 1) % nv).reshape(-1)
    d = (i * nv + (j + 1) % nv).reshape(-1)
    idx = np.concatenate([np.stack([a, b, c], 1), np.stack([a, c, d], 1)]).astype(np.int32)
    return Mesh(p, idx)


CONFIG3_MATERIALS = ["dielectric_smooth", "conductor_rough", "coated_diffuse", None, "dielectric_rough",
                     "conductor_smooth", "coated_diffuse_rough", "conductor_reflectance"]


def write_ply(path, mesh):
    """Binary little-endian PLY of a triangle mesh (float32 x y z; uchar-counted int32 faces), as pbrt's
    `Shape "plymesh"` reads it (util/mesh.cpp TriQuadMesh::ReadPLY)."""
    p = np.ascontiguousarray(mesh.p, dtype="<f4")
    idx = np.ascontiguousarray(mesh.indices, dtype="<i4").reshape(-1, 3)
    face = np.empty(len(idx), dtype=[("n", "u1"), ("i", "<i4", 3)])
    face["n"], face["i"] = 3, idx
    with open(path, "wb") as f:
        f.write((f"ply\nformat binary_little_endian 1.0\nelement vertex {len(p)}\nproperty float x\nproperty float y\n"
                 f"property float z\nelement face {len(idx)}\nproperty list uchar int vertex_indices\nend_header\n").encode())
        f.write(p.tobytes())
        f.write(face.tobytes())


def config3_object_meshes(n_objects, n_instances, grid):
    """The object definitions of config3_pbrt(n_objects, grid, n_instances, ...), in object space: the generator's
    random stream does not depend on `grid`, so the same shapes come out at any tessellation."""
    rng = np.random.default_rng(3000 + n_objects * 7 + n_instances)
    out = []
    for _ in range(n_objects):
        rng.uniform(0.2, 0.8, 3)  # the object's diffuse colour
        out.append(_bumpy_torus(rng, grid, grid))
    return out


def config3_refine(d, n_objects, n_instances, grid):
    """Scene tables `d` dumped from config3_pbrt at a coarse tessellation -> the same scene with every object
    definition tessellated at `grid` (2 * grid^2 triangles each): BASELINE configs[2] at its full size of unique
    triangles without a dump of that size in the repository.  The dump holds an object's vertices in the render
    space of its definition, object space plus one translation, which is read off the coarse meshes and checked."""
    nv, nt = d["mesh_nverts"].astype(np.int64), d["mesh_ntris"].astype(np.int64)
    vo, to = np.concatenate([[0], np.cumsum(nv)]), np.concatenate([[0], np.cumsum(nt)])
    obj = d["mesh_object"]
    coarse_grid = int(round(np.sqrt(nt[obj >= 0][0] / 2)))
    coarse = config3_object_meshes(n_objects, n_instances, coarse_grid)
    fine = config3_object_meshes(n_objects, n_instances, grid)
    P, I, nv2, nt2 = [], [], nv.copy(), nt.copy()
    shift = None
    # the host numbers the definitions in its own (hash map) order; instance i places "obj{i % n_objects}"
    assert n_instances >= n_objects
    index_of = {int(d["instance_object"][i]): i for i in range(n_objects)}
    for m in range(len(nv)):
        if obj[m] < 0:
            P.append(d["P"][vo[m]:vo[m + 1]])
            I.append(d["indices"][to[m]:to[m + 1]])
            continue
        c, f = coarse[index_of[int(obj[m])]], fine[index_of[int(obj[m])]]
        dumped = d["P"][vo[m]:vo[m + 1]]
        if shift is None:
            shift = (dumped[0].astype(np.float64) - c.p[0].astype(np.float64)).astype(np.float32)
        # the reference applies render-from-object as x * 1 + t in float: one rounding
        assert np.array_equal(dumped, c.p + shift) and np.array_equal(d["indices"][to[m]:to[m + 1]], c.indices), \
            "the dump is not config3_pbrt's scene at a coarser tessellation"
        P.append(f.p + shift)
        I.append(f.indices)
        nv2[m], nt2[m] = len(f.p), len(f.indices)
    out = dict(d)
    out["P"] = np.ascontiguousarray(np.concatenate(P), np.float32)
    out["indices"] = np.ascontiguousarray(np.concatenate(I), np.int32)
    out["mesh_nverts"], out["mesh_ntris"] = nv2.astype(d["mesh_nverts"].dtype), nt2.astype(d["mesh_ntris"].dtype)
    return out


def config3_pbrt(path, n_objects=32, grid=32, n_instances=4883, n_lights=1024, res=(1920, 1080), spp=1024,
                 maxdepth=5, out="config3.pfm", seed=0, crop=None, scale=(0.8, 2.2), ply_dir=None):
    """BASELINE configs[2]: procedural instanced scene (n_instances x 2 * grid^2 triangles ~ 10 M at the
    defaults) of dielectric / conductor / coated / diffuse objects over a ground plane, lit by n_lights
    small emissive quads (2 triangles each) through the light-BVH sampler, plus a dim uniform sky.
    Every object definition carries one material; instances are rotation x uniform scale x translation.
    crop = (x0, x1, y0, y1) in NDC renders a window of the frame (for the bounded CPU baseline)."""
    rng = np.random.default_rng(3000 + n_objects * 7 + n_instances)

    def shape(o, mesh):  # inline, or a binary PLY next to the scene file (ply_dir: full-size meshes)
        if ply_dir is None:
            return _mesh_pbrt(mesh)
        write_ply(os.path.join(ply_dir, f"obj{o}.ply"), mesh)
        return f'Shape "plymesh" "string filename" "{os.path.join(ply_dir, f"obj{o}.ply")}"'

    lines = [
        'LookAt 0 14 -48  0 3 0  0 1 0',
        'Camera "perspective" "float fov" 38',
        f'Sampler "zsobol" "integer pixelsamples" {spp} "integer seed" {seed}',
        f'Integrator "path" "integer maxdepth" {maxdepth} "string lightsampler" "bvh"',
        'PixelFilter "box"',
        f'Film "rgb" "string filename" "{out}" "integer xresolution" {res[0]} "integer yresolution" {res[1]} '
        '"bool savefp16" false' + (f' "float cropwindow" [{crop[0]} {crop[1]} {crop[2]} {crop[3]}]' if crop else ''),
        'WorldBegin',
        'LightSource "infinite" "rgb L" [0.03 0.04 0.06]',
        'AttributeBegin', '  Material "diffuse" "rgb reflectance" [0.45 0.42 0.38]',
        '  ' + _mesh_pbrt(_quad([-60, 0, -60], [-60, 0, 60], [60, 0, 60], [60, 0, -60])), 'AttributeEnd',
    ]
    for o in range(n_objects):
        mat = CONFIG3_MATERIALS[o % len(CONFIG3_MATERIALS)]
        c = rng.uniform(0.2, 0.8, 3)
        lines += ['AttributeBegin',
                  '  ' + (MATERIALS[mat] if mat else f'Material "diffuse" "rgb reflectance" [{c[0]:.3f} {c[1]:.3f} {c[2]:.3f}]'),
                  f'  ObjectBegin "obj{o}"', '    ' + shape(o, _bumpy_torus(rng, grid, grid)), '  ObjectEnd', 'AttributeEnd']
    for i in range(n_instances):
        t = (rng.uniform(-30, 30), rng.uniform(0.6, 9.0), rng.uniform(-30, 30))
        axis = rng.normal(size=3)
        axis /= np.linalg.norm(axis)
        lines += ['AttributeBegin', f'  Translate {t[0]:.5f} {t[1]:.5f} {t[2]:.5f}',
                  f'  Rotate {rng.uniform(0, 360):.4f} {axis[0]:.5f} {axis[1]:.5f} {axis[2]:.5f}',
                  f'  Scale {rng.uniform(scale[0], scale[1]):.5f} ' * 1, ]
        sc = lines.pop().split()[1]
        lines += [f'  Scale {sc} {sc} {sc}', f'  ObjectInstance "obj{i % n_objects}"', 'AttributeEnd']
    palette = [rng.uniform(20.0, 200.0) * rng.uniform(0.3, 1.0, 3) for _ in range(32)]  # 32 emission spectra
    for k in range(n_lights):
        c = np.array([rng.uniform(-32, 32), rng.uniform(5.0, 14.0), rng.uniform(-32, 32)])
        n = rng.normal(size=3) * 0.5 + np.array([0.0, -1.0, 0.0])
        n /= np.linalg.norm(n)
        a = np.cross(n, [1.0, 0.0, 0.0])
        a /= np.linalg.norm(a)
        b = np.cross(n, a)
        h = rng.uniform(0.1, 0.3)
        q = [c - h * a - h * b, c + h * a - h * b, c + h * a + h * b, c - h * a + h * b]
        L = palette[k % 32]
        pts = " ".join(f"{float(v):.7g}" for pt in q for v in pt)
        lines += ['AttributeBegin', f'  AreaLightSource "diffuse" "rgb L" [{L[0]:.5g} {L[1]:.5g} {L[2]:.5g}]',
                  '  Material "diffuse" "rgb reflectance" [0 0 0]',
                  f'  Shape "trianglemesh" "point3 P" [{pts}] "integer indices" [0 1 2 0 2 3]', 'AttributeEnd']
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")


# A double-Gauss 50 mm f/2 prescription in pbrt's lens-file units (mm): curvature radius, thickness, index,
# aperture diameter, front element first; radius 0 = the aperture stop (RealisticCamera::Create,
# cameras.cpp:1380-1392; constructor :781-802).
DGAUSS_50MM = [
    (29.475, 3.76, 1.67, 25.2), (84.83, 0.12, 1.0, 25.2), (19.275, 4.025, 1.67, 23.0),
    (40.77, 3.275, 1.699, 23.0), (12.75, 5.705, 1.0, 18.0), (0.0, 4.5, 0.0, 17.1),
    (-14.495, 1.18, 1.603, 17.0), (40.77, 6.065, 1.658, 20.0), (-20.385, 0.19, 1.0, 20.0),
    (437.065, 3.22, 1.717, 20.0), (-39.73, 5.0, 1.0, 20.0),
]


def chart_spectra(n=24, seed=2404):
    """n smooth reflectance spectra on 10 nm knots over 380..720 nm (a Macbeth-like chart: the last six are a
    grey ramp, the rest sums of two sigmoids / a bump), values in [0.03, 0.92]."""
    rng = np.random.default_rng(seed)
    lam = np.arange(380.0, 721.0, 10.0)
    out = []
    for k in range(n):
        if k >= n - 6:
            v = np.full_like(lam, 0.9 * 0.55 ** (k - (n - 6)))
        else:
            c0, c1 = rng.uniform(420, 680, 2)
            w0, w1 = rng.uniform(15, 60, 2)
            a0, a1 = rng.uniform(-0.6, 0.6, 2)
            v = 0.35 + a0 / (1 + np.exp(-(lam - c0) / w0)) + a1 * np.exp(-0.5 * ((lam - c1) / w1) ** 2)
        out.append(np.clip(v, 0.03, 0.92))
    return lam, np.array(out)


def config4_pbrt(path, res=1024, spp=256, maxdepth=5, out="config4.exr", seed=0, nbuckets=31, aperture_mm=12.0,
                 focus=1.9, film="spectral", crop=None, camera="realistic", chromatic=False, microlens=None):
    """BASELINE configs[3]: the Cornell room with a 24-patch spectral reflectance chart on the back wall, seen
    through a RealisticCamera (double-Gauss 50 mm lens file written next to the scene) onto a SpectralFilm
    with `nbuckets` bands over 395..705 nm (the ISET 31-band setting)."""
    sc = cornell_box()
    lens = os.path.splitext(path)[0] + "_lens.dat"
    with open(lens, "w") as f:
        f.write("# double-Gauss 50 mm: radius thickness eta aperture-diameter (mm)\n")
        for e in DGAUSS_50MM:
            f.write(" ".join(repr(float(x)) for x in e) + "\n")
    if camera == "omni":
        # the ISET OmniCamera reads the same prescription as JSON (cameras.cpp:3812-3960): per surface radius,
        # thickness, ior, semi_aperture (mm) and an optional 3x3 + translation transform; two elements are
        # decentred / tilted here so the per-element transforms matter
        import json
        surfaces = []
        for k, (radius, thick, eta, diam) in enumerate(DGAUSS_50MM):
            surf = {"radius": radius, "thickness": thick, "ior": eta if eta else 1.0, "semi_aperture": diam / 2,
                    "conic_constant": 0.0}
            if k == 2:
                surf["transform"] = [[1, 0, 0], [0, 1, 0], [0, 0, 1], [0.15, -0.1, 0.0]]
            if k == 7:
                c, sn = float(np.cos(0.01)), float(np.sin(0.01))
                surf["transform"] = [[c, 0, -sn], [0, 1, 0], [sn, 0, c], [0.0, 0.05, 0.0]]
            surfaces.append(surf)
        lens = os.path.splitext(path)[0] + "_lens.json"
        spec = {"name": "dgauss50", "description": "double-Gauss 50 mm", "surfaces": surfaces}
        if microlens:
            # a microlens array in front of the sensor (light-field camera, cameras.cpp:3259-3327): `microlens` lenses
            # across the film, each a thin biconvex element; per-lens centre offsets in metres (cameras.cpp:3950-3962)
            n = int(microlens)
            rng = np.random.default_rng(5)
            spec["microlens"] = {
                "dimensions": [n, n],
                "offsets": [[float(a), float(b)] for a, b in rng.uniform(-2e-5, 2e-5, (n * n, 2))],
                "surfaces": [
                    {"radius": 1.2, "thickness": 0.3, "ior": 1.5, "semi_aperture": 0.9, "conic_constant": 0.0},
                    {"radius": -1.5, "thickness": 0.6, "ior": 1.0, "semi_aperture": 0.9, "conic_constant": 0.0},
                ]}
        with open(lens, "w") as f:
            json.dump(spec, f)
    white, red, green = (0.73, 0.73, 0.73), (0.65, 0.05, 0.05), (0.12, 0.45, 0.15)
    colours = [white, white, white, red, green, None, white, white]
    cropw = f' "float cropwindow" [{crop[0]} {crop[1]} {crop[2]} {crop[3]}]' if crop else ''
    lines = [
        'LookAt 0.5 0.5 -1.4  0.5 0.5 0.5  0 1 0',
        f'Camera "{camera}" "string lensfile" "{lens}" "float aperturediameter" {aperture_mm} '
        f'"float focusdistance" {focus}' + (' "bool chromaticAberrationEnabled" true' if chromatic else '')
        # with decentred elements the reference's thick-lens focusing diverges: the film distance is given instead
        # (the value FocusThickLens finds for the centred prescription at this focus distance)
        + (' "float filmdistance" 0.0375136' if camera == "omni" else '')
        + (' "float microlenssensoroffset" 0.0006 "integer microlenssimulationradius" 1' if microlens else ''),
        f'Sampler "zsobol" "integer pixelsamples" {spp} "integer seed" {seed}',
        f'Integrator "path" "integer maxdepth" {maxdepth} "string lightsampler" "power"',
        'PixelFilter "box"',
        (f'Film "spectral" "string filename" "{os.path.splitext(out)[0]}.exr" "integer nbuckets" {nbuckets} '
         f'"integer xresolution" {res} "integer yresolution" {res} "bool savefp16" false' if film == "spectral" else
         _gbuffer_film(out, res, res) if film == "gbuffer" else
         f'Film "rgb" "string filename" "{out}" "integer xresolution" {res} "integer yresolution" {res} '
         '"bool savefp16" false') + cropw,
        'WorldBegin',
    ]
    for m, c in zip(sc.meshes, colours):
        lines.append('AttributeBegin')
        if c is None:
            lines.append('  AreaLightSource "diffuse" "rgb L" [17 12 4]')
            lines.append('  Material "diffuse" "rgb reflectance" [0 0 0]')
        else:
            lines.append(f'  Material "diffuse" "rgb reflectance" [{c[0]} {c[1]} {c[2]}]')
        lines.append('  ' + _mesh_pbrt(m))
        lines.append('AttributeEnd')
    lam, spectra = chart_spectra()
    # 6 x 4 patches just in front of the back wall (z = 1), facing the camera
    x0, y0, size, gap, z = 0.14, 0.52, 0.105, 0.018, 0.995
    for k, v in enumerate(spectra):
        i, j = k % 6, k // 6
        xa, ya = x0 + i * (size + gap), y0 + (3 - j) * (size + gap) * 0.9
        xb, yb = xa + size, ya + size * 0.9
        sp = " ".join(f"{l:g} {r:.5f}" for l, r in zip(lam, v))
        lines += ['AttributeBegin', f'  Material "diffuse" "spectrum reflectance" [{sp}]',
                  '  ' + _mesh_pbrt(_quad([xa, ya, z], [xa, yb, z], [xb, yb, z], [xb, ya, z])), 'AttributeEnd']
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return sc


def procedural_density(n=256, seed=7):
    """n^3 density in [0, 1]: three octaves of value noise (hash lattice, trilinear) shaped into a ball that
    fades out towards the grid's faces; x fastest, as SampledGrid stores it (util/containers.h:826)."""
    rng = np.random.default_rng(seed)
    t = (np.arange(n, dtype=np.float32) + 0.5) / n
    out = np.zeros((n, n, n), np.float32)
    amp = 1.0
    for cells in (4, 8, 16):
        lat = rng.random((cells + 1,) * 3, dtype=np.float32)
        x = t * cells
        i = np.minimum(x.astype(np.int32), cells - 1)
        f = (x - i).astype(np.float32)
        f = f * f * (3 - 2 * f)
        def ax(a, axis):
            sh = [1, 1, 1]
            sh[axis] = n
            lo, hi = np.take(a, i, axis=axis), np.take(a, i + 1, axis=axis)
            return lo + (hi - lo) * f.reshape(sh)
        out += amp * ax(ax(ax(lat, 0), 1), 2)
        amp *= 0.5
    out /= 1.75
    z, y, x = np.meshgrid(t, t, t, indexing="ij")
    r = np.sqrt((x - 0.5) ** 2 + (y - 0.5) ** 2 + (z - 0.5) ** 2)
    out = np.clip((out - 0.25) * 2.2, 0, 1) * np.clip((0.5 - r) * 6, 0, 1)
    return np.ascontiguousarray(out.astype(np.float32))


def majorant_grid(density, res=16):
    """MajorantGrid of GridMedium's constructor (media.cpp:240-246) over SampledGrid::MaxValue
    (util/containers.h:836-853): per voxel of a res^3 grid, the max of the density samples whose trilinear
    support overlaps it."""
    nz, ny, nx = density.shape
    out = np.zeros((res, res, res), np.float32)
    def rng_(k, n):
        f32 = np.float32
        lo = int(np.floor(f32(f32(k) / f32(res)) * f32(n) - f32(0.5)))
        hi = int(np.floor(f32(f32(k + 1) / f32(res)) * f32(n) - f32(0.5)))
        return max(lo, 0), min(hi + 1, n - 1)
    for z in range(res):
        z0, z1 = rng_(z, nz)
        for y in range(res):
            y0, y1 = rng_(y, ny)
            for x in range(res):
                x0, x1 = rng_(x, nx)
                out[z, y, x] = density[z0:z1 + 1, y0:y1 + 1, x0:x1 + 1].max()
    return out


_DENSITY_TEXT = {}


def with_density_grid(d, density, medium=0):
    """A copy of .b2rs tables `d` whose grid medium `medium` holds `density` (nz, ny, nx) and the majorant grid
    the host would build for it (majorant_grid)."""
    d = dict(d)
    density = np.ascontiguousarray(density, np.float32)
    d[f"medium_density_{medium}"] = density
    mi = d["media_i"].copy()
    res = int(mi[medium][6])
    mi[medium][3:6] = [density.shape[2], density.shape[1], density.shape[0]]
    d["media_i"] = mi
    d[f"medium_majorants_{medium}"] = majorant_grid(density, res)
    return d


def config5_pbrt(path, res=(1920, 1080), spp=512, maxdepth=5, out="config5.pfm", seed=0, grid=256, density=None,
                 fog=True, crop=None, lightsampler="bvh", patch_cap=False, lights=(), sampler="zsobol", envmap=0,
                 emissive=False):
    """BASELINE configs[4]: VolPathIntegrator in the Cornell room: a thin homogeneous fog fills the room (an
    `interface` quad closes the open front; the camera stays in vacuum: a camera medium trips a race in the
    reference's BasicScene::GetMedium / CreateMedia, scene.cpp:809-850), and a box-shaped `interface` shell
    holds a "uniformgrid" medium with a procedural grid^3 density (procedural_density); ceiling area light."""
    sc = cornell_box()
    white, red, green = (0.73, 0.73, 0.73), (0.65, 0.05, 0.05), (0.12, 0.45, 0.15)
    colours = [white, white, white, red, green, None, white, None]  # the short box stays, the tall one makes room
    dens = procedural_density(grid) if density is None else density
    nz, ny, nx = dens.shape
    cropw = f' "float cropwindow" [{crop[0]} {crop[1]} {crop[2]} {crop[3]}]' if crop else ''
    lines = ['LookAt 0.5 0.5 -1.4  0.5 0.5 0.5  0 1 0']
    lines += [
        'Camera "perspective" "float fov" 39.3',
        f'Sampler "{sampler}" "integer pixelsamples" {spp} "integer seed" {seed}',
        f'Integrator "volpath" "integer maxdepth" {maxdepth} "string lightsampler" "{lightsampler}"',
        'PixelFilter "box"',
        f'Film "rgb" "string filename" "{out}" "integer xresolution" {res[0]} "integer yresolution" {res[1]} '
        '"bool savefp16" false' + cropw,
        'WorldBegin',
    ]
    outside = '"fog"' if fog else '""'
    if fog:
        lines += ['MakeNamedMedium "fog" "string type" "homogeneous" "rgb sigma_a" [0.02 0.02 0.03] '
                  '"rgb sigma_s" [0.25 0.25 0.3] "float scale" 1 "float g" 0.3'
                  # emissive: a faintly glowing fog (HomogeneousMedium's Le, media.h:217-262)
                  + (' "rgb Le" [0.5 0.7 1.0] "float Lescale" 0.4' if emissive else ''),
                  # front cap, normal towards the camera (-z): fog behind it, vacuum in front
                  'AttributeBegin', '  MediumInterface "fog" ""', '  Material "interface"',
                  # patch_cap: the same quad as ONE bilinear patch (medium transitions by a patch's normal)
                  '  ' + ('Shape "bilinearmesh" "integer indices" [0 1 2 3] "point3 P" [0 0 0  0 1 0  1 0 0  1 1 0]'
                          if patch_cap else _mesh_pbrt(_quad([0, 0, 0], [0, 1, 0], [1, 1, 0], [1, 0, 0]))), 'AttributeEnd']
    lo, hi = (0.22, 0.18, 0.25), (0.78, 0.74, 0.81)
    lines.append(f'MakeNamedMedium "cloud" "string type" "uniformgrid" "integer nx" {nx} "integer ny" {ny} "integer nz" {nz} '
                 f'"point3 p0" [{lo[0]} {lo[1]} {lo[2]}] "point3 p1" [{hi[0]} {hi[1]} {hi[2]}] '
                 '"rgb sigma_a" [0.5 0.5 0.5] "rgb sigma_s" [14 14 15] "float scale" 1 "float g" 0.6 '
                 # emissive: the cloud glows where an LeScale grid of its own resolution says so (GridMedium, media.h:301-318);
                 # "half" = no "Lescale" array (the reference then makes a 1 x 1 x 1 grid: a tent over the bounds under its
                 # trilinear lookup with zeros outside)
                 # "blackbody" = a temperature grid (800 .. 4000 on its own scale, with temperaturescale / offset) instead of Le
                 + ('"float temperature" [' + " ".join(f"{v:.6g}" for v in 2400.0 + 1600.0 * np.sin(np.arange(nx * ny * nz) * 0.11))
                    + '] "float temperaturescale" 0.9 "float temperatureoffset" 150 "float Lescale" ['
                    + " ".join(f"{v:.5g}" for v in np.full(nx * ny * nz, 1.5)) + '] ' if emissive == "blackbody" else
                    ('"rgb Le" [1.0 0.55 0.2] ' + ('' if emissive == "half" else '"float Lescale" [' + " ".join(
                        f"{v:.5g}" for v in np.clip(np.sin(np.arange(nx * ny * nz) * 0.37) * 3.0, 0.0, None)) + '] '))
                    if emissive else '')
                 + '"float density" [')
    key = (dens.shape, hash(dens.tobytes()))
    if _DENSITY_TEXT.get("key") != key:  # formatting 256^3 values takes ~10 s: keep the last grid's text
        flat = dens.reshape(-1)
        _DENSITY_TEXT["key"] = key
        _DENSITY_TEXT["text"] = "\n".join(" ".join(f"{v:.6g}" for v in flat[k:k + 65536]) for k in range(0, len(flat), 65536))
    lines.append(_DENSITY_TEXT["text"])
    lines.append(']')
    for m, c in zip(sc.meshes, colours):
        if c is None and m is not sc.meshes[5]:
            continue
        lines.append('AttributeBegin')
        if c is None:
            lines.append('  AreaLightSource "diffuse" "rgb L" [17 12 4]')
            lines.append('  Material "diffuse" "rgb reflectance" [0 0 0]')
        else:
            lines.append(f'  Material "diffuse" "rgb reflectance" [{c[0]} {c[1]} {c[2]}]')
        lines.append('  ' + _mesh_pbrt(m))
        lines.append('AttributeEnd')
    if envmap:  # seen through the front cap (an interface quad) by escaped rays and by light samples
        lines += _envmap_lines(path, int(envmap))
    for name in lights:  # LIGHTS names or LightSource lines; inside the room, so in the fog
        lines += ['AttributeBegin', f'  MediumInterface {outside} {outside}', '  ' + LIGHTS.get(name, name), 'AttributeEnd']
    # the shell: interface material (no BSDF), cloud inside, fog (or vacuum) outside; slightly larger than
    # the grid's box so every ray segment inside the shell sees the whole grid
    e = 0.004
    shell = _box((lo[0] - e, lo[1] - e, lo[2] - e), (hi[0] + e, hi[1] + e, hi[2] + e))
    lines += ['AttributeBegin', f'  MediumInterface "cloud" {outside}', '  Material "interface"',
              '  ' + _mesh_pbrt(shell), 'AttributeEnd']
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return sc


def load_b2rs(path):
    """Reads the named-array container written by a host adaptor (oracle/dumpscene_ref.cpp is the
    test-side one): 'B2RS', u32 count, {u32 nameLen, name, u32 dtype, u32 ndim, u32 dims[], data}."""
    out = {}
    with open(path, "rb") as f:
        magic, count = np.fromfile(f, dtype="<u4", count=2)
        assert magic == 0x53523242, "not a .b2rs file"
        for _ in range(int(count)):
            nl = int(np.fromfile(f, dtype="<u4", count=1)[0])
            name = f.read(nl).decode()
            dtype, nd = (int(x) for x in np.fromfile(f, dtype="<u4", count=2))
            dims = [int(x) for x in np.fromfile(f, dtype="<u4", count=nd)]
            n = int(np.prod(dims)) if dims else 1
            out[name] = np.fromfile(f, dtype=["<f4", "<i4", "<u4"][dtype], count=n).reshape(dims)
    return out
