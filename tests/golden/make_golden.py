"""Regenerates tests/golden/*.npz (ray batches, oracle/_ref/raycast_ref) and the whole-image
fixtures *.b2rs + *_ref.npy (oracle/_ref/dumpscene_ref + pbrt_ref) from the compiled reference.
Run in the build container (needs /root/reference -> `make -f oracle/ref.mk`):
    python tests/golden/make_golden.py
Each fixture stores the ray batch and the reference's BVHAggregate::Intersect / IntersectP
answers for a scene that tests re-create from the same seeds (cases() below)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def cases(S):
    soup = S.random_soup(20000, seed=21, n_meshes=2)
    cb = S.cornell_box()
    sheet = S.Scene([S.grid_sheet(32, 32, jitter=0.0), S.grid_sheet(24, 24, z=0.75, jitter=0.02, seed=9)])
    inst = S.instanced_scene(n_objects=3, tris_per_object=600, n_instances=24, world_tris=1500, seed=41)
    patches = S.patch_scene(6000, 3000, seed=95)
    alpha = S.alpha_scene(seed=43)
    return {
        "alpha_primary": (alpha, S.primary_rays(30000, seed=44)),
        "alpha_adversarial": (alpha, S.adversarial_rays(S.flattened(alpha), 20000, seed=45)),
        "patches_primary": (patches, S.primary_rays(30000, seed=38)),
        "patches_adversarial": (patches, S.adversarial_rays(patches, 20000, seed=39)),
        "instanced_primary": (inst, S.primary_rays(30000, seed=36)),
        "instanced_adversarial": (inst, S.adversarial_rays(S.flattened(inst), 20000, seed=37)),
        "soup20k_primary": (soup, S.primary_rays(40000, seed=31)),
        "soup20k_adversarial": (soup, S.adversarial_rays(soup, 20000, seed=32)),
        "cornell_camera": (cb, S.camera_rays(128, 128, (0.5, 0.5, -1.4), (0.5, 0.5, 0.5), spp=1)),
        "cornell_adversarial": (cb, S.adversarial_rays(cb, 20000, seed=33)),
        "sheets_adversarial": (sheet, S.adversarial_rays(sheet, 20000, seed=34)),
    }


GOLD_DTYPE = np.dtype([("mesh", "<u4"), ("prim", "<u4"), ("instance", "<u4"), ("t", "<f4"), ("uv", "<f4", 2)])


def compact(h):
    out = np.zeros(h.shape[0], dtype=GOLD_DTYPE)
    for f in GOLD_DTYPE.names:
        out[f] = h[f]
    return out


def load(name):
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), name + ".npz"))
    n = z["hits"].shape[0]
    return {"hits": z["hits"], "occ": np.unpackbits(z["occ"])[:n], "shadow_hits": z["shadow_hits"],
            "shadow_occ": np.unpackbits(z["shadow_occ"])[:n]}


# whole-image cases: name -> cornell_pbrt() arguments (config 1 of BASELINE.json, scaled down)
RENDER_CASES = {
    "cornell_r64_s16": dict(res=64, spp=16, maxdepth=5, lightsampler="power", seed=0),
    "cornell_r48_s8_uniform": dict(res=48, spp=8, maxdepth=3, lightsampler="uniform", seed=3),
    "cornell_metal_r64_s16": dict(res=64, spp=16, maxdepth=5, boxes=("conductor_rough", "conductor_smooth")),
    "cornell_metal2_r48_s8": dict(res=48, spp=8, maxdepth=4, boxes=("conductor_reflectance", "dielectric_rough"),
                                  regularize=True),
    "cornell_instanced_r64_s16": dict(res=64, spp=16, maxdepth=5, instanced=True, boxes=("conductor_rough", None)),
    # BVHLightSampler over 202 emitters (the "many-light" mechanism of configs[2]); with a uniform infinite light
    "cornell_manylights_r64_s16": dict(res=64, spp=16, maxdepth=5, lightsampler="bvh", many_lights=200),
    "cornell_infinite_r64_s16": dict(res=64, spp=16, maxdepth=5, lightsampler="bvh", many_lights=20,
                                     infinite=(0.4, 0.5, 0.8)),
    "cornell_infinite_power_r48_s8": dict(res=48, spp=8, maxdepth=4, lightsampler="power", infinite=(0.4, 0.5, 0.8)),
    # SpectralFilm, 31 bands (configs[3]'s film); the reference image has 3 + 31 channels
    "cornell_spectral_r48_s16": dict(res=48, spp=16, maxdepth=5, film="spectral", nbuckets=31),
    # GBufferFilm (the ISET film): RGB + 31 radiance channels + position / ids / normals / variance AOVs, on the
    # instanced Cornell box with smooth-shaded meshes (shading normals differ from the geometric ones, instance ids)
    "cornell_gbuffer_r48_s16": dict(res=48, spp=16, maxdepth=5, film="gbuffer", smooth="instanced"),
    # ... with saveAlbedo: Albedo.R G B from bsdf.rho at the first visible surface (diffuse walls, a rough copper box, a
    # dielectric one whose dispersive index terminates the secondary wavelengths: the film converts with the final lambda)
    "cornell_gbuffer_albedo_r48_s16": dict(res=48, spp=16, maxdepth=5, film="gbuffer_albedo",
                                           boxes=("conductor_rough", "dielectric_glass")),
    "cornell_coated_r64_s16": dict(res=64, spp=16, maxdepth=5, boxes=("coated_diffuse", "coated_diffuse_rough")),
    "cornell_glass_r64_s16": dict(res=64, spp=16, maxdepth=6, boxes=("dielectric_smooth", "dielectric_glass")),
    # bilinear patch meshes shaded (BilinearPatch::InteractionFromIntersection), one under ReverseOrientation
    "cornell_patches_r64_s16": dict(res=64, spp=16, maxdepth=5, patches=True),
    # emissive patches: a rectangle (sampled as a spherical rectangle) and a twisted patch (by area), light BVH
    # per-vertex normals and uv on triangle meshes (shading geometry; an emissive dome oriented by its normals)
    "cornell_smooth_r64_s16": dict(res=64, spp=16, maxdepth=5, smooth=True),
    "cornell_smoothinst_r64_s16": dict(res=64, spp=16, maxdepth=5, smooth="instanced"),
    # per-vertex shading tangents (TriangleMesh::s) under anisotropic metal: with normals and uv, and on their own
    "cornell_tangents_r64_s16": dict(res=64, spp=16, maxdepth=5, smooth="tangents"),
    # alpha cut-outs (constant alpha 0.5 / 0.3 / 0): the test runs inside traversal, for path and shadow rays alike
    "cornell_alpha_r64_s16": dict(res=64, spp=16, maxdepth=5, alpha=True),
    # vertex normals on bilinear patches: shading geometry of a wavy sheet, an emitter oriented by its normals
    "cornell_patchnormals_r64_s16": dict(res=64, spp=16, maxdepth=5, patches="normals"),
    # per-vertex (s, t) on bilinear patches under anisotropic metal: the shading frame follows the mesh's parameterisation
    "cornell_patchuv_r64_s16": dict(res=64, spp=16, maxdepth=5, patches="uv"),
    "cornell_patchlights_r64_s16": dict(res=64, spp=16, maxdepth=5, patches="emissive", lightsampler="bvh"),
}
# Cases whose gate is the Monte-Carlo noise floor (rRMSE(ours, ref) <= 1.25 x rRMSE(ref seed s,
# ref seed s+1)) instead of path-for-path agreement: after a lossless specular bounce the reference's
# Russian-roulette test `rrBeta.MaxComponentValue() < 1` (cpu/integrators.cpp:750) sits on a value
# that is 1 up to rounding, so a last-bit difference in a spectrum lookup decides whether one more
# sampler dimension is consumed; such paths stay valid samples but stop following the reference's.
# For these a second reference image at seed + 1 is stored (<name>_ref2.npy).
# (metal2 follows the reference path for path on the CPU check, which shares glibc's sinf/cosf with the
# reference; on the GPU the last-bit differences of CUDA's sinf/cosf flip a few such roulette tests.)
# (coated: the layered BxDF's random walk is carried as a polynomial in R_k here, csrc/render_layered.h, so its
# internal roulette tests see max_k beta_k with different rounding: most walks are the reference's, a few are not.)
# (alpha: the cut-out decision hashes the ray's bits, HashFloat(ray.o, ray.d); on the CPU check every ray has the
# reference's bits and the image agrees to 6e-6, on the GPU a last-bit difference of sinf / cosf in a sampled direction
# flips decisions further down the path.  The decisions themselves are pinned bit for bit on given rays by the
# alpha_primary / alpha_adversarial traversal fixtures, on the GPU too.)
NOISE_GATED = {"cornell_glass_r64_s16", "cornell_metal2_r48_s8", "cornell_coated_r64_s16", "cornell_alpha_r64_s16"}
# scene tables only (no image): BASELINE.json configs[0] at full size, for bench.py and the
# full-size GPU test, which render the reference image on the spot with oracle/_ref/pbrt_ref
TABLE_CASES = {
    "cornell_r512_s64": dict(res=512, spp=64, maxdepth=5, lightsampler="power", seed=0),
}
# BASELINE.json configs[2] (scenes.config3_pbrt): full size as tables only (bench.py), and in miniature with
# reference images (noise-gated: the scene has smooth glass, see NOISE_GATED)
# Full size: 100 object definitions x 2 * 224^2 = 100 352 triangles (10 M unique), 1000 instances
# (100 M instanced triangles), 10 000 emissive quads = 20 000 lights.  The committed tables are the reference's dump of
# the scene at a coarse tessellation (grid 4); scenes.config3_refine() re-tessellates the object definitions to
# CONFIG3_FULL_GRID with the same generator (checked there, and against a direct dump by tests/test_render_cpu.py).
CONFIG3_FULL = dict(n_objects=100, grid=4, n_instances=1000, n_lights=10000, res=(1920, 1080), spp=1024, maxdepth=5,
                    scale=(2.0, 4.5))
CONFIG3_FULL_GRID = 224
CONFIG3_SMALL = dict(n_objects=8, grid=12, n_instances=300, n_lights=96, res=(96, 54), spp=16, maxdepth=5)
# BASELINE.json configs[3] (scenes.config4_pbrt): RealisticCamera + 31-band SpectralFilm.  Full size as tables
# only (bench.py renders a bounded reference sample on the spot); in miniature with the reference's image.
CONFIG4_FULL = dict(res=1024, spp=256, maxdepth=5)
CONFIG4_SMALL = dict(res=64, spp=16, maxdepth=5)


def make_config4_goldens(P, refrun, out):
    import tempfile
    with tempfile.TemporaryDirectory() as tmp:
        pbrt = os.path.join(tmp, "c4.pbrt")
        # config 4 is quoted on Film "gbuffer" (SURVEY 8d): full-size tables with that film for bench.py; the spectral film
        # keeps its miniature below, the gbuffer film gets one of its own
        P.scenes.config4_pbrt(pbrt, out=os.path.join(tmp, "c4.exr"), film="gbuffer", **CONFIG4_FULL)
        refrun.dump_scene(pbrt, os.path.join(out, "config4_full.b2rs"))
        P.scenes.config4_pbrt(pbrt, out=os.path.join(tmp, "c4.exr"), film="gbuffer", **CONFIG4_SMALL)
        refrun.dump_scene(pbrt, os.path.join(out, "config4_gbuffer_small.b2rs"))
        img, names, secs = refrun.render_channels(pbrt)
        np.save(os.path.join(out, "config4_gbuffer_small_ref.npy"), img.astype(np.float32))
        print("config4_gbuffer_small", img.shape, "mean", img[..., :3].mean(axis=(0, 1)), f"{secs:.2f}s")
        P.scenes.config4_pbrt(pbrt, out=os.path.join(tmp, "c4.exr"), **CONFIG4_SMALL)
        refrun.dump_scene(pbrt, os.path.join(out, "config4_small.b2rs"))
        img, names, secs = refrun.render_channels(pbrt)
        np.save(os.path.join(out, "config4_small_ref.npy"), img.astype(np.float32))
        print("config4_small", img.shape, names[:4], "mean", img[..., :3].mean(axis=(0, 1)), f"{secs:.2f}s")


def make_omni_goldens(P, refrun, out):
    """the config-4 scene through the ISET OmniCamera (JSON prescription, two decentred / tilted
    interfaces), spectral film."""
    import tempfile
    with tempfile.TemporaryDirectory() as tmp:
        pbrt = os.path.join(tmp, "omni.pbrt")
        P.scenes.config4_pbrt(pbrt, out=os.path.join(tmp, "omni.exr"), camera="omni", **CONFIG4_SMALL)
        refrun.dump_scene(pbrt, os.path.join(out, "omni_small.b2rs"))
        img, names, secs = refrun.render_channels(pbrt)
        np.save(os.path.join(out, "omni_small_ref.npy"), img.astype(np.float32))
        print("omni_small", img.shape, "mean", img[..., :3].mean(axis=(0, 1)), f"{secs:.2f}s")
        # the same camera with a 24 x 24 microlens array in front of the sensor (light-field capture)
        P.scenes.config4_pbrt(pbrt, out=os.path.join(tmp, "omni.exr"), camera="omni", microlens=24, **CONFIG4_SMALL)
        refrun.dump_scene(pbrt, os.path.join(out, "omni_microlens_small.b2rs"))
        img, names, secs = refrun.render_channels(pbrt)
        np.save(os.path.join(out, "omni_microlens_small_ref.npy"), img.astype(np.float32))
        print("omni_microlens_small", img.shape, "mean", img[..., :3].mean(axis=(0, 1)), f"{secs:.2f}s")


# BASELINE.json configs[4] (scenes.config5_pbrt): VolPathIntegrator, homogeneous fog + a grid medium inside an
# interface shell.  In miniature (24^3 density) with the reference's image at two seeds; the full-size tables
# (256^3 density, 1920x1080, 512 spp) are derived from the small ones by bench.py (the grid alone is 67 MB).
CONFIG5_FULL = dict(res=(1920, 1080), spp=512, maxdepth=5, grid=256)
CONFIG5_SMALL = dict(res=(64, 48), spp=16, maxdepth=5, grid=24)


def make_config5_goldens(P, refrun, out):
    import tempfile
    with tempfile.TemporaryDirectory() as tmp:
        pbrt, pfm = os.path.join(tmp, "c5.pbrt"), os.path.join(tmp, "c5.pfm")
        for seed, suffix in ((0, "_ref.npy"), (1, "_ref2.npy")):
            P.scenes.config5_pbrt(pbrt, out=pfm, seed=seed, **CONFIG5_SMALL)
            if seed == 0:
                refrun.dump_scene(pbrt, os.path.join(out, "config5_small.b2rs"))
            img, secs = refrun.render(pbrt, pfm)
            np.save(os.path.join(out, "config5_small" + suffix), img.astype(np.float32))
        print("config5_small", img.shape, "mean", img.mean(axis=(0, 1)))
        # the fog's front cap as a bilinear patch: medium transitions decided by a patch's normal
        for seed, suffix in ((0, "_ref.npy"), (1, "_ref2.npy")):
            P.scenes.config5_pbrt(pbrt, out=pfm, seed=seed, patch_cap=True, **CONFIG5_SMALL)
            if seed == 0:
                refrun.dump_scene(pbrt, os.path.join(out, "config5_patchcap_small.b2rs"))
            img, secs = refrun.render(pbrt, pfm)
            np.save(os.path.join(out, "config5_patchcap_small" + suffix), img.astype(np.float32))
        if os.environ.get("B200PT_GOLDEN_PATCHCAP_ONLY"):
            return
        # full-size frame, sampler and integrator tables around the small grid (bench.py swaps the 256^3 grid in)
        P.scenes.config5_pbrt(pbrt, out=pfm, **dict(CONFIG5_FULL, grid=CONFIG5_SMALL["grid"]))
        refrun.dump_scene(pbrt, os.path.join(out, "config5_full.b2rs"))


# The noise-gated cases again where the gate means something (at 16 spp the floor is ~0.35, so 0.44 passed):
# 1024 spp on a small frame, two reference seeds, floor ~0.03-0.05.  A wrong BSDF lobe or a biased roulette shows up as a
# multiple of that.  name -> (generator, arguments)
TIGHT_CASES = {
    "tight_glass": ("cornell", dict(res=32, spp=4096, maxdepth=6, boxes=("dielectric_smooth", "dielectric_glass"))),
    "tight_metal2": ("cornell", dict(res=32, spp=1024, maxdepth=4, boxes=("conductor_reflectance", "dielectric_rough"),
                                     regularize=True)),
    "tight_coated": ("cornell", dict(res=32, spp=1024, maxdepth=5, boxes=("coated_diffuse", "coated_diffuse_rough"))),
    "tight_alpha": ("cornell", dict(res=32, spp=1024, maxdepth=5, alpha=True)),
    # ThinDielectricMaterial: a constant-eta pane and one with a dispersive index (TerminateSecondary)
    "tight_thin": ("cornell", dict(res=32, spp=16384, maxdepth=6, boxes=("thin_dielectric", "thin_dielectric_glass"))),
    # DiffuseTransmissionMaterial: one with the default scale, one whose scale pushes T into the clamp
    "tight_dtrans": ("cornell", dict(res=32, spp=4096, maxdepth=5,
                                     boxes=("diffuse_transmission", "diffuse_transmission_scaled"))),
    # PointLight, SpotLight, DistantLight beside the ceiling emitter: under the BVH light sampler (point and
    # spot in the tree, the distant one in its list of unbounded lights), under the power sampler, and in the fogged
    # room of config 5 (VolPathIntegrator: shadow rays through the fog, the cloud and the front cap)
    "tight_lights": ("cornell", dict(res=32, spp=4096, maxdepth=5, lightsampler="bvh", lights=("point", "spot", "distant"),
                                     boxes=(None, "conductor_rough"))),
    "tight_lights_power": ("cornell", dict(res=32, spp=4096, maxdepth=4, lightsampler="power", lights=("point", "spot", "distant"))),
    "tight_lights_vol": ("config5", dict(CONFIG5_SMALL, res=(32, 24), spp=8192, lights=("point", "spot", "distant"))),
    # IndependentSampler (samplers.h:442-476), a sample count that is no power of two; surfaces with the layered
    # model's own generators, and the volume integrator's
    "tight_independent": ("cornell", dict(res=32, spp=9000, maxdepth=5, sampler="independent",
                                          boxes=("coated_diffuse", "dielectric_rough"))),
    "tight_independent_vol": ("config5", dict(CONFIG5_SMALL, res=(32, 24), spp=9000, sampler="independent")),
    # the filters sampled through a FilterSampler (filters.cpp:92-118): pbrt's default Gaussian (radius 1.5), a Mitchell
    # (negative lobes: negative sample weights), and a Lanczos sinc with unequal radii; 1024 spp on a 24 x 24 frame
    "tight_gaussian": ("cornell", dict(res=24, spp=1024, maxdepth=4, pixelfilter="gaussian")),
    "tight_mitchell": ("cornell", dict(res=24, spp=1024, maxdepth=4, pixelfilter='"mitchell" "float B" 0.2 "float C" 0.4')),
    "tight_triangle": ("cornell", dict(res=24, spp=1024, maxdepth=4, pixelfilter='"triangle" "float xradius" 2 "float yradius" 1.5')),
    "tight_sinc": ("cornell", dict(res=24, spp=1024, maxdepth=4, pixelfilter='"sinc" "float xradius" 3 "float yradius" 2.5')),
    # CoatedConductorMaterial: smooth coat over rough copper + rough coat over a smooth base; and both rough
    # with a dispersive coat (TerminateSecondary)
    "tight_coatcond": ("cornell", dict(res=32, spp=2048, maxdepth=5, boxes=("coated_conductor", "coated_conductor_rough"))),
    "tight_coatcond2": ("cornell", dict(res=32, spp=4096, maxdepth=5, boxes=("coated_conductor_both", "coated_conductor"),
                                        regularize=True)),
    # ImageInfiniteLight: a 32 x 32 equal-area map lights the room through its open front, beside the ceiling
    # emitter; under the BVH and the power light sampler (escaped rays weighted against the light's compensated pdf)
    "tight_envmap": ("cornell", dict(res=32, spp=4096, maxdepth=5, lightsampler="bvh", envmap=32,
                                     boxes=(None, "conductor_smooth"))),
    "tight_envmap_vol": ("config5", dict(CONFIG5_SMALL, res=(32, 24), spp=8192, envmap=32)),
    # emissive media (sigma_a * Le at every tentative collision, cpu/integrators.cpp:990-1001): a glowing fog and a cloud with
    # an LeScale grid of the density's resolution; and with one Lescale number (a 1 x 1 x 1 grid)
    "tight_emissive_vol": ("config5", dict(CONFIG5_SMALL, res=(32, 24), spp=4096, emissive=True)),
    "tight_emissive_vol1": ("config5", dict(CONFIG5_SMALL, res=(32, 24), spp=4096, emissive="half")),
    "tight_blackbody_vol": ("config5", dict(CONFIG5_SMALL, res=(32, 24), spp=4096, emissive="blackbody")),
    "tight_envmap_power": ("cornell", dict(res=32, spp=4096, maxdepth=4, lightsampler="power", envmap=32)),
    "tight_config3": ("config3", dict(CONFIG3_SMALL, res=(32, 18), spp=16384)),
    "tight_config5": ("config5", dict(CONFIG5_SMALL, res=(32, 24), spp=4096)),
}


def make_tight_goldens(P, refrun, out):
    import tempfile
    gen = {"cornell": P.scenes.cornell_pbrt, "config3": P.scenes.config3_pbrt, "config5": P.scenes.config5_pbrt}
    pick = os.environ.get("B200PT_TIGHT_CASES")  # comma-separated names: regenerate those only
    for name, (kind, kw) in TIGHT_CASES.items():
        if pick and name not in pick.split(","):
            continue
        with tempfile.TemporaryDirectory() as tmp:
            pbrt, pfm = os.path.join(tmp, name + ".pbrt"), os.path.join(tmp, name + ".pfm")
            imgs = []
            for seed, suffix in ((0, "_ref.npy"), (1, "_ref2.npy")):
                gen[kind](pbrt, out=pfm, **dict(kw, seed=kw.get("seed", 0) + seed))
                if seed == 0:
                    refrun.dump_scene(pbrt, os.path.join(out, name + ".b2rs"))
                img, secs = refrun.render(pbrt, pfm)
                np.save(os.path.join(out, name + suffix), img.astype(np.float32))
                imgs.append(img.astype(np.float64))
            floor = np.sqrt(np.mean((imgs[1] - imgs[0]) ** 2)) / imgs[0].mean()
            print(name, imgs[0].shape, "mean", imgs[0].mean(), f"noise floor {floor:.4f}", f"{secs:.1f}s")


def make_render_goldens(P, refrun, out):
    import tempfile
    pick = os.environ.get("B200PT_GOLDEN_CASES")  # comma-separated names: regenerate only these image cases
    for name, kw in RENDER_CASES.items():
        if pick and name not in pick.split(","):
            continue
        with tempfile.TemporaryDirectory() as tmp:
            pbrt, pfm = os.path.join(tmp, name + ".pbrt"), os.path.join(tmp, name + ".pfm")
            P.scenes.cornell_pbrt(pbrt, out=pfm, **kw)
            if kw.get("film") in ("spectral", "gbuffer", "gbuffer_albedo"):  # no EXR writer in the oracle build: Film::GetImage read directly
                img, _, secs = refrun.render_channels(pbrt)
            else:
                img, secs = refrun.render(pbrt, pfm)
            refrun.dump_scene(pbrt, os.path.join(out, name + ".b2rs"))
            np.save(os.path.join(out, name + "_ref.npy"), img.astype(np.float32))
            if name in NOISE_GATED:
                P.scenes.cornell_pbrt(pbrt, out=pfm, **dict(kw, seed=kw.get("seed", 0) + 1))
                img2, _ = refrun.render(pbrt, pfm)
                np.save(os.path.join(out, name + "_ref2.npy"), img2.astype(np.float32))
            print(name, img.shape, "mean", img.mean(axis=(0, 1)), f"{secs:.2f}s")
    if pick:
        return
    with tempfile.TemporaryDirectory() as tmp:
        pbrt, pfm = os.path.join(tmp, "c3.pbrt"), os.path.join(tmp, "c3.pfm")
        P.scenes.config3_pbrt(pbrt, out=pfm, **CONFIG3_FULL)
        refrun.dump_scene(pbrt, os.path.join(out, "config3_full.b2rs"), bvh=False)
        print("config3_full tables only")
        for seed, suffix in ((0, "_ref.npy"), (1, "_ref2.npy")):
            P.scenes.config3_pbrt(pbrt, out=pfm, seed=seed, **CONFIG3_SMALL)
            if seed == 0:
                refrun.dump_scene(pbrt, os.path.join(out, "config3_small.b2rs"))
            img, secs = refrun.render(pbrt, pfm)
            np.save(os.path.join(out, "config3_small" + suffix), img.astype(np.float32))
        print("config3_small", img.shape, "mean", img.mean(axis=(0, 1)))
    for name, kw in TABLE_CASES.items():
        with tempfile.TemporaryDirectory() as tmp:
            pbrt = os.path.join(tmp, name + ".pbrt")
            P.scenes.cornell_pbrt(pbrt, out=os.path.join(tmp, name + ".pfm"), **kw)
            refrun.dump_scene(pbrt, os.path.join(out, name + ".b2rs"))
            print(name, "tables only")


if __name__ == "__main__":
    import pbrt_v4_b200 as P
    import refrun
    out = os.path.dirname(os.path.abspath(__file__))
    known = ("--omni-only", "--config4-only", "--config5-only", "--render-only", "--tight-only")
    for a in sys.argv[1:]:
        if a not in known and not a.startswith("--only="):
            sys.exit(f"unknown argument {a}; known: {known}, --only=name,...  (no argument regenerates every fixture)")
    if any(a.startswith("--only=") for a in sys.argv):
        pass
    elif "--config5-only" in sys.argv:
        make_config5_goldens(P, refrun, out)
        sys.exit(0)
    if "--tight-only" in sys.argv:
        make_tight_goldens(P, refrun, out)
        sys.exit(0)
    if "--omni-only" in sys.argv:
        make_omni_goldens(P, refrun, out)
        sys.exit(0)
    if "--config4-only" in sys.argv:
        make_config4_goldens(P, refrun, out)
        sys.exit(0)
    if not any(a.startswith("--only=") for a in sys.argv):
        make_render_goldens(P, refrun, out)
        if not os.environ.get("B200PT_GOLDEN_CASES"):
            make_config4_goldens(P, refrun, out)
            make_omni_goldens(P, refrun, out)
            make_config5_goldens(P, refrun, out)
            make_tight_goldens(P, refrun, out)
    if "--render-only" in sys.argv:
        sys.exit(0)
    only = [a.split("=", 1)[1].split(",") for a in sys.argv if a.startswith("--only=")]
    for name, (sc, rays) in cases(P.scenes).items():
        if only and name not in only[0]:
            continue
        rs = refrun.RefScene(sc)
        hits, _ = rs.closest(rays)
        occ, _ = rs.any(rays)
        sh = P.scenes.shadow_rays(rays, hits, seed=35)
        hits_sh, _ = rs.closest(sh)
        occ_sh, _ = rs.any(sh)
        # rays are re-created from the seeds by the tests; only the answers are stored
        np.savez_compressed(os.path.join(out, name + ".npz"), hits=compact(hits), occ=np.packbits(occ),
                            shadow_hits=compact(hits_sh), shadow_occ=np.packbits(occ_sh))
        print(name, len(rays), "hit", (hits["mesh"] != 0xFFFFFFFF).mean(), "occluded(shadow)", occ_sh.mean())
