"""The restated slab / watertight-triangle arithmetic (csrc/isect.h) compiled for the host,
against the compiled reference on identical ray batches: ids, t and barycentrics bit-exact.
This is the GPU-less proof that the device math follows shapes.cpp:172-273 and
util/vecmath.h:1576-1610 operation for operation (the -m gpu tests repeat it on the device)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import hit_mismatches


def run_hostcheck(binpath, P, rs, rays, layout="pair"):
    """layout "wide" = the production Wide8 walk (traverse_wide.cu), "pair" = the exact binary PairNode walk
    (traverse_persist.cu; irregular rays), "node32" = the parity kernel's."""
    sfx = {"pair": "2", "wide": "8", "node32": ""}[layout]
    rp = os.path.join(rs.dir, "rays_hc.f32")
    np.ascontiguousarray(rays, "<f4").tofile(rp)
    subprocess.check_call([binpath, rs.path, rp, "closest" + sfx, rs.dir + "/hc_hits.bin"])
    subprocess.check_call([binpath, rs.path, rp, "any" + sfx, rs.dir + "/hc_occ.bin"])
    return (np.fromfile(rs.dir + "/hc_hits.bin", dtype=P.capi.HIT_DTYPE),
            np.fromfile(rs.dir + "/hc_occ.bin", dtype=np.uint8))


CASES = {
    "soup_primary": lambda S: (S.random_soup(100000, seed=5, n_meshes=3), "primary"),
    "soup_adversarial": lambda S: (S.random_soup(50000, seed=6), "adversarial"),
    "cornell_camera": lambda S: (S.cornell_box(), "camera"),
    "cornell_adversarial": lambda S: (S.cornell_box(), "adversarial"),
    "sheet_adversarial": lambda S: (S.Scene([S.grid_sheet(64, 64, jitter=0.01)]), "adversarial"),
    "flat_sheet_adversarial": lambda S: (S.Scene([S.grid_sheet(32, 32, jitter=0.0)]), "adversarial"),
    # object instances (TransformedPrimitive, cpu/primitive.cpp:112-131): interval ray transform, t not shifted back
    "instanced_primary": lambda S: (S.instanced_scene(), "primary"),
    "instanced_adversarial": lambda S: (S.instanced_scene(seed=78), "adversarial"),
    # bilinear patches (IntersectBilinearPatch, shapes.h:1279-1347) mixed with triangles: (u, v) and t bit-exact
    "patches_primary": lambda S: (S.patch_scene(), "primary"),
    "patches_adversarial": lambda S: (S.patch_scene(8000, 4000, seed=93), "adversarial"),
    "instanced_only": lambda S: (S.instanced_scene(n_instances=12, world_tris=0, seed=79), "adversarial"),
}


@pytest.mark.parametrize("layout", ["wide", "pair", "node32"])
@pytest.mark.parametrize("case", sorted(CASES))
def test_host_arithmetic_bit_exact(P, refrun, hostcheck_bin, case, layout):
    S = P.scenes
    sc, kind = CASES[case](S)
    if kind == "primary":
        rays = S.primary_rays(150000, seed=1)
    elif kind == "camera":
        rays = S.camera_rays(192, 192, (0.5, 0.5, -1.4), (0.5, 0.5, 0.5), spp=2)
    else:
        rays = S.adversarial_rays(S.flattened(sc), 100000)
    rs = refrun.RefScene(sc)
    ref, _ = rs.closest(rays)
    occ_ref, _ = rs.any(rays)
    mine, occ = run_hostcheck(hostcheck_bin, P, rs, rays, layout)
    assert (ref["mesh"] != 0xFFFFFFFF).mean() > 0.3
    if sc.instances:
        assert (ref["instance"] != 0).mean() > 0.2
    if any(m.kind == 1 for m in sc.meshes):
        kinds = np.array([m.kind for m in sc.meshes] + [0])
        assert (kinds[np.minimum(ref["mesh"], len(sc.meshes))] == 1).mean() > 0.2  # patches are hit
    assert hit_mismatches(mine, ref, sc).sum() == 0
    assert (occ != occ_ref).sum() == 0
    # shadow rays from the hit points (unnormalised d, tMax = 1 - 1e-4)
    sh = S.shadow_rays(rays, mine)
    ref2, _ = rs.closest(sh)
    occ_ref2, _ = rs.any(sh)
    mine2, occ2 = run_hostcheck(hostcheck_bin, P, rs, sh, layout)
    assert hit_mismatches(mine2, ref2, sc).sum() == 0
    assert (occ2 != occ_ref2).sum() == 0
