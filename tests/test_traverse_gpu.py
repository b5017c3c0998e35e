"""Parity tests proper: the CUDA path, called through the C ABI (B200Aggregate -> libb200pt.so),
against (1) the committed golden answers of the reference, (2) the compiled reference re-run on
the same seeded inputs, (3) size-independent properties at BASELINE.json's full config-2 size.
Bar: ids, t and barycentrics bit-exact (0 mismatches)."""
import importlib.util
import os

import numpy as np
import pytest

from conftest import ROOT, hit_mismatches

pytestmark = pytest.mark.gpu

_spec = importlib.util.spec_from_file_location("make_golden", os.path.join(ROOT, "tests", "golden", "make_golden.py"))
G = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(G)
NAMES = ["soup20k_primary", "soup20k_adversarial", "cornell_camera", "cornell_adversarial", "sheets_adversarial",
         "instanced_primary", "instanced_adversarial", "patches_primary", "patches_adversarial",
         "alpha_primary", "alpha_adversarial"]


@pytest.mark.parametrize("name", NAMES)
def test_gpu_matches_golden(P, name):
    sc, rays = G.cases(P.scenes)[name]
    gold = G.load(name)
    agg = P.B200Aggregate(sc)
    hits = agg.IntersectClosest(rays)
    assert hit_mismatches(hits, gold["hits"], sc).sum() == 0
    assert np.array_equal(agg.IntersectShadow(rays), gold["occ"])
    sh = P.scenes.shadow_rays(rays, gold["hits"], seed=35)
    assert hit_mismatches(agg.IntersectClosest(sh), gold["shadow_hits"], sc).sum() == 0
    assert np.array_equal(agg.IntersectShadow(sh), gold["shadow_occ"])
    if not sc.instances:
        assert np.array_equal(agg.Bounds().reshape(-1),
                              np.concatenate([sc.triangle_boxes()[:, :3].min(0), sc.triangle_boxes()[:, 3:].max(0)]))
    else:
        assert (hits["instance"] != 0).mean() > 0.1
    agg.close()


def test_gpu_matches_reference_200k_soup(P, refrun):
    sc = P.scenes.random_soup(200000, seed=5, n_meshes=3)
    rs = refrun.RefScene(sc)
    agg = P.B200Aggregate(sc)
    for rays in (P.scenes.primary_rays(1 << 20, seed=1), P.scenes.adversarial_rays(sc, 1 << 18, seed=2)):
        ref, _ = rs.closest(rays)
        mine = agg.IntersectClosest(rays)
        assert hit_mismatches(mine, ref).sum() == 0
        sh = P.scenes.shadow_rays(rays, mine)
        occ_ref, _ = rs.any(sh)
        assert (agg.IntersectShadow(sh) != occ_ref).sum() == 0
    agg.close()


def test_gpu_instances_match_reference(P, refrun):
    """Object instances: 8 objects x up to 20 K triangles placed 200 times over a world soup,
    own trees and the trees read out of the reference, primary + adversarial + shadow batches."""
    sc = P.scenes.instanced_scene(n_objects=8, tris_per_object=20000, n_instances=200, world_tris=50000, seed=91)
    rs = refrun.RefScene(sc)
    own = P.B200Aggregate(sc)
    given = P.B200Aggregate(sc, bvhs=rs.dump_bvh())
    assert own.stats()["n_instances"] == 200
    for rays in (P.scenes.primary_rays(1 << 19, seed=3), P.scenes.adversarial_rays(P.scenes.flattened(sc), 1 << 17, seed=4)):
        ref, _ = rs.closest(rays)
        mine = own.IntersectClosest(rays)
        assert (ref["instance"] != 0).mean() > 0.2
        assert hit_mismatches(mine, ref).sum() == 0
        assert np.array_equal(mine.view("u1"), given.IntersectClosest(rays).view("u1"))
        occ_ref, _ = rs.any(rays)
        assert (own.IntersectShadow(rays) != occ_ref).sum() == 0
        sh = P.scenes.shadow_rays(rays, mine)
        ref_sh, _ = rs.closest(sh)
        occ_sh, _ = rs.any(sh)
        assert hit_mismatches(own.IntersectClosest(sh), ref_sh).sum() == 0
        assert (own.IntersectShadow(sh) != occ_sh).sum() == 0
    own.close()
    given.close()


def test_gpu_bilinear_patches_match_reference(P, refrun):
    """Bilinear patches: 60 K twisted / planar / axis-aligned patches mixed with 30 K triangles; the
    library's own tree and the tree read out of the reference (prims of kind 2); (u, v) and t bit-exact."""
    sc = P.scenes.patch_scene(60000, 30000, seed=97)
    rs = refrun.RefScene(sc)
    own = P.B200Aggregate(sc)
    given = P.B200Aggregate(sc, bvhs=rs.dump_bvh())
    kinds = np.array([m.kind for m in sc.meshes] + [0])
    for rays in (P.scenes.primary_rays(1 << 19, seed=5), P.scenes.adversarial_rays(sc, 1 << 17, seed=6)):
        ref, _ = rs.closest(rays)
        mine = own.IntersectClosest(rays)
        assert (kinds[np.minimum(ref["mesh"], len(sc.meshes))] == 1).mean() > 0.2
        assert hit_mismatches(mine, ref, sc).sum() == 0
        assert np.array_equal(mine.view("u1"), given.IntersectClosest(rays).view("u1"))
        occ_ref, _ = rs.any(rays)
        assert (own.IntersectShadow(rays) != occ_ref).sum() == 0
        sh = P.scenes.shadow_rays(rays, mine)
        ref_sh, _ = rs.closest(sh)
        occ_sh, _ = rs.any(sh)
        assert hit_mismatches(own.IntersectClosest(sh), ref_sh, sc).sum() == 0
        assert (own.IntersectShadow(sh) != occ_sh).sum() == 0
    own.close()
    given.close()


def test_gpu_with_host_supplied_bvh(P, refrun):
    """Drop-in path: the tree is read out of the reference's BVHAggregate and handed over."""
    sc = P.scenes.random_soup(30000, seed=8)
    rs = refrun.RefScene(sc)
    trees = rs.dump_bvh()
    agg = P.B200Aggregate(sc, bvhs=trees)
    own = P.B200Aggregate(sc)
    rays = P.scenes.primary_rays(100000, seed=4)
    ref, _ = rs.closest(rays)
    a, b = agg.IntersectClosest(rays), own.IntersectClosest(rays)
    assert hit_mismatches(a, ref).sum() == 0
    assert np.array_equal(a.view("u1"), b.view("u1"))
    agg.close()
    own.close()


def test_edge_cases(P):
    sc = P.scenes.cornell_box()
    agg = P.B200Aggregate(sc)
    # empty batch
    assert agg.IntersectClosest(np.zeros((0, 7), np.float32)).shape == (0,)
    assert agg.IntersectShadow(np.zeros((0, 7), np.float32)).shape == (0,)
    # single ray, ragged sizes around the block size
    for n in (1, 31, 127, 129, 1000):
        rays = P.scenes.camera_rays(40, 25, (0.5, 0.5, -1.4), (0.5, 0.5, 0.5))[:n]
        full = agg.IntersectClosest(P.scenes.camera_rays(40, 25, (0.5, 0.5, -1.4), (0.5, 0.5, 0.5)))
        assert np.array_equal(agg.IntersectClosest(rays).view("u1"), full[:n].view("u1"))
    # tMax = 0, negative tMax, NaN direction, zero direction: all misses, no crash
    rays = np.zeros((4, 7), np.float32)
    rays[:, 0:3] = (0.5, 0.5, -1)
    rays[:, 3:6] = (0, 0, 1)
    rays[0, 6], rays[1, 6], rays[2, 6], rays[3, 6] = 0, -1, np.inf, np.inf
    rays[2, 3:6] = np.nan
    rays[3, 3:6] = 0
    h = agg.IntersectClosest(rays)
    assert (h["mesh"] == 0xFFFFFFFF).all()
    assert (agg.IntersectShadow(rays) == 0).all()
    agg.close()


def test_invalid_scene_is_an_error_not_a_crash(P):
    m = P.scenes.Mesh(np.zeros((3, 3), np.float32), np.array([[0, 1, 7]], np.int32))
    with pytest.raises(P.capi.B200PTError) as e:
        P.B200Aggregate(P.scenes.Scene([m]))
    assert e.value.code == P.capi.ERR_INVALID
    with pytest.raises(P.capi.B200PTError):
        P.B200Aggregate(P.scenes.Scene([]))


def test_full_size_properties_1m_triangles(P):
    """BASELINE config 2 at full size (1M triangles, 2^22 rays here): properties that need no oracle.
    (a) any-hit(o,d,tMax) == closest-hit exists; (b) re-tracing with tMax just beyond t_hit returns the
    same record; (c) with tMax just
    below t_hit the first hit is never returned; (d) counters are consistent."""
    sc = P.scenes.random_soup(1_000_000, seed=12111)
    agg = P.B200Aggregate(sc)
    st = agg.stats()
    assert st["n_triangles"] == 1_000_000
    rays = P.scenes.primary_rays(1 << 22, seed=1)
    hits = agg.IntersectClosest(rays, counters=True)
    ts = agg.trace_stats()
    assert ts["rays"] == 1 << 22 and ts["nodes_visited"] > ts["leaves_visited"] > 0 and ts["prims_tested"] > 0
    hit = hits["mesh"] != 0xFFFFFFFF
    assert hit.mean() > 0.99
    assert np.array_equal(agg.IntersectShadow(rays) != 0, hit)
    # re-tracing with tMax a hair beyond t_hit finds the same hit (t itself is not guaranteed to be
    # accepted again: the range test is on tScaled vs tMax*det, shapes.cpp:239-242, and t*det rounds)
    again = rays.copy()
    again[:, 6] = np.where(hit, hits["t"] * np.float32(1.00001), np.inf)
    h2 = agg.IntersectClosest(again)
    assert np.array_equal(h2.view("u1"), hits.view("u1"))
    below = rays.copy()
    below[:, 6] = np.nextafter(hits["t"], np.float32(0))
    h3 = agg.IntersectClosest(below[hit])
    assert ((h3["mesh"] == 0xFFFFFFFF) | (h3["t"] <= hits["t"][hit])).all()
    assert not (agg.IntersectShadow(below[hit]) & ~(h3["mesh"] != 0xFFFFFFFF)).any()
    agg.close()


def test_config2_full_size_bit_exact_vs_reference(P, refrun):
    """BASELINE config 2 at its stated size against the compiled reference on the box's host cores: 1 M triangles,
    2^24 primary + 2^24 shadow + 2^20 adversarial rays, closest-hit records and occlusion bits compared
    with BVHAggregate::Intersect / IntersectP (cpu/aggregates.cpp:529-624) - 0 mismatches in (mesh, prim, instance,
    t bits, b1, b2).  The primary batch goes through in 2^22-ray slices to bound the files handed to raycast_ref."""
    sc = P.scenes.random_soup(1_000_000, seed=12111)
    rs = refrun.RefScene(sc)
    agg = P.B200Aggregate(sc)
    total = 0
    step = 1 << 22
    for k in range(4):
        rays = P.scenes.primary_rays(step, seed=100 + k)
        ref, _ = rs.closest(rays)
        mine = agg.IntersectClosest(rays)
        assert hit_mismatches(mine, ref).sum() == 0
        sh = P.scenes.shadow_rays(rays, mine, seed=200 + k)
        occ_ref, _ = rs.any(sh)
        assert (agg.IntersectShadow(sh) != occ_ref).sum() == 0
        total += 2 * step
    adv = P.scenes.adversarial_rays(sc, 1 << 20, seed=300)
    ref, _ = rs.closest(adv)
    mine = agg.IntersectClosest(adv, counters=True)
    ts = agg.trace_stats()
    assert ts["wide_nodes_visited"] == 0 or ts["exact_rays"] > 0  # wide walk: the axis-aligned rays took the exact binary walk
    assert hit_mismatches(mine, ref).sum() == 0
    occ_ref, _ = rs.any(adv)
    assert (agg.IntersectShadow(adv) != occ_ref).sum() == 0
    total += 2 * len(adv)
    assert total >= (1 << 25) + (1 << 21)
    agg.close()


@pytest.mark.parametrize("walk", ["wide", "binary"])
def test_both_walks_match_reference(P, refrun, monkeypatch, walk):
    """The two production walks - the exact binary pair-node walk (default) and the 8-wide walk with its conservative
    child test + exact leaf-box test (B200PT_TRAVERSAL=wide) - each against the goldens, and on an instanced scene and a
    200 K soup against the compiled reference, adversarial batches included; the wide walk must actually have run
    (wide nodes counted) and must have handed its irregular rays to the binary walk."""
    monkeypatch.setenv("B200PT_TRAVERSAL", walk)
    for name in NAMES:
        sc, rays = G.cases(P.scenes)[name]
        gold = G.load(name)
        agg = P.B200Aggregate(sc)
        hits = agg.IntersectClosest(rays, counters=True)
        ts = agg.trace_stats()
        assert (ts["wide_nodes_visited"] > 0) == (walk == "wide")
        if walk == "wide" and "adversarial" in name:
            assert ts["exact_rays"] > 0
        assert hit_mismatches(hits, gold["hits"], sc).sum() == 0
        assert np.array_equal(agg.IntersectShadow(rays), gold["occ"])
        agg.close()
    sc = P.scenes.instanced_scene(n_objects=6, tris_per_object=15000, n_instances=150, world_tris=40000, seed=191)
    rs = refrun.RefScene(sc)
    agg = P.B200Aggregate(sc)
    for rays in (P.scenes.primary_rays(1 << 18, seed=13), P.scenes.adversarial_rays(P.scenes.flattened(sc), 1 << 16, seed=14)):
        ref, _ = rs.closest(rays)
        mine = agg.IntersectClosest(rays)
        assert hit_mismatches(mine, ref).sum() == 0
        sh = P.scenes.shadow_rays(rays, mine)
        occ, _ = rs.any(sh)
        assert (agg.IntersectShadow(sh) != occ).sum() == 0
    agg.close()
