"""Host SAH builder (b200pt_bvh2_build) against the tree the compiled reference builds
(BVHAggregate::buildRecursive / flattenBVH, cpu/aggregates.cpp:192-387, 505-522): node for node."""
import numpy as np
import pytest


def assert_same_tree(mine, order, ref_nodes, ref_prims):
    assert len(mine) == len(ref_nodes)
    leaf = ref_nodes["nprims"] > 0
    for f in ("bmin", "bmax", "nprims"):
        assert np.array_equal(np.ascontiguousarray(mine[f]).view("u1"), np.ascontiguousarray(ref_nodes[f]).view("u1")), f
    assert np.array_equal(mine["axis"][~leaf], ref_nodes["axis"][~leaf])
    assert np.array_equal(mine["offset"][~leaf], ref_nodes["offset"][~leaf])
    # leaf contents, in leaf order (the reference's slot offsets depend on thread timing above 128K prims)
    ref_tri = ref_prims[:, 2]
    for i in np.nonzero(leaf)[0][:: max(1, leaf.sum() // 20000)]:
        a, n = mine["offset"][i], mine["nprims"][i]
        b = ref_nodes["offset"][i]
        assert np.array_equal(order[a:a + n], ref_tri[b:b + n])


@pytest.mark.parametrize("n,seed,s", [(1, 1, 0.01), (2, 2, 0.01), (3, 3, 0.3), (17, 4, 0.2), (1000, 5, 0.05),
                                      (20000, 6, 0.01), (150000, 7, 0.01)])
def test_soup_tree_matches_reference(P, refrun, n, seed, s):
    sc = P.scenes.random_soup(n, seed=seed, s=s)
    obj, ref_nodes, ref_prims = refrun.RefScene(sc).dump_bvh()[0]
    mine, order = P.capi.bvh2_build(sc.triangle_boxes(), 4)
    assert_same_tree(mine, order, ref_nodes, ref_prims)


def test_cornell_and_sheet_trees_match_reference(P, refrun):
    for sc in (P.scenes.cornell_box(), P.scenes.Scene([P.scenes.grid_sheet(48, 48, jitter=0.0)]),
               P.scenes.Scene([P.scenes.grid_sheet(32, 32, jitter=0.02), P.scenes.grid_sheet(16, 16, z=0.7)])):
        obj, ref_nodes, ref_prims = refrun.RefScene(sc).dump_bvh()[0]
        boxes = sc.triangle_boxes()
        mine, order = P.capi.bvh2_build(boxes, 4)
        assert len(mine) == len(ref_nodes)
        leaf = ref_nodes["nprims"] > 0
        assert np.array_equal(mine["bmin"], ref_nodes["bmin"]) and np.array_equal(mine["bmax"], ref_nodes["bmax"])
        assert np.array_equal(mine["nprims"], ref_nodes["nprims"])
        assert np.array_equal(mine["axis"][~leaf], ref_nodes["axis"][~leaf])
        # global triangle number = position in reference primitive order
        first = np.cumsum([0] + [m.indices.shape[0] for m in sc.meshes])
        ref_global = first[ref_prims[:, 1]] + ref_prims[:, 2]
        assert np.array_equal(order, ref_global)


def test_degenerate_inputs(P):
    # identical boxes -> centroid bounds degenerate -> one leaf with all prims (cpu/aggregates.cpp:224-232)
    boxes = np.tile(np.array([[0, 0, 0, 1, 1, 1]], np.float32), (9, 1))
    nodes, order = P.capi.bvh2_build(boxes, 4)
    assert len(nodes) == 1 and nodes["nprims"][0] == 9 and sorted(order) == list(range(9))
    # zero-area bounds -> leaf (cpu/aggregates.cpp:209)
    boxes = np.array([[0, 0, 0, 0, 0, 0], [0, 0, 0, 0, 0, 0]], np.float32)
    nodes, order = P.capi.bvh2_build(boxes, 4)
    assert len(nodes) == 1 and nodes["nprims"][0] == 2
    # maxPrimsInNode = 1 (object-level trees, scene.cpp:1542) always splits separable prims
    rng = np.random.default_rng(0)
    c = rng.random((64, 3)).astype(np.float32)
    boxes = np.concatenate([c, c + np.float32(0.01)], axis=1)
    nodes, order = P.capi.bvh2_build(boxes, 1)
    assert (nodes["nprims"][nodes["nprims"] > 0] == 1).all() and len(nodes) == 127


def test_parallel_build_is_the_serial_build(P, monkeypatch):
    """The multi-threaded build (serial top, subtrees on worker threads, one depth-first layout) gives the
    single-threaded build's tree byte for byte, whatever the thread count."""
    import time
    sc = P.scenes.random_soup(400000, seed=21, s=0.01)
    boxes = sc.triangle_boxes()
    monkeypatch.setenv("B200PT_BUILD_THREADS", "1")
    t0 = time.perf_counter()
    n1, o1 = P.capi.bvh2_build(boxes, 4)
    t1 = time.perf_counter() - t0
    for threads in ("3", "8"):
        monkeypatch.setenv("B200PT_BUILD_THREADS", threads)
        t0 = time.perf_counter()
        n2, o2 = P.capi.bvh2_build(boxes, 4)
        tp = time.perf_counter() - t0
        assert n1.tobytes() == n2.tobytes() and np.array_equal(o1, o2)
        print(f"build 400K: 1 thread {t1:.3f} s, {threads} threads {tp:.3f} s")
