"""Closest-hit + any-hit launches over the full-size config-3 scene (timing / ncu): python scripts/trav_c3.py [log2 rays] [grid]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import torch
import pbrt_v4_b200 as P
import bench, make_golden
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 22
grid = int(sys.argv[2]) if len(sys.argv) > 2 else make_golden.CONFIG3_FULL_GRID
full = make_golden.CONFIG3_FULL
rsc = P.RenderScene(P.scenes.config3_refine(P.scenes.load_b2rs(os.path.join(ROOT, "tests", "golden", "config3_full.b2rs")),
                                            full["n_objects"], full["n_instances"], grid))
agg = P.B200Aggregate(rsc.geometry, device=0)
n = 1 << lg
prim = bench.config3_rays(P, rsc, n, seed=11)
hits = agg.IntersectClosest(prim)
shad = bench.config3_shadow_rays(P, rsc, prim, hits, seed=1011)
d_prim, d_shad = torch.from_numpy(prim).cuda(), torch.from_numpy(shad).cuda()
d_hits = torch.empty(n * 24, dtype=torch.uint8, device="cuda")
d_occ = torch.empty(n, dtype=torch.uint8, device="cuda")
tc, ta = [], []
for _ in range(int(os.environ.get("REPS", "3"))):
    agg.intersect_closest_ptr(d_prim.data_ptr(), n, d_hits.data_ptr(), P.capi.MEM_DEVICE)
    tc.append(agg.trace_stats()["kernel_ms"])
    agg.intersect_any_ptr(d_shad.data_ptr(), n, d_occ.data_ptr(), P.capi.MEM_DEVICE)
    ta.append(agg.trace_stats()["kernel_ms"])
agg.sync()
ok = np.array_equal(d_hits.cpu().numpy().view(P.capi.HIT_DTYPE).view("u1"), hits.view("u1"))
print(f"config3 grid={grid} mode={os.environ.get('B200PT_TRAVERSAL', 'wide')} nodeMin={os.environ.get('B200PT_NODE_MIN', '-')} "
      f"closest {n / min(tc) * 1e-3:.0f} any {n / min(ta) * 1e-3:.0f} Mrays/s same={ok}")
if os.environ.get("COUNTERS"):
    agg.IntersectClosest(prim, counters=True)
    print(agg.trace_stats(), agg.stats())
