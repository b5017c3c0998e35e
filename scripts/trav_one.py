"""Closest-hit + any-hit launches over the config-2 scene (timing / ncu): python scripts/trav_one.py [log2 rays] [tris]
Prints Mrays/s of the device-resident launches (CUDA events of the library, best of REPS)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pbrt_v4_b200 as P
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 22
tris = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
sc = P.scenes.random_soup(tris, seed=12111)
agg = P.B200Aggregate(sc, device=0)
n = 1 << lg
prim = P.scenes.primary_rays(n, seed=1)
hits = agg.IntersectClosest(prim)
shad = P.scenes.shadow_rays(prim, hits, seed=2)
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
print(f"mode={os.environ.get('B200PT_TRAVERSAL', 'wide')} nodeMin={os.environ.get('B200PT_NODE_MIN', '-')} "
      f"refillMin={os.environ.get('B200PT_REFILL_MIN', '-')} closest {n / min(tc) * 1e-3:.0f} any {n / min(ta) * 1e-3:.0f} Mrays/s same={ok}")
if os.environ.get("COUNTERS"):
    agg.IntersectClosest(prim, counters=True)
    print(agg.trace_stats(), agg.stats())
