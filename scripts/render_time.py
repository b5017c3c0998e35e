"""usage: render_time.py [case] [n_samples]
Times the config-1 render (Cornell box 512x512, 64 spp, maxdepth 5) on cuda:0; the scene tables
are tests/golden/cornell_r512_s64.b2rs (dumped from the reference parser, make_golden.py)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pbrt_v4_b200 as P  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cornell_r512_s64"
rsc = P.RenderScene.load(os.path.join(ROOT, "tests", "golden", name + ".b2rs"))
if name == "config5_full":  # the frame tables hold a small grid: swap the full-size one in (as bench.py does)
    g = int(os.environ.get("B200PT_C5_GRID", "256"))
    rsc = P.RenderScene(P.scenes.with_density_grid(rsc.d, P.scenes.procedural_density(g)))
res, spp = rsc.resolution[0], rsc.spp
nsamp = int(sys.argv[2]) if len(sys.argv) > 2 else spp
r = P.B200Renderer(rsc)
for it in range(3):
    r.clear()
    t = time.time()
    r.render(0, nsamp, 1)
    r.sync()
    dt = time.time() - t
    st = r.stats()
    print(f"iter {it}: wall {dt*1e3:.1f} ms gpu {st['gpu_ms']:.1f} ms  {st['samples']/st['gpu_ms']/1e3:.1f} Msamples/s "
          f"closest {st['closest_rays']} shadow {st['shadow_rays']} launches {st['kernel_launches']} "
          f"-> {(st['closest_rays']+st['shadow_rays'])/st['gpu_ms']/1e3:.1f} Mrays/s")
img = r.image()
print("mean", img.mean(axis=(0, 1)))
if os.environ.get("B200PT_SAVE_IMAGE"):
    np.save(os.path.join(ROOT, "gpurun_out", f"{name}_{nsamp}.npy"), img)
r.close()
