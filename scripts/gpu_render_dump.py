"""Renders every golden render case on cuda:0 and saves the images to gpurun_out/ (debugging aid)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pbrt_v4_b200 as P  # noqa: E402

gold = os.path.join(ROOT, "tests", "golden")
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
for f in sorted(os.listdir(gold)):
    if not f.endswith("_ref.npy"):
        continue
    name = f[:-8]
    rsc = P.RenderScene.load(os.path.join(gold, name + ".b2rs"))
    ref = np.load(os.path.join(gold, f))
    r = P.B200Renderer(rsc)
    r.render()
    img = r.image()
    st = r.stats()
    r.close()
    np.save(os.path.join(ROOT, "gpurun_out", "gpu_" + name + ".npy"), img)
    e = float(np.sqrt(np.mean((img - ref) ** 2)) / ref.mean())
    print(name, "rRMSE", e, "mean", img.mean(), ref.mean(), st)
