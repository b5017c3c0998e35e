"""GPU vs CPU-check on variants of config5_small (debugging aid)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import conftest
import pbrt_v4_b200 as P
rc = conftest.rendercheck.__wrapped__(P)
base = P.scenes.load_b2rs(os.path.join(ROOT, "tests/golden/config5_small.b2rs"))
def variant(name):
    d = {k: v.copy() for k, v in base.items()}
    mi = d["media_i"]
    if name == "nosigma":
        for m in mi:
            d["spectra_values"][m[1]] = 0; d["spectra_values"][m[2]] = 0
    if name == "noscatter":
        for m in mi:
            d["spectra_values"][m[2]] = 0
    if name == "fogonly":
        d["spectra_values"][mi[0][1]] = 0; d["spectra_values"][mi[0][2]] = 0
    if name == "cloudonly":
        d["spectra_values"][mi[1][1]] = 0; d["spectra_values"][mi[1][2]] = 0
    if name == "depth1":
        d["integrator"][0] = 1
    return P.RenderScene(d)
for name in sys.argv[1:] or ["orig", "nosigma", "noscatter", "fogonly", "cloudonly", "depth1"]:
    rsc = variant(name)
    r = P.B200Renderer(rsc)
    r.render(); img = r.image(); st = r.stats(); r.close()
    cpu = P.render.film_to_image(rc(rsc), rsc.d["film_matrix"])
    d = np.abs(img - cpu).sum(axis=2)
    print(name, "rrmse gpu-vs-cpu", conftest.rrmse(img, cpu), "means", img.mean(), cpu.mean(), "bad px", int((d > 1e-3).sum()), st, flush=True)
    np.save(os.path.join(ROOT, "gpurun_out", f"vol_{name}_gpu.npy"), img); np.save(os.path.join(ROOT, "gpurun_out", f"vol_{name}_cpu.npy"), cpu)
