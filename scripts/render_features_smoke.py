"""A few samples per pixel of the goldens that exercise the round-2 render features (finite images, no CUDA error):
   python scripts/render_features_smoke.py [names...]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import pbrt_v4_b200 as P
GOLD = os.path.join(ROOT, "tests", "golden")
names = sys.argv[1:] or ["tight_envmap", "tight_envmap_vol", "tight_coatcond2", "tight_lights", "tight_lights_vol", "tight_dtrans",
                         "tight_thin", "tight_independent", "tight_independent_vol", "tight_mitchell", "cornell_gbuffer_r48_s16",
                         "tight_config3", "config5_small"]
for name in names:
    rsc = P.RenderScene.load(os.path.join(GOLD, name + ".b2rs"))
    r = P.B200Renderer(rsc, device=0)
    r.render(0, min(rsc.spp, 6))
    img = r.image()
    r.close()
    print(name, img.shape, float(img.mean()), bool(np.isfinite(img).all()), flush=True)
print("render_features_smoke done")
