"""diag: films with and without the per-primitive light lists on the small veach-mis test scene, by depth cap"""
import os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import util
from monte_carlo_path_tracing_b200 import api
sc = util.scene_text("veach-mis")
with tempfile.TemporaryDirectory() as td:
    stem = sc.write(td)
    arrays, _ = api.load_scene_files(stem + ".obj", stem + ".xml")
cam = sc.camera
A = api.Scene(arrays, 0)
os.environ["MCPT_LIGHT_LISTS"] = "0"
B = api.Scene(arrays, 0)
print(A.light_lists(), B.light_lists())
for strat, name in ((api.NEE_SPHERICAL, "NEE_SPH"), (api.MIS, "MIS")):
    for md in (1, 2, 5):
        for spp in (32, 256):
            a, sa = A.render(cam, strat, spp, max_depth=md, seed=3)
            b, sb = B.render(cam, strat, spp, max_depth=md, seed=3)
            b2, _ = B.render(cam, strat, spp, max_depth=md, seed=4)
            d = np.abs(a - b).sum(2)
            print(name, "depth", md, "spp", spp, "rmse(a,b) %.3g rmse(b,b2) %.3g mean a %.5g b %.5g" % (util.rmse(a, b), util.rmse(b, b2), a.mean(), b.mean()),
                  "worst pixel", np.unravel_index(d.argmax(), d.shape), float(d.max()), "nan", int(np.isnan(a).sum()),
                  "sweeps", sa["light_sweeps"], sb["light_sweeps"], "evals", sa["light_evals"], sb["light_evals"])
