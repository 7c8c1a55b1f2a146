"""diag: are the rare bright samples of NEE_SPHERICAL (as written) equally frequent with and without the light lists?"""
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
ref, _ = B.render(cam, api.NEE_SPHERICAL, 16384, max_depth=5, seed=999)
refA, _ = A.render(cam, api.NEE_SPHERICAL, 16384, max_depth=5, seed=998)
print("converged: rmse(A,B) %.3g  mean A %.6g B %.6g" % (util.rmse(refA, ref), refA.mean(), ref.mean()))
for name, S in (("lists", A), ("table", B)):
    worst = []
    for seed in range(1, 1201):
        img, _ = S.render(cam, api.NEE_SPHERICAL, 32, max_depth=5, seed=seed)
        worst.append(float(np.abs(img - ref).sum(2).max()))
    w = np.array(worst)
    print(name, "worst-pixel deviation over 1200 seeds: median %.4f  max %.4f  n>0.05: %d" % (np.median(w), w.max(), int((w > 0.05).sum())), np.round(np.sort(w)[-6:], 4))
