"""units behind profiles/r2_flops.json: exact work counts of the launches captured by `ncu --set full` in tools/gpu_r2.sh
(veach-mis, 32 spp per step, seed 1, sample offset 0): sweeps and queue elements of the first vertex and of bounce 1."""
import sys, os, json, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np
from monte_carlo_path_tracing_b200 import api, scenes
sc = scenes.veach_mis(width=1280, height=720)
td = tempfile.mkdtemp(); stem = sc.write(td)
a, _ = api.load_scene_files(stem + ".obj", stem + ".xml")
s = api.Scene(a, 0)
out = {}
for md in (1, 3):
    _, st = s.render(sc.camera, api.MIS, 32, max_depth=md, seed=1)
    out["max_depth_%d" % md] = {k: int(st[k]) for k in ("light_sweeps", "elements", "vertices", "rays_closest", "rays_light", "bounces")}
L = len(a.light_prim)
out["L"] = L; out["Lpad"] = (L + 63) // 64 * 64
out["first_vertex_sweeps"] = out["max_depth_1"]["light_sweeps"]
out["bounce1_sweeps"] = out["max_depth_3"]["light_sweeps"] - out["max_depth_1"]["light_sweeps"]
out["first_elements"] = out["max_depth_1"]["elements"]
print(json.dumps(out))
