"""Where the time of a SMALL render goes (BASELINE.json configs[0] at its own size: cornell-box 512x512, 16 spp, depth 5):
per-bounce device times and host-side issue/sync times printed by the library under MCPT_DEBUG_TIMING."""
import os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
os.environ["MCPT_DEBUG_TIMING"] = "1"
import torch
from monte_carlo_path_tracing_b200 import api, scenes
sc = scenes.cornell_box(width=512, height=512)
with tempfile.TemporaryDirectory() as td:
    a = api.load_scene_files(*[sc.write(td) + e for e in (".obj", ".xml")])[0]
s = api.Scene(a, 0)
film = torch.zeros(512 * 512 * 3, device="cuda")
for k in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    st = s.render_device(sc.camera, api.NEE_UNIFORM, 16, film.data_ptr(), max_depth=5, seed=1, spp_offset=16 * k)
    torch.cuda.synchronize()
    print("render %d: wall %.3f ms, device %.3f ms (bounce %.3f), launches %d" % (k, (time.perf_counter() - t0) * 1e3, st["ms_total"], st["ms_bounce"], st["kernel_launches"]), file=sys.stderr)
