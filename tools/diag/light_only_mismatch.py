"""diagnostic: where does the FP32 light-only trace (trace_fast over lnodes) disagree with the oracle's exhaustive search?"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, tempfile
import util, oracle
from monte_carlo_path_tracing_b200 import api
for name in ("veach-mis", "cornell-box"):
    sc = util.scene_text(name)
    td = tempfile.mkdtemp(); stem = sc.write(td)
    a, cam = api.load_scene_files(stem + ".obj", stem + ".xml")
    s = api.Scene(a, 0); o = oracle.OracleLib(a)
    rng = np.random.default_rng(311)
    ro, rd, ex = util.surface_rays(a, 20000, 310)
    lp = a.light_prim[rng.integers(0, len(a.light_prim), len(ro))]
    b = rng.random(len(ro)); g = rng.random(len(ro)) * (1 - b)
    v = a.positions[lp].astype(np.float64)
    tgt = v[:, 0] * (1 - b - g)[:, None] + v[:, 1] * b[:, None] + v[:, 2] * g[:, None]
    aim = rng.random(len(ro)) < 0.67
    rd[aim] = util.unit(tgt[aim] - ro[aim])
    p32, b32 = s.closest(ro, rd, ex, True, api.FP32)
    p64, b64 = o.closest(ro, rd, ex, 3)
    # the same rays rounded to FP32 first (what the FP32 kernel actually traces)
    ro32, rd32 = ro.astype(np.float32).astype(np.float64), rd.astype(np.float32).astype(np.float64)
    p64r, b64r = o.closest(ro32, rd32, ex, 3)
    mm = p32 != p64
    print(name, "mismatch vs fp64 ray %.4f, vs fp32-rounded ray %.4f" % (mm.mean(), (p32 != p64r).mean()))
    print("  gpu miss / oracle hit:", int(((p32 < 0) & (p64 >= 0)).sum()), " gpu hit / oracle miss:", int(((p32 >= 0) & (p64 < 0)).sum()),
          " both hit, different:", int(((p32 >= 0) & (p64 >= 0) & mm).sum()))
    islight = np.zeros(len(a.material_id), bool); islight[a.light_prim] = True
    print("  mismatches with origin ON a light:", int((mm & islight[ex]).sum()), "of", int(mm.sum()), "; rays with origin on a light:", int(islight[ex].sum()))
    idx = np.nonzero(mm)[0][:8]
    for i in idx:
        print("   ray", i, "ex", ex[i], "light-origin" if islight[ex[i]] else "", "p32", p32[i], b32[i], "p64", p64[i], b64[i])
    s.close()
