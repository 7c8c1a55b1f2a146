#!/usr/bin/env python
"""Small renders of every estimator (+ jitter, + the parity entry points) for compute-sanitizer:
   compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from monte_carlo_path_tracing_b200 import api, scenes  # noqa: E402

for name, kw in (("cornell-box", dict(width=48, height=36)), ("veach-mis", dict(width=40, height=24, sphere_subdiv=1, plate_tess=(4, 2)))):
    sc = scenes.SCENES[name](**kw)
    with tempfile.TemporaryDirectory() as td:
        stem = sc.write(td)
        a, _ = api.load_scene_files(stem + ".obj", stem + ".xml")
    s = api.Scene(a, 0)
    for strat in (api.NEE_UNIFORM, api.NEE_SPHERICAL, api.BRDF_ONLY, api.MIS):
        for flags in (0, api.JITTER):
            img, st = s.render(sc.camera, strat, 5, max_depth=0 if strat == api.MIS else 6, flags=flags, seed=3, batch_spp=3)
            assert np.isfinite(img).all()
            print(name, api.STRATEGY_NAMES[strat], "flags", flags, "vertices", st["vertices"], "launches", st["kernel_launches"], flush=True)
    rng = np.random.default_rng(1)
    lo = a.positions.reshape(-1, 3).min(0); hi = a.positions.reshape(-1, 3).max(0)
    ro = lo + (hi - lo) * rng.random((5000, 3)); rd = rng.normal(size=(5000, 3))
    for prec in (api.FP64, api.FP32):
        s.closest(ro, rd, None, False, prec)
        s.closest(ro, rd, None, True, prec)
    s.close()
soup = scenes.triangle_soup(20000, 8, seed=2, width=64, height=64)
s = api.Scene(soup, 0)
img, st = s.render(soup.camera, api.NEE_UNIFORM, 4, max_depth=5, seed=1)
print("soup", st["vertices"], flush=True)
s.close()
print("sanitize_small done")
