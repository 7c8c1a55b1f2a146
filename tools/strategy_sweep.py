#!/usr/bin/env python
"""BASELINE.json configs[4]: light-sampling strategy sweep on the veach-mis scene.

Strategies: uniform-area NEE, spherical-triangle NEE, BRDF-only, MIS (the reference's path tree) and single-path MIS,
plus the three spherical ones with MCPT_LIGHT_TREE (light hierarchy instead of the O(L) selection loop).  For each: RMSE / relMSE of linear
radiance against a converged reference at equal spp (1..1024) and at equal time (0.1 / 1 / 10 s of device time).
All four run with the F8/F9 fixes (mixture pdf, solid-angle NEE estimator) and no depth cap, so they estimate
the same integral; the reference image is MIS at --ref-spp (default 16384) with a different seed.  The
reference's own experiment (README.md:346-392, 10 spp, wall-clock only) found spherical/MIS sampling less noisy
but far slower per sample; this puts numbers on both axes.

  python tools/strategy_sweep.py --out profiles/r1_strategy_sweep.json        (needs a B200; run under gpurun)
"""
import argparse
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from monte_carlo_path_tracing_b200 import api, scenes  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "strategy_sweep.json"))
    ap.add_argument("--ref-spp", type=int, default=16384)
    ap.add_argument("--width", type=int, default=1280)
    ap.add_argument("--height", type=int, default=720)
    ap.add_argument("--max-spp", type=int, default=1024)
    args = ap.parse_args()
    sc = scenes.veach_mis(width=args.width, height=args.height)
    with tempfile.TemporaryDirectory() as td:
        stem = sc.write(td)
        arrays, _ = api.load_scene_files(stem + ".obj", stem + ".xml")
    cam = sc.camera
    scene = api.Scene(arrays, 0)
    fixes = api.FIX_MIXTURE_PDF | api.FIX_NEE_MEASURE
    T = api.LIGHT_TREE  # "+tree": O(log L) light selection after the first vertex (csrc/light_tree.h) instead of the O(L) loop
    strategies = [("NEE_UNIFORM", (api.NEE_UNIFORM, 0)), ("NEE_SPHERICAL", (api.NEE_SPHERICAL, 0)), ("BRDF_ONLY", (api.BRDF_ONLY, 0)),
                  ("MIS", (api.MIS, 0)), ("MIS_SINGLE", (api.MIS_SINGLE, 0)), ("NEE_SPHERICAL+tree", (api.NEE_SPHERICAL, T)),
                  ("MIS+tree", (api.MIS, T)), ("MIS_SINGLE+tree", (api.MIS_SINGLE, T))]

    def render(strat, spp, seed):
        strat, extra = strat
        # chunks of <= 1024 spp keep int32 sample indices and the film's FP32 sums comfortable
        acc = np.zeros((cam.height, cam.width, 3), np.float64)
        ms = 0.0
        done = 0
        while done < spp:
            n = min(1024, spp - done)
            img, st = scene.render(cam, strat, n, max_depth=0, flags=fixes | extra, seed=seed, spp_offset=done, spp_total=spp)
            acc += img
            ms += st["ms_total"]
            done += n
        return acc, ms

    ref, ref_ms = render((api.MIS, 0), args.ref_spp, seed=1234567)
    scale = float(np.mean(ref))

    def errors(img):
        d = img - ref
        return {"rmse": float(np.sqrt(np.mean(d * d))), "relmse": float(np.mean(d * d / (ref * ref + (0.01 * scale) ** 2))),
                "mean_ratio": float(img.mean() / ref.mean())}

    out = {"scene": "veach-mis", "width": cam.width, "height": cam.height, "lights": int(len(arrays.light_prim)),
           "flags": "FIX_MIXTURE_PDF|FIX_NEE_MEASURE, max_depth=0 (RR 0.6)", "reference": {"strategy": "MIS", "spp": args.ref_spp,
           "device_ms": ref_ms, "mean_radiance": scale}, "equal_spp": {}, "equal_time": {}}
    ms_per_spp = {}
    for name, strat in strategies:
        rows = []
        spp = 1
        while spp <= args.max_spp:
            img, ms = render(strat, spp, seed=42)
            e = errors(img)
            e.update({"spp": spp, "device_ms": ms, "Msamples_per_s": cam.width * cam.height * spp / ms / 1e3})
            rows.append(e)
            spp *= 4
        out["equal_spp"][name] = rows
        ms_per_spp[name] = rows[-1]["device_ms"] / rows[-1]["spp"]
        print(name, " ".join("%d:%.4g" % (r["spp"], r["rmse"]) for r in rows), "ms/spp %.3f" % ms_per_spp[name], flush=True)
    for budget in (0.1, 1.0, 10.0):
        row = {}
        for name, strat in strategies:
            spp = max(1, int(round(budget * 1e3 / ms_per_spp[name])))
            img, ms = render(strat, spp, seed=43)
            e = errors(img)
            e.update({"spp": spp, "device_ms": ms})
            row[name] = e
        out["equal_time"]["%gs" % budget] = row
        print("equal time %gs:" % budget, " ".join("%s spp=%d rmse=%.4g" % (k, v["spp"], v["rmse"]) for k, v in row.items()), flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(out, open(args.out, "w"), indent=1)
    scene.close()


if __name__ == "__main__":
    main()
