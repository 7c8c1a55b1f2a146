"""Generate tests/golden/<scene>.npz from the COMPILED REFERENCE (oracle/_ref/libmcpt_ref.so,
built from the unmodified sources under /root/reference by `make -C oracle ref`).

Run here (where /root/reference exists):   python tests/golden/make_golden.py
The vectors travel with the repo; the GPU box only reads them.  Inputs are regenerated from
tests/util.py with fixed seeds, so the .npz stores outputs (and the flattened scene, which also
pins our own OBJ/MTL/XML loader against tiny_obj_loader + pugixml).
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import util  # noqa: E402
import oracle  # noqa: E402

RENDER = dict(spp=2, max_depth=4, seed=11)
# (strategy, ORACLE_FIX_* flags) of the estimator variants the GPU is compared with: F10 never reproduced (1), the F9 / F8
# switches (2, 4), the uniform-area sampler of main.cpp:297 (strategy 3).  Rendered by ref_render_variant on the
# reference's own objects (oracle/ref_build/ref_driver.cpp).
VARIANTS = [(3, 0), (3, 4), (1, 4), (0, 2), (0, 4), (0, 6), (2, 1), (2, 5)]


def make(name):
    sc = util.scene_text(name)
    with tempfile.TemporaryDirectory() as td:
        stem = sc.write(td)
        ref = oracle.RefLib(stem + ".obj", stem + ".xml")  # one scene per process
        a, gn, area = ref.flatten()
        ref.build_grid(sc.camera.eye, 100000)
        out = dict(positions=a.positions, normals=a.normals, material_id=a.material_id, shape_offset=a.shape_offset,
                   mat_kd=a.mat_kd, mat_ks=a.mat_ks, mat_ns=a.mat_ns, light_prim=a.light_prim, light_class=a.light_class,
                   class_radiance=a.class_radiance, geom_normals=gn, light_area=area)
        # intersection
        ro, rd = util.random_rays(a, 3000, 101)
        out["closest_brute_prim"], out["closest_brute_bgt"] = ref.closest(ro, rd, None, 2)
        out["closest_grid_prim"], out["closest_grid_bgt"] = ref.closest(ro, rd, None, 0)
        out["closest_lbrute_prim"], out["closest_lbrute_bgt"] = ref.closest(ro, rd, None, 3)
        ro, rd, ex = util.surface_rays(a, 3000, 102)
        out["surf_brute_prim"], out["surf_brute_bgt"] = ref.closest(ro, rd, ex, 2)
        out["surf_grid_prim"], out["surf_grid_bgt"] = ref.closest(ro, rd, ex, 0)
        out["surf_lbrute_prim"], out["surf_lbrute_bgt"] = ref.closest(ro, rd, ex, 3)
        out["surf_lgrid_prim"], out["surf_lgrid_bgt"] = ref.closest(ro, rd, ex, 1)
        rng = np.random.default_rng(103)
        prim = rng.integers(0, len(a.material_id), size=3000).astype(np.int32)
        ro, rd = util.aimed_rays(a, prim, 104)
        out["tri_hit"], out["tri_bgt"] = ref.intersect_tri(ro, rd, prim)
        # BRDF
        N, wi, wo, kd, ks, ns, u3 = util.brdf_inputs(3000, 105)
        out["brdf_eval"] = ref.brdf_eval(N, wi, wo, kd, ks, ns)
        out["brdf_sample_dir"], out["brdf_sample_pdf"] = ref.brdf_sample(N, wo, kd, ks, ns, u3)
        out["brdf_pdf"] = ref.brdf_pdf(N, wi, wo, kd, ks, ns)
        # lights
        u4 = util.u53(np.random.default_rng(106), (2000, 4))
        out["lu_prim"], out["lu_point"], out["lu_Le"], out["lu_pdf"] = ref.light_sample_uniform(u4)
        x, Nn, u3, _ = util.shading_points(a, 600, 107)
        query = a.light_prim[np.random.default_rng(108).integers(0, len(a.light_prim), size=600)]
        r = ref.light_sample_spherical(x, Nn, u3, query, staged=1)
        for k, v in r.items():
            out["ls_" + k] = v
        r1 = ref.light_sample_spherical(x, Nn, u3, query, staged=0)  # one-shot variant, Mylight.cpp:163-318
        out["ls1_prim"], out["ls1_point"], out["ls1_pdf"] = r1["prim"], r1["point"], r1["pdf"]
        vals = np.abs(np.random.default_rng(109).normal(size=(200, 3))) * 200
        out["tonemap"] = ref.tonemap(vals)
        # images: the three estimators exactly as checked in (incl. F8, F9, F10)
        for strat in (0, 1, 2):
            img, cnt = ref.render(sc.camera, strat, **RENDER)
            out["img%d" % strat] = img
            out["img%d_vertices" % strat] = np.array([cnt["vertices"], cnt["primary_hits"]], np.int64)
            # closet_ray_intersect / closet_ray_intersect_light_triangle calls, counted by the linker --wrap shims of ref_driver.cpp
            out["img%d_rays" % strat] = np.array([cnt["rays"], cnt["light_rays"]], np.int64)
            # the re-hosted estimator with every switch off IS the reference's own shade* (validates the re-hosting)
            img_v, cnt_v = ref.render_variant(sc.camera, strat, flags=0, **RENDER)
            assert np.array_equal(img, img_v, equal_nan=True) and cnt == cnt_v, ("re-hosted estimator differs from main.cpp", strat)
        for strat, flags in VARIANTS:
            # F10 only shows where recursion #1 changed the survivor set AND the BRDF ray found a light: a few pixels at 24 spp
            img, cnt = ref.render_variant(sc.camera, strat, flags=flags, **dict(RENDER, spp=24 if strat == 2 else RENDER["spp"]))
            out["imgv_s%d_f%d" % (strat, flags)] = img
            out["imgv_s%d_f%d_vertices" % (strat, flags)] = np.array([cnt["vertices"], cnt["primary_hits"]], np.int64)
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print(name, "->", path, "%.1f KB" % (os.path.getsize(path) / 1024))


if __name__ == "__main__":
    if len(sys.argv) > 1:
        make(sys.argv[1])
    else:
        for n in util.SCENE_KW:
            subprocess.check_call([sys.executable, os.path.abspath(__file__), n])
