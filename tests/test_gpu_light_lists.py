"""Per-primitive survivor lists of the spherical-triangle light selection (csrc/light_lists.h).

The lists only leave out lights the reference's two culls (Mylight.cpp:178-197) remove at EVERY vertex of a primitive,
so the estimator is the one of the full O(L) sweep: (1) every light that survives the culls at a point of a primitive
is on that primitive's list (checked in FP64 on random points); (2) films rendered with and without the lists agree to
within Monte Carlo noise, with about half the weight evaluations; (3) the films stay deterministic and independent of
the pass size.
"""
import tempfile

import numpy as np
import pytest

import util
from monte_carlo_path_tracing_b200 import api, scenes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def veach():
    sc = scenes.veach_mis(width=160, height=90)
    with tempfile.TemporaryDirectory() as td:
        stem = sc.write(td)
        arrays, _ = api.load_scene_files(stem + ".obj", stem + ".xml")
    return arrays, sc.camera


@pytest.fixture(scope="module")
def scene_lists(veach):
    s = api.Scene(veach[0], 0)
    yield s
    s.close()


@pytest.fixture(scope="module")
def scene_table(veach):
    import os
    old = os.environ.get("MCPT_LIGHT_LISTS")
    os.environ["MCPT_LIGHT_LISTS"] = "0"
    try:
        s = api.Scene(veach[0], 0)
    finally:
        if old is None:
            del os.environ["MCPT_LIGHT_LISTS"]
        else:
            os.environ["MCPT_LIGHT_LISTS"] = old
    yield s
    s.close()


def test_lists_hold_every_light_that_survives_the_culls(veach, scene_lists):
    arrays, _ = veach
    info = scene_lists.light_lists()
    assert info["lists"] > 0 and 0.3 < info["mean_fraction"] < 0.7, info
    P = arrays.positions.astype(np.float64)
    Nv = arrays.normals.astype(np.float64)
    lp = arrays.light_prim
    L = len(lp)
    ng = scene_lists.geom_normals()[lp]                      # emitting side, FP64
    LV = P[lp]                                               # [L,3,3]
    emissive = np.zeros(len(P), bool)
    emissive[lp] = True
    rng = np.random.default_rng(5)
    eps = 1e-8
    checked = 0
    for prim in np.where(~emissive)[0][::3]:
        on_list = np.zeros(L, bool)
        lst = scene_lists.light_list_of_prim(int(prim))
        assert lst is not None and (np.diff(lst) > 0).all()
        on_list[lst] = True
        b = rng.dirichlet((1, 1, 1), size=12)
        b = np.vstack([b, np.eye(3), [[0.5, 0.5, 0], [0, 0.5, 0.5]]])   # corners and edges too
        x = b @ P[prim]
        n = util.unit(b @ Nv[prim])
        tmp = np.einsum("lc,plc->pl", ng, x[:, None, :] - LV[None, :, 0, :])           # Mylight.cpp:178
        behind = (tmp < 0) | (np.abs(tmp) < eps)
        t = np.einsum("pc,plkc->plk", n, LV[None] - x[:, None, None, :])               # :187-191
        below = ((t < 0) | (np.abs(t) < eps)).all(axis=2)
        survives = ~behind & ~below
        missing = survives & ~on_list[None, :]
        assert not missing.any(), (prim, np.argwhere(missing)[:5])
        checked += survives.sum()
    assert checked > 0
    assert scene_lists.light_list_of_prim(int(lp[0])) is None  # emissive primitives are never sampled from


@pytest.mark.parametrize("strategy,flags", [(api.MIS, 0), (api.NEE_SPHERICAL, api.FIX_NEE_MEASURE | api.FIX_MIXTURE_PDF)],
                         ids=["MIS", "NEE_SPHERICAL"])
def test_lists_leave_the_estimator_alone(veach, scene_lists, scene_table, strategy, flags):
    _, cam = veach
    spp = 64
    assert scene_table.light_lists()["lists"] == 0
    a1, sa = scene_lists.render(cam, strategy, spp, flags=flags, seed=11)
    b1, sb = scene_table.render(cam, strategy, spp, flags=flags, seed=11)
    b2, _ = scene_table.render(cam, strategy, spp, flags=flags, seed=12)
    noise = util.rmse(np.minimum(b1, 4.0), np.minimum(b2, 4.0))
    err = util.rmse(np.minimum(a1, 4.0), np.minimum(b2, 4.0))
    assert noise > 0 and err <= 1.15 * noise, (err, noise)
    # same seed: the same paths up to the handful whose light pick lands on another triangle (FP32 sum order)
    assert abs(a1.mean() - b1.mean()) <= 0.01 * b1.mean(), (a1.mean(), b1.mean())
    assert abs(sa["vertices"] - sb["vertices"]) <= 0.01 * sb["vertices"]
    assert abs(sa["light_sweeps"] - sb["light_sweeps"]) <= 0.01 * sb["light_sweeps"]
    # ... at about half the weight evaluations
    assert sa["light_evals"] < 0.62 * sb["light_evals"], (sa["light_evals"], sb["light_evals"])
    Lpad = (len(veach[0].light_prim) + 63) // 64 * 64
    assert sb["light_evals"] == sb["light_sweeps"] * Lpad


def test_lists_keep_films_deterministic_and_pass_size_independent(veach, scene_lists):
    _, cam = veach
    a, _ = scene_lists.render(cam, api.MIS, 48, flags=api.FIX_MIXTURE_PDF, seed=3)
    b, _ = scene_lists.render(cam, api.MIS, 48, flags=api.FIX_MIXTURE_PDF, seed=3)
    c, _ = scene_lists.render(cam, api.MIS, 48, flags=api.FIX_MIXTURE_PDF, seed=3, batch_spp=7)
    assert np.array_equal(a, b) and np.array_equal(a, c)
    # spp sharding: two halves accumulated = the whole, to FP32 sum order
    h0, _ = scene_lists.render(cam, api.MIS, 24, flags=api.FIX_MIXTURE_PDF, seed=3, spp_offset=0, spp_total=48)
    h1, _ = scene_lists.render(cam, api.MIS, 24, flags=api.FIX_MIXTURE_PDF, seed=3, spp_offset=24, spp_total=48)
    assert np.allclose(h0 + h1, a, rtol=2e-6, atol=1e-7)
