"""Shared helpers for the tests: seeded input generators and scene construction.

Everything is seeded and pure numpy so the same inputs can be regenerated on the GPU box
(where /root/reference does not exist) and compared with tests/golden/*.npz.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from monte_carlo_path_tracing_b200 import scenes  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

# small scene variants used by the parity tests (same generator parameters as golden/make_golden.py)
SCENE_KW = {
    "cornell-box": dict(width=48, height=36),
    "veach-mis": dict(width=40, height=24, sphere_subdiv=1, plate_tess=(4, 2)),
}


def scene_text(name):
    return scenes.SCENES[name](**SCENE_KW[name])


def u53(rng, shape):
    """Uniforms that are exact multiples of 2^-53 (what oracle/stream.h delivers)."""
    m = rng.integers(0, 1 << 53, size=shape, dtype=np.uint64)
    return m.astype(np.float64) * (1.0 / 9007199254740992.0)


def unit(v):
    return v / np.linalg.norm(v, axis=-1, keepdims=True)


def random_rays(arrays, n, seed, inside=True):
    """Rays with origins inside the scene bounds (or on a shell around it) and random directions."""
    rng = np.random.default_rng(seed)
    lo = arrays.positions.reshape(-1, 3).min(0).astype(np.float64)
    hi = arrays.positions.reshape(-1, 3).max(0).astype(np.float64)
    if inside:
        ro = lo + (hi - lo) * rng.random((n, 3))
    else:
        c = 0.5 * (lo + hi)
        ro = c + unit(rng.normal(size=(n, 3))) * np.linalg.norm(hi - lo)
    rd = unit(rng.normal(size=(n, 3)))
    # rd need not be unit for intersect_with_triangle (Myobj.cpp:165): scale a quarter of them
    rd[: n // 4] *= rng.uniform(0.25, 4.0, size=(n // 4, 1))
    return ro, rd


def aimed_rays(arrays, prim, seed):
    """Rays from random origins aimed at (or just beside) triangle prim[i]: exercises the accept/reject
    rules of intersect_with_triangle (Myobj.cpp:185) with a healthy hit rate on any scene."""
    rng = np.random.default_rng(seed)
    n = len(prim)
    ro, _ = random_rays(arrays, n, seed + 1)
    b = rng.uniform(-0.2, 1.0, n); g = rng.uniform(-0.2, 1.0, n) * (1 - np.clip(b, 0, 1))
    v = arrays.positions[prim].astype(np.float64)
    tgt = v[:, 0] * (1 - b - g)[:, None] + v[:, 1] * b[:, None] + v[:, 2] * g[:, None]
    rd = tgt - ro
    rd[: n // 2] = unit(rd[: n // 2])  # half unit-length, half unnormalised (t ~ 1)
    return ro, rd


def surface_rays(arrays, n, seed):
    """Rays leaving random points ON triangles, excluding the origin triangle (the renderer's case)."""
    rng = np.random.default_rng(seed)
    P = len(arrays.material_id)
    prim = rng.integers(0, P, size=n).astype(np.int32)
    b = rng.random(n); g = rng.random(n) * (1 - b)
    v = arrays.positions[prim].astype(np.float64)
    ro = v[:, 0] * (1 - b - g)[:, None] + v[:, 1] * b[:, None] + v[:, 2] * g[:, None]
    rd = unit(rng.normal(size=(n, 3)))
    return ro, rd, prim


def brdf_inputs(n, seed):
    """Randomised (n, wi, wo, kd, ks, ns in {1,10,100,1000,10000})."""
    rng = np.random.default_rng(seed)
    N = unit(rng.normal(size=(n, 3)))
    # directions in the upper hemisphere of N (mostly), some below to exercise the clamps
    def hemi():
        d = unit(rng.normal(size=(n, 3)))
        s = np.sign(np.sum(d * N, axis=1, keepdims=True))
        flip = rng.random((n, 1)) < 0.9
        return np.where(flip, d * s, d)
    wi, wo = hemi(), hemi()
    kd = rng.uniform(0.05, 0.6, size=(n, 3))
    ks = rng.uniform(0.0, 0.4, size=(n, 3))
    ks[rng.random(n) < 0.2] = 0.0  # purely diffuse materials (discrete_distribution(d, 0))
    ns = rng.choice([1.0, 10.0, 100.0, 1000.0, 10000.0], size=n)
    u3 = u53(rng, (n, 3))
    return N, wi, wo, kd, ks, ns, u3


def shading_points(arrays, n, seed):
    """Random non-emissive surface points with their interpolated shading normal."""
    rng = np.random.default_rng(seed)
    is_light = np.zeros(len(arrays.material_id), bool)
    is_light[arrays.light_prim] = True
    cand = np.nonzero(~is_light)[0]
    prim = rng.choice(cand, size=n)
    b = rng.random(n); g = rng.random(n) * (1 - b)
    v = arrays.positions[prim].astype(np.float64)
    nn = arrays.normals[prim].astype(np.float64)
    w = np.stack([1 - b - g, b, g], axis=1)[:, :, None]
    x = (v * w).sum(1)
    N = unit((nn * w).sum(1))
    u3 = u53(rng, (n, 3))
    return x, N, u3, prim.astype(np.int32)


def rel_err(a, b, floor=1e-300):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), floor)


def rmse(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.sqrt(np.mean((a - b) ** 2)))


def rmse_trimmed(a, b, drop=2):
    """RMSE without the `drop` pixels that differ most.  The as-written NEE estimator (F9: area-measure integrand over a
    solid-angle pdf) now and then lands a light sample on the peak of an Ns = 5000 Phong lobe: one sample worth 300x the
    pixel mean, in 1 render out of 36 of the small veach-mis scene at 32 spp -- with or without the light lists, GPU or
    oracle alike (tools/diag/lists_fireflies.py: 33 of 1200 seeds each).  A plain RMSE bound of 1.5 x noise then fails on
    the unlucky seed; the mean-energy check next to it keeps every pixel."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    e = ((a - b) ** 2).reshape(-1, a.shape[-1]).sum(1)
    if drop and e.size > 10 * drop:
        e = np.sort(e)[:-drop]
    return float(np.sqrt(e.sum() / (e.size * a.shape[-1])))


# ---- oracle renders over all host cores -----------------------------------------------------------------
# The oracle is single-threaded like the reference; with the keyed (or hashed) uniform streams a film does not depend on how
# its rows are split, so the full-size parity tests render disjoint row bands in worker processes.  `spawn`, not
# `fork`: the parent holds a CUDA context.
def _oracle_band(args):
    arrays, cam, strategy, spp, kw, rows = args
    import oracle
    o = oracle.OracleLib(arrays)
    o.build_grid(cam.eye, 100000)
    img, cnt = o.render(cam, strategy, spp, rows=rows, **kw)
    return rows, img, cnt


def oracle_render_mp(arrays, cam, strategy, spp, procs=None, **kw):
    """oracle.OracleLib(arrays).render(cam, strategy, spp, **kw) over `procs` processes (row bands)."""
    import concurrent.futures as cf
    import multiprocessing as mp
    procs = procs or min(os.cpu_count() or 1, 32)
    H = cam.height
    nb = max(1, min(H, procs * 4))
    edges = np.linspace(0, H, nb + 1).astype(int)
    jobs = [(arrays, cam, strategy, spp, kw, (int(a), int(b))) for a, b in zip(edges[:-1], edges[1:]) if b > a]
    out = np.zeros((H, cam.width, 3))
    total = {}
    with cf.ProcessPoolExecutor(max_workers=procs, mp_context=mp.get_context("spawn")) as ex:
        for rows, img, cnt in ex.map(_oracle_band, jobs):
            out[rows[0]:rows[1]] = img
            for k, v in cnt.items():
                total[k] = total.get(k, 0) + int(v)
    return out, total


def camera_rays(cam):
    """All primary rays of `cam` (main.cpp:546-553,563-564) in FP64: origins (H*W, 3), unit directions (H*W, 3)."""
    w = np.array(cam.w, np.float64); up = np.array(cam.up, np.float64)
    N = w / np.linalg.norm(w)
    V = np.cross(N, up); V /= np.linalg.norm(V)
    U = np.cross(V, N); U /= np.linalg.norm(U)
    i = np.arange(cam.height, dtype=np.float64)[:, None]; j = np.arange(cam.width, dtype=np.float64)[None, :]
    dx = -cam.pixellen * (i - (cam.height - 1) * 0.5) + 0 * j
    dy = cam.pixellen * (j - (cam.width - 1) * 0.5) + 0 * i
    d = U[None, None, :] * dx[..., None] + V[None, None, :] * dy[..., None] + N[None, None, :] * np.linalg.norm(w)
    d = (d / np.linalg.norm(d, axis=-1, keepdims=True)).reshape(-1, 3)
    o = np.broadcast_to(np.array(cam.eye, np.float64), d.shape).copy()
    return o, d
