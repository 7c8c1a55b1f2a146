"""PARITY render modes of mcpt_render (include/mcpt.h: MCPT_PARITY_FP64, MCPT_LEAF_EXACT) against the oracle.

MCPT_PARITY_FP64 runs the whole estimator in FP64 in the reference's expression order, fed by the renderer's keyed
uniforms; the oracle reads the same table under ORACLE_RNG_KEYED and resolves rays by exhaustive search
(ORACLE_BRUTE_FORCE: the semantics the BVH implements).  Both therefore walk the SAME paths and the films are compared
PIXEL BY PIXEL -- not statistically: every pixel within 1e-9 relative (+ the float32 rounding of the output film), and
identical work counters.  The only arithmetic that differs is the last bit of acos / sin / cos / pow (CUDA vs glibc).

Sizes: the small golden scenes for every estimator and switch, then the configurations bench.py runs -- cornell-box
512x512x16 spp depth 5 (BASELINE.json configs[0]) and the full veach-mis scene with shade_with_mis and NO depth cap
(configs[1], the headline) at a resolution the oracle finishes in about a minute on the host cores.  The FP32 wavefront
(the throughput path) is then held against the PARITY renderer on the GPU, where high spp are cheap.
"""
import tempfile

import numpy as np
import pytest

import oracle
import util
from monte_carlo_path_tracing_b200 import api, scenes

pytestmark = pytest.mark.gpu

OFLAGS = oracle.FIX_STALE_MIS | oracle.BRUTE_FORCE | oracle.RNG_KEYED

# (name, strategy, gpu flags, oracle flags)
CASES = [
    ("NEE_UNIFORM", api.NEE_UNIFORM, 0, 0),
    ("BRDF_ONLY", api.BRDF_ONLY, 0, 0),
    ("BRDF_ONLY_mixture", api.BRDF_ONLY, api.FIX_MIXTURE_PDF, oracle.FIX_MIXTURE_PDF),
    ("NEE_SPHERICAL_as_written", api.NEE_SPHERICAL, 0, 0),
    ("NEE_SPHERICAL_fixed", api.NEE_SPHERICAL, api.FIX_NEE_MEASURE | api.FIX_MIXTURE_PDF, oracle.FIX_NEE_MEASURE | oracle.FIX_MIXTURE_PDF),
    ("MIS_as_written", api.MIS, 0, 0),
    ("MIS_fixed", api.MIS, api.FIX_MIXTURE_PDF, oracle.FIX_MIXTURE_PDF),
]


def assert_same_film(gpu, orc, what, max_bad=0):
    """gpu: float32 film of a PARITY render; orc: float64 oracle film.  Pixel by pixel."""
    orc = np.asarray(orc, np.float64)
    assert np.isfinite(gpu).all() and np.isfinite(orc).all(), what
    tol = 1e-9 * np.abs(orc) + 6e-8 * np.abs(orc) + 1e-12  # 1e-9 relative + half an ulp of the float32 output
    bad = np.abs(gpu.astype(np.float64) - orc) > tol
    nbad = int(np.any(bad, axis=-1).sum())
    assert nbad <= max_bad, (what, nbad, float(np.abs(gpu - orc).max()))
    return nbad


@pytest.mark.parametrize("depth", [5, 0], ids=["depth5", "unlimited"])
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_parity_film_equals_oracle_film(scene_name, case, depth, gpu_scene_of, oracle_of, arrays_of):
    """Every estimator, as written and with the F8 / F9 switches, with and without a depth cap: same film, same counters."""
    _, strat, gflags, oflags = case
    _, cam = arrays_of(scene_name)
    spp = 8
    ref, cnt = oracle_of(scene_name).render(cam, strat, spp, max_depth=depth, flags=OFLAGS | oflags, seed=11)
    img, st = gpu_scene_of(scene_name).render(cam, strat, spp, max_depth=depth, flags=gflags | api.PARITY_FP64, seed=11)
    assert_same_film(img, ref, case[0])
    assert ref.max() > 0
    assert st["vertices"] == cnt["vertices"] and st["primary_hits"] == cnt["primary_hits"]
    assert st["rays_closest"] + st["rays_shadow"] == cnt["rays"] and st["rays_light"] == cnt["light_rays"]


def test_parity_film_is_deterministic_and_shards(scene_name, gpu_scene_of, arrays_of):
    """spp sharding and pass size do not change a PARITY film beyond the float32 output rounding; reruns are identical."""
    _, cam = arrays_of(scene_name)
    s = gpu_scene_of(scene_name)
    kw = dict(max_depth=4, flags=api.PARITY_FP64, seed=5)
    full, _ = s.render(cam, api.MIS, 6, **kw)
    again, _ = s.render(cam, api.MIS, 6, **kw)
    assert np.array_equal(full, again)
    small, _ = s.render(cam, api.MIS, 6, batch_spp=1, **kw)
    assert np.array_equal(full, small)  # FP64 film, samples summed in order whatever the pass size
    lo, _ = s.render(cam, api.MIS, 3, spp_offset=0, spp_total=6, **kw)
    hi, _ = s.render(cam, api.MIS, 3, spp_offset=3, spp_total=6, **kw)
    assert np.allclose(lo + hi, full, rtol=3e-7, atol=1e-9)


def test_parity_rejects_non_reference_estimators(scene_name, gpu_scene_of, arrays_of):
    _, cam = arrays_of(scene_name)
    s = gpu_scene_of(scene_name)
    with pytest.raises(api.McptError):
        s.render(cam, api.MIS_SINGLE, 1, flags=api.PARITY_FP64)
    with pytest.raises(api.McptError):
        s.render(cam, api.MIS, 1, flags=api.PARITY_FP64 | api.LIGHT_TREE)
    with pytest.raises(api.McptError):
        s.render(cam, api.MIS, 1, leaf_mode=7)


# ---- the configurations bench.py runs ------------------------------------------------------------------------------------
def _load(sc):
    with tempfile.TemporaryDirectory() as td:
        stem = sc.write(td)
        a, _ = api.load_scene_files(stem + ".obj", stem + ".xml")
    return a


@pytest.fixture(scope="module")
def cornell512():
    sc = scenes.cornell_box(width=512, height=512)
    a = _load(sc)
    s = api.Scene(a, 0)
    yield s, sc.camera, a
    s.close()


@pytest.fixture(scope="module")
def veach_full():
    """The full veach-mis scene (1282 light triangles, 2.8 k facets) behind a 192x108 camera."""
    sc = scenes.veach_mis(width=192, height=108)
    a = _load(sc)
    s = api.Scene(a, 0)
    yield s, sc.camera, a
    s.close()


def test_cornell_512_all_primary_and_1m_interior_rays_bit_exact(cornell512):
    """hit ids of all 262 144 primary rays of cornell-box 512x512 and of 1 M random interior rays, bit
    exact against the oracle's exhaustive search (Myobj.cpp:194-213 semantics), beta / gamma / t bit-identical too."""
    s, cam, a = cornell512
    o = oracle.OracleLib(a)
    ro, rd = util.camera_rays(cam)
    assert len(ro) == 262144
    p_gpu, b_gpu = s.closest(ro, rd, None, False, api.FP64)
    p_cpu, b_cpu = o.closest(ro, rd, None, 2)
    assert np.array_equal(p_gpu, p_cpu) and np.array_equal(b_gpu, b_cpu)
    assert (p_cpu >= 0).mean() > 0.4
    ro, rd = util.random_rays(a, 1_000_000, seed=77)
    p_gpu, b_gpu = s.closest(ro, rd, None, False, api.FP64)
    p_cpu, b_cpu = o.closest(ro, rd, None, 2)
    assert np.array_equal(p_gpu, p_cpu) and np.array_equal(b_gpu, b_cpu)
    # the wavefront's MCPT_LEAF_EXACT kernel (k_extend_exact) on the FP32 roundings of the same rays: ids identical to the
    # exhaustive search for THOSE rays, beta / gamma / t equal to the float32 rounding the hit record stores
    ro32, rd32 = ro.astype(np.float32).astype(np.float64), rd.astype(np.float32).astype(np.float64)
    p_k, b_k = s.closest(ro32, rd32, None, False, 2)
    p_c, b_c = o.closest(ro32, rd32, None, 2)
    assert np.array_equal(p_k, p_c)
    assert np.array_equal(b_k.astype(np.float32), b_c.astype(np.float32))
    # surface rays with the origin triangle excluded by id (Myobj.cpp:417), the renderer's case
    ro, rd, prim = util.surface_rays(a, 200_000, seed=78)
    p_gpu, b_gpu = s.closest(ro, rd, prim, False, api.FP64)
    p_cpu, b_cpu = o.closest(ro, rd, prim, 2)
    assert np.array_equal(p_gpu, p_cpu) and np.array_equal(b_gpu, b_cpu)


@pytest.mark.parametrize("strat", [api.NEE_UNIFORM, api.MIS], ids=["NEE_UNIFORM", "MIS"])
def test_cornell_512x512x16_depth5_parity_equals_oracle(cornell512, strat):
    """BASELINE.json configs[0] at full size: the PARITY film IS the oracle's film (4.2 M samples, 262 144 pixels)."""
    s, cam, a = cornell512
    ref, cnt = util.oracle_render_mp(a, cam, strat, 16, max_depth=5, flags=OFLAGS, seed=1)
    img, st = s.render(cam, strat, 16, max_depth=5, flags=api.PARITY_FP64, seed=1)
    # a handful of paths may part ways where a last-bit difference of acos / pow flips a discrete decision
    nbad = assert_same_film(img, ref, "cornell 512x512x16", max_bad=3)
    assert abs(st["vertices"] - cnt["vertices"]) <= 8 * max(nbad, 0) and st["primary_hits"] == cnt["primary_hits"]
    assert abs(float(img.mean()) - float(ref.mean())) <= 1e-6 * float(ref.mean())


def test_cornell_512_fast_wavefront_vs_parity(cornell512):
    """The FP32 wavefront against the PARITY renderer at configs[0]'s size.  NEE_UNIFORM / BRDF_ONLY draw the same
    uniforms for the same decisions in both, so the films agree path by path up to FP32 rounding: nearly every pixel
    within 1e-3 relative, the rest (a rounding flipped a discrete choice somewhere along one of its 16 paths) bounded in
    number; mean energy within 0.1 %."""
    s, cam, _ = cornell512
    for strat in (api.NEE_UNIFORM, api.BRDF_ONLY):
        par, _ = s.render(cam, strat, 16, max_depth=5, flags=api.PARITY_FP64, seed=3)
        fast, _ = s.render(cam, strat, 16, max_depth=5, seed=3)
        exact, _ = s.render(cam, strat, 16, max_depth=5, seed=3, leaf_mode=api.LEAF_EXACT)
        for img in (fast, exact):
            lum_p, lum_f = par.sum(-1).astype(np.float64), img.sum(-1).astype(np.float64)
            off = np.abs(lum_f - lum_p) > 1e-3 * lum_p + 1e-5
            assert off.mean() < 0.02, (strat, off.mean())
            assert abs(lum_f.mean() - lum_p.mean()) <= 1e-3 * lum_p.mean()


def test_veach_mis_headline_parity_equals_oracle(veach_full):
    """BASELINE.json configs[1], the headline: full veach-mis scene, shade_with_mis, no depth cap (Russian roulette only,
    main.cpp:429-437), 1282 light triangles swept at every vertex."""
    s, cam, a = veach_full
    ref, cnt = util.oracle_render_mp(a, cam, oracle.MIS, 4, max_depth=0, flags=OFLAGS, seed=2)
    img, st = s.render(cam, api.MIS, 4, max_depth=0, flags=api.PARITY_FP64, seed=2)
    nbad = assert_same_film(img, ref, "veach-mis 192x108x4 unlimited depth", max_bad=2)
    assert abs(st["vertices"] - cnt["vertices"]) <= 16 * nbad and st["primary_hits"] == cnt["primary_hits"]
    assert ref.max() > 0 and abs(float(img.mean()) - float(ref.mean())) <= 1e-5 * float(ref.mean())


def test_veach_mis_headline_fast_wavefront_vs_parity(veach_full):
    """The headline estimator in the FP32 wavefront (exact O(L) light selection, MCPT_LEAF_FAST and MCPT_LEAF_EXACT) against
    the PARITY renderer at 64 spp: RMSE within 1.5 x the PARITY renderer's own seed-to-seed noise, mean energy within 1 %
    (SURVEY.md 8c level 2).  The spherical strategies map a uniform to a light in a different order (lane-major) than
    std::discrete_distribution does, so these films agree in distribution, not path by path."""
    s, cam, _ = veach_full
    spp = 64
    pa, sa = s.render(cam, api.MIS, spp, max_depth=0, flags=api.PARITY_FP64, seed=21)
    pb, _ = s.render(cam, api.MIS, spp, max_depth=0, flags=api.PARITY_FP64, seed=22)
    noise = util.rmse(pa, pb)
    ref = 0.5 * (pa.astype(np.float64) + pb)
    for leaf in (api.LEAF_FAST, api.LEAF_EXACT):
        img, st = s.render(cam, api.MIS, spp, max_depth=0, seed=23, leaf_mode=leaf)
        # against the two-seed mean the expected RMSE is sqrt(1 + 1/2) / sqrt(2) = 0.87 x noise
        assert util.rmse(img, ref) <= 1.5 * noise, (leaf, util.rmse(img, ref), noise)
        assert abs(float(img.mean()) - float(ref.mean())) <= 0.01 * float(ref.mean())
        vpp_f, vpp_p = st["vertices"] / st["paths"], sa["vertices"] / sa["paths"]
        assert abs(vpp_f - vpp_p) <= 0.02 * vpp_p
