"""GPU render parity: images from mcpt_render against the oracle at equal spp, per estimator, plus
size-independent properties at BASELINE.json's full cornell-box size.

Tolerance (SURVEY.md 8c level 2): RMSE(gpu, oracle) <= 1.5 x RMSE(oracle seed A, oracle seed B) at the same
spp -- the two renderers use different random streams, so they can only agree up to Monte Carlo noise --
and the mean image energy within 4 standard errors + 1 %.
"""
import copy

import numpy as np
import pytest

import oracle
import util
from monte_carlo_path_tracing_b200 import api, scenes

pytestmark = pytest.mark.gpu

# (strategy, gpu flags, oracle flags): F10 is never reproduced on the GPU -> oracle runs FIX_STALE_MIS
CASES = [
    ("NEE_UNIFORM", api.NEE_UNIFORM, 0, oracle.FIX_STALE_MIS),
    ("BRDF_ONLY", api.BRDF_ONLY, 0, oracle.FIX_STALE_MIS),
    ("NEE_SPHERICAL_as_written", api.NEE_SPHERICAL, 0, oracle.FIX_STALE_MIS),
    ("NEE_SPHERICAL_fixed", api.NEE_SPHERICAL, api.FIX_NEE_MEASURE | api.FIX_MIXTURE_PDF,
     oracle.FIX_STALE_MIS | oracle.FIX_NEE_MEASURE | oracle.FIX_MIXTURE_PDF),
    ("MIS_as_written", api.MIS, 0, oracle.FIX_STALE_MIS),
    ("MIS_fixed", api.MIS, api.FIX_MIXTURE_PDF, oracle.FIX_STALE_MIS | oracle.FIX_MIXTURE_PDF),
]
SPP = {"cornell-box": 128, "veach-mis": 32}
DEPTH = 5


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_image_matches_oracle(scene_name, case, gpu_scene_of, oracle_of, arrays_of):
    _, strat, gflags, oflags = case
    _, cam = arrays_of(scene_name)
    spp = SPP[scene_name]
    o = oracle_of(scene_name)
    a, _ = o.render(cam, strat, spp, max_depth=DEPTH, flags=oflags, seed=1)
    b, cnt = o.render(cam, strat, spp, max_depth=DEPTH, flags=oflags, seed=2)
    img, st = gpu_scene_of(scene_name).render(cam, strat, spp, max_depth=DEPTH, flags=gflags, seed=3)
    assert np.isfinite(img).all()
    a = np.nan_to_num(a); b = np.nan_to_num(b)
    noise = util.rmse_trimmed(a, b)
    ref = 0.5 * (a + b)
    err = util.rmse_trimmed(img, a)
    assert noise > 0
    # + the float32 rounding of the output image (directly visible emitters are ~1e2 brighter than the rest)
    assert err <= 1.5 * noise + 2e-7 * ref.max(), (err, noise)
    # mean energy: difference of means vs its standard error over pixels
    diff = (img.astype(np.float64) - ref).mean(axis=2).ravel()
    se = diff.std() / np.sqrt(diff.size)
    assert abs(diff.mean()) <= 4 * se + 0.01 * ref.mean(), (diff.mean(), se, ref.mean())
    # same control flow -> the same work, up to noise: sample-bounces per path within 3 %
    vpp_gpu = st["vertices"] / st["paths"]
    vpp_orc = cnt["vertices"] / (cam.width * cam.height * spp)
    assert abs(vpp_gpu - vpp_orc) <= 0.03 * vpp_orc + 0.01, (vpp_gpu, vpp_orc)
    assert st["primary_hits"] == cnt["primary_hits"]


def test_primary_hits_match_oracle(scene_name, gpu_scene_of, oracle_of, arrays_of):
    """Camera rays (main.cpp:563-564) + closest hit: emission-only image (max_depth=1, BRDF_ONLY) is
    deterministic, so it must equal the oracle's pixel for pixel (FP32 rays: allow border pixels)."""
    _, cam = arrays_of(scene_name)
    a, _ = oracle_of(scene_name).render(cam, oracle.BRDF_ONLY, 1, max_depth=1, seed=1)
    img, _ = gpu_scene_of(scene_name).render(cam, api.BRDF_ONLY, 1, max_depth=1, seed=1)
    mism = np.mean(np.any(np.abs(img - a) > 1e-4 * (1 + np.abs(a)), axis=2))
    assert mism < 0.01, mism


@pytest.fixture(scope="module")
def cornell_full():
    sc = scenes.cornell_box(width=512, height=512)
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        stem = sc.write(td)
        a, _ = api.load_scene_files(stem + ".obj", stem + ".xml")
    s = api.Scene(a, 0)
    yield s, sc.camera, a
    s.close()


def test_full_size_properties(cornell_full):
    """BASELINE.json config 1 size (512x512, 16 spp, depth 5): determinism, batch-size independence,
    spp sharding (the multi-GPU decomposition), linearity in the emitted radiance."""
    s, cam, a = cornell_full
    kw = dict(max_depth=5, seed=9)
    full, st = s.render(cam, api.NEE_UNIFORM, 16, **kw)
    again, _ = s.render(cam, api.NEE_UNIFORM, 16, **kw)
    assert np.array_equal(full, again)  # deterministic: counter-based RNG, atomic-free accumulation
    small_batch, st_small = s.render(cam, api.NEE_UNIFORM, 16, batch_spp=3, **kw)
    assert np.array_equal(full, small_batch)  # the film adds its samples one by one in sample order, whatever the pass size
    # the visibility-ray queue is filled through per-warp reservations whose unused tails are holes: they are counted and
    # taken off, so the ray counters do not depend on how many passes (and reservations) a render was made of
    for k in ("rays_shadow", "rays_closest", "vertices", "primary_hits"):
        assert st[k] == st_small[k], (k, st[k], st_small[k])
    assert 0 < st["rays_shadow"] <= st["vertices"]
    # spp sharding: samples [0,8) and [8,16) rendered separately, each normalised by the global 16
    lo, _ = s.render(cam, api.NEE_UNIFORM, 8, spp_offset=0, spp_total=16, **kw)
    hi, _ = s.render(cam, api.NEE_UNIFORM, 8, spp_offset=8, spp_total=16, **kw)
    assert np.allclose(lo + hi, full, rtol=1e-5, atol=1e-6)
    assert st["paths"] == 512 * 512 * 16
    assert st["vertices"] <= st["paths"] * 5 and st["vertices"] >= st["primary_hits"] * 0.5
    assert np.isfinite(full).all() and full.mean() > 0
    # different seeds differ, same expectation
    other, _ = s.render(cam, api.NEE_UNIFORM, 16, max_depth=5, seed=10)
    assert not np.array_equal(full, other)
    assert abs(other.mean() - full.mean()) < 0.02 * full.mean()
    # linearity: doubling every emitted radiance doubles the image exactly (power of two, one light class)
    b = copy.copy(a)
    b.class_radiance = a.class_radiance * 2.0
    s2 = api.Scene(b, 0)
    try:
        dbl, _ = s2.render(cam, api.NEE_UNIFORM, 16, **kw)
    finally:
        s2.close()
    assert np.array_equal(dbl, full * 2.0)


def test_mis_tree_full_size(cornell_full):
    """The branching estimator (shade_with_mis recurses twice, F6): atomic-free run sums are deterministic,
    independent of batch size, and the queue never overflows at the default capacity."""
    s, cam, _ = cornell_full
    a, st = s.render(cam, api.MIS, 8, max_depth=5, seed=4)
    b, _ = s.render(cam, api.MIS, 8, max_depth=5, seed=4)
    assert np.array_equal(a, b)
    c, _ = s.render(cam, api.MIS, 8, max_depth=5, seed=4, batch_spp=2)
    assert np.allclose(a, c, rtol=1e-5, atol=1e-6)
    assert st["queue_peak"] <= 3 * 512 * 512 * 8
    assert st["rays_light"] > 0 and np.isfinite(a).all()


def test_unlimited_depth_and_jitter(gpu_scene_of, arrays_of):
    """max_depth <= 0 is the reference's behaviour (RR only, F5); jitter is an extension (F4)."""
    _, cam = arrays_of("cornell-box")
    s = gpu_scene_of("cornell-box")
    a, st = s.render(cam, api.NEE_UNIFORM, 64, max_depth=0, seed=5)
    b, _ = s.render(cam, api.NEE_UNIFORM, 64, max_depth=12, seed=5)
    assert np.isfinite(a).all()
    clip = lambda im: np.minimum(im, 1.0).mean()  # the few directly visible emitter pixels dominate a plain mean
    assert abs(clip(a) - clip(b)) < 0.02 * clip(b)  # depth > 12 carries 0.6^12 of the energy
    assert st["bounces"] > 12
    c, _ = s.render(cam, api.NEE_UNIFORM, 64, max_depth=5, seed=5)
    j, _ = s.render(cam, api.NEE_UNIFORM, 64, max_depth=5, seed=5, flags=api.JITTER)
    assert np.isfinite(j).all() and abs(clip(j) - clip(c)) < 0.05 * clip(c)


def test_axis_aligned_cornell_is_finite(tmp_path):
    """The un-rotated Cornell box has a wall normal of exactly (-1,0,0), which NaNs the reference's tangent
    frame (F11, BRDF.cpp:60-67).  The GPU frame falls back to e_y there: the image must be finite."""
    sc = scenes.cornell_box(width=64, height=64, rot=None, name="cornell-aa")
    stem = sc.write(str(tmp_path))
    a, _ = api.load_scene_files(stem + ".obj", stem + ".xml")
    s = api.Scene(a, 0)
    try:
        for strat in (api.NEE_UNIFORM, api.MIS):
            img, _ = s.render(sc.camera, strat, 32, max_depth=5, seed=1)
            assert np.isfinite(img).all() and img.mean() > 0
    finally:
        s.close()


def test_cpp_driver_end_to_end(scene_files, tmp_path, gpu_scene_of, arrays_of):
    """mcpt_render: the C++ replacement of main() -- load, render through the C ABI, tone-map, write PPM/PFM.
    Same seed and parameters as a ctypes render -> the PFM holds exactly the same linear film."""
    import os
    import subprocess
    drv = os.path.join(util.ROOT, "monte_carlo_path_tracing_b200", "mcpt_render")
    _, stem = scene_files("cornell-box")
    _, cam = arrays_of("cornell-box")
    ppm, pfm = tmp_path / "o.ppm", tmp_path / "o.pfm"
    r = subprocess.run([drv, stem + ".obj", stem + ".xml", str(ppm), "--spp", "8", "--strategy", "nee_uniform",
                        "--max-depth", "4", "--seed", "7", "--pfm", str(pfm)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    raw = ppm.read_bytes()
    hdr = ("P6\n%d %d\n255\n" % (cam.width, cam.height)).encode()
    assert raw.startswith(hdr) and len(raw) == len(hdr) + cam.width * cam.height * 3
    data = pfm.read_bytes()
    h2 = ("PF\n%d %d\n-1.0\n" % (cam.width, cam.height)).encode()
    film = np.frombuffer(data[len(h2):], np.float32).reshape(cam.height, cam.width, 3)[::-1]
    img, _ = gpu_scene_of("cornell-box").render(cam, api.NEE_UNIFORM, 8, max_depth=4, seed=7)
    assert np.array_equal(film, img)
    rgb8 = np.frombuffer(raw[len(hdr):], np.uint8).reshape(cam.height, cam.width, 3)
    assert np.array_equal(rgb8, gpu_scene_of("cornell-box").tonemap(img))
    # the reference's saveimage() format: 24-bit BMP, bottom-up, BGR, rows padded to 4 bytes
    bmp = tmp_path / "o.bmp"
    r = subprocess.run([drv, stem + ".obj", stem + ".xml", str(bmp), "--spp", "8", "--strategy", "nee_uniform",
                        "--max-depth", "4", "--seed", "7"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    b = bmp.read_bytes()
    row = (cam.width * 3 + 3) // 4 * 4
    assert b[:2] == b"BM" and int.from_bytes(b[10:14], "little") == 54 and len(b) == 54 + row * cam.height
    assert int.from_bytes(b[18:22], "little") == cam.width and int.from_bytes(b[22:26], "little") == cam.height
    pix = np.frombuffer(b[54:], np.uint8).reshape(cam.height, row)[:, :cam.width * 3].reshape(cam.height, cam.width, 3)
    assert np.array_equal(pix[::-1, :, ::-1], rgb8)


@pytest.mark.gpu
def test_cpp_driver_checkpoint_resume(scene_files, tmp_path, arrays_of):
    """Progressive rendering: 12 spp in batches of 4 with a checkpoint after each batch; a run
    stopped at 4 spp and resumed gives bit-for-bit the film of the uninterrupted run (samples are keyed by their global
    index), and a checkpoint written with other parameters is refused."""
    import os
    import subprocess
    drv = os.path.join(util.ROOT, "monte_carlo_path_tracing_b200", "mcpt_render")
    _, stem = scene_files("cornell-box")
    _, cam = arrays_of("cornell-box")
    base = [drv, stem + ".obj", stem + ".xml"]
    opts = ["--strategy", "mis", "--seed", "11", "--every", "4"]

    def film_of(path):
        data = path.read_bytes()
        h2 = ("PF\n%d %d\n-1.0\n" % (cam.width, cam.height)).encode()
        return np.frombuffer(data[len(h2):], np.float32).reshape(cam.height, cam.width, 3)[::-1]

    full = tmp_path / "full.pfm"
    r = subprocess.run(base + [str(tmp_path / "a.ppm"), "--spp", "12", "--pfm", str(full)] + opts, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "12 / 12 spp" in r.stdout and "mean relative error" in r.stdout
    ck = tmp_path / "render.ckpt"
    r = subprocess.run(base + [str(tmp_path / "b.ppm"), "--spp", "4", "--checkpoint", str(ck)] + opts, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and ck.exists(), r.stderr
    part = tmp_path / "part.pfm"
    r = subprocess.run(base + [str(tmp_path / "c.ppm"), "--spp", "12", "--checkpoint", str(ck), "--resume", "--pfm", str(part)] + opts,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "resumed" in r.stdout and "at 4 spp" in r.stdout, r.stdout + r.stderr
    assert np.array_equal(film_of(full), film_of(part))
    assert film_of(full).mean() > 0
    r = subprocess.run(base + [str(tmp_path / "d.ppm"), "--spp", "12", "--checkpoint", str(ck), "--resume", "--seed", "12",
                               "--strategy", "mis", "--every", "4"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 2 and "different" in r.stderr
    # adaptive stop: a loose target ends the render before the spp budget is used up
    r = subprocess.run(base + [str(tmp_path / "e.ppm"), "--spp", "4096", "--target-rel-error", "0.5"] + opts, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "4096 / 4096" not in r.stdout


def test_single_path_mis_converges_to_the_same_image(scene_name, gpu_scene_of, oracle_of, arrays_of):
    """MCPT_MIS_SINGLE (not a reference estimator): textbook single-path MIS.  Same expectation as
    the corrected reference estimators: compared with the ORACLE's shade() with both fixes (solid-angle NEE, mixture
    pdf) at equal spp -- error within the oracle's own seed-to-seed noise, mean energy within 4 standard errors -- and
    with the GPU's path-tree MIS on block averages.  No path tree: at most one child and one visibility ray per vertex."""
    _, cam = arrays_of(scene_name)
    spp = SPP[scene_name]
    o = oracle_of(scene_name)
    of = oracle.FIX_STALE_MIS | oracle.FIX_NEE_MEASURE | oracle.FIX_MIXTURE_PDF
    a, _ = o.render(cam, oracle.NEE_SPHERICAL, spp, max_depth=DEPTH, flags=of, seed=1)
    b, _ = o.render(cam, oracle.NEE_SPHERICAL, spp, max_depth=DEPTH, flags=of, seed=2)
    a = np.nan_to_num(a); b = np.nan_to_num(b)
    s = gpu_scene_of(scene_name)
    img, st = s.render(cam, api.MIS_SINGLE, spp, max_depth=DEPTH, flags=api.FIX_MIXTURE_PDF, seed=3)
    assert np.isfinite(img).all()
    noise, ref = util.rmse(a, b), 0.5 * (a + b)
    assert util.rmse(img, a) <= 1.5 * noise + 2e-7 * ref.max(), (util.rmse(img, a), noise)
    diff = (img.astype(np.float64) - ref).mean(axis=2).ravel()
    se = diff.std() / np.sqrt(diff.size)
    assert abs(diff.mean()) <= 4 * se + 0.01 * ref.mean(), (diff.mean(), se, ref.mean())
    assert st["rays_shadow"] > 0 and st["rays_light"] > 0 and st["queue_peak"] <= cam.width * cam.height * spp
    # against the path-tree MIS on the GPU, 8x8 block means of a 4x longer render
    tree, _ = s.render(cam, api.MIS, 4 * spp, max_depth=DEPTH, flags=api.FIX_MIXTURE_PDF, seed=4)
    single, _ = s.render(cam, api.MIS_SINGLE, 4 * spp, max_depth=DEPTH, flags=api.FIX_MIXTURE_PDF, seed=5)
    H, W = cam.height // 8 * 8, cam.width // 8 * 8
    blk = lambda im: np.minimum(im[:H, :W], 4.0).reshape(H // 8, 8, W // 8, 8, 3).mean(axis=(1, 3))
    bt, bs = blk(tree), blk(single)
    rel = np.linalg.norm(bt - bs) / np.linalg.norm(bt)
    assert rel < 0.05, rel
    # deterministic and batch-size independent like every other strategy
    again, _ = s.render(cam, api.MIS_SINGLE, spp, max_depth=DEPTH, flags=api.FIX_MIXTURE_PDF, seed=3)
    assert np.array_equal(img, again)


@pytest.mark.parametrize("strategy,flags", [(api.MIS, api.FIX_MIXTURE_PDF), (api.NEE_SPHERICAL, api.FIX_NEE_MEASURE | api.FIX_MIXTURE_PDF),
                                            (api.MIS_SINGLE, api.FIX_MIXTURE_PDF)], ids=["MIS", "NEE_SPHERICAL", "MIS_SINGLE"])
def test_light_tree_converges_to_the_same_image(scene_name, strategy, flags, gpu_scene_of, arrays_of):
    """MCPT_LIGHT_TREE: picking the light through the hierarchy instead of the O(L) loop changes
    the variance, not the expectation.  Compared with the exact selection on 8x8 block means of long renders and on
    the mean image energy; deterministic and independent of the pass size like every other mode."""
    _, cam = arrays_of(scene_name)
    spp = 4 * SPP[scene_name]
    s = gpu_scene_of(scene_name)
    exact, _ = s.render(cam, strategy, spp, max_depth=DEPTH, flags=flags, seed=11)
    tree, st = s.render(cam, strategy, spp, max_depth=DEPTH, flags=flags | api.LIGHT_TREE, seed=12)
    assert np.isfinite(tree).all() and not np.array_equal(exact, tree)
    H, W = cam.height // 8 * 8, cam.width // 8 * 8
    blk = lambda im: np.minimum(im[:H, :W], 4.0).reshape(H // 8, 8, W // 8, 8, 3).mean(axis=(1, 3))
    be, bt = blk(exact), blk(tree)
    rel = np.linalg.norm(be - bt) / np.linalg.norm(be)
    assert rel < 0.05, rel
    diff = (tree.astype(np.float64) - exact).mean(axis=2).ravel()
    se = diff.std() / np.sqrt(diff.size)
    assert abs(diff.mean()) <= 4 * se + 0.01 * exact.mean(), (diff.mean(), se, exact.mean())
    again, _ = s.render(cam, strategy, spp, max_depth=DEPTH, flags=flags | api.LIGHT_TREE, seed=12)
    assert np.array_equal(tree, again)
    small, _ = s.render(cam, strategy, spp, max_depth=DEPTH, flags=flags | api.LIGHT_TREE, seed=12, batch_spp=5)
    assert np.allclose(tree, small, rtol=1e-5, atol=1e-6)
    assert st["kernel_launches"] > 0 and st["vertices"] > 0
    # with sub-pixel jitter the first vertex is walked through the tree as well (no per-pixel sweep to share)
    jit, _ = s.render(cam, strategy, spp, max_depth=DEPTH, flags=flags | api.LIGHT_TREE | api.JITTER, seed=13)
    jit_exact, _ = s.render(cam, strategy, spp, max_depth=DEPTH, flags=flags | api.JITTER, seed=14)
    bj, bje = blk(jit), blk(jit_exact)
    assert np.isfinite(jit).all() and np.linalg.norm(bj - bje) / np.linalg.norm(bje) < 0.05


# ---- edge cases -----------------------------------------------------------------------------------------------
def _tiny_scene(num_tris, with_light):
    """num_tris unit triangles facing the camera at z = 0 (x = 2k .. 2k+1), the last one emissive if with_light."""
    from monte_carlo_path_tracing_b200.scenes import Camera, SceneArrays
    pos = np.zeros((num_tris, 3, 3), np.float32)
    for k in range(num_tris):
        pos[k] = [[2 * k, 0, 0], [2 * k + 1, 0, 0], [2 * k, 1, 0]]
    nor = np.zeros_like(pos); nor[..., 2] = 1.0
    mat = np.zeros(num_tris, np.int32)
    lights = np.array([num_tris - 1], np.int32) if with_light else np.zeros(0, np.int32)
    if with_light:
        mat[-1] = 1
    cam = Camera(eye=(num_tris, 0.4, 6.0), lookat=(num_tris, 0.4, 0.0), up=(0, 1, 0), fovy_deg=60.0, width=64, height=32)
    return SceneArrays(pos, nor, mat, np.array([0, num_tris], np.uint32), np.array([[0.5, 0.5, 0.5], [0, 0, 0]], np.float32),
                       np.zeros((2, 3), np.float32), np.ones(2, np.float32), lights, np.zeros(len(lights), np.int32),
                       np.array([[3.0, 2.0, 1.0]]) if with_light else np.zeros((0, 3)), cam)


@pytest.mark.parametrize("build", [api.BUILD_HOST_SAH, api.BUILD_GPU_LBVH], ids=["host-sah", "gpu-lbvh"])
@pytest.mark.parametrize("num_tris", [1, 2, 3, 5])
def test_tiny_scenes(num_tris, build):
    """One to five triangles (a lone root leaf gets a wrapper node; the GPU builder falls back to the host below two
    triangles): closest hits agree with the oracle bit for bit, and the emissive triangle is seen with its radiance."""
    a = _tiny_scene(num_tris, True)
    s = api.Scene(a, 0, build=build)
    try:
        assert s.bvh_check() == 0
        o = oracle.OracleLib(a)
        rng = np.random.default_rng(num_tris)
        ro = np.column_stack([rng.uniform(-1, 2 * num_tris + 1, 2000), rng.uniform(-0.5, 1.5, 2000), np.full(2000, 3.0)])
        rd = util.unit(np.column_stack([rng.normal(0, 0.2, 2000), rng.normal(0, 0.2, 2000), np.full(2000, -1.0)]))
        p1, b1 = s.closest(ro, rd, None, False, api.FP64)
        p0, b0 = o.closest(ro, rd, None, 2)
        assert np.array_equal(p1, p0) and np.array_equal(b1, b0) and (p0 >= 0).any()
        for strat in (api.NEE_UNIFORM, api.NEE_SPHERICAL, api.MIS, api.MIS_SINGLE, api.BRDF_ONLY):
            img, st = s.render(a.camera, strat, 4, max_depth=3, seed=1)
            assert np.isfinite(img).all() and st["primary_hits"] > 0
            assert np.isclose(img.max(axis=(0, 1)), [3.0, 2.0, 1.0]).all()  # the light seen directly
    finally:
        s.close()


def test_light_tree_edge_cases():
    """MCPT_LIGHT_TREE on a one-light scene (the root is a leaf: probability 1, same film as the exact selection up to
    rounding of w / (w / 1)), with sub-pixel jitter (tree walks at the first vertex too) and on a strategy that never
    selects a light spherically (flag ignored)."""
    a = _tiny_scene(5, True)
    s = api.Scene(a, 0)
    try:
        cam = a.camera
        for strat in (api.MIS, api.NEE_SPHERICAL, api.MIS_SINGLE):
            exact, _ = s.render(cam, strat, 16, max_depth=4, seed=5, flags=api.FIX_MIXTURE_PDF)
            tree, _ = s.render(cam, strat, 16, max_depth=4, seed=5, flags=api.FIX_MIXTURE_PDF | api.LIGHT_TREE)
            assert np.isfinite(tree).all() and np.allclose(exact, tree, rtol=1e-4, atol=1e-6)
            jit, _ = s.render(cam, strat, 16, max_depth=4, seed=5, flags=api.FIX_MIXTURE_PDF | api.LIGHT_TREE | api.JITTER)
            assert np.isfinite(jit).all() and abs(jit.mean() - exact.mean()) < 0.25 * exact.mean() + 1e-6
        u0, _ = s.render(cam, api.NEE_UNIFORM, 8, max_depth=4, seed=6)
        u1, _ = s.render(cam, api.NEE_UNIFORM, 8, max_depth=4, seed=6, flags=api.LIGHT_TREE)
        assert np.array_equal(u0, u1)
        r = s.light_tree(np.array([[2.0, 0.3, 3.0]]), np.array([[0.0, 0.0, -1.0]]), np.array([0.3]), np.array([0], np.int32))
        assert r["light"][0] == 0 and r["prob"][0] == 1.0 and r["query_prob"][0] == 1.0
    finally:
        s.close()


def test_scene_without_lights():
    """BRDF_ONLY needs no light (black image, paths still traced); every light-sampling strategy refuses with INVALID."""
    a = _tiny_scene(3, False)
    s = api.Scene(a, 0)
    try:
        img, st = s.render(a.camera, api.BRDF_ONLY, 4, max_depth=3, seed=1)
        assert not img.any() and st["primary_hits"] > 0 and st["vertices"] > 0
        for strat in (api.NEE_UNIFORM, api.NEE_SPHERICAL, api.MIS, api.MIS_SINGLE):
            with pytest.raises(api.McptError) as e:
                s.render(a.camera, strat, 1)
            assert e.value.code == -1
    finally:
        s.close()


def test_veach_full_size_sharding_and_determinism():
    """BASELINE.json configs[1] geometry at full resolution (1280x720, shade_with_mis, RR only): deterministic,
    independent of the pass size, and the two halves of an spp-sharded render (what two GPUs do) add up to the whole."""
    sc = scenes.veach_mis(width=1280, height=720)
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        stem = sc.write(td)
        a, _ = api.load_scene_files(stem + ".obj", stem + ".xml")
    s = api.Scene(a, 0)
    try:
        cam = sc.camera
        full, st = s.render(cam, api.MIS, 6, seed=21)
        again, _ = s.render(cam, api.MIS, 6, seed=21)
        assert np.array_equal(full, again)
        small, _ = s.render(cam, api.MIS, 6, seed=21, batch_spp=4)
        assert np.allclose(full, small, rtol=1e-5, atol=1e-6)
        lo, _ = s.render(cam, api.MIS, 3, seed=21, spp_offset=0, spp_total=6)
        hi, _ = s.render(cam, api.MIS, 3, seed=21, spp_offset=3, spp_total=6)
        assert np.allclose(lo + hi, full, rtol=1e-5, atol=1e-6)
        assert st["paths"] == 1280 * 720 * 6 and st["bounces"] > 10 and np.isfinite(full).all() and full.mean() > 0
    finally:
        s.close()


def _read_pfm(path, cam):
    data = path.read_bytes()
    hdr = ("PF\n%d %d\n-1.0\n" % (cam.width, cam.height)).encode()
    assert data.startswith(hdr)
    return np.frombuffer(data[len(hdr):], np.float32).reshape(cam.height, cam.width, 3)[::-1]


@pytest.mark.parametrize("strategy", ["mis", "nee_uniform"])
def test_reference_side_binding_is_a_drop_in(scene_name, strategy, scene_files, tmp_path, arrays_of):
    """oracle/_ref/ref_gpu_main = the reference's own main() objects -- Myobj / Mylight loading the OBJ / MTL / XML with
    tiny_obj_loader and pugixml, unmodified -- with the pixel loop handed to libmcpt.so (oracle/ref_build/ref_gpu_main.cpp,
    quoted in INTEGRATION.md).  Its film equals, bit for bit, what the repo's own driver renders from the same files."""
    import os
    import subprocess
    exe = os.path.join(util.ROOT, "oracle", "_ref", "ref_gpu_main")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/ref_gpu_main was not built (needs /root/reference at build time)")
    drv = os.path.join(util.ROOT, "monte_carlo_path_tracing_b200", "mcpt_render")
    _, stem = scene_files(scene_name)
    _, cam = arrays_of(scene_name)
    a_pfm, b_pfm, b_ppm, a_ppm = tmp_path / "ref.pfm", tmp_path / "own.pfm", tmp_path / "own.ppm", tmp_path / "ref.ppm"
    common = ["--spp", "8", "--strategy", strategy, "--seed", "3", "--max-depth", "6"]
    # bytes, not text: the reference's loaders print GBK-encoded banners (Myobj.cpp:12,26-27)
    r = subprocess.run([exe, stem + ".obj", stem + ".xml", str(a_pfm), "--ppm", str(a_ppm)] + common, capture_output=True, timeout=300)
    assert r.returncode == 0, (r.stdout + r.stderr).decode(errors="replace")
    r = subprocess.run([drv, stem + ".obj", stem + ".xml", str(b_ppm), "--pfm", str(b_pfm), "--every", "8"] + common,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    fa, fb = _read_pfm(a_pfm, cam), _read_pfm(b_pfm, cam)
    assert fa.max() > 0 and np.array_equal(fa, fb)
    # RadianceRGB::tone_mapping (the reference's own, RadianceRGB.cpp:51-67) vs mcpt_tonemap on the device
    hdr = ("P6\n%d %d\n255\n" % (cam.width, cam.height)).encode()
    pa = np.frombuffer(a_ppm.read_bytes()[len(hdr):], np.uint8).astype(int)
    pb = np.frombuffer(b_ppm.read_bytes()[len(hdr):], np.uint8).astype(int)
    # same formula, FP64 pow from two libraries (glibc / CUDA): a value within an ulp of a rounding boundary may land on
    # either side of floor(255 x + 0.5)
    assert np.abs(pa - pb).max() <= 1 and np.mean(pa != pb) < 1e-3


def test_multi_abi_on_one_device_equals_render(scene_name, gpu_scene_of, arrays_of):
    """mcpt_multi_* with a single device is mcpt_render: same film bit for bit, same counters."""
    a, cam = arrays_of(scene_name)
    m = api.MultiScene(a, [0])
    try:
        img_m, st_m = m.render(cam, api.MIS, 6, max_depth=5, seed=8)
        m.upload()
        again, _ = m.render(cam, api.MIS, 6, max_depth=5, seed=8)
    finally:
        m.close()
    img, st = gpu_scene_of(scene_name).render(cam, api.MIS, 6, max_depth=5, seed=8)
    assert np.array_equal(img_m, img) and np.array_equal(again, img)
    for k in ("paths", "vertices", "primary_hits", "rays_light"):
        assert st_m[k] == st[k], k


def test_multi_gpu_render_behind_the_c_abi(scene_name, scene_files, tmp_path, gpu_scene_of, arrays_of):
    """spp sharded over the devices of ONE process inside the library, one ncclReduce of the films.
    Needs >= 2 GPUs (gpurun --gpus 2).  The film differs from the one-device film only by the FP32 order of one sum."""
    import os
    import subprocess
    if api.load().mcpt_device_count() < 2:
        pytest.skip("needs two GPUs")
    a, cam = arrays_of(scene_name)
    one, st1 = gpu_scene_of(scene_name).render(cam, api.MIS, 9, max_depth=5, seed=8)
    m = api.MultiScene(a, [0, 1])
    try:
        two, st2 = m.render(cam, api.MIS, 9, max_depth=5, seed=8)   # 5 + 4 samples
        lone, _ = m.render(cam, api.MIS, 1, max_depth=5, seed=8)    # fewer samples than devices: device 1 idles
    finally:
        m.close()
    assert np.allclose(two, one, rtol=2e-6, atol=1e-7)
    assert st2["paths"] == st1["paths"] and st2["vertices"] == st1["vertices"]
    ref1, _ = gpu_scene_of(scene_name).render(cam, api.MIS, 1, max_depth=5, seed=8)
    assert np.array_equal(lone, ref1)
    # the C++ driver, no Python in the loop
    drv = os.path.join(util.ROOT, "monte_carlo_path_tracing_b200", "mcpt_render")
    _, stem = scene_files(scene_name)
    films = []
    for g in ("1", "2"):
        pfm = tmp_path / ("g%s.pfm" % g)
        r = subprocess.run([drv, stem + ".obj", stem + ".xml", str(tmp_path / "o.ppm"), "--pfm", str(pfm), "--spp", "9", "--every", "9",
                            "--strategy", "mis", "--seed", "8", "--max-depth", "5", "--gpus", g], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        films.append(_read_pfm(pfm, cam))
    assert np.array_equal(films[0], one) and np.allclose(films[1], one, rtol=2e-6, atol=1e-7)


def test_mis_queue_overflow_is_retried_inside_the_library(gpu_scene_of, arrays_of, monkeypatch):
    """shade_with_mis grows a path tree (F6).  When it outgrows its queue (here the first attempt is told its queue ends
    after 1000 entries, MCPT_TEST_OVERFLOW_ABOVE, testing only) the library abandons the pass, halves it, doubles the
    room per path slot and renders again -- the caller sees the same film as without the overflow, bit for bit
    (nothing depends on the pass size)."""
    _, cam = arrays_of("cornell-box")
    s = gpu_scene_of("cornell-box")
    ref, st0 = s.render(cam, api.MIS, 16, max_depth=6, seed=12)
    monkeypatch.setenv("MCPT_TEST_OVERFLOW_ABOVE", "1000")
    img, st = s.render(cam, api.MIS, 16, max_depth=6, seed=12)
    monkeypatch.delenv("MCPT_TEST_OVERFLOW_ABOVE")
    assert np.array_equal(img, ref)
    assert st["vertices"] == st0["vertices"] and st["passes"] > st0["passes"]


def test_single_path_mis_without_flags_is_unbiased(scene_name, gpu_scene_of, arrays_of):
    """MCPT_MIS_SINGLE with flags = 0 (the default of `mcpt_render --strategy mis_single`): the BRDF branch always uses
    the mixture pdf -- single-lobe weights (F8) have no meaning for an estimator the reference does not have -- so the
    film is the same with and without MCPT_FIX_MIXTURE_PDF, and it converges to the corrected path-tree MIS image."""
    _, cam = arrays_of(scene_name)
    s = gpu_scene_of(scene_name)
    a, _ = s.render(cam, api.MIS_SINGLE, 64, max_depth=5, seed=31)
    b, _ = s.render(cam, api.MIS_SINGLE, 64, max_depth=5, seed=31, flags=api.FIX_MIXTURE_PDF)
    assert np.array_equal(a, b)
    tree, _ = s.render(cam, api.MIS, 256, max_depth=5, seed=32, flags=api.FIX_MIXTURE_PDF)
    big, _ = s.render(cam, api.MIS_SINGLE, 256, max_depth=5, seed=33)
    assert abs(float(big.mean()) - float(tree.mean())) <= 0.02 * float(tree.mean())
