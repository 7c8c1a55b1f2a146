"""mcpt_visibility_noalloc (include/mcpt.h): the NEE visibility test with its scratch in a fixed device array inside
the library -- no cudaMalloc / cudaFree in the call.  No counterpart in the reference.
Collected last on purpose.

The new entry point runs the same kernels as mcpt_visibility_batch on the same FP32 rays, so the answers must be EQUAL
(integers: the tolerance the visibility tests use for same-kernel comparisons is exact equality) -- on two calls in a row
with different rays, the second of which finds the first one's contents in the shared scratch.
"""
import numpy as np
import pytest

import util
from monte_carlo_path_tracing_b200 import api

pytestmark = pytest.mark.gpu


def _rays(a, n, seed):
    """FP32 rays from surface points towards random points on random light triangles (what k_shade queues)."""
    rng = np.random.default_rng(seed)
    x, _, _, prim = util.shading_points(a, n, seed + 1)
    lp = a.light_prim[rng.integers(0, len(a.light_prim), n)].astype(np.int32)
    b = rng.random(n); g = rng.random(n) * (1 - b)
    v = a.positions[lp].astype(np.float64)
    tgt = v[:, 0] * (1 - b - g)[:, None] + v[:, 1] * b[:, None] + v[:, 2] * g[:, None]
    ro = x.astype(np.float32)
    rd = util.unit(tgt - ro.astype(np.float64)).astype(np.float32)
    return ro, rd, prim.astype(np.int32), lp


def _device_inputs(torch, ro, rd, prim, lp):
    n = len(ro)
    org = np.zeros((n, 4), np.float32); org[:, :3] = ro
    dr = np.zeros((n, 4), np.float32); dr[:, :3] = rd
    ids = np.stack([prim, lp], axis=1).astype(np.int32)
    dev = torch.device("cuda:0")
    return (torch.from_numpy(org).to(dev).contiguous(), torch.from_numpy(dr).to(dev).contiguous(),
            torch.from_numpy(ids).to(dev).contiguous(), torch.full((n,), -7, dtype=torch.int32, device=dev))


# `which` rather than the session-scoped `scene_name` fixture: pytest regroups tests around parametrised session fixtures,
# which would pull these in between the other files' tests
@pytest.mark.parametrize("leaf_mode", [api.LEAF_FAST, api.LEAF_EXACT], ids=["fast", "exact"])
@pytest.mark.parametrize("which", ["cornell-box", "veach-mis"])
def test_noalloc_visibility_equals_the_allocating_entry_point(which, gpu_scene_of, arrays_of, leaf_mode):
    import torch
    scene_name = which
    a, _ = arrays_of(scene_name)
    s = gpu_scene_of(scene_name)
    stream = torch.cuda.current_stream().cuda_stream
    seen = set()
    # two calls in a row, different sizes and rays: the second call finds the first one's queue, accumulators and counters
    for call, (n, seed) in enumerate(((20000, 11), (7001, 12))):
        ro, rd, prim, lp = _rays(a, n, seed)
        expect = s.visibility(ro.astype(np.float64), rd.astype(np.float64), prim, lp, leaf_mode)  # mcpt_visibility_batch
        org, dr, ids, vis = _device_inputs(torch, ro, rd, prim, lp)
        free_before = torch.cuda.mem_get_info()[0]
        s.visibility_noalloc(org.data_ptr(), dr.data_ptr(), ids.data_ptr(), n, vis.data_ptr(), leaf_mode, stream)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before or call == 0  # the first call may load the module lazily
        got = vis.cpu().numpy()
        assert np.array_equal(got, expect), (call, int((got != expect).sum()))
        seen |= set(np.unique(got).tolist())
    assert {1, 0} <= seen  # visible and occluded rays both occur


def test_noalloc_limit_and_empty_batch(gpu_scene_of, arrays_of):
    import torch
    a, _ = arrays_of("cornell-box")
    s = gpu_scene_of("cornell-box")
    n = api.NOALLOC_MAX_RAYS
    ro, rd, prim, lp = _rays(a, n, 21)
    org, dr, ids, vis = _device_inputs(torch, ro, rd, prim, lp)
    # exactly the maximum works and equals the allocating entry point
    s.visibility_noalloc(org.data_ptr(), dr.data_ptr(), ids.data_ptr(), n, vis.data_ptr(), api.LEAF_FAST, 0)
    torch.cuda.synchronize()
    assert np.array_equal(vis.cpu().numpy(), s.visibility(ro.astype(np.float64), rd.astype(np.float64), prim, lp, api.LEAF_FAST))
    # one more is refused with a clear error, and nothing is launched
    vis.fill_(-7)
    with pytest.raises(api.McptError) as e:
        s.visibility_noalloc(org.data_ptr(), dr.data_ptr(), ids.data_ptr(), n + 1, vis.data_ptr(), api.LEAF_FAST, 0)
    assert "MCPT_NOALLOC_MAX_RAYS" in str(e.value) and str(n) in str(e.value)
    torch.cuda.synchronize()
    assert int((vis != -7).sum()) == 0
    s.visibility_noalloc(0, 0, 0, 0, 0, api.LEAF_FAST, 0)  # empty batch: a no-op, null pointers allowed
    with pytest.raises(api.McptError):
        s.visibility_noalloc(org.data_ptr(), dr.data_ptr(), ids.data_ptr(), 16, vis.data_ptr(), 7, 0)  # unknown leaf_mode
