#!/usr/bin/env python
"""bench.py -- headline benchmark of the path-tracing hot path (BASELINE.json).

  python bench.py --gpus N --steps K --warmup W            our CUDA path (one rank per GPU under torchrun)
  python bench.py --impl reference --gpus N --steps K ...  the reference's own CPU code (oracle/_ref) on the
                                                           host cores, same workload/metric

Workload (BASELINE.json configs[1]): veach-mis look-alike, 1280x720, shade_with_mis (spherical-triangle +
BRDF sampling, balance heuristic), no depth cap (the reference has none; Russian roulette p = 0.6 ends
paths).  One "step" = `--spp-per-step` samples of every pixel on every rank (weak scaling: per-GPU work is
fixed) followed, for N > 1, by the NCCL reduce of the W*H*3 float film onto rank 0.
metric = primary Msamples/s = W*H*spp_per_step*N / step time.  Inputs (scene, BVH) are resident in HBM
for `value`; `e2e` times the host-buffer C ABI per step: scene upload (pinned H2D) + render + film D2H -- mcpt_scene_upload
+ mcpt_render at N = 1, and at N > 1 mcpt_multi_upload + mcpt_multi_render called by rank 0 alone (one process, spp sharded
over the N devices INSIDE the library, one ncclReduce; the other ranks have released their GPUs and wait).
`extra` carries, from the same run: cornell-box config 1 exactly (512x512, 16 spp, depth 5) and at 512 spp, veach-mis
with MCPT_JITTER, the FAST-vs-EXACT hit-id mismatch rate, and config 3 strong-scaled (interior, 4096 spp / N).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from monte_carlo_path_tracing_b200 import scenes  # noqa: E402

WORKLOADS = {
    # name: (scene generator kwargs, strategy name, max_depth)
    "veach-mis": dict(scene="veach-mis", kw=dict(width=1280, height=720), strategy="MIS", max_depth=0, spp=128),
    "cornell-box": dict(scene="cornell-box", kw=dict(width=512, height=512), strategy="NEE_UNIFORM", max_depth=5, spp=512),
    # configs[2]: stand-in for the unavailable bathroom scene: generated interior, ~185 k triangles
    "interior": dict(scene="interior", kw=dict(width=1280, height=720), strategy="MIS", max_depth=0, spp=128),
    # configs[3]: 1 M random triangles + 64 emissive triangles, 1024x1024, uniform-area NEE (BVH traversal stress)
    "soup-1m": dict(scene="soup", kw=dict(num_tris=1_000_000, num_lights=64, width=1024, height=1024), strategy="NEE_UNIFORM",
                    max_depth=5, spp=128),
}
STRAT = {"NEE_SPHERICAL": 0, "BRDF_ONLY": 1, "MIS": 2, "NEE_UNIFORM": 3}
# Compulsory wavefront state traffic per sample-bounce of the split pipeline: k_extend reads the
# ray (32 B) and writes the hit (16); k_shade reads ray + throughput (48, +8 tree node for MIS) + hit (16) and stages
# ~1 child (48, +8); k_gather moves it (2 x 48); the slot accumulator is read and written (32); NEE adds one visibility
# record written and read (2 x 56); MIS stages / moves / extends a second child and writes its contribution word.
ALGO_BYTES_PER_VERTEX = {"MIS": 576, "NEE_UNIFORM": 400, "NEE_SPHERICAL": 400, "BRDF_ONLY": 288}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)), "measured"
        except Exception:
            pass
    return {"hbm_gbs": 6650.0}, "fallback"


# ---- clocks during the timed region ------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.stop_flag, self.gpu = [], False, gpu_index
        self.thread = threading.Thread(target=self._run, daemon=True)
        # NVML in-process (nvidia_ml_py).  nvmlInit / the first query take the driver lock for tens of ms, so they
        # happen HERE, before the timed region; inside it one query costs microseconds.  (Spawning nvidia-smi every
        # 200 ms inside the timed region measurably slowed the host-driven bounce loop.)
        self.nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.gpu)
            self.nv = (pynvml, h, pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            self._query()
            self.rows.clear()
        except Exception:
            self.nv = None

    def _query(self):
        pynvml, h, mx = self.nv
        bits = [(pynvml.nvmlClocksEventReasonHwSlowdown, 0), (pynvml.nvmlClocksEventReasonHwThermalSlowdown, 1),
                (pynvml.nvmlClocksEventReasonSwThermalSlowdown, 2), (pynvml.nvmlClocksEventReasonSwPowerCap, 3)]
        sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
        try:
            r = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)
        except Exception:
            r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
        row = [str(sm), str(mx)] + ["Not Active"] * 4
        for bit, k in bits:
            if r & bit:
                row[2 + k] = "Active"
        self.rows.append(row)

    def _run(self):
        if self.nv is not None:
            try:
                while not self.stop_flag:
                    self._query()
                    time.sleep(0.05)
                return
            except Exception:
                pass
        while not self.stop_flag:
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([x.strip() for x in o.splitlines()[0].split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def start(self):
        self.thread.start()

    def stop(self):
        self.stop_flag = True
        self.thread.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for r in self.rows for k in range(4) if len(r) > 2 + k and r[2 + k].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---- the reference's CPU code on the host cores ----------------------------------------------------------------
def _cpu_worker(args):
    """One process = one copy of the reference (it is not thread-safe: global caches and Mylight scratch,
    SURVEY.md 8b).  Renders interleaved rows of the workload at 1 spp until the deadline; returns samples done."""
    stem, cam_d, strategy, max_depth, wid, nw, budget_s, seed, use_ref = args
    sys.path.insert(0, ROOT)
    import oracle
    from monte_carlo_path_tracing_b200.scenes import Camera
    cam = Camera(**cam_d)
    if isinstance(stem, dict):  # array scene (triangle soup): no OBJ files, so the C restatement renders it
        from monte_carlo_path_tracing_b200 import scenes as _sc
        lib = oracle.OracleLib(_sc.triangle_soup(**stem))
        lib.build_grid(cam.eye, 100000)
        render = lambda r0, r1, k: lib.render(cam, STRAT[strategy], 1, max_depth=max_depth, seed=seed, rows=(r0, r1), spp_offset=k, spp_total=1)
    elif use_ref:
        lib = oracle.RefLib(stem + ".obj", stem + ".xml")
        lib.build_grid(cam.eye, 100000)
        strat = {"NEE_SPHERICAL": 0, "BRDF_ONLY": 1, "MIS": 2}[strategy]
        render = lambda r0, r1, k: lib.render(cam, strat, 1, max_depth=max_depth, seed=seed, rows=(r0, r1), spp_offset=k, spp_total=1)
    else:
        from monte_carlo_path_tracing_b200 import api
        arrays, _ = api.load_scene_files(stem + ".obj", stem + ".xml")
        lib = oracle.OracleLib(arrays)
        lib.build_grid(cam.eye, 100000)
        render = lambda r0, r1, k: lib.render(cam, STRAT[strategy], 1, max_depth=max_depth, seed=seed, rows=(r0, r1), spp_offset=k, spp_total=1)
    t0 = time.time()
    samples = vertices = rays = 0
    k = 0
    rows = list(range(wid, cam.height, nw))
    # spread the rows this worker visits over the whole frame
    rng = np.random.default_rng(1234 + wid)
    rng.shuffle(rows)
    while True:
        for r in rows:
            _, cnt = render(r, r + 1, k)
            samples += cam.width
            vertices += cnt["vertices"]
            rays += cnt.get("rays", 0) + cnt.get("light_rays", 0)
            if time.time() - t0 >= budget_s:
                return samples, vertices, time.time() - t0, rays
        k += 1


def cpu_reference_run(stem, cam, strategy, max_depth, budget_s, seed=1, procs=None, row_stride=None):
    """All host cores (or `procs`), one reference process each, for ~budget_s seconds.  Returns dict(value Msamples/s, ...).
    Process w renders rows w, w + row_stride, ... (row_stride defaults to the process count)."""
    import multiprocessing as mp
    import oracle
    use_ref = oracle.ref_available() and strategy in ("NEE_SPHERICAL", "BRDF_ONLY", "MIS") and not isinstance(stem, dict)
    if not os.path.exists(oracle.ORACLE_SO):
        oracle.build(ref=False)
    nproc = procs or os.cpu_count() or 1
    cam_d = dict(eye=cam.eye, lookat=cam.lookat, up=cam.up, fovy_deg=cam.fovy_deg, width=cam.width, height=cam.height)
    ctx = mp.get_context("spawn")
    t0 = time.time()
    with ctx.Pool(nproc) as pool:
        res = pool.map(_cpu_worker, [(stem, cam_d, strategy, max_depth, w, row_stride or nproc, budget_s, seed, use_ref) for w in range(nproc)])
    wall = time.time() - t0
    samples = sum(r[0] for r in res)
    vertices = sum(r[1] for r in res)
    busy = max(r[2] for r in res)
    rays = sum(r[3] for r in res)
    return {"value": samples / busy / 1e6, "unit": "Msamples/s", "cores": nproc, "kind": "reference" if use_ref else "port",
            "sample": "%d primary samples (interleaved full-width rows at 1 spp, %.1f s on %d processes; wall incl. load %.1f s)" % (samples, busy, nproc, wall),
            "sample_bounces_per_s": vertices / busy, "Mrays_per_s": rays / busy / 1e6,
            "rays_counted": "closet_ray_intersect + closet_ray_intersect_light_triangle calls (linker --wrap shims in oracle/ref_build/ref_driver.cpp)",
            "samples": samples, "seconds": busy}


# ---- arms -------------------------------------------------------------------------------------------------------
def write_scene(wl, td):
    """(scene object with .camera, stem of the OBJ/XML files) -- for the array scene the 'stem' is the generator kwargs."""
    if wl["scene"] == "soup":
        return scenes.triangle_soup(**wl["kw"]), dict(wl["kw"])
    sc = scenes.SCENES[wl["scene"]](**wl["kw"])
    return sc, sc.write(td)


def load_arrays(sc, stem):
    from monte_carlo_path_tracing_b200 import api
    if isinstance(stem, dict):
        return sc
    return api.load_scene_files(stem + ".obj", stem + ".xml")[0]


def base_line(args, wl_name, wl, cam, spp_per_step):
    return {"metric": "primary Msamples/s", "unit": "Msamples/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": "%s %dx%d %s max_depth=%s spp_per_step=%d per GPU" % (
                wl_name, cam.width, cam.height, wl["strategy"], wl["max_depth"] or "none(RR 0.6)", spp_per_step),
                "scene": wl["scene"], "width": cam.width, "height": cam.height, "strategy": wl["strategy"],
                "max_depth": wl["max_depth"], "spp_per_step": spp_per_step, "parallelism": "spp-shard x%d + 1 NCCL reduce" % args.gpus,
                "l2_policy": "per-step working set (ray queues + film, >400 MB) exceeds the 126 MB L2"}}


def run_reference(args, wl_name, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # the CPU reference runs once, on rank 0's host cores
    with tempfile.TemporaryDirectory() as td:
        sc, stem = write_scene(wl, td)
        cam = sc.camera
        for _ in range(min(args.warmup, 1)):
            cpu_reference_run(stem, cam, wl["strategy"], wl["max_depth"], 1.0)
        per = max(2.0, min(20.0, 60.0 / max(args.steps, 1)))
        runs = [cpu_reference_run(stem, cam, wl["strategy"], wl["max_depth"], per, seed=1 + k) for k in range(args.steps)]
    samples = sum(r["samples"] for r in runs)
    secs = sum(r["seconds"] for r in runs)
    v = samples / secs / 1e6
    line = base_line(args, wl_name, wl, cam, args.spp_per_step)
    line.update({"impl": "reference", "value": v, "ms_per_step": 1e3 * secs / max(args.steps, 1), "n_gpus": args.gpus,
                 "dtype": "f64", "gpu_launches": 0,
                 "cpu_baseline": {"value": v, "unit": "Msamples/s", "cores": runs[0]["cores"], "kind": runs[0]["kind"],
                                  "sample": "%d steps x %s" % (args.steps, runs[0]["sample"])},
                 "e2e": {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    print(json.dumps(line), flush=True)


def shade_flop_constants():
    """FP32 flops per unit of work of the shading kernels, from the committed ncu capture (profiles/r3b_flops.json:
    thread-level SASS instruction counts by opcode from `ncu --set full --import-source on`, FFMA = 2 flops, FFMA2 = 4,
    FADD / FMUL = 1, FADD2 / FMUL2 = 2, MUFU = 1, divided by the units the captured launch processed).  Falls back to
    the hand count of lights.cuh (81 flops + 4 MUFU + ~5 min/max per light and vertex; ~320 per shaded vertex)."""
    p = os.path.join(ROOT, "profiles", "r3b_flops.json")
    try:
        j = json.load(open(p))
        return float(j["flops_per_light_vertex"]), float(j["flops_per_shade_element"]), "ncu: profiles/r3b_flops.json"
    except Exception:
        return 90.0, 320.0, "hand count (lights.cuh)"


def extra_measurements(args, api, torch, rank, world, local, dist):
    """Side measurements the verdict asks for in the same driver run.  Each is a short, separately timed render; none of
    them touches the headline numbers.  Returns a dict (rank 0) or None."""
    out = {}
    with tempfile.TemporaryDirectory() as td:
        if world == 1:
            # cornell-box, BASELINE.json configs[0]: 512x512, 16 spp, depth 5 -- exactly, and at 512 spp per step
            sc = scenes.cornell_box(width=512, height=512)
            a = api.load_scene_files(*[sc.write(td) + e for e in (".obj", ".xml")])[0]
            s = api.Scene(a, local)
            film = torch.zeros(512 * 512 * 3, dtype=torch.float32, device="cuda")
            for spp, reps in ((16, 20), (512, 3)):
                for k in range(3):
                    s.render_device(sc.camera, api.NEE_UNIFORM, spp, film.data_ptr(), max_depth=5, seed=args.seed, spp_offset=k * spp)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                v = 0
                for k in range(reps):
                    v += s.render_device(sc.camera, api.NEE_UNIFORM, spp, film.data_ptr(), max_depth=5, seed=args.seed, spp_offset=(3 + k) * spp)["vertices"]
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1)
                out["cornell_box_512x512_%dspp_depth5" % spp] = {"Gsample_bounces_per_s": v / ms / 1e6, "Msamples_per_s": 512 * 512 * spp * reps / ms / 1e3,
                                                                "ms_per_render": ms / reps, "renders": reps}
            # FAST vs EXACT leaf test: hit-id mismatch rate on 1 M rays leaving random surface points (cornell-box)
            rng = np.random.default_rng(5)
            n = 1_000_000
            prim = rng.integers(0, len(a.material_id), n).astype(np.int32)
            b = rng.random(n); g = rng.random(n) * (1 - b)
            v3 = a.positions[prim].astype(np.float64)
            ro = v3[:, 0] * (1 - b - g)[:, None] + v3[:, 1] * b[:, None] + v3[:, 2] * g[:, None]
            rd = rng.normal(size=(n, 3)); rd /= np.linalg.norm(rd, axis=1, keepdims=True)
            pf, _ = s.closest(ro, rd, prim, False, api.FP32)
            px, _ = s.closest(ro, rd, prim, False, api.FP64)
            out["fast_vs_exact_hit_id_mismatch"] = {"rate": float(np.mean(pf != px)), "rays": n, "scene": "cornell-box",
                                                    "what": "k_extend (FP32 Moeller-Trumbore, 4-wide BVH) vs the FP64 Cramer test of Myobj.cpp:165-192"}
            s.close()
            del film
    return out


def run_ours(args, wl_name, wl):
    import torch
    import torch.distributed as dist
    from monte_carlo_path_tracing_b200 import api
    from monte_carlo_path_tracing_b200 import dist as mdist
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    strategy = STRAT[wl["strategy"]]
    S = args.spp_per_step
    td = tempfile.mkdtemp(prefix="mcpt_bench_")
    sc, stem = write_scene(wl, td)
    cam = sc.camera
    arrays = load_arrays(sc, stem)
    scene = api.Scene(arrays, local, build=api.BUILD_GPU_LBVH if args.bvh == "gpu" else api.BUILD_HOST_SAH)
    bvh_extra = dict(scene.build_info(), sah=scene.bvh_info()["sah"], nodes=scene.bvh_info()["nodes"])
    npix = cam.width * cam.height
    film = torch.zeros(npix * 3, dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    kw = dict(max_depth=wl["max_depth"], seed=args.seed, batch_spp=args.batch_spp)
    if args.light_tree:
        kw["flags"] = api.LIGHT_TREE

    def step(k):
        # global sample range of this step: [k*world*S, (k+1)*world*S), rank r owns S of them
        st = scene.render_device(cam, strategy, S, film.data_ptr(), stream=stream, spp_offset=(k * world + rank) * S,
                                 spp_total=S * world, **kw)
        mdist.reduce_film(film, 0)
        return st

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # A second, host-side group: while rank 0 drives all N devices through mcpt_multi_* (the e2e leg) the other ranks must
    # wait on the CPU.  An NCCL barrier is a kernel that spins on the waiting rank's GPU, and a GPU shared by two processes is
    # time-sliced between them: measured at N = 2, it halved the rate of the device rank 0 was rendering on.
    cpu_group = dist.new_group(backend="gloo") if world > 1 else None

    def host_barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(group=cpu_group)

    for k in range(args.warmup):
        step(k)
    sync()
    bvh_extra["trace_layout"] = scene.trace_layout()  # settled on the first real bounce of the warm-up
    bvh_extra["light_lists"] = scene.light_lists()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stats = []
    sync()
    e0.record()
    host_t = [time.perf_counter()]
    for k in range(args.steps):
        stats.append(step(args.warmup + k))
        host_t.append(time.perf_counter())
    e1.record()
    sync()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    clocks = sampler.stop() if sampler else None

    # ---- the same workload with MCPT_JITTER (extension of F4): no per-pixel sharing of the primary hit and of the first
    # vertex's light sweep -- the number the no-jitter headline should be read beside
    side = {}
    if wl_name == "veach-mis" and not args.no_extras:
        jkw = dict(kw, flags=kw.get("flags", 0) | api.JITTER)
        scene.render_device(cam, strategy, S, film.data_ptr(), stream=stream, spp_offset=0, spp_total=S * world, **jkw)
        sync()
        j0, j1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        j0.record()
        jst = scene.render_device(cam, strategy, S, film.data_ptr(), stream=stream, spp_offset=(world + rank) * S, spp_total=S * world, **jkw)
        mdist.reduce_film(film, 0)
        j1.record()
        sync()
        jms = torch.tensor([j0.elapsed_time(j1)], device="cuda")
        if world > 1:
            dist.all_reduce(jms, op=dist.ReduceOp.MAX)
        side["veach_mis_with_MCPT_JITTER"] = {"Msamples_per_s": npix * S * world / float(jms.item()) / 1e3, "ms_per_step": float(jms.item()),
                                              "sample_bounces_per_path": jst["vertices"] / max(jst["paths"], 1)}

    # ---- e2e through the host-buffer C ABI: scene upload (pinned H2D) + render + film D2H, every step
    h2d_bytes = scene.device_bytes() * world
    pinned = torch.empty(npix * 3, dtype=torch.float32).pin_memory()
    pinned_np = pinned.numpy()  # the host film buffer the ABI call fills
    e2e_how = "mcpt_scene_upload + mcpt_render"
    if world == 1:
        scene.upload(); scene.render(cam, strategy, min(S, 2), **kw)
        sync()
        t0 = time.perf_counter()
        for k in range(args.steps):
            scene.upload()  # pinned H2D of the whole scene blob (geometry, BVHs, lights, materials)
            scene.render(cam, strategy, S, spp_offset=k * S, spp_total=S, out=pinned_np, **kw)  # film D2H inside
        e2e_s = time.perf_counter() - t0
    else:
        # N GPUs behind the C ABI (mcpt_multi_*): ONE process drives all devices -- rank 0 calls, the others have released
        # their GPUs (scene + 80 GB workspace) and wait at the barrier.  spp sharded inside the library, one ncclReduce.
        e2e_how = "mcpt_multi_upload + mcpt_multi_render (rank 0 drives all %d devices)" % world
        scene.close()
        scene = None
        del film
        torch.cuda.empty_cache()
        host_barrier()
        e2e_s = 0.0
        if rank == 0:
            multi = api.MultiScene(arrays, list(range(world)), build=api.BUILD_GPU_LBVH if args.bvh == "gpu" else api.BUILD_HOST_SAH)
            multi.upload(); multi.render(cam, strategy, min(S, 2) * world, **kw)
            multi.render(cam, strategy, S * world, out=pinned_np, **kw)
            t0 = time.perf_counter()
            for k in range(args.steps):
                multi.upload()
                multi.render(cam, strategy, S * world, spp_offset=k * S * world, spp_total=S * world, out=pinned_np, **kw)
            e2e_s = time.perf_counter() - t0
            multi.close()
        host_barrier()

    # ---- config 3 strong-scaled (BASELINE.json configs[2]): interior stand-in, 1280x720, 4096 spp split over the N GPUs
    if not args.no_extras and wl_name == "veach-mis":
        with tempfile.TemporaryDirectory() as td3:
            sc3 = scenes.interior(width=1280, height=720)
            a3 = api.load_scene_files(*[sc3.write(td3) + e for e in (".obj", ".xml")])[0]
        s3 = api.Scene(a3, local)
        film3 = torch.zeros(npix * 3, dtype=torch.float32, device="cuda")
        off3, cnt3 = mdist.shard_spp(4096, world, rank)
        s3.render_device(sc3.camera, api.MIS, 8, film3.data_ptr(), stream=stream, seed=args.seed)
        sync()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        s3.render_device(sc3.camera, api.MIS, cnt3, film3.data_ptr(), stream=stream, seed=args.seed, spp_offset=off3, spp_total=4096)
        mdist.reduce_film(film3, 0)
        c1.record()
        sync()
        cms = torch.tensor([c0.elapsed_time(c1)], device="cuda")
        if world > 1:
            dist.all_reduce(cms, op=dist.ReduceOp.MAX)
        side["config3_interior_1280x720_4096spp_strong"] = {"seconds": float(cms.item()) / 1e3, "spp_per_gpu": cnt3, "n_gpus": world,
                                                            "Msamples_per_s": npix * 4096 / float(cms.item()) / 1e3, "triangles": len(a3.material_id),
                                                            "scaling": "strong: total work fixed, one NCCL reduce at the end"}
        s3.close()
        del film3
    if not args.no_extras and rank == 0:
        side.update(extra_measurements(args, api, torch, rank, world, local, dist) or {})

    tot = {k: sum(s[k] for s in stats) for k in ("paths", "vertices", "rays_closest", "rays_shadow", "rays_light", "kernel_launches", "bounces")}
    ms_bounce = sum(s["ms_bounce"] for s in stats)
    if rank == 0:
        samples = npix * S * world * args.steps
        value = samples / ms / 1e3
        peaks, which = measured_peaks()
        algo_b = ALGO_BYTES_PER_VERTEX[wl["strategy"]]
        launches_bounce = max(tot["bounces"], 1)
        # Dominant kernel of the step, timed live with CUDA events on the launch stream inside mcpt_render_device
        # (McptStats.ms_shade / ms_trace): k_shade when the light sweeps dominate, else the trace kernels.
        # Algorithmic bytes per launch = per-record constants x the record counts the library reports.
        tree = wl["strategy"] == "MIS"
        nee = wl["strategy"] in ("NEE_UNIFORM", "NEE_SPHERICAL")
        ms_shade = sum(s["ms_shade"] for s in stats)
        ms_trace = sum(s["ms_trace"] for s in stats)
        elements = sum(s["elements"] for s in stats)
        passes_n = sum(s["passes"] for s in stats)
        children = max(0, tot["rays_closest"] - npix * passes_n)   # rays k_extend traced = entries staged and gathered
        rec = 56 if tree else 48
        sph = wl["strategy"] in ("MIS", "NEE_SPHERICAL")
        later_el = max(0, elements - npix * passes_n)              # elements of the later bounces
        # k_shade: ray + hit in, slot accumulator read-modify-write, children staged, visibility records out;
        # k_lightsel (spherical strategies, later bounces): ray direction + hit in, selection out, read again by k_shade
        shade_bytes = elements * (rec + 16 + 16) + children * rec + tot["rays_shadow"] * 56 + (later_el * (32 + 16 + 16) if sph else 0)
        trace_bytes = children * (32 + 16 + 2 * rec) + tot["rays_shadow"] * (56 + 32) + (elements * 32 if tree else 0)
        if ms_shade >= ms_trace:
            dom = ("k_lightsel + k_shade<%s>" if sph else "k_shade<%s>") % wl["strategy"]
            dom_ms, dom_bytes, dom_launches = ms_shade, shade_bytes, launches_bounce
        else:
            dom, dom_ms, dom_bytes = "trace kernels k_gather + k_extend + k_shadow", ms_trace, trace_bytes
            dom_launches = launches_bounce
        achieved = dom_bytes / (dom_ms * 1e-3) / 1e9
        # dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel(s) of ONE bounce, from the committed
        # `ncu --set full` capture (profiles/ncu_traffic.json); compare with `algorithmic_bytes_captured_launch` there
        traffic, traffic_detail = None, None
        try:
            t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))[wl_name]
            if t.get("kernel") == dom:
                traffic, traffic_detail = t["dram_bytes_per_launch"], t
                if t.get("dram_bytes_per_element"):   # captured launch -> average launch of this run
                    traffic = t["dram_bytes_per_element"] * elements / dom_launches
        except Exception:
            pass
        L = len(arrays.light_prim)
        # FP32 work actually executed (estimate): ~320 flops of shading per vertex, plus ~90 flops per light per sweep of
        # the light table; without jitter the first vertex of a pixel is swept once per pass (k_lightprep), later
        # vertices once each -- for MIS only the 60 % that survive the roulette taken before sampling (main.cpp:429-437)
        passes = sum(s["passes"] for s in stats)
        later = max(0, tot["vertices"] - sum(s["primary_hits"] for s in stats))
        sweeps = (npix * passes + later * (0.6 if wl["strategy"] == "MIS" else 1.0)) if (sph and not args.light_tree) else 0
        # per light and vertex (lights.cuh light_weight_pair): 10 FADD + 17 FMUL + 27 FFMA (eight dot products, three norms, the
        # Van Oosterom-Strackee denominator, the 8-term atan polynomial) = 81 flops + 4 MUFU + ~5 min/max/select
        f_light, f_shade, f_src = shade_flop_constants()
        counted_sweeps = sum(s.get("light_sweeps", 0) for s in stats)   # McptStats.light_sweeps: exact, counted on the device
        if counted_sweeps and sph and not args.light_tree:
            sweeps = counted_sweeps
        Lpad = (L + 63) // 64 * 64
        # (light, vertex) weights the sweeps evaluated: counted on the device (McptStats.light_evals) -- the padded table per
        # sweep, or the per-primitive survivor lists (csrc/light_lists.h) where the scene has them
        light_evals = sum(s.get("light_evals", 0) for s in stats) if (sph and not args.light_tree) else 0
        if not light_evals:
            light_evals = Lpad * sweeps
        flops_sweep = f_light * light_evals
        flops = f_shade * elements + flops_sweep
        flops_per_vertex = flops / max(tot["vertices"], 1)
        try:
            fp32_scalar, fp32_packed = api.measure_fp32(local)
            fp32_peak, fp32_src = max(fp32_scalar, fp32_packed) * 1e12, "measured: mcpt_measure_fp32 (register-resident FFMA2 chains) %.1f packed / %.1f scalar TFLOP/s" % (fp32_packed, fp32_scalar)
        except Exception:
            fp32_peak, fp32_src = 148 * 128 * 2 * 1.965e9, "computed: 148 SMs x 128 lanes x 2 x 1.965 GHz"
        line = base_line(args, wl_name, wl, cam, S)
        if args.light_tree:
            line["config"]["light_selection"] = "MCPT_LIGHT_TREE: O(log L) hierarchy instead of the reference's O(L) loop (not a reference estimator)"
            line["config"]["workload"] += " light-tree"
        line.update({
            "value": value, "ms_per_step": ms / args.steps,
            "gpu_launches": int(tot["kernel_launches"]),
            "clocks": clocks,
            "e2e": {"value": samples / e2e_s / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": npix * 12, "through": e2e_how},
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": achieved / peaks["hbm_gbs"], "traffic": traffic, "traffic_detail": traffic_detail, "peak_source": which,
                         "algorithmic_bytes_per_launch": dom_bytes / dom_launches, "avg_launch_ms": dom_ms / dom_launches,
                         "launches": dom_launches, "kernel_share_of_step": dom_ms / ms,
                         "pipeline": {"ms_shade": ms_shade, "ms_trace": ms_trace, "ms_bounce": ms_bounce, "share_of_step": ms_bounce / ms,
                                      "algorithmic_bytes_per_sample_bounce": algo_b,
                                      "achieved_GBps": algo_b * tot["vertices"] / (ms_bounce * 1e-3) / 1e9},
                         "note": "latency/issue-bound by design (SURVEY.md 8d): HBM fraction shown because the contract asks; "
                                 "the fp32 object is the bound that applies to the light-sampling strategies"},
            "fp32": {"achieved_tflops": flops / (ms * 1e-3) / 1e12, "peak_tflops": fp32_peak / 1e12,
                     "frac": flops / (ms * 1e-3) / fp32_peak, "executed_flops_per_sample_bounce": flops_per_vertex, "lights": L,
                     "light_table_sweeps": sweeps, "light_weight_evaluations": light_evals,
                     "evaluations_per_sweep": light_evals / max(sweeps, 1), "flops_per_light_vertex": f_light, "flops_per_shade_element": f_shade,
                     "flop_constants": f_src, "peak_source": fp32_src, "note": "rank 0, whole step; sweeps counted on the device (McptStats.light_sweeps)"},
            "extra": {"Msample_bounces_per_s": tot["vertices"] * world / ms / 1e3,
                      "Mrays_per_s": (tot["rays_closest"] + tot["rays_shadow"] + tot["rays_light"]) * world / ms / 1e3,
                      "sample_bounces_per_path": tot["vertices"] / max(tot["paths"], 1),
                      "time_to_1024spp_s": 1024.0 / (S * world) * (ms / args.steps) / 1e3,
                      "bvh": bvh_extra, "device_ms_per_step": [round(float(x["ms_total"]), 3) for x in stats],
                      "host_ms_per_step": [round(1e3 * (b - a), 3) for a, b in zip(host_t[:-1], host_t[1:])]},
        })
        if sph and not args.light_tree and ms_shade >= ms_trace:
            # the dominant kernels are the light sweeps + shading: FP32-issue-bound (ncu: FMA pipe 66 % busy, 252 GB/s of DRAM), so the
            # roofline that applies is FP32, against the MEASURED multiply-add ceiling; the HBM reading is kept beside it
            hbm = dict(line["roofline"])
            # The dominant kernels are the light sweeps: k_lightsel_sorted (k_lightsel without survivor lists) at the later vertices
            # and k_lightprep at the first one.  Both are timed alone with CUDA events on the launch stream inside the library
            # (McptStats.ms_sweep, ms_prep); their flops = evaluations counted on the device x the per-evaluation constant of the
            # committed ncu capture, which was derived from exactly these kernels.
            ms_sweep = sum(s.get("ms_sweep", 0.0) for s in stats)
            ms_prep = sum(s.get("ms_prep", 0.0) for s in stats)
            sweep_launches = sum(s["bounces"] for s in stats)  # one sweep launch per bounce after the first, one k_lightprep per pass
            dom_ms2 = ms_sweep + ms_prep
            ach = flops_sweep / (dom_ms2 * 1e-3) / 1e12
            # the same selection without the per-primitive survivor lists evaluates the whole padded table at every sweep:
            # what the executed work stands for in the reference's terms (Mylight.cpp:169-259 loops over all L lights)
            full_table = f_light * Lpad * sweeps
            group_ms = ms_shade + ms_prep
            line["roofline"] = {"bound": "fp32", "kernel": "light sweeps: k_lightsel_sorted + k_lightprep", "achieved": ach, "peak": fp32_peak / 1e12,
                                "unit": "TFLOP/s", "frac": ach * 1e12 / fp32_peak, "traffic": traffic, "peak_source": fp32_src,
                                "algorithmic_flops_per_launch": flops_sweep / max(sweep_launches, 1), "avg_launch_ms": dom_ms2 / max(sweep_launches, 1),
                                "launches": sweep_launches, "kernel_share_of_step": dom_ms2 / ms, "flop_constants": f_src,
                                "ms_sweep": ms_sweep, "ms_prep": ms_prep,
                                "full_table_equivalent": {"tflops": full_table / (dom_ms2 * 1e-3) / 1e12, "frac": full_table / (dom_ms2 * 1e-3) / fp32_peak,
                                                          "note": "flops the same sweeps would execute over the whole light table (no survivor lists): "
                                                                  "the lists skip lights whose weight is exactly zero"},
                                "shade_group": {"kernels": dom + " + k_lightprep (selection incl. its sort, sweeps, shading)", "achieved": flops / (group_ms * 1e-3) / 1e12,
                                                "frac": flops / (group_ms * 1e-3) / fp32_peak, "ms": group_ms, "share_of_step": group_ms / ms},
                                "hbm": hbm,
                                "note": "bound = fp32: executed FP32 flops of the light-sweep kernels over their own device-timed duration (CUDA events "
                                        "around each launch), against the measured FFMA ceiling of this GPU; `shade_group` = the whole selection + shading "
                                        "side of the bounce; `hbm` holds the byte-roofline view"}
        line["extra"].update(side)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_reference_run(stem, cam, wl["strategy"], wl["max_depth"], args.cpu_seconds)
            # the reference as designed is one thread (README.md:418): same rows, one process, a quarter of the budget
            one = cpu_reference_run(stem, cam, wl["strategy"], wl["max_depth"], max(2.0, args.cpu_seconds / 4), procs=1,
                                    row_stride=line["cpu_baseline"]["cores"])
            line["cpu_baseline"]["value_1thread"] = one["value"]
            for k in ("samples", "seconds"):
                line["cpu_baseline"].pop(k, None)
        print(json.dumps(line), flush=True)
    if scene is not None:
        scene.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="veach-mis", choices=sorted(WORKLOADS))
    ap.add_argument("--spp-per-step", type=int, default=0, help="samples per pixel per GPU per step (0 = the workload's default)")
    ap.add_argument("--batch-spp", type=int, default=0)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the side measurements (cornell config 1, jitter, config 3, id mismatch)")
    ap.add_argument("--light-tree", action="store_true", help="MCPT_LIGHT_TREE light selection (SURVEY.md 8(f) row 4; not the headline)")
    ap.add_argument("--bvh", default="host", choices=["host", "gpu"], help="scene BVH builder: host SAH (default) or device LBVH")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.spp_per_step <= 0:
        args.spp_per_step = wl["spp"]
    if args.impl == "reference":
        run_reference(args, args.workload, wl)
    else:
        run_ours(args, args.workload, wl)


if __name__ == "__main__":
    main()
