"""Probe: do two renders on two streams overlap usefully on one B200?  Two replicas of veach-mis, each rendering half of a
128-spp step from its own host thread on its own stream (MCPT_GRID_DIV=2 halves the resident blocks of every persistent
kernel so that both fit), with the second thread started `stagger` ms after the first.  Compared with the same two halves
back to back and with one 128-spp render.  Usage: python tools/diag/overlap_probe.py [stagger_ms ...]"""
import os, sys, tempfile, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from monte_carlo_path_tracing_b200 import api, scenes

sc = scenes.veach_mis(width=1280, height=720)
with tempfile.TemporaryDirectory() as td:
    stem = sc.write(td)
    arrays, _ = api.load_scene_files(stem + ".obj", stem + ".xml")
cam = sc.camera
A, B = api.Scene(arrays, 0), api.Scene(arrays, 0)
filmA = torch.zeros(720 * 1280 * 3, device="cuda"); filmB = torch.zeros_like(filmA)
sA, sB = torch.cuda.Stream(), torch.cuda.Stream()
S = 128

def run(scene, film, stream, spp, off, out):
    t0 = time.perf_counter()
    st = scene.render_device(cam, api.MIS, spp, film.data_ptr(), stream=stream.cuda_stream, spp_offset=off, spp_total=S, seed=1)
    out.append((time.perf_counter() - t0) * 1e3); out.append(st)

def wall(fn, n=3):
    best = 1e9
    for _ in range(n):
        torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize()
        best = min(best, (time.perf_counter() - t0) * 1e3)
    return best

for _ in range(2):
    run(A, filmA, sA, S, 0, []); run(B, filmB, sB, S // 2, 0, [])
print("one 128-spp render            : %.2f ms" % wall(lambda: run(A, filmA, sA, S, 0, [])))
print("two 64-spp renders, in series : %.2f ms" % wall(lambda: (run(A, filmA, sA, S // 2, 0, []), run(B, filmB, sB, S // 2, S // 2, []))))
for stagger in [float(x) for x in (sys.argv[1:] or ["0", "3", "6", "10"])]:
    def both():
        oa, ob = [], []
        ta = threading.Thread(target=run, args=(A, filmA, sA, S // 2, 0, oa))
        tb = threading.Thread(target=run, args=(B, filmB, sB, S // 2, S // 2, ob))
        ta.start(); time.sleep(stagger * 1e-3); tb.start(); ta.join(); tb.join()
        both.last = (oa[0], ob[0], oa[1]["ms_shade"], oa[1]["ms_trace"], ob[1]["ms_shade"], ob[1]["ms_trace"])
    w = wall(both)
    print("two 64-spp renders, 2 threads, stagger %4.1f ms: %.2f ms  (A %.1f B %.1f; A shade %.1f trace %.1f, B shade %.1f trace %.1f)" % ((stagger, w) + both.last))
