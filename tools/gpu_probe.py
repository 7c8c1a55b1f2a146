"""Quick GPU probe: render each configuration briefly and print throughput from McptStats."""
import sys, time, os, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from monte_carlo_path_tracing_b200 import api, scenes

def load(sc):
    with tempfile.TemporaryDirectory() as td:
        stem = sc.write(td)
        return api.load_scene_files(stem + '.obj', stem + '.xml')[0]

def run(tag, scene, cam, strat, spp, **kw):
    scene.render(cam, strat, min(spp, 4), **kw)  # warm-up
    t = time.time(); img, st = scene.render(cam, strat, spp, **kw); wall = time.time() - t
    ms = st['ms_total']
    print('%-28s %-14s spp %4d  %8.1f ms (bounce %7.1f, wall %7.1f)  %8.1f Msamples/s  %8.1f Mvert/s  %8.1f Mrays/s  v/path %.2f  launches %d bounces %d peak %d mean %.4f' % (
        tag, api.STRATEGY_NAMES[strat], spp, ms, st['ms_bounce'], wall * 1e3, st['paths'] / ms / 1e3, st['vertices'] / ms / 1e3,
        (st['rays_closest'] + st['rays_shadow'] + st['rays_light']) / ms / 1e3, st['vertices'] / st['paths'], st['kernel_launches'], st['bounces'], st['queue_peak'], img.mean()), flush=True)

which = sys.argv[1:] or ['cornell', 'veach', 'interior', 'soup']
if 'cornell' in which:
    sc = scenes.cornell_box(512, 512); s = api.Scene(load(sc))
    for strat in (api.NEE_UNIFORM, api.BRDF_ONLY, api.NEE_SPHERICAL, api.MIS):
        run('cornell 512x512 d5', s, sc.camera, strat, 64, max_depth=5)
    run('cornell 512x512 d5 b64', s, sc.camera, api.NEE_UNIFORM, 256, max_depth=5, batch_spp=64)
    s.close()
if 'veach' in which:
    sc = scenes.veach_mis(1280, 720); s = api.Scene(load(sc))
    for strat in (api.NEE_UNIFORM, api.BRDF_ONLY, api.NEE_SPHERICAL, api.MIS):
        run('veach 1280x720 d5', s, sc.camera, strat, 8, max_depth=5)
    s.close()
if 'interior' in which:
    sc = scenes.interior(1280, 720); t = time.time(); a = load(sc); print('interior load %.1fs' % (time.time() - t))
    t = time.time(); s = api.Scene(a); print('interior scene_create %.1fs' % (time.time() - t), s.bvh_info())
    for strat in (api.NEE_UNIFORM, api.MIS):
        run('interior 1280x720 d5', s, sc.camera, strat, 8, max_depth=5)
    s.close()
if 'soup' in which:
    a = scenes.triangle_soup(); t = time.time(); s = api.Scene(a); print('soup scene_create %.1fs' % (time.time() - t), s.bvh_info())
    run('soup 1M 1024x1024 d5', s, a.camera, api.NEE_UNIFORM, 8, max_depth=5)
    s.close()
