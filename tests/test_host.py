"""CPU-only tests of the host side: the C ABI surface, the host loaders, the SAH BVH builder, the scene
generators, the spp sharding and its reduce (2 ranks over gloo), and the no-GPU error path.  No compute
entry point is called here (there is no GPU and no CPU fallback)."""
import ctypes as C
import os
import re
import socket
import subprocess

import numpy as np
import pytest

import util
from monte_carlo_path_tracing_b200 import api, dist as mdist, scenes

ROOT = util.ROOT


def test_abi_exports_every_declared_symbol():
    """libmcpt.so loads and exports every function include/mcpt.h declares; api.SYMBOLS lists them all."""
    hdr = open(os.path.join(ROOT, "include", "mcpt.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = sorted(set(re.findall(r"\b(mcpt_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 25
    lib = api.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert sorted(api.SYMBOLS) == declared
    assert lib.mcpt_version() == 100


def test_abi_struct_layouts_match_header(tmp_path):
    """ctypes mirrors of the PODs have the sizes AND field offsets the C compiler gives include/mcpt.h."""
    assert C.sizeof(api.CameraC) == 10 * 8
    assert C.sizeof(api.RenderParams) == 4 * 8 + 8 + 4 + 4 + 4 + 4
    assert C.sizeof(api.Stats) == 6 * 8 + 4 * 4 + 5 * 4 + 4 + 8 + 8 + 8 + 2 * 4  # 5 floats, 4 bytes of padding, elements, light_sweeps, light_evals, ms_sweep, ms_prep
    assert C.sizeof(api.SceneDesc) == 5 * 4 + 4 + 10 * 8
    structs = {"McptCamera": api.CameraC, "McptRenderParams": api.RenderParams, "McptStats": api.Stats, "McptSceneDesc": api.SceneDesc}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "mcpt.h"', 'int main(void){']
    for cname, ct in structs.items():
        lines.append('printf("%s %%zu\\n", sizeof(%s));' % (cname, cname))
        for f, _ in ct._fields_:
            lines.append('printf("%s.%s %%zu\\n", offsetof(%s, %s));' % (cname, f, cname, f))
    lines.append("return 0;}")
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)])
    got = dict(l.split() for l in subprocess.check_output([str(exe)], text=True).splitlines())
    for cname, ct in structs.items():
        assert int(got[cname]) == C.sizeof(ct), cname
        for f, _ in ct._fields_:
            assert int(got["%s.%s" % (cname, f)]) == getattr(ct, f).offset, (cname, f)


def test_integration_md_quotes_the_binding_verbatim():
    """INTEGRATION.md section 2 IS oracle/ref_build/ref_gpu_main.cpp, the reference-side binding that is compiled against
    the unmodified reference objects and run by tests/test_gpu_render.py -- not a sketch that can rot."""
    md = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    src = open(os.path.join(ROOT, "oracle", "ref_build", "ref_gpu_main.cpp")).read()
    a = md.index("<!-- BEGIN ref_gpu_main.cpp (verbatim) -->")
    b = md.index("<!-- END ref_gpu_main.cpp -->")
    quoted = md[a:b].split("```cpp\n", 1)[1].rsplit("```", 1)[0]
    assert quoted == src


@pytest.mark.skipif(os.path.exists("/dev/nvidiactl"), reason="a GPU is present")
def test_no_gpu_fails_loudly(arrays_of):
    """No CUDA device -> MCPT_ERR_NO_DEVICE (or a CUDA error), never a silent CPU path."""
    a, _ = arrays_of("cornell-box")
    with pytest.raises(api.McptError) as e:
        api.Scene(a, 0)
    assert e.value.code in (-3, -2)


def test_loader_errors_do_not_exit(tmp_path):
    """The reference exit(1)s on a missing file (Myobj.cpp:15-20, Mylight.cpp:16-20); the ABI returns a code."""
    with pytest.raises(api.McptError) as e:
        api.load_scene_files(str(tmp_path / "nope.obj"), str(tmp_path / "nope.xml"))
    assert e.value.code == -4
    (tmp_path / "a.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2//1 3//1\n")
    with pytest.raises(api.McptError):
        api.load_scene_files(str(tmp_path / "a.obj"), str(tmp_path / "missing.xml"))


def test_loader_camera_and_light_order(scene_files):
    """<camera> node parsed (README.md:339-343); lights ordered by class (mtlname ascending) then prim."""
    sc, stem = scene_files("veach-mis")
    a, cam = api.load_scene_files(stem + ".obj", stem + ".xml")
    assert cam is not None and (cam.width, cam.height) == (sc.camera.width, sc.camera.height)
    assert np.allclose(cam.eye, sc.camera.eye) and abs(cam.fovy_deg - sc.camera.fovy_deg) < 1e-12
    assert np.all(np.diff(a.light_class) >= 0)
    for c in np.unique(a.light_class):
        assert np.all(np.diff(a.light_prim[a.light_class == c]) > 0)
    assert a.shape_offset[0] == 0 and a.shape_offset[-1] == len(a.material_id)
    assert a.material_id.min() >= 0 and a.material_id.max() < len(a.mat_ns)


@pytest.mark.parametrize("name", ["cornell-box", "veach-mis"])
def test_host_bvh_is_valid(name, arrays_of):
    a, _ = arrays_of(name)
    for lights_only in (False, True):
        r = api.host_bvh_check(a, lights_only)
        n = len(a.light_prim) if lights_only else len(a.material_id)
        assert r["errors"] == 0 and r["max_depth"] < 64
        assert r["leaves"] >= (n + 7) // 8 and r["nodes"] >= r["leaves"] - 1
        assert r["sah"] > 0


def test_host_bvh_triangle_soup():
    """C4-style soup at 50 k triangles: structurally valid, depth far below the traversal stack."""
    a = scenes.triangle_soup(50_000, 16, seed=3)
    r = api.host_bvh_check(a, False)
    assert r["errors"] == 0 and r["max_depth"] <= 48
    r = api.host_bvh_check(a, True)
    assert r["errors"] == 0


def test_host_bvh_degenerate_inputs():
    """All-coincident centroids and zero-area triangles must not break the builder."""
    a = scenes.triangle_soup(64, 2, seed=4)
    a.positions[:40] = a.positions[0]        # 40 identical triangles
    a.positions[40:50, 1] = a.positions[40:50, 0]  # zero-area
    r = api.host_bvh_check(a, False)
    assert r["errors"] == 0


@pytest.mark.parametrize("name", ["cornell-box", "veach-mis"])
def test_host_light_tree_is_valid(name, arrays_of):
    """Light hierarchy of MCPT_LIGHT_TREE (csrc/light_tree.h): 2L-1 nodes, every light in exactly one leaf, every
    ancestor's bounding sphere holds the light's vertices and every ancestor's cone its normal, powers add up."""
    a, _ = arrays_of(name)
    r = api.host_light_tree_check(a)
    L = len(a.light_prim)
    assert r["errors"] == 0 and r["lights"] == L and r["nodes"] == 2 * L - 1
    assert r["max_depth"] <= 4 * max(1, int(np.ceil(np.log2(max(L, 2)))))


def test_host_light_tree_soup_and_degenerate():
    """Full-size veach-mis light set (1282 triangles on four spheres), random lights, coincident lights, one light."""
    sc = scenes.veach_mis()
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        stem = sc.write(td)
        a, _ = api.load_scene_files(stem + ".obj", stem + ".xml")
    r = api.host_light_tree_check(a)
    assert r["errors"] == 0 and r["nodes"] == 2 * 1282 - 1 and r["max_depth"] <= 24
    a = scenes.triangle_soup(2000, 500, seed=5)
    r = api.host_light_tree_check(a)
    assert r["errors"] == 0 and r["nodes"] == 999 and r["max_depth"] <= 30
    a.positions[a.light_prim[:300]] = a.positions[a.light_prim[0]]  # 300 coincident lights: split by order
    r = api.host_light_tree_check(a)
    assert r["errors"] == 0 and r["nodes"] == 999
    a = scenes.triangle_soup(100, 1, seed=6)
    r = api.host_light_tree_check(a)
    assert r["errors"] == 0 and r["nodes"] == 1 and r["max_depth"] == 0


def test_scene_generators():
    c = scenes.cornell_box(width=32, height=32)
    assert "usemtl" in c.obj and "vn " in c.obj and "<light" in c.xml and "<camera" in c.xml
    v = scenes.veach_mis(width=32, height=18)
    assert v.obj.count("\nf ") > 3000 * 0.8  # ~3 k facets like the reference's scene (README.md:212)
    s = scenes.triangle_soup(1000, 8)
    assert s.positions.shape == (1008, 3, 3) and len(s.light_prim) == 8


def test_shard_spp_partitions():
    for total in (1, 7, 16, 1024, 4096):
        for world in (1, 2, 3, 4, 8):
            parts = [mdist.shard_spp(total, world, r) for r in range(world)]
            assert sum(c for _, c in parts) == total
            off = 0
            for o, c in parts:
                assert o == off
                off += c
            counts = [c for _, c in parts]
            assert max(counts) - min(counts) <= 1


# ---- 2 ranks over gloo: the N>1 path of dist.render_sharded with the oracle standing in for the device ----------
class _OracleScene:
    """Test double with Scene.render_device's signature: fills the rank's film from the CPU oracle."""

    def __init__(self, orc, film):
        self.orc, self.film = orc, film

    def render_device(self, cam, strategy, spp, d_out_ptr, stream=0, spp_offset=0, spp_total=0, **kw):
        import torch
        img, cnt = self.orc.render(cam, strategy, spp, spp_offset=spp_offset, spp_total=spp_total, **kw)
        self.film.copy_(torch.from_numpy(img.reshape(-1).astype(np.float32)))
        return dict(vertices=cnt["vertices"])


def _gloo_worker(rank, world, port, total_spp, out_path):
    import torch
    import torch.distributed as td
    import oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    td.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import tempfile
        sc = util.scene_text("cornell-box")
        with tempfile.TemporaryDirectory() as d:
            stem = sc.write(d)
            a, cam = api.load_scene_files(stem + ".obj", stem + ".xml")
        orc = oracle.OracleLib(a)
        orc.build_grid(cam.eye, 100000)
        film = torch.zeros(cam.width * cam.height * 3, dtype=torch.float32)
        st = mdist.render_sharded(_OracleScene(orc, film), cam, oracle.NEE_UNIFORM, total_spp, rank, world, film,
                                  max_depth=3, seed=5, flags=oracle.FIX_STALE_MIS)
        assert st["vertices"] > 0
        if rank == 0:
            np.save(out_path, film.numpy())
    finally:
        td.destroy_process_group()


def test_spp_sharding_two_ranks_gloo(tmp_path, oracle_of, arrays_of):
    """World size 2: each rank renders its half of the global sample range (disjoint counter-RNG streams keyed
    by the GLOBAL sample index), one reduce sums the films on rank 0 -> equals the 1-rank render."""
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = str(tmp_path / "film.npy")
    total = 6
    mp.spawn(_gloo_worker, args=(2, port, total, out), nprocs=2, join=True)
    got = np.load(out)
    import oracle
    _, cam = arrays_of("cornell-box")
    ref, _ = oracle_of("cornell-box").render(cam, oracle.NEE_UNIFORM, total, max_depth=3, seed=5, flags=oracle.FIX_STALE_MIS)
    assert np.allclose(got, ref.reshape(-1), rtol=1e-5, atol=1e-6)
    assert got.max() > 0


def _driver():
    p = os.path.join(ROOT, "monte_carlo_path_tracing_b200", "mcpt_render")
    if not os.path.exists(p):
        api.build()
    return p


@pytest.mark.skipif(os.path.exists("/dev/nvidiactl"), reason="a GPU is present")
def test_cpp_driver_fails_loudly_without_gpu(scene_files, tmp_path):
    """mcpt_render (main.cpp minus EasyX, C++ over the C ABI) loads the scene on the host, then stops with the
    library's message -- it never falls back to a CPU renderer and never exit()s from inside the library."""
    import subprocess
    _, stem = scene_files("cornell-box")
    r = subprocess.run([_driver(), stem + ".obj", stem + ".xml", str(tmp_path / "o.ppm"), "--spp", "1"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 1
    assert "light triangles" in r.stdout and "no CPU fallback" in r.stderr
    assert not (tmp_path / "o.ppm").exists()


# ---- loader edge cases (what tiny_obj_loader does with the files users of the reference have) --------
_MTL = "newmtl white\nKd 0.7 0.6 0.5\nKs 0.1 0.2 0.3\nNs 25\nnewmtl lamp\nKd 0 0 0\n"
_XML = '<?xml version="1.0"?>\n<camera type="perspective" width="64" height="48" fovy="35">\n <eye x="0" y="1" z="5"/>\n <lookat x="0" y="0" z="0"/>\n <up x="0" y="1" z="0"/>\n</camera>\n<light mtlname="lamp" radiance="10,20,30"/>\n'


def _write(tmp_path, obj, xml=_XML, mtl=_MTL):
    (tmp_path / "s.obj").write_text("mtllib s.mtl\n" + obj)
    (tmp_path / "s.mtl").write_text(mtl)
    (tmp_path / "s.xml").write_text(xml)
    return str(tmp_path / "s.obj"), str(tmp_path / "s.xml")


def test_loader_quads_ngons_negative_indices(tmp_path):
    """tinyobj triangulates (tiny_obj_loader.h:509-531): quads split along the shorter diagonal, polygons as fans;
    negative indices count back from the current end of the vertex list; o / g start new shapes."""
    obj = ("v 0 0 0\nv 2 0 0\nv 2 1 0\nv 0 1 0\nvn 0 0 1\n"
           "o floor\nusemtl white\nf 1//1 2//1 3//1 4//1\n"           # quad: diagonals equal -> (0,1,3),(1,2,3) branch
           "v 0 0 1\nv 1 0 1\nv 1.5 1 1\nv 0.5 2 1\nv -0.5 1 1\n"
           "g lampshape\nusemtl lamp\nf -5//1 -4//1 -3//1 -2//1 -1//1\n")  # pentagon by negative indices -> 3 triangles
    a, cam = api.load_scene_files(*_write(tmp_path, obj))
    assert len(a.material_id) == 5 and list(a.shape_offset) == [0, 2, 5]
    assert a.material_id.tolist() == [0, 0, 1, 1, 1]
    assert a.light_prim.tolist() == [2, 3, 4] and a.light_class.tolist() == [0, 0, 0]
    assert np.allclose(a.class_radiance, [[10, 20, 30]])
    assert np.allclose(a.mat_kd[0], [0.7, 0.6, 0.5]) and np.allclose(a.mat_ks[0], [0.1, 0.2, 0.3]) and a.mat_ns[0] == 25
    # the quad's two triangles cover its area exactly once
    v = a.positions[:2].astype(np.float64)
    area = 0.5 * np.linalg.norm(np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0]), axis=1)
    assert abs(area.sum() - 2.0) < 1e-6
    # pentagon fan: all triangles share the first vertex, z = 1
    assert np.all(a.positions[2:, 0] == a.positions[2, 0]) and np.all(a.positions[2:, :, 2] == 1.0)
    assert (cam.width, cam.height, cam.fovy_deg) == (64, 48, 35.0) and cam.eye == (0.0, 1.0, 5.0)
    r = api.host_bvh_check(a, False)
    assert r["errors"] == 0


def test_loader_rejects_what_the_reference_reads_out_of_bounds(tmp_path):
    """A face without vn makes the reference index attrib.normals[-3..] (Myobj.cpp:671-673), a face without usemtl
    makes it call materials.at(-1) (main.cpp:292): both are reported as errors instead."""
    with pytest.raises(api.McptError) as e:
        api.load_scene_files(*_write(tmp_path, "v 0 0 0\nv 1 0 0\nv 0 1 0\nusemtl white\nf 1 2 3\n"))
    assert e.value.code == -4 and "normal" in str(e.value)
    with pytest.raises(api.McptError) as e:
        api.load_scene_files(*_write(tmp_path, "v 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2//1 3//1\n"))
    assert e.value.code == -4 and "material" in str(e.value)
    with pytest.raises(api.McptError):
        api.load_scene_files(*_write(tmp_path, "v 0 0 0\nv 1 0 0\nvn 0 0 1\nusemtl white\nf 1//1 2//1 7//1\n"))


def test_loader_light_without_triangles_and_no_camera(tmp_path):
    """<light> whose material no face uses: lightsTriangles.at() throws in the reference (Mylight.cpp:126); here the
    scene loads and scene_create rejects it.  A file without <camera> reports has_camera = 0."""
    obj = "v 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nusemtl white\nf 1//1 2//1 3//1\n"
    a, cam = api.load_scene_files(*_write(tmp_path, obj, xml='<light mtlname="lamp" radiance="1,1,1"/>\n'))
    assert cam is None and len(a.light_prim) == 0


@pytest.mark.parametrize("name", ["course", "course2"])
def test_loader_matches_the_reference_loaders_on_a_dcc_style_export(name):
    """The host loader (csrc/host_scene.cpp) against the reference's own tiny_obj_loader + pugixml + Myobj / Mylight accessors
    (Myobj.cpp:10-28, Mylight.cpp:11-100) on files the scene generators of this repo never write: smoothing groups, vt
    indices, map_Kd / map_Bump / illum / Ni / d / Tr lines, v/vt/vn, v//vn and relative indices mixed, usemtl and g in
    mid-object, quads and a pentagon; `course2` with CRLF line ends, tabs, vp / l / p statements, exponent notation and no
    final newline.  Golden arrays from the compiled reference: tests/golden/make_course_golden.py."""
    d = os.path.join(ROOT, "tests", "golden", "course-style")
    g = np.load(os.path.join(d, name + ".npz"))
    a, cam = api.load_scene_files(os.path.join(d, name + ".obj"), os.path.join(d, name + ".xml"))
    for f in ("positions", "normals", "material_id", "shape_offset", "mat_kd", "mat_ks", "mat_ns", "light_prim", "light_class", "class_radiance"):
        assert np.array_equal(getattr(a, f), g[f]), f
    assert len(a.material_id) == 13 and a.light_prim.tolist() == [10, 11, 12]
    assert (cam.width, cam.height, cam.fovy_deg) == (96, 64, 40.0)
