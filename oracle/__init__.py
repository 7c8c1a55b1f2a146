"""oracle -- TEST INFRASTRUCTURE (ctypes front-ends), not product code.

``OracleLib``  : oracle/liboracle.so, the plain-C FP64 restatement (oracle/mcpt_oracle.c).
``RefLib``     : oracle/_ref/libmcpt_ref.so, the reference's own sources compiled unmodified
                 (oracle/Makefile target ``ref``); only present where it was built from
                 /root/reference (this container) or travelled prebuilt (GPU box).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libmcpt_ref.so")

NEE_SPHERICAL, BRDF_ONLY, MIS, NEE_UNIFORM = 0, 1, 2, 3
FIX_STALE_MIS, FIX_NEE_MEASURE, FIX_MIXTURE_PDF, BRUTE_FORCE, RNG_KEYED = 1, 2, 4, 8, 16

_d3 = C.c_double * 3


def build(ref: bool = True) -> None:
    """Compile the restatement and, when /root/reference is present, oracle/_ref."""
    subprocess.check_call(["make", "-s", "-C", HERE, "restatement"])
    if ref:
        subprocess.check_call(["make", "-s", "-C", HERE, "-j8", "ref"])


def ref_available() -> bool:
    return os.path.exists(REF_SO)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a if shape is None else a.reshape(shape)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def camera_args(cam):
    return _d3(*cam.eye), _d3(*cam.w), _d3(*cam.up), C.c_double(cam.pixellen)


class _Probes:
    """Batch probes shared by both libraries (same C signatures modulo the scene handle)."""
    _pre = ""
    _hs = ()

    def _fn(self, name):
        return getattr(self.lib, self._pre + name)

    def intersect_tri(self, ro, rd, prim):
        ro, rd, prim = _f64(ro), _f64(rd), _i32(prim)
        n = len(prim)
        hit = np.zeros(n, np.int32); bgt = np.zeros((n, 3))
        self._fn("intersect_tri")(*self._hs, C.c_int(n), _p(ro), _p(rd), _p(prim), _p(hit), _p(bgt))
        return hit, bgt

    def closest(self, ro, rd, exclude=None, mode=2):
        ro, rd = _f64(ro), _f64(rd)
        n = len(ro)
        ex = _i32(exclude) if exclude is not None else np.full(n, -1, np.int32)
        prim = np.zeros(n, np.int32); bgt = np.zeros((n, 3))
        self._fn("closest")(*self._hs, C.c_int(n), _p(ro), _p(rd), _p(ex), C.c_int(mode), _p(prim), _p(bgt))
        return prim, bgt

    def brdf_eval(self, N, wi, wr, kd, ks, ns):
        N, wi, wr, kd, ks, ns = map(_f64, (N, wi, wr, kd, ks, ns))
        out = np.zeros_like(N)
        self._fn("brdf_eval")(C.c_int(len(ns)), _p(N), _p(wi), _p(wr), _p(kd), _p(ks), _p(ns), _p(out))
        return out

    def brdf_sample(self, N, wr, kd, ks, ns, u3):
        N, wr, kd, ks, ns, u3 = map(_f64, (N, wr, kd, ks, ns, u3))
        d = np.zeros_like(N); pdf = np.zeros(len(ns))
        self._fn("brdf_sample")(C.c_int(len(ns)), _p(N), _p(wr), _p(kd), _p(ks), _p(ns), _p(u3), _p(d), _p(pdf))
        return d, pdf

    def brdf_pdf(self, N, wi, wr, kd, ks, ns):
        N, wi, wr, kd, ks, ns = map(_f64, (N, wi, wr, kd, ks, ns))
        pdf = np.zeros(len(ns))
        self._fn("brdf_pdf")(C.c_int(len(ns)), _p(N), _p(wi), _p(wr), _p(kd), _p(ks), _p(ns), _p(pdf))
        return pdf

    def light_sample_uniform(self, u4):
        u4 = _f64(u4)
        n = len(u4)
        prim = np.zeros(n, np.int32); pt = np.zeros((n, 3)); Le = np.zeros((n, 3)); pdf = np.zeros(n)
        self._fn("light_sample_uniform")(*self._hs, C.c_int(n), _p(u4), _p(prim), _p(pt), _p(Le), _p(pdf))
        return prim, pt, Le, pdf

    def light_sample_spherical(self, x1, N, u3, query=None, staged=1):
        x1, N, u3 = map(_f64, (x1, N, u3))
        n = len(x1)
        q = _i32(query) if query is not None else np.zeros(n, np.int32)
        prim = np.zeros(n, np.int32); pt = np.zeros((n, 3)); Le = np.zeros((n, 3)); pdf = np.zeros(n)
        wsum = np.zeros(n); nsurv = np.zeros(n, np.int32); qpdf = np.zeros(n)
        self._fn("light_sample_spherical")(*self._hs, C.c_int(n), C.c_int(staged), _p(x1), _p(N), _p(u3), _p(q),
                                           _p(prim), _p(pt), _p(Le), _p(pdf), _p(wsum), _p(nsurv), _p(qpdf))
        return dict(prim=prim, point=pt, Le=Le, pdf=pdf, wsum=wsum, nsurv=nsurv, query_pdf=qpdf)

    def tonemap(self, rgb, max_radiance=380.0, gamma=0.25):
        rgb = _f64(rgb).reshape(-1, 3)
        out = np.zeros(rgb.shape, np.int32)
        self._fn("tonemap")(C.c_int(len(rgb)), _p(rgb), C.c_double(max_radiance), C.c_double(gamma), _p(out))
        return out


class RefLib(_Probes):
    """The compiled reference.  ONE scene per process (the reference never clears its caches)."""
    _pre = "ref_"

    def __init__(self, obj_path: str, xml_path: str):
        if not ref_available():
            raise FileNotFoundError(REF_SO + " (run `make -C oracle ref` where /root/reference exists)")
        self.lib = C.CDLL(REF_SO)
        self.lib.ref_vertices.restype = C.c_uint64
        self.lib.ref_load(obj_path.encode(), xml_path.encode())
        c = [C.c_int() for _ in range(5)]
        self.lib.ref_counts(*[C.byref(x) for x in c])
        self.P, self.S, self.M, self.L, self.C = (x.value for x in c)

    def build_grid(self, eye, n0=100000):
        self.lib.ref_build_grid(_d3(*eye), C.c_int(n0))

    def flatten(self):
        from monte_carlo_path_tracing_b200.scenes import SceneArrays
        P, S, M, L, Cn = self.P, self.S, self.M, self.L, self.C
        a = SceneArrays(np.zeros((P, 3, 3), np.float32), np.zeros((P, 3, 3), np.float32), np.zeros(P, np.int32),
                        np.zeros(S + 1, np.uint32), np.zeros((M, 3), np.float32), np.zeros((M, 3), np.float32),
                        np.zeros(M, np.float32), np.zeros(L, np.int32), np.zeros(L, np.int32), np.zeros((Cn, 3)))
        gn = np.zeros((P, 3)); area = np.zeros(L)
        self.lib.ref_flatten(_p(a.positions), _p(a.normals), _p(a.material_id), _p(a.shape_offset), _p(a.mat_kd),
                             _p(a.mat_ks), _p(a.mat_ns), _p(a.light_prim), _p(a.light_class), _p(a.class_radiance),
                             _p(gn), _p(area))
        return a, gn, area

    def render(self, cam, strategy, spp, max_depth=0, seed=1, rows=None, cols=None, spp_offset=0, spp_total=None):
        W, H = cam.width, cam.height
        i0, i1 = rows or (0, H)
        j0, j1 = cols or (0, W)
        out = np.zeros((i1 - i0, j1 - j0, 3)); cnt = (C.c_uint64 * 4)()
        e, w, u, pl = camera_args(cam)
        self.lib.ref_render(e, w, u, pl, W, H, i0, i1, j0, j1, spp, spp_offset, spp_total or spp, strategy, max_depth,
                            C.c_uint64(seed), _p(out), cnt)
        return out, dict(vertices=cnt[0], primary_hits=cnt[1], rays=cnt[2], light_rays=cnt[3])


    def render_variant(self, cam, strategy, spp, max_depth=0, flags=0, seed=1, rows=None, cols=None, spp_offset=0, spp_total=None):
        """The estimator re-hosted on the reference's own objects with the ORACLE_FIX_* switches (ref_driver.cpp:
        ref_render_variant); strategy 3 = uniform-area NEE.  flags == 0, strategy 0..2 equals ``render`` bit for bit."""
        W, H = cam.width, cam.height
        i0, i1 = rows or (0, H)
        j0, j1 = cols or (0, W)
        out = np.zeros((i1 - i0, j1 - j0, 3)); cnt = (C.c_uint64 * 4)()
        e, w, u, pl = camera_args(cam)
        self.lib.ref_render_variant(e, w, u, pl, W, H, i0, i1, j0, j1, spp, spp_offset, spp_total or spp, strategy, max_depth,
                                    flags, C.c_uint64(seed), _p(out), cnt)
        return out, dict(vertices=cnt[0], primary_hits=cnt[1], rays=cnt[2], light_rays=cnt[3])


class OracleLib(_Probes):
    """The plain-C restatement on a flat scene (monte_carlo_path_tracing_b200.scenes.SceneArrays)."""
    _pre = "oracle_"

    def __init__(self, arrays):
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        self.lib = C.CDLL(ORACLE_SO)
        self.lib.oracle_scene_create.restype = C.c_void_p
        a = arrays
        self.P, self.L = len(a.material_id), len(a.light_prim)
        self._keep = [np.ascontiguousarray(x) for x in (a.positions, a.normals, a.material_id, a.shape_offset, a.mat_kd,
                                                        a.mat_ks, a.mat_ns, a.light_prim, a.light_class,
                                                        a.class_radiance)]
        k = self._keep
        self.scene = C.c_void_p(self.lib.oracle_scene_create(
            C.c_int(self.P), C.c_int(len(a.shape_offset) - 1), C.c_int(len(a.mat_ns)), C.c_int(self.L),
            C.c_int(len(a.class_radiance)), *[_p(x) for x in k]))
        self._hs = (self.scene,)

    def __del__(self):
        try:
            self.lib.oracle_scene_destroy(self.scene)
        except Exception:
            pass

    def build_grid(self, eye, n0=100000):
        return self.lib.oracle_build_grid(self.scene, _d3(*eye), C.c_int(n0))

    def geom_normals(self):
        out = np.zeros((self.P, 3))
        self.lib.oracle_geom_normals(self.scene, _p(out))
        return out

    def light_areas(self):
        out = np.zeros(self.L)
        self.lib.oracle_light_areas(self.scene, _p(out))
        return out

    def keyed_uniforms(self, seed, pixel, sample, node, call):
        out = (C.c_double * 4)()
        self.lib.oracle_keyed_uniforms(C.c_uint64(seed), C.c_uint32(pixel), C.c_uint32(sample), C.c_uint64(node), C.c_uint32(call), out)
        return np.array(out[:])

    def render(self, cam, strategy, spp, max_depth=0, flags=0, seed=1, rows=None, cols=None, spp_offset=0,
               spp_total=None):
        W, H = cam.width, cam.height
        i0, i1 = rows or (0, H)
        j0, j1 = cols or (0, W)
        out = np.zeros((i1 - i0, j1 - j0, 3)); cnt = (C.c_uint64 * 4)()
        e, w, u, pl = camera_args(cam)
        self.lib.oracle_render(self.scene, e, w, u, pl, W, H, i0, i1, j0, j1, spp, spp_offset, spp_total or spp,
                               strategy, max_depth, flags, C.c_uint64(seed), _p(out), cnt)
        return out, dict(vertices=cnt[0], primary_hits=cnt[1], rays=cnt[2], light_rays=cnt[3])
