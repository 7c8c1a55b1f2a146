"""ctypes binding of the C ABI in include/mcpt.h (libmcpt.so, CUDA sm_100a).

There is no CPU fallback: if the library is missing, ``load()`` raises; if there is no CUDA
device, ``Scene(...)`` raises ``McptError`` (MCPT_ERR_NO_DEVICE).  The host loaders
(``load_scene_files``) work without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from .scenes import Camera, SceneArrays

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MCPT_LIB") or os.path.join(HERE, "libmcpt.so")  # MCPT_LIB: an A/B build variant (tools/)
CSRC = os.path.join(HERE, "csrc")

NEE_SPHERICAL, BRDF_ONLY, MIS, NEE_UNIFORM, MIS_SINGLE = 0, 1, 2, 3, 4
STRATEGY_NAMES = {0: "NEE_SPHERICAL", 1: "BRDF_ONLY", 2: "MIS", 3: "NEE_UNIFORM"}
FIX_NEE_MEASURE, FIX_MIXTURE_PDF, JITTER, ACCUMULATE, LIGHT_TREE, PARITY_FP64 = 2, 4, 16, 64, 128, 256
LEAF_FAST, LEAF_EXACT = 0, 1
FP64, FP32 = 0, 1
BUILD_HOST_SAH, BUILD_GPU_LBVH = 0, 1

# every symbol include/mcpt.h declares (tests check the .so exports all of them)
SYMBOLS = [
    "mcpt_scene_create", "mcpt_scene_destroy", "mcpt_scene_device_bytes", "mcpt_scene_upload", "mcpt_render",
    "mcpt_render_device", "mcpt_tonemap", "mcpt_intersect_tri_batch", "mcpt_intersect_batch", "mcpt_phong_eval_batch",
    "mcpt_phong_sample_batch", "mcpt_phong_pdf_batch", "mcpt_light_sample_uniform_batch",
    "mcpt_light_sample_spherical_batch", "mcpt_scene_geom_normals", "mcpt_scene_light_areas", "mcpt_scene_bvh_info",
    "mcpt_last_error", "mcpt_version", "mcpt_device_count", "mcpt_host_scene_load", "mcpt_host_scene_desc",
    "mcpt_host_scene_camera", "mcpt_host_scene_free", "mcpt_host_bvh_check", "mcpt_scene_create_ex",
    "mcpt_scene_build_info", "mcpt_scene_bvh_check", "mcpt_host_light_tree_check", "mcpt_light_tree_batch",
    "mcpt_scene_light_tree_info", "mcpt_scene_light_tree_nodes", "mcpt_visibility_batch",
    "mcpt_multi_create", "mcpt_multi_device_count", "mcpt_multi_upload", "mcpt_multi_render", "mcpt_multi_destroy",
    "mcpt_render_multi", "mcpt_multi_scene", "mcpt_measure_fp32", "mcpt_scene_trace_layout", "mcpt_scene_light_lists", "mcpt_scene_light_list_of_prim", "mcpt_host_light_lists", "mcpt_visibility_noalloc",
]


NOALLOC_MAX_RAYS = 65536  # include/mcpt.h MCPT_NOALLOC_MAX_RAYS


class McptError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("mcpt error %d: %s" % (code, msg))
        self.code = code


class SceneDesc(C.Structure):
    _fields_ = [("num_prims", C.c_uint32), ("num_shapes", C.c_uint32), ("num_materials", C.c_uint32),
                ("num_lights", C.c_uint32), ("num_classes", C.c_uint32),
                ("positions", C.c_void_p), ("normals", C.c_void_p), ("material_id", C.c_void_p),
                ("shape_offset", C.c_void_p), ("mat_kd", C.c_void_p), ("mat_ks", C.c_void_p), ("mat_ns", C.c_void_p),
                ("light_prim", C.c_void_p), ("light_class", C.c_void_p), ("class_radiance", C.c_void_p)]


class CameraC(C.Structure):
    _fields_ = [("eye", C.c_double * 3), ("w", C.c_double * 3), ("up", C.c_double * 3), ("pixellen", C.c_double)]


class RenderParams(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("spp_offset", C.c_int32),
                ("spp_total", C.c_int32), ("max_depth", C.c_int32), ("strategy", C.c_int32), ("flags", C.c_uint32),
                ("seed", C.c_uint64), ("p_rr", C.c_float), ("batch_spp", C.c_int32), ("leaf_mode", C.c_int32),
                ("reserved", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("primary_hits", C.c_uint64), ("vertices", C.c_uint64),
                ("rays_closest", C.c_uint64), ("rays_shadow", C.c_uint64), ("rays_light", C.c_uint64),
                ("passes", C.c_uint32), ("bounces", C.c_uint32), ("kernel_launches", C.c_uint32),
                ("queue_peak", C.c_uint32), ("ms_total", C.c_float), ("ms_bounce", C.c_float), ("ms_other", C.c_float),
                ("ms_shade", C.c_float), ("ms_trace", C.c_float), ("elements", C.c_uint64), ("light_sweeps", C.c_uint64), ("light_evals", C.c_uint64),
                ("ms_sweep", C.c_float), ("ms_prep", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


_lib = None


def build(verbose: bool = False) -> str:
    """Compile csrc/ into libmcpt.so for sm_100a (nvcc cross-compiles without a GPU)."""
    subprocess.check_call(["make", "-C", CSRC, "-j8"] + ([] if verbose else ["-s"]))
    return LIB_PATH


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(LIB_PATH + " is missing: run __graft_entry__.build() (there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    lib.mcpt_last_error.restype = C.c_char_p
    lib.mcpt_scene_device_bytes.restype = C.c_size_t
    lib.mcpt_scene_device_bytes.argtypes = [C.c_void_p]
    lib.mcpt_host_scene_desc.restype = C.POINTER(SceneDesc)
    lib.mcpt_host_scene_desc.argtypes = [C.c_void_p]
    lib.mcpt_host_scene_free.argtypes = [C.c_void_p]
    lib.mcpt_scene_destroy.argtypes = [C.c_void_p]
    lib.mcpt_multi_destroy.argtypes = [C.c_void_p]
    lib.mcpt_multi_upload.argtypes = [C.c_void_p]
    lib.mcpt_multi_device_count.argtypes = [C.c_void_p]
    lib.mcpt_multi_scene.restype = C.c_void_p
    lib.mcpt_multi_scene.argtypes = [C.c_void_p, C.c_int32]
    _lib = lib
    return lib


def _check(rc):
    if rc != 0:
        raise McptError(rc, load().mcpt_last_error().decode(errors="replace"))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def camera_c(cam: Camera) -> CameraC:
    c = CameraC()
    c.eye[:] = cam.eye
    c.w[:] = cam.w
    c.up[:] = cam.up
    c.pixellen = cam.pixellen
    return c


def make_desc(a: SceneArrays):
    """SceneDesc pointing into (contiguous copies of) the arrays; returns (desc, keepalive)."""
    keep = [np.ascontiguousarray(a.positions, np.float32), np.ascontiguousarray(a.normals, np.float32),
            _i32(a.material_id), np.ascontiguousarray(a.shape_offset, np.uint32),
            np.ascontiguousarray(a.mat_kd, np.float32), np.ascontiguousarray(a.mat_ks, np.float32),
            np.ascontiguousarray(a.mat_ns, np.float32), _i32(a.light_prim), _i32(a.light_class),
            _f64(a.class_radiance)]
    d = SceneDesc()
    d.num_prims = len(keep[2]); d.num_shapes = len(keep[3]) - 1; d.num_materials = len(keep[6])
    d.num_lights = len(keep[7]); d.num_classes = len(keep[9].reshape(-1, 3))
    (d.positions, d.normals, d.material_id, d.shape_offset, d.mat_kd, d.mat_ks, d.mat_ns, d.light_prim, d.light_class,
     d.class_radiance) = [x.ctypes.data for x in keep]
    return d, keep


def load_scene_files(obj_path: str, xml_path: str):
    """Host loaders (mirror of Myobj::read / Mylight::read + gather_light_triangles).
    Returns (SceneArrays, Camera or None).  Needs no GPU."""
    lib = load()
    h = C.c_void_p()
    _check(lib.mcpt_host_scene_load(obj_path.encode(), xml_path.encode(), C.byref(h)))
    try:
        d = lib.mcpt_host_scene_desc(h).contents
        P, S, M, L, Cn = d.num_prims, d.num_shapes, d.num_materials, d.num_lights, d.num_classes

        def arr(ptr, ctype, shape):
            n = int(np.prod(shape))
            if n == 0:
                return np.zeros(shape, ctype)
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(np.ctypeslib.as_ctypes_type(ctype))), (n,)).reshape(shape).copy()
        a = SceneArrays(arr(d.positions, np.float32, (P, 3, 3)), arr(d.normals, np.float32, (P, 3, 3)),
                        arr(d.material_id, np.int32, (P,)), arr(d.shape_offset, np.uint32, (S + 1,)),
                        arr(d.mat_kd, np.float32, (M, 3)), arr(d.mat_ks, np.float32, (M, 3)), arr(d.mat_ns, np.float32, (M,)),
                        arr(d.light_prim, np.int32, (L,)), arr(d.light_class, np.int32, (L,)),
                        arr(d.class_radiance, np.float64, (Cn, 3)))
        cc = CameraC(); w = C.c_int32(); hh = C.c_int32(); fov = C.c_double()
        cam = None
        if lib.mcpt_host_scene_camera(h, C.byref(cc), C.byref(w), C.byref(hh), C.byref(fov)):
            eye = tuple(cc.eye)
            cam = Camera(eye, tuple(e + x for e, x in zip(eye, cc.w)), tuple(cc.up), fov.value, w.value, hh.value)
        a.camera = cam
        return a, cam
    finally:
        lib.mcpt_host_scene_free(h)


def host_bvh_check(a: SceneArrays, lights_only: bool = False):
    """Build the SAH BVH on the host and validate it structurally (no GPU needed)."""
    lib = load()
    d, keep = make_desc(a)
    out = (C.c_uint32 * 4)(); sah = C.c_double()
    _check(lib.mcpt_host_bvh_check(C.byref(d), C.c_int(int(lights_only)), out, C.byref(sah)))
    return dict(nodes=out[0], leaves=out[1], max_depth=out[2], errors=out[3], sah=sah.value)


def host_light_tree_check(a: SceneArrays):
    """Build the light hierarchy of LIGHT_TREE on the host and validate it structurally (no GPU needed)."""
    lib = load()
    d, keep = make_desc(a)
    out = (C.c_uint32 * 4)()
    _check(lib.mcpt_host_light_tree_check(C.byref(d), out))
    return dict(nodes=out[0], max_depth=out[1], errors=out[2], lights=out[3])


def host_light_lists(a: SceneArrays):
    """Per-primitive survivor lists of the light sweeps (csrc/light_lists.h) built on the host, as a [P, L] membership
    matrix (no GPU needed); None when the builder declines (few lights, too many primitives)."""
    lib = load()
    d, keep = make_desc(a)
    mask = np.zeros((len(a.positions), len(a.light_prim)), np.uint8)
    out = (C.c_uint32 * 3)()
    _check(lib.mcpt_host_light_lists(C.byref(d), _p(mask), out))
    if out[0] == 0:
        return None
    return dict(mask=mask.astype(bool), lists=out[0], max_len=out[1], mean_fraction=out[2] / 1000.0)


def measure_fp32(device: int = 0):
    """(scalar FFMA, packed FFMA2) FP32 ceilings of the device in TFLOP/s, measured (mcpt_measure_fp32)."""
    a, b = C.c_double(), C.c_double()
    _check(load().mcpt_measure_fp32(C.c_int(device), C.byref(a), C.byref(b)))
    return a.value, b.value


class MultiScene:
    """A scene replicated on several GPUs of this process (mcpt_multi_create): spp sharded inside the library, one NCCL
    reduce of the films, host film out."""

    def __init__(self, arrays: SceneArrays, devices, build: int = BUILD_HOST_SAH):
        self.lib = load()
        d, self._keep = make_desc(arrays)
        devs = (C.c_int32 * len(devices))(*devices)
        self.handle = C.c_void_p()
        self.n = len(devices)
        _check(self.lib.mcpt_multi_create(C.byref(d), devs, C.c_int32(len(devices)), C.c_uint32(build), C.byref(self.handle)))

    def close(self):
        if getattr(self, "handle", None):
            self.lib.mcpt_multi_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self):
        _check(self.lib.mcpt_multi_upload(self.handle))

    def render(self, cam: Camera, strategy: int, spp: int, max_depth: int = 0, flags: int = 0, seed: int = 1,
               spp_offset: int = 0, spp_total: int = 0, batch_spp: int = 0, p_rr: float = 0.0, out=None,
               leaf_mode: int = LEAF_FAST):
        if out is None:
            out = np.zeros((cam.height, cam.width, 3), np.float32)
        st = Stats()
        p = Scene._params(None, cam, strategy, spp, max_depth, flags, seed, spp_offset, spp_total, batch_spp, p_rr, leaf_mode)
        cc = camera_c(cam)
        _check(self.lib.mcpt_multi_render(self.handle, C.byref(cc), C.byref(p), _p(out), C.byref(st)))
        return out, st.as_dict()


class Scene:
    """A scene resident on one GPU (mcpt_scene_create)."""

    def __init__(self, arrays: SceneArrays, device: int = 0, build: int = BUILD_HOST_SAH):
        self.lib = load()
        self.arrays = arrays
        d, self._keep = make_desc(arrays)
        self.P, self.L = d.num_prims, d.num_lights
        self.handle = C.c_void_p()
        _check(self.lib.mcpt_scene_create_ex(C.byref(d), C.c_int(device), C.c_uint32(build), C.byref(self.handle)))
        self.device = device

    def close(self):
        if getattr(self, "handle", None):
            self.lib.mcpt_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- render -------------------------------------------------------------------------------
    def _params(self, cam, strategy, spp, max_depth, flags, seed, spp_offset, spp_total, batch_spp, p_rr, leaf_mode=LEAF_FAST):
        p = RenderParams()
        p.width, p.height = cam.width, cam.height
        p.spp, p.spp_offset, p.spp_total = spp, spp_offset, spp_total or spp
        p.max_depth, p.strategy, p.flags, p.seed, p.p_rr, p.batch_spp = max_depth, strategy, flags, seed, p_rr, batch_spp
        p.leaf_mode, p.reserved = leaf_mode, 0
        return p

    def render(self, cam: Camera, strategy: int, spp: int, max_depth: int = 0, flags: int = 0, seed: int = 1,
               spp_offset: int = 0, spp_total: int = 0, batch_spp: int = 0, p_rr: float = 0.0, out=None,
               leaf_mode: int = LEAF_FAST):
        """mcpt_render: host image out (H, W, 3) float32 linear radiance + stats dict.  ``out``: an existing C-contiguous
        float32 array of that shape (e.g. a view of pinned memory) to receive the film instead of a fresh one."""
        if out is None:
            out = np.zeros((cam.height, cam.width, 3), np.float32)
        elif out.dtype != np.float32 or out.size != cam.height * cam.width * 3 or not out.flags.c_contiguous:
            raise ValueError("out must be a C-contiguous float32 array of H*W*3 elements")
        st = Stats()
        p = self._params(cam, strategy, spp, max_depth, flags, seed, spp_offset, spp_total, batch_spp, p_rr, leaf_mode)
        cc = camera_c(cam)
        _check(self.lib.mcpt_render(self.handle, C.byref(cc), C.byref(p), _p(out), C.byref(st)))
        return out, st.as_dict()

    def render_device(self, cam: Camera, strategy: int, spp: int, d_out_ptr: int, stream: int = 0, max_depth: int = 0,
                      flags: int = 0, seed: int = 1, spp_offset: int = 0, spp_total: int = 0, batch_spp: int = 0,
                      p_rr: float = 0.0, leaf_mode: int = LEAF_FAST):
        """mcpt_render_device: writes W*H*3 floats at device pointer ``d_out_ptr`` on CUDA stream ``stream``."""
        st = Stats()
        p = self._params(cam, strategy, spp, max_depth, flags, seed, spp_offset, spp_total, batch_spp, p_rr, leaf_mode)
        cc = camera_c(cam)
        _check(self.lib.mcpt_render_device(self.handle, C.byref(cc), C.byref(p), C.c_void_p(d_out_ptr),
                                           C.c_void_p(stream), C.byref(st)))
        return st.as_dict()

    def upload(self):
        _check(self.lib.mcpt_scene_upload(self.handle))

    def device_bytes(self) -> int:
        return int(self.lib.mcpt_scene_device_bytes(self.handle))

    def tonemap(self, rgb, max_radiance=380.0, gamma=0.25):
        rgb = np.ascontiguousarray(rgb, np.float32)
        out = np.zeros(rgb.shape, np.uint8)
        _check(self.lib.mcpt_tonemap(self.handle, _p(rgb), C.c_size_t(rgb.size // 3), C.c_double(max_radiance),
                                     C.c_double(gamma), _p(out)))
        return out

    # ---- introspection ------------------------------------------------------------------------
    def geom_normals(self):
        out = np.zeros((self.P, 3))
        _check(self.lib.mcpt_scene_geom_normals(self.handle, _p(out)))
        return out

    def light_areas(self):
        out = np.zeros(self.L)
        _check(self.lib.mcpt_scene_light_areas(self.handle, _p(out)))
        return out

    def build_info(self):
        """ms spent on the scene BVH tree / on the whole scene_create, and the builder used (BUILD_*)."""
        t = C.c_double(); a = C.c_double(); b = C.c_int32()
        _check(self.lib.mcpt_scene_build_info(self.handle, C.byref(t), C.byref(a), C.byref(b)))
        return dict(tree_ms=t.value, total_ms=a.value, builder=b.value)

    def trace_layout(self):
        """Wide BVH layout the trace kernels use for this scene (4 or 8) and the calibration times (ms) of both."""
        w = C.c_int32(); ms = (C.c_float * 2)(); rs = (C.c_float * 2)()
        _check(self.lib.mcpt_scene_trace_layout(self.handle, C.byref(w), ms, rs))
        return dict(width=w.value, calib_ms_bvh4=ms[0], calib_ms_bvh8=ms[1], first_bounce_ms_bvh4=rs[0], first_bounce_ms_bvh8=rs[1])

    def light_lists(self):
        """Per-primitive survivor lists of the light sweeps (csrc/light_lists.h); lists == 0: not in use for this scene."""
        n = C.c_uint32(); e = C.c_uint64(); f = C.c_double(); ms = C.c_double()
        _check(self.lib.mcpt_scene_light_lists(self.handle, C.byref(n), C.byref(e), C.byref(f), C.byref(ms)))
        return dict(lists=n.value, entries=e.value, mean_fraction=f.value, build_ms=ms.value)

    def light_list_of_prim(self, prim: int):
        """light indices on the survivor list of `prim`, or None when it has no list"""
        out = np.zeros(max(1, len(self.arrays.light_prim)), np.int32); n = C.c_int32()
        _check(self.lib.mcpt_scene_light_list_of_prim(self.handle, C.c_uint32(prim), _p(out), C.byref(n)))
        return None if n.value < 0 else out[:n.value].copy()

    def bvh_check(self) -> int:
        """violations found in the scene BVH this scene holds (0 = sound)"""
        e = C.c_uint32()
        _check(self.lib.mcpt_scene_bvh_check(self.handle, C.byref(e)))
        return e.value

    def bvh_info(self, lights=False):
        n = C.c_uint32(); l = C.c_uint32(); d = C.c_uint32(); s = C.c_double()
        _check(self.lib.mcpt_scene_bvh_info(self.handle, int(lights), C.byref(n), C.byref(l), C.byref(d), C.byref(s)))
        return dict(nodes=n.value, leaves=l.value, max_depth=d.value, sah=s.value)

    # ---- parity entry points (same signatures as oracle.OracleLib probes) ---------------------
    def intersect_tri(self, ro, rd, prim, precision=FP64):
        ro, rd, prim = _f64(ro), _f64(rd), _i32(prim)
        n = len(prim)
        hit = np.zeros(n, np.int32); bgt = np.zeros((n, 3))
        _check(self.lib.mcpt_intersect_tri_batch(self.handle, C.c_size_t(n), _p(ro), _p(rd), _p(prim), precision, _p(hit), _p(bgt)))
        return hit, bgt

    def closest(self, ro, rd, exclude=None, light_only=False, precision=FP64):
        ro, rd = _f64(ro), _f64(rd)
        n = len(ro)
        ex = _i32(exclude) if exclude is not None else np.full(n, -1, np.int32)
        prim = np.zeros(n, np.int32); bgt = np.zeros((n, 3))
        _check(self.lib.mcpt_intersect_batch(self.handle, C.c_size_t(n), _p(ro), _p(rd), _p(ex), int(light_only), precision,
                                             _p(prim), _p(bgt)))
        return prim, bgt

    def visibility(self, ro, rd, exclude, light_prim, leaf_mode=LEAF_FAST):
        """mcpt_visibility_batch: the renderer's NEE visibility kernels on explicit rays (1 visible, 0 occluded,
        -1 the ray misses its light triangle)."""
        ro, rd = _f64(ro), _f64(rd)
        n = len(ro)
        ex = _i32(exclude) if exclude is not None else np.full(n, -1, np.int32)
        lp = _i32(light_prim)
        vis = np.zeros(n, np.int32)
        _check(self.lib.mcpt_visibility_batch(self.handle, C.c_size_t(n), _p(ro), _p(rd), _p(ex), _p(lp), int(leaf_mode), _p(vis)))
        return vis

    def visibility_noalloc(self, d_org: int, d_dir: int, d_ids: int, n: int, d_visible: int, leaf_mode=LEAF_FAST, stream: int = 0):
        """mcpt_visibility_noalloc: the same test on the caller's DEVICE buffers (raw pointers: float4 org, float4 dir,
        int2 exclude | light prim, int32 visible out), scratch in a fixed array inside the library -- no allocation,
        at most NOALLOC_MAX_RAYS rays, launched on ``stream`` without a synchronise."""
        _check(self.lib.mcpt_visibility_noalloc(self.handle, C.c_size_t(n), C.c_void_p(d_org), C.c_void_p(d_dir), C.c_void_p(d_ids),
                                                int(leaf_mode), C.c_void_p(d_visible), C.c_void_p(stream)))

    def brdf_eval(self, N, wi, wr, kd, ks, ns, precision=FP64):
        N, wi, wr, kd, ks, ns = map(_f64, (N, wi, wr, kd, ks, ns))
        out = np.zeros_like(N)
        _check(self.lib.mcpt_phong_eval_batch(self.handle, C.c_size_t(len(ns)), _p(N), _p(wi), _p(wr), _p(kd), _p(ks), _p(ns),
                                              precision, _p(out)))
        return out

    def brdf_sample(self, N, wr, kd, ks, ns, u3, precision=FP64):
        N, wr, kd, ks, ns, u3 = map(_f64, (N, wr, kd, ks, ns, u3))
        d = np.zeros_like(N); pdf = np.zeros(len(ns))
        _check(self.lib.mcpt_phong_sample_batch(self.handle, C.c_size_t(len(ns)), _p(N), _p(wr), _p(kd), _p(ks), _p(ns), _p(u3),
                                                precision, _p(d), _p(pdf)))
        return d, pdf

    def brdf_pdf(self, N, wi, wr, kd, ks, ns, precision=FP64):
        N, wi, wr, kd, ks, ns = map(_f64, (N, wi, wr, kd, ks, ns))
        pdf = np.zeros(len(ns))
        _check(self.lib.mcpt_phong_pdf_batch(self.handle, C.c_size_t(len(ns)), _p(N), _p(wi), _p(wr), _p(kd), _p(ks), _p(ns),
                                             precision, _p(pdf)))
        return pdf

    def light_sample_uniform(self, u4, precision=FP64):
        u4 = _f64(u4)
        n = len(u4)
        prim = np.zeros(n, np.int32); pt = np.zeros((n, 3)); Le = np.zeros((n, 3)); pdf = np.zeros(n)
        _check(self.lib.mcpt_light_sample_uniform_batch(self.handle, C.c_size_t(n), _p(u4), precision, _p(prim), _p(pt), _p(Le), _p(pdf)))
        return prim, pt, Le, pdf

    def light_sample_spherical(self, x1, N, u3, query=None, precision=FP64):
        x1, N, u3 = map(_f64, (x1, N, u3))
        n = len(x1)
        q = _i32(query) if query is not None else np.full(n, -1, np.int32)
        prim = np.zeros(n, np.int32); pt = np.zeros((n, 3)); Le = np.zeros((n, 3)); pdf = np.zeros(n)
        wsum = np.zeros(n); nsurv = np.zeros(n, np.int32); qpdf = np.zeros(n)
        _check(self.lib.mcpt_light_sample_spherical_batch(self.handle, C.c_size_t(n), _p(x1), _p(N), _p(u3), _p(q), precision,
                                                          _p(prim), _p(pt), _p(Le), _p(pdf), _p(wsum), _p(nsurv), _p(qpdf)))
        return dict(prim=prim, point=pt, Le=Le, pdf=pdf, wsum=wsum, nsurv=nsurv, query_pdf=qpdf)

    def light_tree(self, x1, N, u, query_light=None):
        """LIGHT_TREE selection at vertices (x1, N): one draw per vertex with uniform u, plus the selection
        probability and FP32 spherical-triangle weight of light ``query_light`` (index into the light list)."""
        x1, N, u = map(_f64, (x1, N, u))
        n = len(x1)
        q = _i32(query_light) if query_light is not None else np.full(n, -1, np.int32)
        light = np.zeros(n, np.int32); prob = np.zeros(n); qprob = np.zeros(n); qw = np.zeros(n)
        _check(self.lib.mcpt_light_tree_batch(self.handle, C.c_size_t(n), _p(x1), _p(N), _p(u), _p(q), _p(light), _p(prob),
                                              _p(qprob), _p(qw)))
        return dict(light=light, prob=prob, query_prob=qprob, query_weight=qw)

    def light_tree_arrays(self):
        """The light hierarchy as built for this scene: nodes [n, 12] float32 (csrc/light_tree.h layout, the three
        child / parent words are int32 bit patterns) and the leaf node of every light."""
        n = C.c_uint32(); d = C.c_uint32()
        _check(self.lib.mcpt_scene_light_tree_info(self.handle, C.byref(n), C.byref(d)))
        nodes = np.zeros((n.value, 12), np.float32)
        leaf = np.zeros(len(self.arrays.light_prim), np.int32)
        _check(self.lib.mcpt_scene_light_tree_nodes(self.handle, _p(nodes), _p(leaf)))
        return nodes, leaf, d.value
