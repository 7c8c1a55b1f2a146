"""Procedural scene generators (OBJ + MTL + XML text, or flat arrays).

The reference's example scenes (``example-scenes-cg23.zip``) are not in the
checkout (/root/reference/.MISSING_LARGE_BLOBS:1), so every scene used by the
tests and the bench is generated here, deterministically, in the exact file
formats the reference loads:

* OBJ/MTL as read by tiny_obj_loader through ``Myobj::read`` (Myobj.cpp:10-28):
  every face carries ``vn`` normals and a ``usemtl`` because the reference
  indexes ``attrib.normals`` / ``materials`` unchecked (Myobj.cpp:671-673,
  main.cpp:292).
* XML with document-level ``<light mtlname= radiance="r,g,b"/>`` nodes as
  iterated by ``Mylight::read`` (Mylight.cpp:21-25) plus the ``<camera>`` node
  shown in README.md:339-343 (the reference hard-codes the camera, F7).

Scenes:
  cornell_box()   C1  classic Cornell data, small baked rotation (F11)
  veach_mis()     C2  four glossy Phong plates + four icosphere lights + fill
  interior()      C3  stand-in for the unavailable bathroom scene (>=100k tris)
  triangle_soup() C4  1M random triangles + 64 emissive triangles (flat arrays)
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass, field

import numpy as np


# --------------------------------------------------------------------------
# containers
# --------------------------------------------------------------------------
@dataclass
class Camera:
    """Pin-hole camera in the reference's own parametrisation (main.cpp:546-553).

    ``eye`` and ``w`` (= lookat - eye, NOT normalised: |w| enters pixellen) and
    ``up``; ``pixellen = tan(fovy/2) * |w| / (height/2)`` (radians done right;
    the reference's literal ``tan(20.1143/360)`` is F7).
    """
    eye: tuple
    lookat: tuple
    up: tuple
    fovy_deg: float
    width: int
    height: int

    @property
    def w(self):
        return tuple(l - e for l, e in zip(self.lookat, self.eye))

    @property
    def pixellen(self):
        wn = math.sqrt(sum(c * c for c in self.w))
        return math.tan(math.radians(self.fovy_deg) * 0.5) * wn / (self.height * 0.5)


@dataclass
class Material:
    name: str
    kd: tuple = (0.0, 0.0, 0.0)
    ks: tuple = (0.0, 0.0, 0.0)
    ns: float = 1.0


@dataclass
class SceneText:
    """A scene as the three text files the reference reads."""
    name: str
    obj: str
    mtl: str
    xml: str
    camera: Camera
    lights: dict = field(default_factory=dict)  # mtlname -> (r,g,b)

    def write(self, directory: str) -> str:
        """Write <dir>/<name>.{obj,mtl,xml}; return the path stem."""
        os.makedirs(directory, exist_ok=True)
        stem = os.path.join(directory, self.name)
        with open(stem + ".obj", "w") as f:
            f.write(self.obj)
        with open(stem + ".mtl", "w") as f:
            f.write(self.mtl)
        with open(stem + ".xml", "w") as f:
            f.write(self.xml)
        return stem


# --------------------------------------------------------------------------
# mesh builder
# --------------------------------------------------------------------------
class _Mesh:
    """Accumulates ``o`` groups of triangles with per-face-vertex normals."""

    def __init__(self):
        self.v = []      # list of (x,y,z)
        self.vn = []     # list of (x,y,z)
        self.groups = []  # (name, mtl, [(v0,v1,v2,n0,n1,n2)]) 0-based indices

    def begin(self, name, mtl):
        self.groups.append((name, mtl, []))

    def _add_v(self, p):
        self.v.append((float(p[0]), float(p[1]), float(p[2])))
        return len(self.v) - 1

    def _add_n(self, n):
        l = math.sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2])
        self.vn.append((n[0] / l, n[1] / l, n[2] / l))
        return len(self.vn) - 1

    def tri(self, a, b, c, na=None, nb=None, nc=None, toward=None, away=None):
        """Add a triangle.  Without explicit normals the face normal is used,
        oriented toward point ``toward`` / away from point ``away`` if given,
        else by the winding."""
        a = np.asarray(a, float); b = np.asarray(b, float); c = np.asarray(c, float)
        if na is None:
            n = np.cross(b - a, c - a)
            cen = (a + b + c) / 3.0
            if toward is not None and np.dot(n, np.asarray(toward, float) - cen) < 0:
                n = -n
            if away is not None and np.dot(n, np.asarray(away, float) - cen) > 0:
                n = -n
            na = nb = nc = n
        ia, ib, ic = self._add_v(a), self._add_v(b), self._add_v(c)
        ja, jb, jc = self._add_n(na), self._add_n(nb), self._add_n(nc)
        self.groups[-1][2].append((ia, ib, ic, ja, jb, jc))

    def quad(self, p0, p1, p2, p3, **kw):
        self.tri(p0, p1, p2, **kw)
        self.tri(p0, p2, p3, **kw)

    def grid_quad(self, p0, p1, p2, p3, nu, nv, **kw):
        """Bilinear patch p0->p1 (u) / p0->p3 (v) tessellated nu x nv."""
        p0, p1, p2, p3 = (np.asarray(p, float) for p in (p0, p1, p2, p3))

        def P(u, v):
            return (p0 * (1 - u) * (1 - v) + p1 * u * (1 - v) + p2 * u * v + p3 * (1 - u) * v)
        for i in range(nu):
            for j in range(nv):
                u0, u1 = i / nu, (i + 1) / nu
                v0, v1 = j / nv, (j + 1) / nv
                self.quad(P(u0, v0), P(u1, v0), P(u1, v1), P(u0, v1), **kw)

    def box(self, lo, hi, **kw):
        x0, y0, z0 = lo; x1, y1, z1 = hi
        cen = ((x0 + x1) / 2, (y0 + y1) / 2, (z0 + z1) / 2)
        q = [
            ((x0, y0, z0), (x1, y0, z0), (x1, y1, z0), (x0, y1, z0)),
            ((x0, y0, z1), (x1, y0, z1), (x1, y1, z1), (x0, y1, z1)),
            ((x0, y0, z0), (x0, y0, z1), (x0, y1, z1), (x0, y1, z0)),
            ((x1, y0, z0), (x1, y0, z1), (x1, y1, z1), (x1, y1, z0)),
            ((x0, y0, z0), (x1, y0, z0), (x1, y0, z1), (x0, y0, z1)),
            ((x0, y1, z0), (x1, y1, z0), (x1, y1, z1), (x0, y1, z1)),
        ]
        for f in q:
            self.quad(*f, away=cen, **kw)

    def icosphere(self, center, radius, subdiv, smooth=True):
        verts, faces = _icosphere(subdiv)
        c = np.asarray(center, float)
        for f in faces:
            a, b, cc = (verts[i] for i in f)
            pa, pb, pc = c + radius * a, c + radius * b, c + radius * cc
            if smooth:
                self.tri(pa, pb, pc, na=a, nb=b, nc=cc)
            else:
                self.tri(pa, pb, pc, away=c)

    # ---- transforms / output
    def transform(self, R):
        R = np.asarray(R, float)
        self.v = [tuple(R @ np.asarray(p)) for p in self.v]
        self.vn = [tuple(R @ np.asarray(n)) for n in self.vn]

    def num_tris(self):
        return sum(len(g[2]) for g in self.groups)

    def to_obj(self, mtllib):
        out = ["# generated by monte_carlo_path_tracing_b200.scenes", "mtllib %s" % mtllib]
        # %.9g round-trips an IEEE float32 exactly; tinyobj's real_t is float (F2)
        f32 = np.float32
        out += ["v %.9g %.9g %.9g" % tuple(f32(c) for c in p) for p in self.v]
        out += ["vn %.9g %.9g %.9g" % tuple(f32(c) for c in n) for n in self.vn]
        for name, mtl, tris in self.groups:
            out.append("o %s" % name)
            out.append("usemtl %s" % mtl)
            out += ["f %d//%d %d//%d %d//%d" % (a + 1, ja + 1, b + 1, jb + 1, c + 1, jc + 1)
                    for (a, b, c, ja, jb, jc) in tris]
        return "\n".join(out) + "\n"


def _icosphere(subdiv):
    t = (1.0 + math.sqrt(5.0)) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t),
         (0, -1, -t), (0, 1, -t), (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    v = [np.asarray(p, float) / np.linalg.norm(p) for p in v]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4),
         (11, 10, 2), (10, 7, 6), (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8),
         (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    for _ in range(subdiv):
        cache = {}
        nf = []

        def mid(i, j):
            key = (min(i, j), max(i, j))
            if key not in cache:
                m = v[i] + v[j]
                v.append(m / np.linalg.norm(m))
                cache[key] = len(v) - 1
            return cache[key]
        for (a, b, c) in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    return v, f


def _mtl_text(materials):
    out = ["# generated by monte_carlo_path_tracing_b200.scenes"]
    for m in materials:
        out += ["newmtl %s" % m.name,
                "Kd %.9g %.9g %.9g" % tuple(np.float32(c) for c in m.kd),
                "Ks %.9g %.9g %.9g" % tuple(np.float32(c) for c in m.ks),
                "Ns %.9g" % np.float32(m.ns), ""]
    return "\n".join(out) + "\n"


def _xml_text(cam: Camera, lights: dict):
    out = ['<?xml version="1.0" encoding="utf-8"?>',
           '<camera type="perspective" width="%d" height="%d" fovy="%.17g">' % (cam.width, cam.height, cam.fovy_deg),
           '\t<eye x="%.17g" y="%.17g" z="%.17g"/>' % tuple(cam.eye),
           '\t<lookat x="%.17g" y="%.17g" z="%.17g"/>' % tuple(cam.lookat),
           '\t<up x="%.17g" y="%.17g" z="%.17g"/>' % tuple(cam.up),
           '</camera>']
    for name, rad in lights.items():
        out.append('<light mtlname="%s" radiance="%.17g,%.17g,%.17g"/>' % ((name,) + tuple(rad)))
    return "\n".join(out) + "\n"


def _rot(ay_deg, az_deg):
    ay, az = math.radians(ay_deg), math.radians(az_deg)
    Ry = np.array([[math.cos(ay), 0, math.sin(ay)], [0, 1, 0], [-math.sin(ay), 0, math.cos(ay)]])
    Rz = np.array([[math.cos(az), -math.sin(az), 0], [math.sin(az), math.cos(az), 0], [0, 0, 1]])
    return Rz @ Ry


def _finish(name, mesh, materials, cam, lights, rot):
    if rot is not None:
        R = _rot(*rot)
        mesh.transform(R)
        cam = Camera(tuple(R @ np.asarray(cam.eye)), tuple(R @ np.asarray(cam.lookat)),
                     tuple(R @ np.asarray(cam.up)), cam.fovy_deg, cam.width, cam.height)
    return SceneText(name, mesh.to_obj(name + ".mtl"), _mtl_text(materials), _xml_text(cam, lights), cam, dict(lights))


# --------------------------------------------------------------------------
# C1: cornell box
# --------------------------------------------------------------------------
def cornell_box(width=512, height=512, rot=(0.5, 0.5), glossy_block=True, name="cornell-box") -> SceneText:
    """Classic Cornell-box geometry (32 triangles incl. 2 emissive), camera literal
    from main.cpp:512-513.  ``rot`` bakes a small rotation (degrees about y, z)
    into geometry and camera so no shading normal is exactly (-1,0,0), which NaNs
    the reference's tangent frame (F11, BRDF.cpp:60-67).  ``rot=None`` gives the
    axis-aligned variant (GPU-only robustness test)."""
    m = _Mesh()
    room_c = (278.0, 274.4, 279.6)
    m.begin("floor", "white")
    m.quad((552.8, 0, 0), (0, 0, 0), (0, 0, 559.2), (549.6, 0, 559.2), toward=room_c)
    m.begin("ceiling", "white")
    m.quad((556, 548.8, 0), (556, 548.8, 559.2), (0, 548.8, 559.2), (0, 548.8, 0), toward=room_c)
    m.begin("backwall", "white")
    m.quad((549.6, 0, 559.2), (0, 0, 559.2), (0, 548.8, 559.2), (556, 548.8, 559.2), toward=room_c)
    m.begin("rightwall", "green")
    m.quad((0, 0, 559.2), (0, 0, 0), (0, 548.8, 0), (0, 548.8, 559.2), toward=room_c)
    m.begin("leftwall", "red")
    m.quad((552.8, 0, 0), (549.6, 0, 559.2), (556, 548.8, 559.2), (556, 548.8, 0), toward=room_c)
    m.begin("light", "light")
    m.quad((343, 548.3, 227), (343, 548.3, 332), (213, 548.3, 332), (213, 548.3, 227), toward=room_c)
    sc = (186.0, 82.5, 169.0)
    m.begin("shortblock", "white")
    for q in [((130, 165, 65), (82, 165, 225), (240, 165, 272), (290, 165, 114)),
              ((290, 0, 114), (290, 165, 114), (240, 165, 272), (240, 0, 272)),
              ((130, 0, 65), (130, 165, 65), (290, 165, 114), (290, 0, 114)),
              ((82, 0, 225), (82, 165, 225), (130, 165, 65), (130, 0, 65)),
              ((240, 0, 272), (240, 165, 272), (82, 165, 225), (82, 0, 225))]:
        m.quad(*q, away=sc)
    tc = (368.5, 165.0, 351.0)
    m.begin("tallblock", "glossy" if glossy_block else "white")
    for q in [((423, 330, 247), (265, 330, 296), (314, 330, 456), (472, 330, 406)),
              ((423, 0, 247), (423, 330, 247), (472, 330, 406), (472, 0, 406)),
              ((472, 0, 406), (472, 330, 406), (314, 330, 456), (314, 0, 456)),
              ((314, 0, 456), (314, 330, 456), (265, 330, 296), (265, 0, 296)),
              ((265, 0, 296), (265, 330, 296), (423, 330, 247), (423, 0, 247))]:
        m.quad(*q, away=tc)
    mats = [Material("white", (0.725, 0.71, 0.68)),
            Material("red", (0.63, 0.065, 0.05)),
            Material("green", (0.14, 0.45, 0.091)),
            Material("glossy", (0.35, 0.35, 0.35), (0.4, 0.4, 0.4), 40.0),
            Material("light", (0.78, 0.78, 0.78))]
    cam = Camera((278.0, 273.0, -800.0), (278.0, 273.0, -799.0), (0, 1, 0), 39.3, width, height)
    lights = {"light": (34.0, 24.0, 8.0)}
    return _finish(name, m, mats, cam, lights, rot)


# --------------------------------------------------------------------------
# C2: veach-mis look-alike
# --------------------------------------------------------------------------
def veach_mis(width=1280, height=720, sphere_subdiv=2, plate_tess=(16, 4), rot=(0.37, 0.23),
              name="veach-mis") -> SceneText:
    """Look-alike of Veach's MIS scene with the README camera (README.md:339-343):
    four glossy Phong plates (Ns 10/50/500/5000, ks-dominant) under four icosphere
    lights of radii 0.03/0.1/0.3/0.9 with equal power, plus a fill light; floor
    and back wall tessellated so the facet count (~2.8k, 1282 emissive) is close
    to the 3092 facets of the course scene (README.md:212)."""
    m = _Mesh()
    up_c = (2.0, 6.0, 0.0)
    m.begin("floor", "diffuse")
    m.grid_quad((-8, 0, -12), (14, 0, -12), (14, 0, 12), (-8, 0, 12), 16, 16, toward=up_c)
    m.begin("backwall", "diffuse")
    m.grid_quad((-8, 0, -12), (-8, 0, 12), (-8, 14, 12), (-8, 14, -12), 16, 16, toward=up_c)
    # plates: (centre x, centre y, tilt degrees toward camera (+x), Ns)
    plates = [(6.2, 0.55, 14.0, 10.0), (3.9, 1.05, 25.0, 50.0), (1.7, 1.85, 37.0, 500.0), (-0.3, 2.95, 50.0, 5000.0)]
    half_w, half_d = 4.6, 0.95
    for k, (cx, cy, tilt, ns) in enumerate(plates):
        t = math.radians(tilt)
        # plate normal leans from +y toward +x; in-plane depth axis d is orthogonal to z and n
        n = np.array([math.sin(t), math.cos(t), 0.0])
        d = np.array([math.cos(t), -math.sin(t), 0.0])
        c = np.array([cx, cy, 0.0])
        z = np.array([0.0, 0.0, 1.0])
        p0 = c - half_d * d - half_w * z
        p1 = c + half_d * d - half_w * z
        p2 = c + half_d * d + half_w * z
        p3 = c - half_d * d + half_w * z
        m.begin("plate%d" % (k + 1), "plate%d" % (k + 1))
        m.grid_quad(p0, p1, p2, p3, plate_tess[1], plate_tess[0], toward=c + n)
    radii = [0.03, 0.1, 0.3, 0.9]
    zs = [-3.75, -1.25, 1.25, 3.75]
    for k, (r, z) in enumerate(zip(radii, zs)):
        m.begin("sphere%d" % (k + 1), "light%d" % (k + 1))
        m.icosphere((-1.8, 6.3, z), r, sphere_subdiv, smooth=True)
    m.begin("fill", "light0")
    m.quad((9.0, 13.0, -3.0), (15.0, 13.0, -3.0), (15.0, 13.0, 3.0), (9.0, 13.0, 3.0), toward=(4.0, 0.0, 0.0))
    mats = [Material("diffuse", (0.4, 0.4, 0.4))]
    for k, (_, _, _, ns) in enumerate(plates):
        mats.append(Material("plate%d" % (k + 1), (0.07, 0.09, 0.13), (0.6, 0.6, 0.6), ns))
    lights = {"light0": (6.0, 6.0, 6.0)}
    cols = [(1.0, 0.35, 0.35), (1.0, 1.0, 0.35), (0.35, 1.0, 0.35), (0.35, 0.45, 1.0)]
    for k, r in enumerate(radii):
        mats.append(Material("light%d" % (k + 1), (0.0, 0.0, 0.0)))
        s = 1.0 / (r * r)
        lights["light%d" % (k + 1)] = tuple(round(c * s, 4) for c in cols[k])
    mats.append(Material("light0", (0.0, 0.0, 0.0)))
    cam = Camera((28.2792, 5.2, 1.23612e-06), (0.0, 2.8, 0.0), (0.0, 1.0, 0.0), 20.1143, width, height)
    return _finish(name, m, mats, cam, lights, rot)


# --------------------------------------------------------------------------
# C3: large interior stand-in
# --------------------------------------------------------------------------
def interior(width=1280, height=720, grid=(7, 5), sphere_subdiv=4, rot=(0.41, 0.29), name="interior") -> SceneText:
    """Stand-in for the unavailable bathroom scene (main.cpp:21,516-517): a room
    with a grid of smooth-shaded icospheres and boxes on pedestals, tessellated
    walls and 8 emissive ceiling panels (16 emissive triangles).  Default
    parameters give ~190k triangles."""
    m = _Mesh()
    X, Y, Z = 24.0, 9.0, 18.0
    rc = (X / 2, Y / 2, Z / 2)
    m.begin("floor", "tile")
    m.grid_quad((0, 0, 0), (X, 0, 0), (X, 0, Z), (0, 0, Z), 48, 36, toward=rc)
    m.begin("ceiling", "wall")
    m.grid_quad((0, Y, 0), (X, Y, 0), (X, Y, Z), (0, Y, Z), 24, 18, toward=rc)
    m.begin("walls", "wall")
    m.grid_quad((0, 0, 0), (0, 0, Z), (0, Y, Z), (0, Y, 0), 18, 9, toward=rc)
    m.grid_quad((X, 0, 0), (X, 0, Z), (X, Y, Z), (X, Y, 0), 18, 9, toward=rc)
    m.grid_quad((0, 0, Z), (X, 0, Z), (X, Y, Z), (0, Y, Z), 24, 9, toward=rc)
    m.grid_quad((0, 0, 0), (X, 0, 0), (X, Y, 0), (0, Y, 0), 24, 9, toward=rc)
    gx, gz = grid
    k = 0
    for i in range(gx):
        for j in range(gz):
            cx = 2.5 + i * (X - 5.0) / max(gx - 1, 1)
            cz = 4.0 + j * (Z - 6.0) / max(gz - 1, 1)
            m.begin("ped%d" % k, "wood")
            m.box((cx - 0.7, 0, cz - 0.7), (cx + 0.7, 1.0, cz + 0.7))
            mat = ("ceramic", "chrome", "tile")[k % 3]
            m.begin("orb%d" % k, mat)
            m.icosphere((cx, 1.0 + 0.85, cz), 0.85, sphere_subdiv, smooth=True)
            k += 1
    m.begin("panels", "panel")
    for i in range(4):
        for j in range(2):
            x0 = 2.5 + i * 5.5; z0 = 4.0 + j * 7.0
            m.quad((x0, Y - 0.02, z0), (x0 + 2.5, Y - 0.02, z0), (x0 + 2.5, Y - 0.02, z0 + 3.0), (x0, Y - 0.02, z0 + 3.0),
                   toward=rc)
    mats = [Material("tile", (0.55, 0.6, 0.62), (0.15, 0.15, 0.15), 60.0),
            Material("wall", (0.7, 0.68, 0.62)),
            Material("wood", (0.45, 0.3, 0.16), (0.05, 0.05, 0.05), 20.0),
            Material("ceramic", (0.6, 0.6, 0.65), (0.3, 0.3, 0.3), 200.0),
            Material("chrome", (0.05, 0.05, 0.05), (0.8, 0.8, 0.8), 2000.0),
            Material("panel", (0.0, 0.0, 0.0))]
    cam = Camera((1.2, 4.6, 0.9), (14.0, 1.8, 10.5), (0, 1, 0), 55.0, width, height)
    lights = {"panel": (18.0, 17.0, 15.0)}
    return _finish(name, m, mats, cam, lights, rot)


# --------------------------------------------------------------------------
# C4: synthetic triangle soup as flat arrays
# --------------------------------------------------------------------------
@dataclass
class SceneArrays:
    """A scene in the flat layout ``mcpt_scene_create`` takes (include/mcpt.h)."""
    positions: np.ndarray    # [P,3,3] float32
    normals: np.ndarray      # [P,3,3] float32
    material_id: np.ndarray  # [P] int32
    shape_offset: np.ndarray  # [S+1] uint32
    mat_kd: np.ndarray       # [M,3] float32
    mat_ks: np.ndarray       # [M,3] float32
    mat_ns: np.ndarray       # [M] float32
    light_prim: np.ndarray   # [L] int32, reference order (class asc, then prim asc)
    light_class: np.ndarray  # [L] int32
    class_radiance: np.ndarray  # [C,3] float64
    camera: Camera = None


def triangle_soup(num_tris=1_000_000, num_lights=64, seed=12345, width=1024, height=1024) -> SceneArrays:
    """``num_tris`` triangles with vertices c + U(-0.5,0.5)^3 * 0.5,
    c ~ U(0,100)^3, vn = face normal, diffuse grey kd=0.5, plus ``num_lights``
    emissive triangles (scale 2, radiance 50) at U(0,100)^3; camera at
    (50,50,-150) looking +z, fovy 40 deg.  BVH-traversal stress scene."""
    rng = np.random.Generator(np.random.PCG64(seed))
    c = rng.uniform(0.0, 100.0, size=(num_tris, 1, 3))
    v = c + rng.uniform(-0.5, 0.5, size=(num_tris, 3, 3)) * 0.5
    cl = rng.uniform(0.0, 100.0, size=(num_lights, 1, 3))
    vl = cl + rng.uniform(-0.5, 0.5, size=(num_lights, 3, 3)) * 2.0
    pos = np.concatenate([v, vl], axis=0).astype(np.float32)
    p64 = pos.astype(np.float64)
    n = np.cross(p64[:, 1] - p64[:, 0], p64[:, 2] - p64[:, 0])
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    n = n / np.maximum(ln, 1e-300)
    # orient every face toward the camera side (-z) so primary hits are front-facing
    flip = n[:, 2] > 0
    n[flip] = -n[flip]
    nor = np.repeat(n[:, None, :], 3, axis=1).astype(np.float32)
    P = num_tris + num_lights
    mat = np.zeros(P, np.int32)
    mat[num_tris:] = 1
    cam = Camera((50.0, 50.0, -150.0), (50.0, 50.0, -149.0), (0, 1, 0), 40.0, width, height)
    return SceneArrays(
        positions=pos, normals=nor, material_id=mat,
        shape_offset=np.array([0, num_tris, P], np.uint32),
        mat_kd=np.array([[0.5, 0.5, 0.5], [0, 0, 0]], np.float32),
        mat_ks=np.zeros((2, 3), np.float32), mat_ns=np.ones(2, np.float32),
        light_prim=np.arange(num_tris, P, dtype=np.int32),
        light_class=np.zeros(num_lights, np.int32),
        class_radiance=np.array([[50.0, 50.0, 50.0]], np.float64), camera=cam)


SCENES = {"cornell-box": cornell_box, "veach-mis": veach_mis, "interior": interior}


def write_scene(name: str, directory: str, **kw) -> str:
    """Generate scene ``name`` into ``directory`` (cached by file existence); return stem."""
    stem = os.path.join(directory, name)
    if not all(os.path.exists(stem + e) for e in (".obj", ".mtl", ".xml")) or kw:
        SCENES[name](**kw).write(directory)
    return stem
