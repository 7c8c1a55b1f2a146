"""tests/golden/course-style/*.npz: the flattened arrays the COMPILED REFERENCE (tiny_obj_loader + pugixml + Myobj / Mylight
accessors, oracle/_ref/libmcpt_ref.so) produces from two hand-written scenes in the style of a DCC export -- smoothing
groups, vt indices, map_Kd / map_Bump / illum / Ni / d / Tr lines, v/vt/vn, v//vn and relative indices mixed, usemtl and g
in mid-object, quads and a pentagon; the second one with CRLF line ends, tabs, vp / l / p statements, exponent notation and
no final newline.  Run here (where /root/reference exists): python tests/golden/make_course_golden.py <name>
(one scene per process: the reference never clears its caches)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle  # noqa: E402

FIELDS = ("positions", "normals", "material_id", "shape_offset", "mat_kd", "mat_ks", "mat_ns", "light_prim", "light_class", "class_radiance")

if __name__ == "__main__":
    name = sys.argv[1]
    d = os.path.join(HERE, "course-style")
    ref = oracle.RefLib(os.path.join(d, name + ".obj"), os.path.join(d, name + ".xml"))
    a, gn, area = ref.flatten()
    np.savez_compressed(os.path.join(d, name + ".npz"), geom_normals=gn, light_area=area, **{f: getattr(a, f) for f in FIELDS})
    print(name, "prims", len(a.material_id), "lights", len(a.light_prim))
