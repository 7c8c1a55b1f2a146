"""monte_carlo_path_tracing_b200 -- B200 (sm_100a) wavefront path tracer behind a C ABI.

Drop-in for the render loop of luotong96/Monte_Carlo_Path_Tracing (main.cpp:557-588): see
include/mcpt.h for the boundary, csrc/ for the CUDA kernels, api.py for the ctypes binding,
scenes.py for the procedural scene files, dist.py for spp sharding across GPUs.
"""
from . import scenes  # noqa: F401
from .api import (BRDF_ONLY, MIS, NEE_SPHERICAL, NEE_UNIFORM, McptError, Scene, build, host_bvh_check,  # noqa: F401
                  load, load_scene_files)
