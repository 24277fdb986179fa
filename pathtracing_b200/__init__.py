"""pathtracing_b200 — B200-native wavefront path tracer behind the C-ABI of include/pt_b200.h.

The product is the shared library `libptb200.so` (hand-written sm_100a CUDA + C++ host
code, pathtracing_b200/csrc).  The Python modules here are thin plumbing: ctypes
bindings (capi), a scene-description recorder (scene) and seeded synthetic inputs (synth).
"""
from . import capi  # noqa: F401
from .capi import PtError  # noqa: F401
from .scene import GpuScene, SceneDesc, camera_desc, render_params  # noqa: F401
