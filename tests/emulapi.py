"""TEST INFRASTRUCTURE ONLY — ctypes binding of tests/emul/libptemul.so: the product's
packed scene + device traversal header compiled for the host (see tests/emul/emul.cpp)."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from pathtracing_b200 import capi
from pathtracing_b200.capi import pt_light_desc, pt_material_desc, pt_medium_desc, pt_mesh_desc, pt_shape_desc, pt_texture_desc

_DIR = Path(__file__).resolve().parent / "emul"
_lib = None
_libs = {}


def load(name="libptemul.so"):
    """name: libptemul.so, or libptemul_stack20.so (the same code with a 20-row traversal stack)."""
    global _lib
    if name in _libs:
        return _libs[name]
    subprocess.run(["make", "-s", "-C", str(_DIR)], check=True)
    lib = C.CDLL(str(_DIR / name))
    vp, i, sz = C.c_void_p, C.c_int, C.c_size_t
    lib.emul_scene_create.restype = vp
    lib.emul_scene_destroy.argtypes = [vp]
    lib.emul_last_error.restype = C.c_char_p
    for name, desc in (("emul_scene_add_texture", pt_texture_desc), ("emul_scene_add_material", pt_material_desc),
                       ("emul_scene_add_medium", pt_medium_desc), ("emul_scene_add_mesh", pt_mesh_desc),
                       ("emul_scene_add_light", pt_light_desc)):
        f = getattr(lib, name)
        f.restype = i
        f.argtypes = [vp, C.POINTER(desc)]
    lib.emul_scene_add_shape.restype = i
    lib.emul_scene_add_shape.argtypes = [vp, C.POINTER(pt_shape_desc), C.POINTER(C.c_float)]
    lib.emul_scene_animate.restype = i
    lib.emul_scene_animate.argtypes = [vp, i, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.emul_mat4_inverse.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.emul_scene_add_blas.restype = i
    lib.emul_scene_add_blas.argtypes = [vp, C.POINTER(C.c_int32), i]
    lib.emul_scene_add_instance.restype = i
    lib.emul_scene_add_instance.argtypes = [vp, i, C.POINTER(C.c_float)]
    lib.emul_scene_commit.restype = i
    lib.emul_scene_commit.argtypes = [vp, i]
    lib.emul_scene_bounds.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.emul_node_count.argtypes = [vp]
    lib.emul_flatten_check.restype = C.c_longlong
    lib.emul_flatten_check.argtypes = [C.c_uint, C.c_uint, i, C.c_uint, i, C.POINTER(C.c_int)]
    lib.emul_stack_overflows.restype = C.c_ulonglong
    lib.emul_stack_overflows.argtypes = [vp]
    lib.emul_bvh4q_check.restype = i
    lib.emul_bvh4q_check.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_float)]
    lib.emul_trace_closest.restype = i
    lib.emul_trace_closest.argtypes = [vp, vp, sz, vp, vp, vp]
    lib.emul_trace_any.restype = i
    lib.emul_trace_any.argtypes = [vp, vp, sz, vp]
    _libs[name] = lib
    if name == "libptemul.so":
        _lib = lib
    return lib


class EmulScene:
    def __init__(self, desc, bvh_width=4, lib_name="libptemul.so"):
        self.lib = load(lib_name)
        self.h = self.lib.emul_scene_create()
        self.desc = desc
        desc.replay(self)
        self._ck(self.lib.emul_scene_commit(self.h, bvh_width))

    def _ck(self, rc):
        if rc < 0:
            raise RuntimeError(f"emul error {rc}: {self.lib.emul_last_error().decode()}")
        return rc

    def add_texture(self, d):
        return self._ck(self.lib.emul_scene_add_texture(self.h, C.byref(d)))

    def add_material(self, d):
        return self._ck(self.lib.emul_scene_add_material(self.h, C.byref(d)))

    def add_medium(self, d):
        return self._ck(self.lib.emul_scene_add_medium(self.h, C.byref(d)))

    def add_mesh(self, d):
        return self._ck(self.lib.emul_scene_add_mesh(self.h, C.byref(d)))

    def add_blas(self, ids):
        return self._ck(self.lib.emul_scene_add_blas(self.h, ids.ctypes.data_as(C.POINTER(C.c_int32)), len(ids)))

    def add_instance(self, blas_id, xf):
        p = xf.ctypes.data_as(C.POINTER(C.c_float)) if xf is not None else None
        return self._ck(self.lib.emul_scene_add_instance(self.h, blas_id, p))

    def add_shape(self, d, xf=None):
        p = xf.ctypes.data_as(C.POINTER(C.c_float)) if xf is not None else None
        return self._ck(self.lib.emul_scene_add_shape(self.h, C.byref(d), p))

    def animate(self, prim_id, direction, time_bounds):
        return self._ck(self.lib.emul_scene_animate(self.h, prim_id, (C.c_float * 3)(*direction), (C.c_float * 2)(*time_bounds)))

    def add_light(self, d):
        return self._ck(self.lib.emul_scene_add_light(self.h, C.byref(d)))

    def set_medium(self, m):
        return 0

    def bounds(self):
        lo, hi = (C.c_float * 3)(), (C.c_float * 3)()
        self.lib.emul_scene_bounds(self.h, lo, hi)
        return np.array(lo[:], np.float32), np.array(hi[:], np.float32)

    def node_count(self):
        return self.lib.emul_node_count(self.h)

    def stack_overflows(self):
        return int(self.lib.emul_stack_overflows(self.h))

    def bvh4q_check(self):
        n, worst = C.c_int(0), C.c_float(0)
        bad = self.lib.emul_bvh4q_check(self.h, C.byref(n), C.byref(worst))
        return bad, n.value, worst.value

    def trace_closest(self, rays, want_inst=True):
        rays = np.ascontiguousarray(rays, dtype=capi.RAY_DTYPE)
        hits = np.empty(len(rays), dtype=capi.HIT_DTYPE)
        inst = np.empty(len(rays), dtype=np.uint32) if want_inst else None
        stats = np.zeros(4, np.uint64)
        self._ck(self.lib.emul_trace_closest(self.h, rays.ctypes.data, len(rays), hits.ctypes.data,
                                             inst.ctypes.data if want_inst else None, stats.ctypes.data))
        self.last_stats = dict(zip(("tlas_nodes", "blas_nodes", "instances", "triangles"), (int(x) for x in stats)))
        return hits, inst

    def trace_any(self, rays):
        rays = np.ascontiguousarray(rays, dtype=capi.RAY_DTYPE)
        occ = np.empty(len(rays), dtype=np.uint8)
        self._ck(self.lib.emul_trace_any(self.h, rays.ctypes.data, len(rays), occ.ctypes.data))
        return occ

    def close(self):
        if self.h:
            self.lib.emul_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
