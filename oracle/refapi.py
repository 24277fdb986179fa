"""TEST INFRASTRUCTURE ONLY — ctypes binding of oracle/_ref/libptref_{strict,fast}.so,
the unmodified reference sources behind oracle/ref_driver.cpp.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; the product (pathtracing_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

from pathtracing_b200 import capi
from pathtracing_b200.capi import (pt_camera_desc, pt_light_desc, pt_material_desc, pt_medium_desc, pt_mesh_desc, pt_shape_desc,
                                   pt_render_params, pt_texture_desc)

_DIR = Path(__file__).resolve().parent
_libs: dict[str, C.CDLL] = {}


class ref_render_out(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("seconds", C.c_double)]


def available(flavour="strict") -> bool:
    return (_DIR / "_ref" / f"libptref_{flavour}.so").exists()


def load(flavour="strict") -> C.CDLL:
    """flavour: 'strict' (-ffp-contract=off, bit-level parity) | 'fast' (reference Release flags)."""
    if flavour in _libs:
        return _libs[flavour]
    p = _DIR / "_ref" / f"libptref_{flavour}.so"
    if not p.exists():
        raise RuntimeError(f"{p} missing: run `make -C oracle ref` where /root/reference is mounted")
    # RTLD_LOCAL: both flavours define the same symbols
    lib = C.CDLL(str(p), mode=os.RTLD_LOCAL | os.RTLD_NOW)
    vp, i, sz = C.c_void_p, C.c_int, C.c_size_t
    lib.ref_scene_create.restype = vp
    lib.ref_scene_destroy.argtypes = [vp]
    lib.ref_scene_destroy.restype = None
    lib.ref_hardware_threads.restype = i
    for name, desc in (("ref_scene_add_texture", pt_texture_desc), ("ref_scene_add_material", pt_material_desc),
                       ("ref_scene_add_medium", pt_medium_desc), ("ref_scene_add_mesh", pt_mesh_desc),
                       ("ref_scene_add_light", pt_light_desc)):
        f = getattr(lib, name)
        f.restype = i
        f.argtypes = [vp, C.POINTER(desc)]
    lib.ref_scene_add_shape.restype = i
    lib.ref_scene_add_shape.argtypes = [vp, C.POINTER(pt_shape_desc), C.POINTER(C.c_float)]
    lib.ref_scene_animate.restype = i
    lib.ref_scene_animate.argtypes = [vp, i, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.ref_scene_add_blas.restype = i
    lib.ref_scene_add_blas.argtypes = [vp, C.POINTER(C.c_int32), i]
    lib.ref_scene_add_instance.restype = i
    lib.ref_scene_add_instance.argtypes = [vp, i, C.POINTER(C.c_float)]
    lib.ref_scene_set_medium.restype = i
    lib.ref_scene_set_medium.argtypes = [vp, i]
    lib.ref_scene_commit.restype = i
    lib.ref_scene_commit.argtypes = [vp, i, i, i, i]
    lib.ref_scene_bounds.restype = i
    lib.ref_scene_bounds.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.ref_trace_closest.restype = i
    lib.ref_trace_closest.argtypes = [vp, vp, sz, vp, vp, i, C.POINTER(C.c_double)]
    lib.ref_trace_any.restype = i
    lib.ref_trace_any.argtypes = [vp, vp, sz, vp, i, C.POINTER(C.c_double)]
    lib.ref_render.restype = i
    lib.ref_render.argtypes = [vp, C.POINTER(pt_camera_desc), C.POINTER(pt_render_params), i, vp, vp, C.POINTER(ref_render_out)]
    lib.ref_render_stock.restype = i
    lib.ref_render_stock.argtypes = [vp, C.POINTER(pt_camera_desc), C.POINTER(pt_render_params), i, C.POINTER(C.c_double)]
    lib.ref_hdr_roundtrip.restype = i
    lib.ref_hdr_roundtrip.argtypes = [vp, i, i, i, vp]
    _libs[flavour] = lib
    return lib


class RefScene:
    """The reference's own Scene/BLASx/TLASx objects built from a SceneDesc."""

    def __init__(self, desc, bvh_width=4, light_sampler=1, id_mode=False, count_rays=False, flavour="strict"):
        self.lib = load(flavour)
        self.h = self.lib.ref_scene_create()
        self.desc = desc
        desc.replay(self)
        rc = self.lib.ref_scene_commit(self.h, bvh_width, light_sampler, int(id_mode), int(count_rays))
        if rc != 0:
            raise RuntimeError(f"ref_scene_commit failed: {rc}")
        self.id_mode = id_mode
        self.threads = self.lib.ref_hardware_threads()

    def _ck(self, rc):
        if rc < 0:
            raise RuntimeError(f"ref error {rc}")
        return rc

    def add_texture(self, d):
        return self._ck(self.lib.ref_scene_add_texture(self.h, C.byref(d)))

    def add_material(self, d):
        return self._ck(self.lib.ref_scene_add_material(self.h, C.byref(d)))

    def add_medium(self, d):
        return self._ck(self.lib.ref_scene_add_medium(self.h, C.byref(d)))

    def add_mesh(self, d):
        return self._ck(self.lib.ref_scene_add_mesh(self.h, C.byref(d)))

    def add_blas(self, ids):
        return self._ck(self.lib.ref_scene_add_blas(self.h, ids.ctypes.data_as(C.POINTER(C.c_int32)), len(ids)))

    def add_instance(self, blas_id, xf):
        p = xf.ctypes.data_as(C.POINTER(C.c_float)) if xf is not None else None
        return self._ck(self.lib.ref_scene_add_instance(self.h, blas_id, p))

    def add_shape(self, d, xf=None):
        p = xf.ctypes.data_as(C.POINTER(C.c_float)) if xf is not None else None
        return self._ck(self.lib.ref_scene_add_shape(self.h, C.byref(d), p))

    def animate(self, prim_id, direction, time_bounds):
        return self._ck(self.lib.ref_scene_animate(self.h, prim_id, (C.c_float * 3)(*direction), (C.c_float * 2)(*time_bounds)))

    def add_light(self, d):
        return self._ck(self.lib.ref_scene_add_light(self.h, C.byref(d)))

    def set_medium(self, m):
        return self._ck(self.lib.ref_scene_set_medium(self.h, m))

    def bounds(self):
        lo, hi = (C.c_float * 3)(), (C.c_float * 3)()
        self._ck(self.lib.ref_scene_bounds(self.h, lo, hi))
        return np.array(lo[:], np.float32), np.array(hi[:], np.float32)

    def trace_closest(self, rays, want_inst=False, threads=None):
        rays = np.ascontiguousarray(rays, dtype=capi.RAY_DTYPE)
        hits = np.empty(len(rays), dtype=capi.HIT_DTYPE)
        inst = np.empty(len(rays), dtype=np.uint32) if want_inst else None
        sec = C.c_double()
        self._ck(self.lib.ref_trace_closest(self.h, rays.ctypes.data, len(rays), hits.ctypes.data,
                                            inst.ctypes.data if want_inst else None,
                                            threads or self.threads, C.byref(sec)))
        self.last_seconds = sec.value
        return hits, inst

    def trace_any(self, rays, threads=None):
        rays = np.ascontiguousarray(rays, dtype=capi.RAY_DTYPE)
        occ = np.empty(len(rays), dtype=np.uint8)
        sec = C.c_double()
        self._ck(self.lib.ref_trace_any(self.h, rays.ctypes.data, len(rays), occ.ctypes.data,
                                        threads or self.threads, C.byref(sec)))
        self.last_seconds = sec.value
        return occ

    def render(self, cam, params, threads=None, want_sumsq=True):
        h, w = params.height, params.width
        rgbw = np.zeros((h, w, 4), np.float64)
        sumsq = np.zeros((h, w, 3), np.float64) if want_sumsq else None
        out = ref_render_out()
        self._ck(self.lib.ref_render(self.h, C.byref(cam), C.byref(params), threads or self.threads,
                                     rgbw.ctypes.data, sumsq.ctypes.data if want_sumsq else None, C.byref(out)))
        stats = {k: getattr(out, k) for k, _ in ref_render_out._fields_}
        return rgbw, sumsq, stats

    def close(self):
        if self.h:
            self.lib.ref_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
