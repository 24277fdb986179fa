"""ctypes binding of the C-ABI in include/pt_b200.h (libptb200.so).

This is plumbing for tests / bench.py; the product is the shared library.  There
is no CPU fallback: loading fails loudly when the CUDA extension has not been
built, and scene creation fails when no CUDA device is present.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_PKG = Path(__file__).resolve().parent
LIB_PATH_SMALL_STACK = Path(__file__).resolve().parent / "libptb200_stack20.so"    # test-only twin, csrc/Makefile
LIB_PATH = _PKG / "libptb200.so"
if os.environ.get("PTB_LIB"):            # kernel experiments: an alternative build of the same C-ABI
    LIB_PATH = Path(os.environ["PTB_LIB"])

PT_MISS = 0xFFFFFFFF

# ---- POD descriptors (mirror include/pt_b200.h) ------------------------------------


class pt_texture_desc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("color", C.c_float * 3), ("color_scale", C.c_float * 3),
                ("width", C.c_int32), ("height", C.c_int32), ("channels", C.c_int32),
                ("pixels", C.c_void_p), ("tex_a", C.c_int32), ("tex_b", C.c_int32),
                ("uv_scale", C.c_float * 2)]


class pt_material_desc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("tex", C.c_int32), ("norm", C.c_int32), ("rough", C.c_int32),
                ("metal", C.c_int32), ("alpha", C.c_int32), ("ri", C.c_float), ("albedo", C.c_float * 3),
                ("alpha_mode", C.c_int32), ("alpha_cutoff", C.c_float)]


class pt_medium_desc(C.Structure):
    _fields_ = [("sigma_a", C.c_float * 3), ("sigma_s", C.c_float * 3), ("g", C.c_float),
                ("density", C.c_float), ("Le", C.c_float * 3), ("le_density", C.c_float)]


class pt_mesh_desc(C.Structure):
    _fields_ = [("positions", C.c_void_p), ("normals", C.c_void_p), ("texcoords", C.c_void_p),
                ("tangents", C.c_void_p), ("indices", C.c_void_p),
                ("n_vertices", C.c_uint32), ("n_triangles", C.c_uint32),
                ("material", C.c_int32), ("emissive_tex", C.c_int32), ("medium", C.c_int32)]


class pt_shape_desc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("p", C.c_float * 3), ("u", C.c_float * 3), ("v", C.c_float * 3),
                ("radius", C.c_float), ("material", C.c_int32), ("medium", C.c_int32),
                ("emissive_tex", C.c_int32), ("one_sided", C.c_int32)]


class pt_light_desc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("color", C.c_float * 3), ("color_b", C.c_float * 3),
                ("scale", C.c_float), ("dir_or_pos", C.c_float * 3), ("tex", C.c_int32),
                ("power_func", C.c_int32)]


class pt_camera_desc(C.Structure):
    _fields_ = [("look_from", C.c_float * 3), ("look_at", C.c_float * 3), ("fov", C.c_float),
                ("focus_angle", C.c_float), ("focus_distance", C.c_float), ("medium", C.c_int32),
                ("shutter_start", C.c_float), ("shutter_end", C.c_float)]


class pt_render_params(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("max_depth", C.c_int32),
                ("integrator", C.c_int32), ("sampler", C.c_int32), ("strata_x", C.c_int32),
                ("strata_y", C.c_int32), ("filter", C.c_int32), ("filter_radius", C.c_float * 2),
                ("filter_param", C.c_float * 2), ("seed", C.c_uint64), ("pass_begin", C.c_int32),
                ("pass_stride", C.c_int32), ("adaptive", C.c_int32)]


class pt_render_stats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("gpu_ms", C.c_double), ("nonfinite_samples", C.c_uint64)]


class pt_trace_stats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("tlas_nodes", C.c_uint64), ("blas_nodes", C.c_uint64),
                ("instances", C.c_uint64), ("triangles", C.c_uint64), ("hits", C.c_uint64)]


class pt_build_stats(C.Structure):
    _fields_ = [("total_s", C.c_double), ("pack_s", C.c_double), ("bvh2_s", C.c_double), ("flatten_s", C.c_double),
                ("upload_s", C.c_double), ("shade_tables_s", C.c_double), ("gpu_build_ms", C.c_float),
                ("gpu_upload_ms", C.c_float), ("gpu_download_ms", C.c_float), ("builder", C.c_int32),
                ("gpu_levels", C.c_int32), ("gpu_subtrees", C.c_uint32), ("gpu_items", C.c_uint64)]


ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p)     # pt_allreduce_fn

BVH2_NODE_DTYPE = np.dtype([("lo", "<f4", 3), ("hi", "<f4", 3), ("left", "<u4"), ("right", "<u4"), ("count", "<u4"), ("axis", "<u4")])
assert BVH2_NODE_DTYPE.itemsize == 40
BUILDER_HOST, BUILDER_GPU = range(2)
LAYOUT_EXACT, LAYOUT_FAST = range(2)

RAY_DTYPE = np.dtype([("o", "<f4", 3), ("tmax", "<f4"), ("d", "<f4", 3), ("time", "<f4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("prim", "<u4")])
RAY24_DTYPE = np.dtype([("o", "<f4", 3), ("d", "<f4", 3)])       # pt_ray24: tmax = inf, time = 0 implied
HIT8_DTYPE = np.dtype([("t", "<f4"), ("prim", "<u4")])              # pt_hit8
assert RAY_DTYPE.itemsize == 32 and HIT_DTYPE.itemsize == 16 and RAY24_DTYPE.itemsize == 24 and HIT8_DTYPE.itemsize == 8

# enums
TEX_SOLID, TEX_IMAGE_U8, TEX_IMAGE_F32, TEX_CHECKER, TEX_UV, TEX_NORMAL = range(6)
MAT_DIFFUSE, MAT_DIELECTRIC, MAT_THIN_DIELECTRIC, MAT_CONDUCTOR = range(4)
ALPHA_OPAQUE, ALPHA_BLEND, ALPHA_MASK = range(3)
LIGHT_UNIFORM_INFINITE, LIGHT_GRADIENT_INFINITE, LIGHT_TEXTURE_INFINITE, LIGHT_DISTANT, LIGHT_POINT = range(5)
INTEGRATOR_SIMPLE, INTEGRATOR_PATH, INTEGRATOR_VOLPATH = range(3)
SHAPE_SPHERE, SHAPE_QUAD = range(2)
STATIC_MAX_RAYS = 65536            # PT_STATIC_MAX_RAYS
FILTER_BOX, FILTER_GAUSSIAN, FILTER_MITCHELL, FILTER_LANCZOS = range(4)

EXPORTS = [
    "pt_device_count", "pt_scene_create", "pt_scene_destroy", "pt_last_error",
    "pt_scene_add_texture", "pt_scene_add_material", "pt_scene_add_medium", "pt_scene_add_mesh",
    "pt_scene_add_blas", "pt_scene_add_instance", "pt_scene_add_shape", "pt_scene_add_shape_xf", "pt_scene_animate", "pt_scene_add_light", "pt_scene_set_medium",
    "pt_scene_commit", "pt_scene_set_builder", "pt_scene_set_layout", "pt_scene_set_allreduce", "pt_scene_build_stats", "pt_build_bvh2", "pt_scene_bounds", "pt_trace_closest", "pt_trace_closest_compact", "pt_trace_any",
    "pt_trace_closest_dev", "pt_trace_any_dev", "pt_trace_closest_static", "pt_scene_stack_overflows", "pt_trace_closest_stats", "pt_scene_set_counting", "pt_scene_render_counts", "pt_render",
    "pt_film_device_ptr", "pt_film_download", "pt_film_download_sumsq", "pt_film_upload",
    "pt_scene_replicate", "pt_multi_create", "pt_multi_destroy", "pt_multi_gpus", "pt_multi_scene", "pt_multi_render",
    "pt_dist_unique_id", "pt_dist_init", "pt_dist_render", "pt_dist_finalize",
    "pt_unit_shade", "pt_unit_light_count", "pt_unit_light_sample", "pt_unit_hg", "pt_unit_table",
]

_lib = None


class PtError(RuntimeError):
    pass


def load_library(path: os.PathLike | None = None) -> C.CDLL:
    """dlopen libptb200.so and declare the signatures.  Raises if it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = Path(path) if path else LIB_PATH
    if not p.exists():
        raise PtError(f"{p} not found: build the CUDA extension first (python -c 'import __graft_entry__ as g; g.build()')")
    lib = C.CDLL(str(p))
    vp, i, sz = C.c_void_p, C.c_int, C.c_size_t
    lib.pt_device_count.restype = i
    lib.pt_scene_create.restype = vp
    lib.pt_scene_create.argtypes = [i]
    lib.pt_scene_destroy.restype = None
    lib.pt_scene_destroy.argtypes = [vp]
    lib.pt_last_error.restype = C.c_char_p
    for name, desc in (("pt_scene_add_texture", pt_texture_desc), ("pt_scene_add_material", pt_material_desc),
                       ("pt_scene_add_medium", pt_medium_desc), ("pt_scene_add_mesh", pt_mesh_desc),
                       ("pt_scene_add_light", pt_light_desc), ("pt_scene_add_shape", pt_shape_desc)):
        f = getattr(lib, name)
        f.restype = i
        f.argtypes = [vp, C.POINTER(desc)]
    lib.pt_scene_add_shape_xf.restype = i
    lib.pt_scene_add_shape_xf.argtypes = [vp, C.POINTER(pt_shape_desc), C.POINTER(C.c_float)]
    lib.pt_scene_animate.restype = i
    lib.pt_scene_animate.argtypes = [vp, i, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.pt_scene_add_blas.restype = i
    lib.pt_scene_add_blas.argtypes = [vp, C.POINTER(C.c_int32), i]
    lib.pt_scene_add_instance.restype = i
    lib.pt_scene_add_instance.argtypes = [vp, i, C.POINTER(C.c_float)]
    lib.pt_scene_set_medium.restype = i
    lib.pt_scene_set_medium.argtypes = [vp, i]
    lib.pt_scene_commit.restype = i
    lib.pt_scene_commit.argtypes = [vp, i, i]
    lib.pt_scene_set_builder.restype = i
    lib.pt_scene_set_builder.argtypes = [vp, i]
    lib.pt_scene_set_layout.restype = i
    lib.pt_scene_set_layout.argtypes = [vp, i]
    lib.pt_scene_set_allreduce.restype = i
    lib.pt_scene_set_allreduce.argtypes = [vp, ALLREDUCE_FN, vp]
    lib.pt_scene_build_stats.restype = i
    lib.pt_scene_build_stats.argtypes = [vp, C.POINTER(pt_build_stats)]
    lib.pt_build_bvh2.restype = i
    lib.pt_build_bvh2.argtypes = [i, i, vp, C.c_uint32, vp, C.c_uint32, C.POINTER(C.c_uint32), vp, C.POINTER(C.c_float)]
    lib.pt_scene_bounds.restype = i
    lib.pt_scene_bounds.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.pt_trace_closest.restype = i
    lib.pt_trace_closest.argtypes = [vp, vp, sz, vp, vp]
    lib.pt_trace_any.restype = i
    lib.pt_trace_any.argtypes = [vp, vp, sz, vp]
    lib.pt_scene_stack_overflows.restype = i
    lib.pt_scene_stack_overflows.argtypes = [vp, C.POINTER(C.c_uint64)]
    lib.pt_trace_closest_static.restype = i
    lib.pt_trace_closest_static.argtypes = [vp, vp, sz, vp, vp]
    lib.pt_trace_closest_dev.restype = i
    lib.pt_trace_closest_dev.argtypes = [vp, vp, sz, vp, vp, vp]
    lib.pt_trace_any_dev.restype = i
    lib.pt_trace_any_dev.argtypes = [vp, vp, sz, vp, vp]
    lib.pt_scene_set_counting.restype = i
    lib.pt_scene_set_counting.argtypes = [vp, i]
    lib.pt_scene_render_counts.restype = i
    lib.pt_scene_render_counts.argtypes = [vp, C.POINTER(pt_trace_stats), C.POINTER(pt_trace_stats)]
    lib.pt_trace_closest_stats.restype = i
    lib.pt_trace_closest_stats.argtypes = [vp, vp, sz, C.POINTER(pt_trace_stats)]
    lib.pt_render.restype = i
    lib.pt_render.argtypes = [vp, C.POINTER(pt_camera_desc), C.POINTER(pt_render_params), i, C.POINTER(pt_render_stats)]
    lib.pt_film_device_ptr.restype = i
    lib.pt_film_device_ptr.argtypes = [vp, C.POINTER(vp), C.POINTER(sz)]
    lib.pt_film_download.restype = i
    lib.pt_film_download.argtypes = [vp, vp]
    lib.pt_film_download_sumsq.restype = i
    lib.pt_film_download_sumsq.argtypes = [vp, vp]
    lib.pt_film_upload.restype = i
    lib.pt_film_upload.argtypes = [vp, vp, i, i]
    lib.pt_scene_replicate.restype = i
    lib.pt_scene_replicate.argtypes = [vp, i, C.POINTER(vp)]
    lib.pt_multi_create.restype = vp
    lib.pt_multi_create.argtypes = [vp, C.POINTER(C.c_int), i]
    lib.pt_multi_destroy.restype = None
    lib.pt_multi_destroy.argtypes = [vp]
    lib.pt_multi_gpus.restype = i
    lib.pt_multi_gpus.argtypes = [vp]
    lib.pt_multi_scene.restype = vp
    lib.pt_multi_scene.argtypes = [vp, i]
    lib.pt_multi_render.restype = i
    lib.pt_multi_render.argtypes = [vp, C.POINTER(pt_camera_desc), C.POINTER(pt_render_params), i, C.POINTER(pt_render_stats)]
    lib.pt_dist_unique_id.restype = i
    lib.pt_dist_unique_id.argtypes = [vp]
    lib.pt_dist_init.restype = i
    lib.pt_dist_init.argtypes = [vp, vp, i, i]
    lib.pt_dist_render.restype = i
    lib.pt_dist_render.argtypes = [vp, C.POINTER(pt_camera_desc), C.POINTER(pt_render_params), i, C.POINTER(pt_render_stats)]
    lib.pt_dist_finalize.restype = i
    lib.pt_dist_finalize.argtypes = [vp]
    lib.pt_trace_closest_compact.restype = i
    lib.pt_trace_closest_compact.argtypes = [vp, vp, sz, vp, vp]
    lib.pt_unit_shade.restype = i
    lib.pt_unit_shade.argtypes = [vp, vp, sz, vp, vp, i]
    lib.pt_unit_light_count.restype = i
    lib.pt_unit_light_count.argtypes = [vp]
    lib.pt_unit_light_sample.restype = i
    lib.pt_unit_light_sample.argtypes = [vp, i, sz, vp, vp, i, C.c_float]
    lib.pt_unit_table.restype = i
    lib.pt_unit_table.argtypes = [vp, i, vp, sz, C.POINTER(sz)]
    lib.pt_unit_hg.restype = i
    lib.pt_unit_hg.argtypes = [i, C.c_float, sz, vp, vp]
    if path is None:
        _lib = lib
    return lib


def build_bvh2(boxes: np.ndarray, builder: int = BUILDER_HOST, device: int = 0):
    """pt_build_bvh2: binary binned-SAH build over (n, 6) float32 boxes {lo, hi}.
    Returns (nodes[BVH2_NODE_DTYPE] in DFS preorder, order[u32], gpu_build_ms)."""
    lib = load_library()
    boxes = np.ascontiguousarray(boxes, dtype=np.float32).reshape(-1, 6)
    n = len(boxes)
    cap = max(2 * n - 1, 1)
    nodes = np.zeros(cap, dtype=BVH2_NODE_DTYPE)
    order = np.zeros(max(n, 1), dtype=np.uint32)
    nn = C.c_uint32(0)
    ms = C.c_float(0)
    rc = lib.pt_build_bvh2(device, builder, boxes.ctypes.data, n, nodes.ctypes.data, cap, C.byref(nn), order.ctypes.data, C.byref(ms))
    if rc < 0:
        raise PtError(f"pt error {rc}: {lib.pt_last_error().decode()}")
    return nodes[:nn.value], order[:n], ms.value
