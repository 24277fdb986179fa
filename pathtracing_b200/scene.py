"""Backend-neutral scene description + the GPU scene object.

`SceneDesc` records the same calls a user of the reference makes
(examples/example_1.cpp:17-101: textures -> materials -> meshes -> BLAS ->
instances -> lights) as POD descriptors.  `SceneDesc.replay(backend)` pushes
them through any object exposing the add_* entry points: the product's C-ABI
(`GpuScene`) or the reference harness used by the tests (tests/refapi.py).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np

from . import capi
from .capi import (PtError, pt_camera_desc, pt_light_desc, pt_material_desc, pt_medium_desc, pt_mesh_desc, pt_shape_desc,
                   pt_render_params, pt_render_stats, pt_texture_desc, pt_trace_stats)


def _f3(dst, v):
    v = np.broadcast_to(np.asarray(v, dtype=np.float32), (3,))
    for k in range(3):
        dst[k] = float(v[k])


@dataclass
class SceneDesc:
    textures: list = field(default_factory=list)
    materials: list = field(default_factory=list)
    media: list = field(default_factory=list)
    meshes: list = field(default_factory=list)
    blas: list = field(default_factory=list)
    instances: list = field(default_factory=list)
    shape_xf: dict = field(default_factory=dict)    # instance index -> column-major transform of a TransformedPrimitive over a shape
    animated: dict = field(default_factory=dict)    # instance index -> (direction, timeBounds) of its AnimatedPrimitive
    lights: list = field(default_factory=list)
    scene_medium: int = -1
    _keep: list = field(default_factory=list)   # numpy arrays referenced by descriptors

    # -- textures ---------------------------------------------------------------------
    def _tex(self, kind, color=(0, 0, 0), scale=(1, 1, 1), pixels=None, tex_a=-1, tex_b=-1, uv_scale=(1, 1)):
        d = pt_texture_desc()
        d.kind = kind
        _f3(d.color, color)
        _f3(d.color_scale, scale)
        d.tex_a, d.tex_b = tex_a, tex_b
        d.uv_scale[0], d.uv_scale[1] = float(uv_scale[0]), float(uv_scale[1])
        if pixels is not None:
            px = np.ascontiguousarray(pixels)
            if px.ndim == 2:
                px = px[:, :, None]
            d.height, d.width, d.channels = px.shape
            d.pixels = px.ctypes.data
            self._keep.append(px)
        self.textures.append(d)
        return len(self.textures) - 1

    def solid(self, color, scale=(1, 1, 1)):
        return self._tex(capi.TEX_SOLID, color=color, scale=scale)

    def image_u8(self, pixels, scale=(1, 1, 1)):
        assert pixels.dtype == np.uint8
        return self._tex(capi.TEX_IMAGE_U8, pixels=pixels, scale=scale)

    def image_f32(self, pixels, scale=(1, 1, 1)):
        assert pixels.dtype == np.float32
        return self._tex(capi.TEX_IMAGE_F32, pixels=pixels, scale=scale)

    def uv_texture(self, scale=(1, 1, 1)):
        return self._tex(capi.TEX_UV, scale=scale)

    def normal_texture(self, scale=(1, 1, 1)):
        return self._tex(capi.TEX_NORMAL, scale=scale)

    def checker(self, tex_a, tex_b, uv_scale, scale=(1, 1, 1)):
        return self._tex(capi.TEX_CHECKER, tex_a=tex_a, tex_b=tex_b, uv_scale=uv_scale, scale=scale)

    # -- materials --------------------------------------------------------------------
    def _mat(self, kind, tex=-1, norm=-1, rough=-1, metal=-1, alpha=-1, ri=1.0, albedo=(0, 0, 0),
             alpha_mode=capi.ALPHA_BLEND, alpha_cutoff=0.5):
        d = pt_material_desc()
        d.kind, d.tex, d.norm, d.rough, d.metal, d.alpha = kind, tex, norm, rough, metal, alpha
        d.ri = ri
        _f3(d.albedo, albedo)
        d.alpha_mode, d.alpha_cutoff = alpha_mode, alpha_cutoff
        self.materials.append(d)
        return len(self.materials) - 1

    def diffuse(self, tex, norm=-1, rough=-1, metal=-1, alpha=-1, alpha_mode=capi.ALPHA_BLEND, alpha_cutoff=0.5):
        """MicrofacetDiffuse (Material.hpp:200-361)."""
        return self._mat(capi.MAT_DIFFUSE, tex, norm, rough, metal, alpha, alpha_mode=alpha_mode, alpha_cutoff=alpha_cutoff)

    def dielectric(self, ri, tex, norm=-1, rough=-1, alpha=-1, alpha_mode=capi.ALPHA_BLEND):
        """MicrofacetDielectric (Material.hpp:363-598)."""
        return self._mat(capi.MAT_DIELECTRIC, tex, norm, rough, -1, alpha, ri=ri, alpha_mode=alpha_mode)

    def thin_dielectric(self, ri, tex=-1):
        return self._mat(capi.MAT_THIN_DIELECTRIC, tex, ri=ri)

    def conductor(self, albedo):
        return self._mat(capi.MAT_CONDUCTOR, albedo=albedo)

    # -- media ------------------------------------------------------------------------
    def medium(self, sigma_a, sigma_s, g, density=1.0, Le=(0, 0, 0), le_density=1.0):
        d = pt_medium_desc()
        _f3(d.sigma_a, sigma_a)
        _f3(d.sigma_s, sigma_s)
        _f3(d.Le, Le)
        d.g, d.density, d.le_density = g, density, le_density
        self.media.append(d)
        return len(self.media) - 1

    # -- geometry ---------------------------------------------------------------------
    def mesh(self, positions, indices, normals=None, texcoords=None, tangents=None, material=-1,
             emissive_tex=-1, medium=-1):
        pos = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 3)
        idx = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1, 3)
        if normals is None:
            normals = vertex_normals(pos, idx)
        nrm = np.ascontiguousarray(normals, dtype=np.float32).reshape(-1, 3)
        if texcoords is None:
            texcoords = np.zeros((len(pos), 2), np.float32)
        uv = np.ascontiguousarray(texcoords, dtype=np.float32).reshape(-1, 2)
        assert len(nrm) == len(pos) and len(uv) == len(pos)
        d = pt_mesh_desc()
        d.positions, d.normals, d.texcoords, d.indices = pos.ctypes.data, nrm.ctypes.data, uv.ctypes.data, idx.ctypes.data
        keep = [pos, nrm, uv, idx]
        if tangents is not None:
            tan = np.ascontiguousarray(tangents, dtype=np.float32).reshape(-1, 3)
            d.tangents = tan.ctypes.data
            keep.append(tan)
        d.n_vertices, d.n_triangles = len(pos), len(idx)
        d.material, d.emissive_tex, d.medium = material, emissive_tex, medium
        self._keep.extend(keep)
        self.meshes.append(d)
        return len(self.meshes) - 1

    def add_blas(self, mesh_ids):
        self.blas.append(np.asarray(mesh_ids, dtype=np.int32).reshape(-1))
        return len(self.blas) - 1

    def instance(self, blas_id, transform=None):
        """transform: 4x4 (row-major numpy, i.e. M @ [x,y,z,1]) or None."""
        if transform is not None:
            m = np.asarray(transform, dtype=np.float32).reshape(4, 4)
            transform = np.ascontiguousarray(m.T).reshape(16)   # -> column-major (glm)
        self.instances.append((blas_id, transform))
        return len(self.instances) - 1

    def animate(self, prim_id, direction, time_bounds=(0.0, 1.0)):
        """AnimatedPrimitive(primitive, direction, timeBounds) around an already added top-level primitive"""
        self.animated[prim_id] = (tuple(float(x) for x in direction), tuple(float(x) for x in time_bounds))
        return prim_id

    def _shape(self, kind, p, u, v, radius, material, medium, emissive_tex, one_sided, transform=None):
        if transform is not None:   # TransformedPrimitive(GeometricPrimitive(shape), M)
            m = np.asarray(transform, dtype=np.float32).reshape(4, 4)
            self.shape_xf[len(self.instances)] = np.ascontiguousarray(m.T).reshape(16)
        d = pt_shape_desc()
        d.kind = kind
        _f3(d.p, p)
        _f3(d.u, u)
        _f3(d.v, v)
        d.radius, d.material, d.medium, d.emissive_tex, d.one_sided = radius, material, medium, emissive_tex, int(one_sided)
        self.instances.append((-1, d))
        return len(self.instances) - 1

    def sphere(self, center, radius, material=-1, medium=-1, emissive_tex=-1, one_sided=False, transform=None):
        """GeometricPrimitive(SphereShape(center, radius), material, AreaLight(shape, emissive_tex, one_sided), medium),
        inside a TransformedPrimitive when `transform` (4x4, row-major numpy) is given"""
        return self._shape(capi.SHAPE_SPHERE, center, (0, 0, 0), (0, 0, 0), radius, material, medium, emissive_tex, one_sided, transform)

    def quad(self, q, u, v, material=-1, medium=-1, emissive_tex=-1, one_sided=False, transform=None):
        """GeometricPrimitive(QuadShape(Q, u, v), ...)"""
        return self._shape(capi.SHAPE_QUAD, q, u, v, 0.0, material, medium, emissive_tex, one_sided, transform)

    # -- lights -----------------------------------------------------------------------
    def light(self, kind, color=(0, 0, 0), color_b=(0, 0, 0), scale=1.0, v=(0, 0, 0), tex=-1, power_func=0):
        d = pt_light_desc()
        d.kind = kind
        _f3(d.color, color)
        _f3(d.color_b, color_b)
        _f3(d.dir_or_pos, v)
        d.scale, d.tex, d.power_func = scale, tex, power_func
        self.lights.append(d)
        return len(self.lights) - 1

    # -- replay -----------------------------------------------------------------------
    def replay(self, b):
        for d in self.textures:
            b.add_texture(d)
        for d in self.materials:
            b.add_material(d)
        for d in self.media:
            b.add_medium(d)
        for d in self.meshes:
            b.add_mesh(d)
        for ids in self.blas:
            b.add_blas(ids)
        for k, (blas_id, xf) in enumerate(self.instances):
            if blas_id < 0:
                b.add_shape(xf, self.shape_xf.get(k))
            else:
                b.add_instance(blas_id, xf)
            if k in self.animated:
                b.animate(k, *self.animated[k])
        for d in self.lights:
            b.add_light(d)
        if self.scene_medium >= 0:
            b.set_medium(self.scene_medium)


def vertex_normals(pos, idx):
    """Area-weighted vertex normals (synthetic-scene helper, not on the hot path)."""
    p0, p1, p2 = pos[idx[:, 0]], pos[idx[:, 1]], pos[idx[:, 2]]
    fn = np.cross(p1 - p0, p2 - p0)
    n = np.zeros_like(pos)
    for k in range(3):
        np.add.at(n, idx[:, k], fn)
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    ln[ln == 0] = 1
    return (n / ln).astype(np.float32)


def camera_desc(look_from, look_at, fov, focus_angle=0.0, focus_distance=0.0, medium=-1, shutter=(0.0, 1.0)):
    d = pt_camera_desc()
    _f3(d.look_from, look_from)
    _f3(d.look_at, look_at)
    d.fov, d.focus_angle, d.focus_distance, d.medium = fov, focus_angle, focus_distance, medium
    d.shutter_start, d.shutter_end = shutter
    return d


def render_params(width, height, spp, max_depth, integrator=capi.INTEGRATOR_PATH, sampler=0, strata=(1, 1),
                  filter=capi.FILTER_BOX, filter_radius=(0.5, 0.5), filter_param=(0.0, 0.0), seed=1,
                  pass_begin=0, pass_stride=1, adaptive=0):
    d = pt_render_params()
    d.width, d.height, d.spp, d.max_depth, d.integrator = width, height, spp, max_depth, integrator
    d.sampler, d.strata_x, d.strata_y = sampler, strata[0], strata[1]
    d.filter = filter
    d.filter_radius[0], d.filter_radius[1] = filter_radius
    d.filter_param[0], d.filter_param[1] = filter_param
    d.seed, d.pass_begin, d.pass_stride, d.adaptive = seed, pass_begin, pass_stride, adaptive
    return d


class GpuScene:
    """A committed scene on one B200, through the C-ABI only."""

    def __init__(self, desc: SceneDesc, bvh_width=4, light_sampler=1, device=0, builder=None, layout=None, lib=None):
        self.lib = lib if lib is not None else capi.load_library()      # lib: a second build of the library (tests: the small-stack twin)
        self.h = self.lib.pt_scene_create(device)
        if not self.h:
            raise PtError(self._err())
        self.desc = desc
        desc.replay(self)
        if builder is not None:                      # library default: GPU builder (PTB_BUILDER=host overrides)
            self._ck(self.lib.pt_scene_set_builder(self.h, builder))
        if layout is None:      # experiments: PTB_LAYOUT=fast turns the quantised layout on for every width-4 scene
            layout = capi.LAYOUT_FAST if (os.environ.get("PTB_LAYOUT") == "fast" and bvh_width == 4) else capi.LAYOUT_EXACT
        self.layout = layout
        self._ck(self.lib.pt_scene_set_layout(self.h, layout))
        self._ck(self.lib.pt_scene_commit(self.h, bvh_width, light_sampler))
        self.bvh_width = bvh_width

    def build_stats(self) -> dict:
        st = capi.pt_build_stats()
        self._ck(self.lib.pt_scene_build_stats(self.h, C.byref(st)))
        return {k: getattr(st, k) for k, _ in st._fields_}

    def _err(self):
        return self.lib.pt_last_error().decode()

    def _ck(self, rc):
        if rc < 0:
            raise PtError(f"pt error {rc}: {self._err()}")
        return rc

    # replay backend
    def add_texture(self, d):
        return self._ck(self.lib.pt_scene_add_texture(self.h, C.byref(d)))

    def add_material(self, d):
        return self._ck(self.lib.pt_scene_add_material(self.h, C.byref(d)))

    def add_medium(self, d):
        return self._ck(self.lib.pt_scene_add_medium(self.h, C.byref(d)))

    def add_mesh(self, d):
        return self._ck(self.lib.pt_scene_add_mesh(self.h, C.byref(d)))

    def add_blas(self, ids):
        return self._ck(self.lib.pt_scene_add_blas(self.h, ids.ctypes.data_as(C.POINTER(C.c_int32)), len(ids)))

    def add_instance(self, blas_id, xf):
        p = xf.ctypes.data_as(C.POINTER(C.c_float)) if xf is not None else None
        return self._ck(self.lib.pt_scene_add_instance(self.h, blas_id, p))

    def add_shape(self, d, xf=None):
        if xf is None:
            return self._ck(self.lib.pt_scene_add_shape(self.h, C.byref(d)))
        return self._ck(self.lib.pt_scene_add_shape_xf(self.h, C.byref(d), xf.ctypes.data_as(C.POINTER(C.c_float))))

    def animate(self, prim_id, direction, time_bounds):
        a, b = (C.c_float * 3)(*direction), (C.c_float * 2)(*time_bounds)
        return self._ck(self.lib.pt_scene_animate(self.h, prim_id, a, b))

    def add_light(self, d):
        return self._ck(self.lib.pt_scene_add_light(self.h, C.byref(d)))

    def set_medium(self, m):
        return self._ck(self.lib.pt_scene_set_medium(self.h, m))

    # queries
    def bounds(self):
        lo, hi = (C.c_float * 3)(), (C.c_float * 3)()
        self._ck(self.lib.pt_scene_bounds(self.h, lo, hi))
        return np.array(lo[:], np.float32), np.array(hi[:], np.float32)

    # ray casting, host buffers (numpy structured arrays of RAY_DTYPE)
    def trace_closest(self, rays, want_inst=True):
        rays = np.ascontiguousarray(rays, dtype=capi.RAY_DTYPE)
        hits = np.empty(len(rays), dtype=capi.HIT_DTYPE)
        inst = np.empty(len(rays), dtype=np.uint32) if want_inst else None
        self._ck(self.lib.pt_trace_closest(self.h, rays.ctypes.data, len(rays), hits.ctypes.data,
                                           inst.ctypes.data if want_inst else None))
        return hits, inst

    def table(self, which):
        """pt_unit_table: one of the committed traversal tables as bytes (0 clusters, 1 triangles, 2 quantised clusters, 3 padded triangles, 4 instances)."""
        n = C.c_size_t(0)
        self._ck(self.lib.pt_unit_table(self.h, which, None, 0, C.byref(n)))
        out = np.empty(n.value, np.uint8)
        if n.value:
            self._ck(self.lib.pt_unit_table(self.h, which, out.ctypes.data, n.value, C.byref(n)))
        return out

    def trace_closest_compact(self, rays24, want_inst=True):
        """pt_trace_closest_compact: 24 B rays {o, d} in (tmax = inf, time = 0), 8 B hits {t, prim} out."""
        rays24 = np.ascontiguousarray(rays24, dtype=capi.RAY24_DTYPE)
        hits = np.empty(len(rays24), dtype=capi.HIT8_DTYPE)
        inst = np.empty(len(rays24), dtype=np.uint32) if want_inst else None
        self._ck(self.lib.pt_trace_closest_compact(self.h, rays24.ctypes.data, len(rays24), hits.ctypes.data,
                                                   inst.ctypes.data if want_inst else None))
        return hits, inst

    def trace_any(self, rays):
        rays = np.ascontiguousarray(rays, dtype=capi.RAY_DTYPE)
        occ = np.empty(len(rays), dtype=np.uint8)
        self._ck(self.lib.pt_trace_any(self.h, rays.ctypes.data, len(rays), occ.ctypes.data))
        return occ

    # ray casting, device buffers (raw pointers, e.g. torch tensors' data_ptr())
    def trace_closest_dev(self, d_rays, n, d_hits, d_inst=None, stream=None):
        self._ck(self.lib.pt_trace_closest_dev(self.h, d_rays, n, d_hits, d_inst, stream))

    def trace_any_dev(self, d_rays, n, d_occ, stream=None):
        self._ck(self.lib.pt_trace_any_dev(self.h, d_rays, n, d_occ, stream))

    def trace_stats(self, d_rays, n):
        st = pt_trace_stats()
        self._ck(self.lib.pt_trace_closest_stats(self.h, d_rays, n, C.byref(st)))
        return {k: int(getattr(st, k)) for k, _ in pt_trace_stats._fields_}

    # rendering
    def render(self, cam, params, clear=True):
        st = pt_render_stats()
        self._ck(self.lib.pt_render(self.h, C.byref(cam), C.byref(params), 1 if clear else 0, C.byref(st)))
        self._film_shape = (params.height, params.width)
        return {k: getattr(st, k) for k, _ in pt_render_stats._fields_}

    def film(self):
        h, w = self._film_shape
        out = np.empty((h, w, 4), np.float64)
        self._ck(self.lib.pt_film_download(self.h, out.ctypes.data))
        return out

    def film_sumsq(self):
        h, w = self._film_shape
        out = np.empty((h, w, 3), np.float64)
        self._ck(self.lib.pt_film_download_sumsq(self.h, out.ctypes.data))
        return out

    def film_device_ptr(self):
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self.lib.pt_film_device_ptr(self.h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def counted_render(self, cam, params):
        """One frame with the counting build of the traversal kernels: (render stats, closest-ray visits, shadow-ray visits)."""
        self._ck(self.lib.pt_scene_set_counting(self.h, 1))
        try:
            st = self.render(cam, params)
            c, s = pt_trace_stats(), pt_trace_stats()
            self._ck(self.lib.pt_scene_render_counts(self.h, C.byref(c), C.byref(s)))
        finally:
            self._ck(self.lib.pt_scene_set_counting(self.h, 0))
        d = lambda x: {k: int(getattr(x, k)) for k, _ in pt_trace_stats._fields_}
        return st, d(c), d(s)

    def stack_overflows(self) -> int:
        n = C.c_uint64(0)
        self._ck(self.lib.pt_scene_stack_overflows(self.h, C.byref(n)))
        return int(n.value)

    # one process per GPU (pt_dist_*): the NCCL communicator lives inside libptb200.so; `dist` (torch.distributed,
    # any backend) only carries the 128-byte id from rank 0 to the others
    def dist_init(self, dist=None):
        world = dist.get_world_size() if dist is not None and dist.is_initialized() else 1
        rank = dist.get_rank() if world > 1 else 0
        ident = None
        if world > 1:
            import torch  # noqa: F401  (loads PyTorch's NCCL before ours is looked for: one NCCL per process)
            box = [None]
            if rank == 0:
                buf = (C.c_ubyte * 128)()
                self._ck(self.lib.pt_dist_unique_id(buf))
                box[0] = bytes(buf)
            dist.broadcast_object_list(box, src=0)
            ident = (C.c_ubyte * 128).from_buffer_copy(box[0])
        self._ck(self.lib.pt_dist_init(self.h, ident, rank, world))
        self.rank, self.world = rank, world

    def dist_render(self, cam, params, clear=True):
        """This rank's passes of the frame + the one film reduce to rank 0 (pt_dist_render)."""
        st = pt_render_stats()
        self._ck(self.lib.pt_dist_render(self.h, C.byref(cam), C.byref(params), 1 if clear else 0, C.byref(st)))
        self._film_shape = (params.height, params.width)
        return {k: getattr(st, k) for k, _ in pt_render_stats._fields_}

    def close(self):
        if self.h:
            self.lib.pt_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class MultiGpu:
    """pt_multi_*: one process, one host thread per GPU inside libptb200.so, the frame in the primary scene's film.
    devices=None: every visible GPU; [0, 0, ...]: virtual ranks on one GPU (the sharding logic without NCCL)."""

    def __init__(self, primary: GpuScene, devices=None):
        self.primary, self.lib = primary, primary.lib
        if devices is None:
            self.h = self.lib.pt_multi_create(primary.h, None, 0)
        else:
            arr = (C.c_int * len(devices))(*devices)
            self.h = self.lib.pt_multi_create(primary.h, arr, len(devices))
        if not self.h:
            raise PtError(primary._err())

    @property
    def gpus(self) -> int:
        return self.lib.pt_multi_gpus(self.h)

    def render(self, cam, params, clear=True):
        st = pt_render_stats()
        rc = self.lib.pt_multi_render(self.h, C.byref(cam), C.byref(params), 1 if clear else 0, C.byref(st))
        if rc < 0:
            raise PtError(f"pt error {rc}: {self.primary._err()}")
        self.primary._film_shape = (params.height, params.width)
        return {k: getattr(st, k) for k, _ in pt_render_stats._fields_}

    def close(self):
        if self.h:
            self.lib.pt_multi_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
