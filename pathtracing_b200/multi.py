"""Multi-GPU plumbing: scene replicated per rank, sample passes sharded
(rank r renders passes r, r+G, ...), ONE reduce(sum) of the film accumulators to rank 0.

The product path is csrc/multi.cu (pt_multi_* in one process, pt_dist_* with one process per GPU: NCCL inside
libptb200.so).  What is left here: render_sharded (the call bench.py makes under torchrun), a torch view of the
device film, and the same sharding / reduce / all-reduce glue over torch.distributed for the gloo tests that run
without a GPU.  The reference has no counterpart: its only shared state is the
atomic<double> film that worker threads merge into (Film.hpp:125-132, 244-249).
"""
from __future__ import annotations

import numpy as np


def shard_passes(spp: int, world: int, rank: int):
    """Sample-pass indices rank `rank` of `world` renders: rank, rank+world, ... < spp."""
    return list(range(rank, spp, world))


def pass_range(params, world: int, rank: int):
    """Return a copy of pt_render_params restricted to this rank's passes."""
    import copy
    p = copy.copy(params)
    p.pass_begin = rank
    p.pass_stride = world
    return p


class _DevArray:
    """__cuda_array_interface__ view of a raw device pointer (the film the C-ABI exposes)."""

    def __init__(self, ptr: int, n: int, typestr="<f8"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 3}


def film_tensor(scene):
    """torch view (no copy) of the scene's device film: float64[4*W*H]."""
    import torch
    ptr, n = scene.film_device_ptr()
    return torch.as_tensor(_DevArray(ptr, n), device=f"cuda:{torch.cuda.current_device()}")


def reduce_film(film, dist, dst: int = 0):
    """The single collective of the path: sum of RGBW accumulators to rank `dst` (in place)."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM)
    return film


def make_allreduce(dist, cuda: bool = True):
    """pt_allreduce_fn over torch.distributed: sums n doubles at a raw pointer over all ranks, in place.
    Used by adaptive renders whose passes are sharded (every rank must see the same per-pixel statistics,
    Integrators.cpp:82-85).  cuda=False wraps a host pointer instead (gloo tests of this glue).
    Keep the returned object alive for as long as the scene may call it."""
    import ctypes as C

    from .capi import ALLREDUCE_FN

    def _sum(user, ptr, n, stream):
        try:
            import torch
            if cuda:
                t = torch.as_tensor(_DevArray(ptr, n), device=f"cuda:{torch.cuda.current_device()}")
                torch.cuda.synchronize()              # the scene's stream is not torch's
                dist.all_reduce(t, op=dist.ReduceOp.SUM)
                torch.cuda.synchronize()
            else:
                a = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(n,))
                t = torch.from_numpy(a)
                dist.all_reduce(t, op=dist.ReduceOp.SUM)
            return 0
        except Exception as e:  # never unwind through the C caller
            print(f"[pathtracing_b200] all-reduce callback failed: {e!r}", flush=True)
            return 1

    return ALLREDUCE_FN(_sum)


def render_sharded(scene, cam, params, dist=None):
    """Render this rank's passes and reduce the film to rank 0 -- all inside libptb200.so (pt_dist_render: NCCL
    communicator, pass sharding, the film reduce and the per-round all-reduce of adaptive renders live behind the
    C-ABI; `dist` only carries the 128-byte NCCL id once, GpuScene.dist_init).  Returns (stats, film tensor);
    the tensor is a view of the device film and holds the complete image on rank 0 only."""
    if getattr(scene, "world", None) is None:
        scene.dist_init(dist)
    stats = scene.dist_render(cam, params, clear=True)
    return stats, film_tensor(scene)


def resolve(rgbw: np.ndarray):
    """Film pixel value = sum(RGB*w) / sum(w)  (Film.hpp:161)."""
    w = np.maximum(rgbw[..., 3:4], 1e-300)
    return rgbw[..., :3] / w
