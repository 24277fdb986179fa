"""GPU tier: multi-GPU rendering behind the C-ABI (pt_multi_* / pt_dist_*, csrc/multi.cu) -- the replacement of
the reference's worker threads + Film::Merge (Integrators.cpp:23-41,117-123, Film.hpp:125-132).

The sample set of a frame does not depend on how its passes are sharded (streams are keyed by sample id), so the
G-rank film must equal the 1-rank film up to the order of the fp64 sums: allclose 1e-9.  On a 1-GPU box the ranks are
"virtual" (all on device 0, reduced by an add kernel: NCCL refuses two ranks on one GPU); with >= 2 GPUs the same
test runs over NCCL with one host thread per GPU."""
import ctypes as C

import numpy as np
import pytest

from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene, MultiGpu, render_params

pytestmark = pytest.mark.gpu


def _n_gpus():
    return capi.load_library().pt_device_count()


def _single(desc, cam, p, width=4):
    gpu = GpuScene(desc, width)
    st = gpu.render(cam, p)
    film, sq = gpu.film().copy(), gpu.film_sumsq().copy()
    gpu.close()
    return st, film, sq


def _check_equal(one, many):
    (st1, f1, q1), (stg, fg, qg) = one, many
    assert stg["samples"] == st1["samples"] and stg["closest_rays"] == st1["closest_rays"] and stg["shadow_rays"] == st1["shadow_rays"]
    assert np.array_equal(f1[..., 3], fg[..., 3])
    assert np.allclose(f1[..., :3], fg[..., :3], rtol=1e-9, atol=1e-12)
    assert np.allclose(q1, qg, rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("ranks", [2, 3, 5])
def test_virtual_ranks_on_one_gpu_equal_the_single_gpu_film(ranks):
    desc, cam = synth.cornell_scene()
    p = render_params(64, 48, 13, 6, seed=4)          # 13 passes over 2 / 3 / 5 ranks: uneven shares, 5 ranks > some shares
    one = _single(desc, cam, p)
    gpu = GpuScene(desc, 4)
    m = MultiGpu(gpu, devices=[0] * ranks)
    assert m.gpus == ranks
    for _ in range(2):                               # twice: the primary's film is cleared, not accumulated
        st = m.render(cam, p)
        _check_equal(one, (st, gpu.film(), gpu.film_sumsq()))
    m.close()
    # the primary scene still works on its own afterwards
    st = gpu.render(cam, p)
    _check_equal(one, (st, gpu.film(), gpu.film_sumsq()))
    gpu.close()


@pytest.mark.parametrize("ranks,spp", [(4, 2), (3, 5)])
def test_tile_sharding_for_frames_with_fewer_passes_than_ranks(ranks, spp, monkeypatch):
    """with fewer sample passes than ranks the frame is split by interleaved 32 x 32 tiles (rank = tile index mod G)
    and the film reduce adds the disjoint tile sets: same film as one GPU.  (3 ranks, 5 spp: forced with PTB_SHARD=tiles.)"""
    desc, cam = synth.cornell_scene()
    p = render_params(100, 70, spp, 5, seed=6)        # 4 x 3 tiles, the last column / row partial
    one = _single(desc, cam, p)
    if spp >= ranks:
        monkeypatch.setenv("PTB_SHARD", "tiles")
    gpu = GpuScene(desc, 4)
    m = MultiGpu(gpu, devices=[0] * ranks)
    st = m.render(cam, p)
    _check_equal(one, (st, gpu.film(), gpu.film_sumsq()))
    m.close(); gpu.close()


def test_replica_traces_and_renders_like_the_original():
    """pt_scene_replicate: device-to-device copy of every table (here onto the same GPU), no second build."""
    desc = synth.instanced_scene(3, 24)
    gpu = GpuScene(desc, 4)
    rep = C.c_void_p()
    assert gpu.lib.pt_scene_replicate(gpu.h, 0, C.byref(rep)) == 0, gpu._err()
    lo, hi = gpu.bounds()
    rays = synth.random_rays(50000, lo, hi, seed=3)
    want_h, want_i = gpu.trace_closest(rays)
    hits = np.empty(len(rays), capi.HIT_DTYPE)
    inst = np.empty(len(rays), np.uint32)
    assert gpu.lib.pt_trace_closest(rep, rays.ctypes.data, len(rays), hits.ctypes.data, inst.ctypes.data) == 0, gpu._err()
    assert np.array_equal(hits.view(np.uint8), want_h.view(np.uint8)) and np.array_equal(inst, want_i)
    gpu.lib.pt_scene_destroy(rep)
    gpu.close()


def test_dist_entry_points_with_one_rank():
    """pt_dist_init / pt_dist_render at world size 1: no NCCL, same film as pt_render."""
    desc, cam = synth.cornell_scene()
    p = render_params(48, 48, 8, 5, seed=2)
    one = _single(desc, cam, p)
    gpu = GpuScene(desc, 4)
    gpu.dist_init(None)
    st = gpu.dist_render(cam, p)
    _check_equal(one, (st, gpu.film(), gpu.film_sumsq()))
    gpu.close()


def test_multi_rejects_bad_device_lists():
    desc, cam = synth.cornell_scene()
    gpu = GpuScene(desc, 4)
    lib = gpu.lib
    bad = (C.c_int * 2)(1, 0) if _n_gpus() > 1 else (C.c_int * 2)(0, 7)
    assert not lib.pt_multi_create(gpu.h, bad, 2)
    assert lib.pt_last_error()
    if _n_gpus() > 2:
        mixed = (C.c_int * 3)(0, 1, 1)
        assert not lib.pt_multi_create(gpu.h, mixed, 3)
    gpu.close()


@pytest.mark.skipif(_n_gpus() < 2, reason="needs two GPUs (NCCL refuses two ranks on one device)")
@pytest.mark.parametrize("adaptive", [0, 1])
def test_nccl_ranks_equal_the_single_gpu_film(adaptive):
    desc, cam = synth.cornell_scene()
    p = render_params(96, 64, 12, 6, seed=8, adaptive=adaptive)
    one = _single(desc, cam, p)
    gpu = GpuScene(desc, 4)
    m = MultiGpu(gpu)                                 # every visible GPU
    assert m.gpus == _n_gpus()
    for _ in range(2):
        st = m.render(cam, p)
        if adaptive:
            # the stopping rule sees the same per-pixel sums on every rank (all-reduce per round): same samples per pixel
            f = gpu.film()
            assert np.array_equal(one[1][..., 3], f[..., 3])
            assert np.allclose(one[1][..., :3], f[..., :3], rtol=1e-9, atol=1e-12)
        else:
            _check_equal(one, (st, gpu.film(), gpu.film_sumsq()))
    m.close()
    gpu.close()
