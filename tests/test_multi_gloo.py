"""CPU tier, world_size 2 over gloo: the host logic of the multi-GPU path (pass sharding + the one
film reduce) without a GPU.  Each 'rank' accumulates a deterministic per-pass film; the reduced
film on rank 0 must equal the single-rank film exactly (integer-valued doubles: order-free)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pathtracing_b200 import multi
from pathtracing_b200.scene import render_params

W, H, SPP = 16, 12, 7


def fake_pass_film(k):
    """what one sample pass adds: integer-valued RGB*w and w=1 per pixel"""
    y, x = np.mgrid[0:H, 0:W]
    f = np.zeros((H, W, 4), np.float64)
    f[..., 0] = (x * 3 + y * 5 + k * 7) % 11
    f[..., 1] = (x + k) % 5
    f[..., 2] = (y * k) % 3
    f[..., 3] = 1.0
    return f


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    p = multi.pass_range(render_params(W, H, SPP, 4), world, rank)
    assert (p.pass_begin, p.pass_stride) == (rank, world)
    passes = list(range(p.pass_begin, p.spp, p.pass_stride))
    assert passes == multi.shard_passes(SPP, world, rank)
    film = torch.zeros(4 * W * H, dtype=torch.float64)
    for k in passes:
        film += torch.from_numpy(fake_pass_film(k).reshape(-1))
    multi.reduce_film(film, dist)
    if rank == 0:
        np.save(out, film.numpy())
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.timeout(120)
def test_two_rank_reduce_equals_single_rank(tmp_path):
    out = str(tmp_path / "film.npy")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = np.load(out).reshape(H, W, 4)
    want = sum(fake_pass_film(k) for k in range(SPP))
    assert np.array_equal(got, want)
    assert np.array_equal(got[..., 3], np.full((H, W), SPP))
    img = multi.resolve(got)
    assert np.allclose(img, want[..., :3] / SPP)


def _allreduce_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cb = multi.make_allreduce(dist, cuda=False)        # the pt_allreduce_fn pt_render calls between adaptive rounds
    stats = np.arange(6 * 5, dtype=np.float64) * (rank + 1)
    rc = cb(None, stats.ctypes.data, stats.size, None)
    assert rc == 0
    if rank == 0:
        np.save(out, stats)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_adaptive_round_allreduce_callback(tmp_path):
    """the per-round sum of the per-pixel estimator statistics (sharded adaptive renders), through the
    same ctypes callback type the C-ABI takes"""
    out = str(tmp_path / "stats.npy")
    mp.spawn(_allreduce_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert np.array_equal(np.load(out), np.arange(30, dtype=np.float64) * 3)


def test_shards_partition_the_passes():
    for spp in (1, 2, 7, 16, 1024):
        for world in (1, 2, 3, 4, 8):
            allp = sorted(k for r in range(world) for k in multi.shard_passes(spp, world, r))
            assert allp == list(range(spp))
            sizes = [len(multi.shard_passes(spp, world, r)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
