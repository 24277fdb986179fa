"""Binary SAH build (BVHBase::BuildBaseThreaded, BVH.hpp:290-390) through pt_build_bvh2.
CPU tier: structural invariants of the host builder (its equality with the reference's tree is what
tests/test_raycast_emul.py pins, ray by ray).  GPU tier: the device builder (csrc/bvh_build_gpu.cu) must
return the SAME tree — every node, split axis, box bit and leaf order — as the host builder, and a scene
committed with it must trace bit-identically."""
import numpy as np
import pytest

from pathtracing_b200 import capi, synth


def _tri_boxes(pos, idx):
    p = pos[idx.reshape(-1, 3)]
    return np.concatenate([p.min(axis=1), p.max(axis=1)], axis=1).astype(np.float32)


def _boxes(kind, n, seed=1):
    rng = np.random.default_rng(seed)
    if kind == "soup":
        pos, idx = synth.soup(n, seed=seed)
        return _tri_boxes(pos, idx)
    if kind == "terrain":
        nx = max(2, int(np.sqrt(n / 2)) + 1)
        pos, idx, _ = synth.terrain(nx, nx, seed=seed)
        return _tri_boxes(pos, idx)[:n]
    if kind == "lattice":          # many exactly equal centroids and costs: tie-breaking and degenerate partitions
        c = rng.integers(0, 6, size=(n, 3)).astype(np.float32)
        h = rng.integers(1, 3, size=(n, 3)).astype(np.float32) * 0.25
        return np.concatenate([c - h, c + h], axis=1)
    if kind == "same":             # all items identical: no axis is splittable
        return np.tile(np.array([[0, 0, 0, 1, 1, 1]], dtype=np.float32), (n, 1))
    if kind == "line":             # extent on one axis only, very uneven density
        x = (rng.random(n) ** 8).astype(np.float32) * 100
        z = np.zeros(n, dtype=np.float32)
        return np.stack([x, z, z, x + 0.01, z + 1, z + 1], axis=1)
    raise ValueError(kind)


def _check_tree(nodes, order, boxes):
    n = len(boxes)
    assert sorted(order.tolist()) == list(range(n))
    assert len(nodes) <= 2 * n - 1
    covered = np.zeros(n, dtype=np.int32)
    stack = [0]
    visited = 0
    while stack:
        k = stack.pop()
        nd = nodes[k]
        visited += 1
        if nd["count"]:
            items = order[nd["right"]:nd["right"] + nd["count"]]
            covered[nd["right"]:nd["right"] + nd["count"]] += 1
            b = boxes[items]
            assert np.array_equal(nd["lo"], b[:, :3].min(axis=0)) and np.array_equal(nd["hi"], b[:, 3:].max(axis=0))
        else:
            l, r = nodes[nd["left"]], nodes[nd["right"]]
            assert nd["left"] == k + 1 and nd["right"] > nd["left"]            # DFS preorder
            assert np.array_equal(nd["lo"], np.minimum(l["lo"], r["lo"])) and np.array_equal(nd["hi"], np.maximum(l["hi"], r["hi"]))
            assert nd["axis"] < 3
            stack += [nd["right"], nd["left"]]
    assert visited == len(nodes) and np.all(covered == 1)


@pytest.mark.parametrize("kind,n", [("soup", 1), ("soup", 2), ("soup", 3), ("soup", 777), ("terrain", 20000), ("lattice", 5000), ("same", 100), ("line", 3000)])
def test_host_builder_invariants(kind, n):
    boxes = _boxes(kind, n)
    nodes, order, _ = capi.build_bvh2(boxes, capi.BUILDER_HOST)
    _check_tree(nodes, order, boxes)
    if kind == "same":
        assert len(nodes) == 1 and nodes[0]["count"] == n


def test_gpu_builder_needs_a_device_on_cpu_box():
    lib = capi.load_library()
    if lib.pt_device_count() > 0:
        pytest.skip("GPU present")
    with pytest.raises(capi.PtError):
        capi.build_bvh2(_boxes("soup", 10), capi.BUILDER_GPU)


GPU_CASES = [("soup", 1), ("soup", 2), ("soup", 3), ("soup", 31), ("soup", 32), ("soup", 33), ("soup", 64), ("soup", 1000), ("soup", 2048), ("soup", 2049),
             ("soup", 5000), ("soup", 100000), ("soup", 1000000), ("terrain", 300000), ("lattice", 2000), ("lattice", 60000), ("same", 100), ("same", 5000),
             ("line", 3000), ("line", 200000)]


@pytest.mark.gpu
@pytest.mark.parametrize("kind,n", GPU_CASES)
def test_gpu_builder_equals_host_builder(kind, n):
    boxes = _boxes(kind, n, seed=3)
    hn, ho, _ = capi.build_bvh2(boxes, capi.BUILDER_HOST)
    gn, go, ms = capi.build_bvh2(boxes, capi.BUILDER_GPU)
    assert len(hn) == len(gn)
    assert np.array_equal(ho, go)
    for f in ("left", "right", "count", "axis"):
        assert np.array_equal(hn[f], gn[f]), f
    assert np.array_equal(hn["lo"], gn["lo"]) and np.array_equal(hn["hi"], gn["hi"])     # numerically (+0 == -0)


@pytest.mark.gpu
@pytest.mark.parametrize("width", [2, 4, 8])
def test_scene_committed_with_gpu_builder_traces_identically(width):
    from pathtracing_b200.scene import GpuScene
    for desc in (synth.soup_scene(150000), synth.instanced_scene(5, 12)):
        a = GpuScene(desc, width, builder=capi.BUILDER_HOST)
        b = GpuScene(desc, width, builder=capi.BUILDER_GPU)
        st = b.build_stats()
        assert st["builder"] == capi.BUILDER_GPU and st["gpu_items"] > 0 and st["gpu_build_ms"] > 0
        lo, hi = a.bounds()
        rays = synth.random_rays(300000, lo, hi, seed=9)
        ah, ai = a.trace_closest(rays)
        bh, bi = b.trace_closest(rays)
        assert np.array_equal(ah.view(np.uint8), bh.view(np.uint8)) and np.array_equal(ai, bi)
        assert np.array_equal(a.trace_any(rays), b.trace_any(rays))
        a.close(); b.close()


@pytest.mark.parametrize("width", [2, 4, 8])
@pytest.mark.parametrize("n,threads,clustered", [(5000, 3, 0), (200000, 7, 0), (150000, 16, 1)])
def test_threaded_flatten_gives_the_one_thread_arrays(width, n, threads, clustered):
    """flatten_bvh hands subtrees to worker threads and stitches their arrays back into DFS preorder
    (csrc/bvh_build.cpp flatten_threaded): byte-identical BVH2 / BVH4 / BVH8 node arrays, leaf marks and root."""
    import ctypes as C
    import emulapi
    nodes = C.c_int(0)
    bad = emulapi.load().emul_flatten_check(n, 11 + width, width, threads, clustered, C.byref(nodes))
    assert bad == 0 and nodes.value > n // 8, (bad, nodes.value)


@pytest.mark.gpu
@pytest.mark.parametrize("width,layout", [(4, 0), (4, 1), (2, 0), (8, 0)])
def test_blas_packed_on_the_device_gives_the_host_arrays(width, layout):
    """PT_BUILDER_GPU packs every large BLAS (BVH2, BVH4 or BVH8) on the device (csrc/scene_pack_gpu.cu: boxes, binary build, collapse, triangle records,
    quantised clusters): the scene-wide tables must be, byte for byte, the ones the host path packs (pt_unit_table) -- multi-mesh BLAS, several
    BLASes of which one is small enough to stay on the host, instances, and a material with an alpha test among them."""
    from pathtracing_b200.scene import GpuScene
    cases = [synth.soup_scene(150000), synth.instanced_scene(5, 48), synth.c3_scene(subdiv=4, n_instances=9)[0], synth.c5_scene(grid=300, n_objects=6)[0],
             synth.alpha_scene()[0] if hasattr(synth, "alpha_scene") else synth.terrain_scene(200, 150)]
    for k, desc in enumerate(cases):
        a = GpuScene(desc, width, builder=capi.BUILDER_HOST, layout=layout)
        b = GpuScene(desc, width, builder=capi.BUILDER_GPU, layout=layout)
        st = b.build_stats()
        assert st["gpu_items"] > 0 or k == 4          # the last scene is all small BLASes: everything stays on the host
        for t in ((0, 1, 4) if layout == 0 else (0, 1, 2, 3, 4)) + (5, 6, 7, 8, 9, 10, 11):
            ta, tb = a.table(t), b.table(t)
            assert len(ta) == len(tb) and len(ta) > 0, (t, len(ta), len(tb))
            diff = int((ta != tb).sum())
            assert diff == 0, (t, diff, len(ta))
        a.close(); b.close()
