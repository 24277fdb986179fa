"""GPU tier: the CUDA ray-cast kernels, called through the C-ABI, against the
UNMODIFIED reference (oracle/_ref).  Integer work (primitive / instance ids) is
bit-exact; hit distances are bit-exact against the -ffp-contract=off build of the
reference and within 4 ulp against its default (FMA-contracting) Release build."""
import numpy as np
import pytest

from conftest import needs_ref
from pathtracing_b200 import synth
from pathtracing_b200.scene import GpuScene, SceneDesc

pytestmark = pytest.mark.gpu


def _cmp(desc, width, n_rays=200000, tmax=np.inf, seed=7):
    from oracle.refapi import RefScene
    ref = RefScene(desc, width, id_mode=True)
    gpu = GpuScene(desc, width)
    lo, hi = ref.bounds()
    glo, ghi = gpu.bounds()
    assert np.array_equal(lo, glo) and np.array_equal(hi, ghi)
    rays = synth.random_rays(n_rays, lo, hi, seed=seed, tmax=tmax)
    rh, ri = ref.trace_closest(rays, want_inst=True)
    gh, gi = gpu.trace_closest(rays)
    assert np.array_equal(rh["prim"], gh["prim"])
    assert np.array_equal(rh["t"].view(np.uint32), gh["t"].view(np.uint32))
    assert np.array_equal(ri, gi)
    assert np.array_equal(ref.trace_any(rays), gpu.trace_any(rays))
    gpu.close()
    return rh


@needs_ref
@pytest.mark.parametrize("width", [2, 4])
@pytest.mark.parametrize("opaque", [False, True])
def test_soup_bit_exact(width, opaque):
    rh = _cmp(synth.soup_scene(100000, opaque=opaque), width)
    assert (rh["prim"] != 0xFFFFFFFF).mean() > 0.2


@needs_ref
@pytest.mark.parametrize("width", [2, 4])
def test_terrain_bit_exact(width):
    _cmp(synth.terrain_scene(301, 201), width)


@needs_ref
@pytest.mark.parametrize("width", [2, 4])
def test_instances_bit_exact(width):
    d = synth.instanced_scene(3, 24)
    _cmp(d, width, n_rays=100000)
    _cmp(d, width, n_rays=50000, tmax=2.0, seed=11)


@needs_ref
@pytest.mark.parametrize("width", [2, 4])
@pytest.mark.parametrize("n", [1, 2, 3, 5])
def test_tiny_blas(width, n):
    _cmp(synth.soup_scene(n, seed=5), width, n_rays=3000)


@needs_ref
@pytest.mark.parametrize("width", [2, 4])
def test_against_default_release_build(width):
    """Reference compiled with its own Release flags contracts mul+add into FMA; ids must
    still agree except for near-ties, t within a few ulp."""
    from oracle.refapi import RefScene
    d = synth.soup_scene(100000)
    ref = RefScene(d, width, id_mode=True, flavour="fast")
    gpu = GpuScene(d, width)
    lo, hi = ref.bounds()
    rays = synth.random_rays(200000, lo, hi)
    rh, _ = ref.trace_closest(rays)
    gh, _ = gpu.trace_closest(rays, want_inst=False)
    same = rh["prim"] == gh["prim"]
    assert same.mean() > 0.9999
    hit = same & (rh["prim"] != 0xFFFFFFFF)
    ulp = np.abs(rh["t"][hit].view(np.int32).astype(np.int64) - gh["t"][hit].view(np.int32).astype(np.int64))
    # grazing hits amplify the rounding difference of a contracted determinant
    assert np.percentile(ulp, 99) <= 8 and np.percentile(ulp, 99.9) <= 32 and ulp.max() <= 1 << 14
    gpu.close()


def test_empty_and_ragged_batches():
    d = synth.soup_scene(1000)
    gpu = GpuScene(d, 4)
    lo, hi = gpu.bounds()
    h, i = gpu.trace_closest(synth.random_rays(0, lo, hi))
    assert len(h) == 0
    for n in (1, 31, 129, 1025):
        rays = synth.random_rays(n, lo, hi)
        h, i = gpu.trace_closest(rays)
        h2, i2 = gpu.trace_closest(rays[::-1].copy())
        assert np.array_equal(h, h2[::-1]) and np.array_equal(i, i2[::-1])
    gpu.close()


def test_chunked_host_path_equals_single_launch():
    """the host-buffer entry point cuts the batch into chunks over two streams"""
    d = synth.soup_scene(20000)
    gpu = GpuScene(d, 4)
    lo, hi = gpu.bounds()
    rays = synth.random_rays((1 << 21) + 12345, lo, hi)
    h, i = gpu.trace_closest(rays)
    k = 1 << 20
    h2, i2 = gpu.trace_closest(rays[k - 7:k + 9])
    assert np.array_equal(h[k - 7:k + 9], h2) and np.array_equal(i[k - 7:k + 9], i2)
    occ = gpu.trace_any(rays)
    assert np.array_equal(occ != 0, h["prim"] != 0xFFFFFFFF)   # tmax = inf: occluded iff something is hit
    gpu.close()


def test_bvh2_and_bvh4_agree_at_full_size():
    """size-independent property at BASELINE's C2 size: two different hierarchies over the same
    1M triangles return identical primitive ids and bit-identical distances (ties excepted)."""
    d = synth.soup_scene(1_000_000)
    g2, g4 = GpuScene(d, 2), GpuScene(d, 4)
    lo, hi = g2.bounds()
    rays = synth.random_rays(1 << 21, lo, hi)
    h2, _ = g2.trace_closest(rays, want_inst=False)
    h4, _ = g4.trace_closest(rays, want_inst=False)
    diff = h2["prim"] != h4["prim"]
    assert diff.mean() < 1e-5
    assert np.array_equal(h2["t"][~diff].view(np.uint32), h4["t"][~diff].view(np.uint32))
    # hits lie inside the scene box and t reproduces a point on the reported triangle's plane
    hit = h4["prim"] != 0xFFFFFFFF
    p = rays["o"][hit] + h4["t"][hit, None] * rays["d"][hit]
    assert (p >= lo - 1e-4).all() and (p <= hi + 1e-4).all()
    g2.close()
    g4.close()


def test_error_paths():
    from pathtracing_b200 import capi
    lib = capi.load_library()
    d = SceneDesc()
    with pytest.raises(capi.PtError):
        GpuScene(d, 4)            # no instances
    d = synth.soup_scene(10)
    with pytest.raises(capi.PtError):
        GpuScene(d, 3)            # bad width
    g = GpuScene(d, 4)
    assert lib.pt_scene_commit(g.h, 4, 1) == -4   # PT_ERR_STATE: already committed
    g.close()
