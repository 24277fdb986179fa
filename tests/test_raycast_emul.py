"""CPU tier: the product's host BVH build + flattening + the device traversal header
compiled for the host (tests/emul) against the UNMODIFIED reference (oracle/_ref,
-ffp-contract=off build).  Bit-exact primitive ids, instance ids and hit distances."""
import numpy as np
import pytest

from conftest import needs_ref
from pathtracing_b200 import synth


def _cmp(desc, width, n_rays=40000, tmax=np.inf, seed=7):
    from emulapi import EmulScene
    from oracle.refapi import RefScene
    ref = RefScene(desc, width, id_mode=True)
    em = EmulScene(desc, width)
    lo, hi = ref.bounds()
    elo, ehi = em.bounds()
    assert np.array_equal(lo, elo) and np.array_equal(hi, ehi)
    rays = synth.random_rays(n_rays, lo, hi, seed=seed, tmax=tmax)
    rh, ri = ref.trace_closest(rays, want_inst=True)
    eh, ei = em.trace_closest(rays)
    assert np.array_equal(rh["prim"], eh["prim"])
    assert np.array_equal(rh["t"].view(np.uint32), eh["t"].view(np.uint32))
    assert np.array_equal(ri, ei)
    assert np.array_equal(ref.trace_any(rays), em.trace_any(rays))
    return (rh["prim"] != 0xFFFFFFFF).mean()


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
@pytest.mark.parametrize("opaque", [False, True])
def test_soup_bit_exact(width, opaque):
    assert _cmp(synth.soup_scene(20000, opaque=opaque), width) > 0.05


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
def test_terrain_bit_exact(width):
    assert _cmp(synth.terrain_scene(101, 81), width) > 0.2


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
def test_instances_bit_exact(width):
    d = synth.instanced_scene(2, 16)
    assert _cmp(d, width) > 0.1
    _cmp(d, width, tmax=1.5, seed=11)


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
@pytest.mark.parametrize("n", [1, 2, 3, 5])
def test_tiny_blas(width, n):
    """root is a leaf (n<=2) or a partially filled cluster"""
    _cmp(synth.soup_scene(n, seed=5), width, n_rays=2000)
    d = synth.soup_scene(n, seed=5)
    d.instances.clear()
    d.instance(0, np.diag([2.0, 0.5, 1.5, 1.0]))
    _cmp(d, width, n_rays=2000)


@needs_ref
def test_degenerate_identical_triangles():
    """many coincident centroids: SAH cannot split, big leaves"""
    pos = np.tile(np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32), (40, 1))
    idx = np.arange(120, dtype=np.uint32).reshape(-1, 3)
    from pathtracing_b200.scene import SceneDesc
    d = SceneDesc()
    m = d.diffuse(d.solid((1, 1, 1)))
    d.instance(d.add_blas([d.mesh(pos, idx, material=m)]))
    from emulapi import EmulScene
    from oracle.refapi import RefScene
    for width in (2, 4, 8):
        ref, em = RefScene(d, width, id_mode=True), EmulScene(d, width)
        rays = synth.random_rays(2000, (-0.5, -0.5, -1), (1.5, 1.5, 1))
        rh, _ = ref.trace_closest(rays)
        eh, _ = em.trace_closest(rays)
        # all 40 triangles tie exactly: the LAST one tested wins (t == max is accepted)
        assert np.array_equal(rh["prim"], eh["prim"])
        assert np.array_equal(rh["t"].view(np.uint32), eh["t"].view(np.uint32))


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
def test_shapes_bit_exact(width):
    """SphereShape / QuadShape GeometricPrimitives in the TLAS next to BLAS instances (Shape.cpp:3-56, 316-359)"""
    d, _ = synth.shapes_scene()
    from emulapi import EmulScene
    from oracle.refapi import RefScene
    ref, em = RefScene(d, width, id_mode=True), EmulScene(d, width)
    lo, hi = ref.bounds()
    elo, ehi = em.bounds()
    assert np.array_equal(lo, elo) and np.array_equal(hi, ehi)
    for seed, tmax, box in ((7, np.inf, ((-3, -0.3, -2.5), (3, 2.2, 1.5))), (9, 2.0, ((-3, -0.3, -2.5), (3, 2.2, 1.5))), (11, np.inf, (lo, hi))):
        rays = synth.random_rays(30000, box[0], box[1], seed=seed, tmax=tmax)
        rh, ri = ref.trace_closest(rays, want_inst=True)
        eh, ei = em.trace_closest(rays)
        assert np.array_equal(ri, ei)
        assert np.array_equal(rh["prim"], eh["prim"])
        assert np.array_equal(rh["t"].view(np.uint32), eh["t"].view(np.uint32))
        shape_hit = np.isin(ri, [k for k, (b, _) in enumerate(d.instances) if b < 0])
        assert np.array_equal(rh["u"][shape_hit].view(np.uint32), eh["u"][shape_hit].view(np.uint32))
        assert np.array_equal(rh["v"][shape_hit].view(np.uint32), eh["v"][shape_hit].view(np.uint32))
        assert np.array_equal(ref.trace_any(rays), em.trace_any(rays))
        if seed == 7:
            assert shape_hit.mean() > 0.3 and (~shape_hit & (rh["prim"] != 0xFFFFFFFF)).mean() > 0.02


@needs_ref
def test_shapes_only_scene():
    """no BLAS at all: the TLAS holds only analytic shapes; also a single-shape scene (TLAS root is a leaf)"""
    from emulapi import EmulScene
    from oracle.refapi import RefScene
    d, _ = synth.shapes_scene(mesh=False)
    from pathtracing_b200.scene import SceneDesc
    one = SceneDesc()
    one.sphere((0.1, 0.2, 0.3), 0.7, material=one.diffuse(one.solid((1, 1, 1))))
    for desc, box in ((d, ((-3, -0.3, -2.5), (3, 2.2, 1.5))), (one, ((-1, -1, -1), (1, 1, 1)))):
        for width in (2, 4, 8):
            ref, em = RefScene(desc, width, id_mode=True), EmulScene(desc, width)
            rays = synth.random_rays(8000, box[0], box[1], seed=3)
            rh, ri = ref.trace_closest(rays, want_inst=True)
            eh, ei = em.trace_closest(rays)
            assert np.array_equal(ri, ei) and np.array_equal(rh["t"].view(np.uint32), eh["t"].view(np.uint32))
            assert np.array_equal(ref.trace_any(rays), em.trace_any(rays))


@needs_ref
def test_translation_inverse_is_exact():
    """AnimatedPrimitive builds TransformedPrimitive(translate(dir * t)) per ray (Primitive.cpp:82-89); trace.cuh
    assumes glm::inverse of that matrix is the identity with column 3 = -dir*t, bit for bit"""
    import ctypes as C
    from emulapi import load
    lib = load()
    rng = np.random.default_rng(0)
    for _ in range(2000):
        v = (rng.standard_normal(3) * 10.0 ** rng.integers(-6, 4)).astype(np.float32)
        m = np.eye(4, dtype=np.float32)
        m[3, :3] = v                      # column-major storage: column 3
        out = np.zeros(16, np.float32)
        lib.emul_mat4_inverse(m.reshape(16).ctypes.data_as(C.POINTER(C.c_float)), out.ctypes.data_as(C.POINTER(C.c_float)))
        want = np.eye(4, dtype=np.float32)
        want[3, :3] = -v
        assert np.array_equal(out.reshape(4, 4), want)      # == treats -0 and +0 alike, as every use of the matrix does


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
def test_wrapped_primitives_bit_exact(width):
    """TransformedPrimitive over sphere / quad GeometricPrimitives and AnimatedPrimitive over shapes, transformed
    shapes, BLASes and transformed BLASes (Primitive.cpp:32-96), rays carrying a time"""
    d, _ = synth.wrapped_scene()
    from emulapi import EmulScene
    from oracle.refapi import RefScene
    ref, em = RefScene(d, width, id_mode=True), EmulScene(d, width)
    lo, hi = ref.bounds()
    elo, ehi = em.bounds()
    assert np.array_equal(lo, elo) and np.array_equal(hi, ehi)
    wrapped = sorted(set(d.shape_xf) | set(d.animated))
    seen = set()
    for seed, tmax, box in ((7, np.inf, ((-3, -0.3, -2.5), (3, 2.2, 1.5))), (9, 2.0, ((-3, -0.3, -2.5), (3, 2.2, 1.5))), (11, np.inf, (lo, hi))):
        rays = synth.with_times(synth.random_rays(40000, box[0], box[1], seed=seed, tmax=tmax), seed=seed)
        rh, ri = ref.trace_closest(rays, want_inst=True)
        eh, ei = em.trace_closest(rays)
        assert np.array_equal(ri, ei)
        assert np.array_equal(rh["prim"], eh["prim"])
        assert np.array_equal(rh["t"].view(np.uint32), eh["t"].view(np.uint32))
        shape_hit = np.isin(ri, [k for k, (b, _) in enumerate(d.instances) if b < 0])
        assert np.array_equal(rh["u"][shape_hit].view(np.uint32), eh["u"][shape_hit].view(np.uint32))
        assert np.array_equal(rh["v"][shape_hit].view(np.uint32), eh["v"][shape_hit].view(np.uint32))
        assert np.array_equal(ref.trace_any(rays), em.trace_any(rays))
        seen |= set(np.unique(ri[np.isin(ri, wrapped)]).tolist())
    assert seen == set(wrapped), (seen, wrapped)      # every wrapped primitive was actually hit


# ---- PT_LAYOUT_FAST (quantised 64 B clusters, csrc/trace.cuh node_step4q) ---------------------------------------------

def _cmp_fast(desc, n_rays=60000, tmax=np.inf, seed=7, allow=0):
    """The fast layout walks the reference's BVH4 tree with boxes that CONTAIN the exact ones: it may test more
    candidates than the reference, never fewer.  Same closest hit (ids and the bits of t) except where candidates tie
    or the reference's own box test rounds a grazing hit away -- `allow` bounds those, and each one must be a hit the
    fast layout found IN FRONT of (or instead of a miss of) the reference's."""
    from emulapi import EmulScene
    from oracle.refapi import RefScene
    ref = RefScene(desc, 4, id_mode=True)
    exact = EmulScene(desc, 4)
    fast = EmulScene(desc, 44)
    lo, hi = ref.bounds()
    rays = synth.random_rays(n_rays, lo, hi, seed=seed, tmax=tmax)
    rh, ri = ref.trace_closest(rays, want_inst=True)
    fh, fi = fast.trace_closest(rays)
    diff = (rh["prim"] != fh["prim"]) | (rh["t"].view(np.uint32) != fh["t"].view(np.uint32)) | (ri != fi)
    assert diff.sum() <= allow, (int(diff.sum()), rh[diff][:5], fh[diff][:5])
    if diff.any():
        assert (fh["t"][diff] <= rh["t"][diff]).all()
    occ_r, occ_f = ref.trace_any(rays), fast.trace_any(rays)
    d2 = occ_r != occ_f
    assert d2.sum() <= allow and (occ_f[d2] == 1).all()
    # and it does visit a superset: never fewer node / triangle visits than the exact layout, and not many more
    exact.trace_closest(rays)
    se, sf = exact.last_stats, fast.last_stats
    for k in ("tlas_nodes", "blas_nodes", "triangles"):
        assert sf[k] >= se[k], (k, se, sf)
    assert sf["blas_nodes"] <= 1.15 * se["blas_nodes"] + 100 and sf["triangles"] <= 1.15 * se["triangles"] + 100, (se, sf)
    return (rh["prim"] != 0xFFFFFFFF).mean()


@needs_ref
@pytest.mark.parametrize("opaque", [False, True])
def test_fast_layout_soup(opaque):
    assert _cmp_fast(synth.soup_scene(20000, opaque=opaque)) > 0.05


@needs_ref
def test_fast_layout_terrain_instances_tiny():
    assert _cmp_fast(synth.terrain_scene(101, 81)) > 0.2
    d = synth.instanced_scene(2, 16)
    assert _cmp_fast(d) > 0.1
    _cmp_fast(d, tmax=1.5, seed=11)
    for n in (1, 2, 3, 5):
        _cmp_fast(synth.soup_scene(n, seed=5), n_rays=2000)


@needs_ref
def test_fast_layout_shapes_and_wrappers():
    _cmp_fast(synth.shapes_scene()[0], n_rays=30000)
    d, _ = synth.wrapped_scene()
    lo, hi = (-3, -0.3, -2.5), (3, 2.2, 1.5)
    from emulapi import EmulScene
    from oracle.refapi import RefScene
    ref, fast = RefScene(d, 4, id_mode=True), EmulScene(d, 44)
    rays = synth.with_times(synth.random_rays(30000, lo, hi, seed=8), seed=2)
    rh, ri = ref.trace_closest(rays, want_inst=True)
    fh, fi = fast.trace_closest(rays)
    assert np.array_equal(rh["prim"], fh["prim"]) and np.array_equal(ri, fi) and np.array_equal(rh["t"].view(np.uint32), fh["t"].view(np.uint32))


def test_fast_layout_boxes_contain_the_exact_ones():
    """Decode every quantised cluster the way a reader of the format would (org + q * 2^(ex-127), float arithmetic)
    and check containment with at least a quarter step of margin on the low side (emul_bvh4q_check)."""
    from emulapi import EmulScene
    for desc in (synth.soup_scene(30000), synth.terrain_scene(121, 91), synth.instanced_scene(3, 24)):
        e = EmulScene(desc, 44)
        bad, nodes, worst = e.bvh4q_check()
        assert nodes > 0 and bad == 0, (bad, nodes)
        assert worst <= 1.25 / 127.0 * 1.05, worst         # a box grows by at most 1.25 steps per side, a step is at most 1/127 of the cluster extent (power-of-two steps)


def test_stack_overflow_flag_on_a_small_stack():
    """The host build of the traversal with a 20-row stack (libptemul_stack20.so): the deep-chain scene in BVH8 overflows on the two
    rays along +x and they come back wrong -- and flagged; the 96-row build answers all four like the reference.  GPU twin:
    test_raycast_gpu.py::test_stack_overflow_is_a_counted_error."""
    from emulapi import EmulScene
    from test_raycast_gpu import _deep_rays
    desc, rays = synth.deep_chain_scene(230), _deep_rays()
    full, small = EmulScene(desc, 8), EmulScene(desc, 8, lib_name="libptemul_stack20.so")
    want, _ = full.trace_closest(rays)
    assert full.stack_overflows() == 0 and (want["prim"] != 0xFFFFFFFF).all()
    got, _ = small.trace_closest(rays)
    assert small.stack_overflows() == 2
    assert np.array_equal(got[2:], want[2:]) and not np.array_equal(got[:2]["prim"], want[:2]["prim"])
    from oracle import refapi
    if refapi.available("strict"):
        from oracle.refapi import RefScene
        r = RefScene(desc, 8, id_mode=True)
        rh, _ = r.trace_closest(rays, want_inst=True)
        assert np.array_equal(rh["prim"], want["prim"]) and np.array_equal(rh["t"].view(np.uint32), want["t"].view(np.uint32))
        r.close()
