"""GPU tier: the CUDA ray-cast kernels, called through the C-ABI, against the
UNMODIFIED reference (oracle/_ref).  Integer work (primitive / instance ids) is
bit-exact; hit distances are bit-exact against the -ffp-contract=off build of the
reference and within 4 ulp against its default (FMA-contracting) Release build."""
import numpy as np
import pytest

from conftest import needs_ref
from pathtracing_b200 import capi, synth
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
@pytest.mark.parametrize("width", [2, 4, 8])
@pytest.mark.parametrize("opaque", [False, True])
def test_soup_bit_exact(width, opaque):
    rh = _cmp(synth.soup_scene(100000, opaque=opaque), width)
    assert (rh["prim"] != 0xFFFFFFFF).mean() > 0.2


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
def test_terrain_bit_exact(width):
    _cmp(synth.terrain_scene(301, 201), width)


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
def test_instances_bit_exact(width):
    d = synth.instanced_scene(3, 24)
    _cmp(d, width, n_rays=100000)
    _cmp(d, width, n_rays=50000, tmax=2.0, seed=11)


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
@pytest.mark.parametrize("n", [1, 2, 3, 5])
def test_tiny_blas(width, n):
    _cmp(synth.soup_scene(n, seed=5), width, n_rays=3000)


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
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


# ---- BASELINE.json C2 at its real size, against the reference itself (oracle/_ref), not GPU-vs-GPU ----------------------
# 1 000 000 triangles, the two meshes SURVEY.md section 8d names (triangle soup, displaced-grid "surface"), 2 Mi incoherent rays
# (a prefix-sized sample of the 16 Mi-ray set the bench times: the reference does ~3.5 Mrays/s on the box's host threads),
# BVH2 / BVH4 / BVH8: primitive ids, the bits of t and the any-hit flags.  Reference side: BVH.hpp:503-548, 1111-1211.

def _full_size_scene(mesh):
    return synth.soup_scene(1_000_000) if mesh == "soup" else synth.terrain_scene(1001, 501)      # 1000 x 500 x 2 triangles


@needs_ref
@pytest.mark.parametrize("mesh", ["soup", "surface"])
@pytest.mark.parametrize("width", [4, 2, 8])
def test_c2_full_size_against_the_reference(mesh, width):
    from oracle.refapi import RefScene
    d = _full_size_scene(mesh)
    ref = RefScene(d, width, id_mode=True)
    gpu = GpuScene(d, width, layout=capi.LAYOUT_EXACT)
    lo, hi = ref.bounds()
    rays = synth.random_rays(1 << 21, lo, hi, seed=7)
    rh, _ = ref.trace_closest(rays)
    gh, _ = gpu.trace_closest(rays, want_inst=False)
    assert np.array_equal(rh["prim"], gh["prim"])
    assert np.array_equal(rh["t"].view(np.uint32), gh["t"].view(np.uint32))
    assert (rh["prim"] != 0xFFFFFFFF).mean() > 0.2
    k = 1 << 20                                        # any-hit on half of them, with finite and infinite tmax
    sh = rays[:k].copy()
    sh["tmax"][: k // 2] = 0.35
    assert np.array_equal(ref.trace_any(sh), gpu.trace_any(sh))
    assert gpu.stack_overflows() == 0
    gpu.close(); ref.close()


# ---- PT_LAYOUT_FAST: quantised clusters, same tree ------------------------------------------------------------------------

def _cmp_fast(desc, n_rays=200000, tmax=np.inf, seed=7, allow=0):
    """Same closest hit as the reference (ids, bits of t, any-hit flag).  `allow` bounds the rays where the two may differ: ties (two
    triangles hit at bit-equal t: the fast closest-hit kernel visits children by entry distance, not in the reference's per-octant
    order, so the other one can win) and hits that the reference's own box arithmetic rounds away by an ulp or two and the fast layout,
    whose boxes contain the reference's, finds -- never the reverse: the fast kernel's t is never larger.  Measured at full size
    (scratch/r02_tie_check.py, 2 x 16 Mi rays): 12 rays, 10 ties and 2 hits that are 1 and 2 ulp closer."""
    from oracle.refapi import RefScene
    ref = RefScene(desc, 4, id_mode=True)
    gpu = GpuScene(desc, 4, layout=capi.LAYOUT_FAST)
    lo, hi = ref.bounds()
    rays = synth.random_rays(n_rays, lo, hi, seed=seed, tmax=tmax)
    rh, ri = ref.trace_closest(rays, want_inst=True)
    gh, gi = gpu.trace_closest(rays)
    diff = (rh["prim"] != gh["prim"]) | (rh["t"].view(np.uint32) != gh["t"].view(np.uint32)) | (ri != gi)
    assert diff.sum() <= allow, (int(diff.sum()), rh[diff][:4], gh[diff][:4])
    assert (gh["t"][diff] <= rh["t"][diff]).all()
    ro, go = ref.trace_any(rays), gpu.trace_any(rays)
    d2 = ro != go
    assert d2.sum() <= allow and (go[d2] == 1).all()
    gpu.close(); ref.close()
    return rh


@needs_ref
def test_fast_layout_small_scenes():
    rh = _cmp_fast(synth.soup_scene(100000))
    assert (rh["prim"] != 0xFFFFFFFF).mean() > 0.2
    _cmp_fast(synth.soup_scene(100000, opaque=True))
    _cmp_fast(synth.terrain_scene(301, 201))
    d = synth.instanced_scene(3, 24)
    _cmp_fast(d, n_rays=100000)
    _cmp_fast(d, n_rays=50000, tmax=2.0, seed=11)
    for n in (1, 2, 3, 5):
        _cmp_fast(synth.soup_scene(n, seed=5), n_rays=3000)
    _cmp_fast(synth.shapes_scene()[0], n_rays=50000)


@needs_ref
@pytest.mark.parametrize("mesh", ["soup", "surface"])
def test_fast_layout_c2_full_size_against_the_reference(mesh):
    _cmp_fast(_full_size_scene(mesh), n_rays=1 << 21, allow=4)


def test_fast_layout_needs_width_4_and_streams_like_the_exact_one():
    d = synth.soup_scene(20000)
    with pytest.raises(capi.PtError):
        GpuScene(d, 8, layout=capi.LAYOUT_FAST)
    fast, exact = GpuScene(d, 4, layout=capi.LAYOUT_FAST), GpuScene(d, 4)
    lo, hi = fast.bounds()
    rays = synth.random_rays((1 << 21) + 777, lo, hi)          # large enough for the streamed host-buffer path
    hf, i_f = fast.trace_closest(rays)
    he, ie = exact.trace_closest(rays)
    assert np.array_equal(hf.view(np.uint8), he.view(np.uint8)) and np.array_equal(i_f, ie)
    assert np.array_equal(fast.trace_any(rays), exact.trace_any(rays))
    fast.close(); exact.close()


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


@pytest.mark.parametrize("width,layout", [(4, 0), (4, 1), (2, 0), (8, 0)])
def test_compact_transfer_form_equals_the_full_one(width, layout):
    """pt_trace_closest_compact (24 B rays in, 8 B hits out): same t bits, primitive and instance ids as pt_trace_closest on the
    expanded rays -- in the chunked form (small batch; a scene with wrappers) and in the streamed one (>= 1 Mi rays, plain scene)."""
    for desc, n in ((synth.instanced_scene(3, 24), 3000), (synth.instanced_scene(3, 24), (1 << 20) + 12345), (synth.wrapped_scene()[0], 50000)):
        g = GpuScene(desc, width, layout=layout)
        lo, hi = g.bounds()
        rays = synth.random_rays(n, lo, hi, seed=width + n)
        r24 = np.zeros(n, capi.RAY24_DTYPE)
        r24["o"], r24["d"] = rays["o"], rays["d"]
        want_h, want_i = g.trace_closest(rays)
        got_h, got_i = g.trace_closest_compact(r24)
        assert np.array_equal(got_h["t"].view(np.uint32), want_h["t"].view(np.uint32)) and np.array_equal(got_h["prim"], want_h["prim"])
        assert np.array_equal(got_i, want_i) and (want_h["prim"] != 0xFFFFFFFF).mean() > 0.05
        got_h2, none = g.trace_closest_compact(r24, want_inst=False)
        assert none is None and np.array_equal(got_h2, got_h)
        g.close()
    with pytest.raises(capi.PtError):
        g2 = GpuScene(synth.soup_scene(10), 4)
        try:
            g2._ck(g2.lib.pt_trace_closest_compact(g2.h, None, 5, None, None))
        finally:
            g2.close()


def _deep_rays():
    r = np.zeros(4, dtype=capi.RAY_DTYPE)
    r["tmax"] = np.inf
    r["o"][:2], r["o"][2:] = (-1, 0, 0), (2, 0, 0)
    for k, d in enumerate(((1, 0, 0), (1, 1e-3, 1e-3), (-1, 0, 0), (-1, 1e-3, 0))):
        r["d"][k] = np.asarray(d, np.float32) / np.linalg.norm(d)
    return r


def test_stack_overflow_is_a_counted_error():
    """A ray that needs more traversal-stack rows than there are used to lose a push silently (a possibly wrong hit, no error).
    Now the ray is flagged, counted, and every entry point that traced it returns PT_ERR_OVERFLOW.  No scene a SAH builder makes
    comes near the product's 96 rows (the deepest test scene needs 22), so the error path runs against the library's test-only twin
    built with 20 rows (csrc/Makefile, PTB_STACK=20): synth.deep_chain_scene in BVH8 needs more for rays along +x."""
    from pathtracing_b200 import capi
    small = capi.load_library(capi.LIB_PATH_SMALL_STACK)
    desc, rays = synth.deep_chain_scene(230), _deep_rays()
    good = GpuScene(desc, 8)
    want, _ = good.trace_closest(rays)
    assert good.stack_overflows() == 0
    g = GpuScene(desc, 8, lib=small)
    hits = np.empty(len(rays), capi.HIT_DTYPE)
    rc = small.pt_trace_closest(g.h, rays.ctypes.data, len(rays), hits.ctypes.data, None)
    assert rc == -6 and b"stack overflow on 2 rays" in small.pt_last_error(), (rc, small.pt_last_error())      # PT_ERR_OVERFLOW
    assert g.stack_overflows() == 2
    assert np.array_equal(hits[2:], want[2:]) and not np.array_equal(hits[:2]["prim"], want[:2]["prim"])        # the two flagged rays are the ones that went wrong
    occ = np.empty(len(rays), np.uint8)
    assert small.pt_trace_any(g.h, rays.ctypes.data, len(rays), occ.ctypes.data) in (0, -6)                     # any-hit stops at the first hit: may or may not get that deep
    # the same rays in BVH4 fit in 20 rows: no error, right answer, and the count does not move
    g4 = GpuScene(desc, 4, lib=small)
    h4 = np.empty(len(rays), capi.HIT_DTYPE)
    assert small.pt_trace_closest(g4.h, rays.ctypes.data, len(rays), h4.ctypes.data, None) == 0
    assert np.array_equal(h4["prim"], want["prim"]) and g4.stack_overflows() == 0
    for s in (good, g, g4):
        s.close()


def _alpha_rays(n, seed=3):
    """Rays from in front of the Cornell opening towards the cards (most of them hit a card first)."""
    rng = np.random.default_rng(seed)
    o = np.stack([rng.uniform(0.05, 0.95, n), rng.uniform(0.1, 0.9, n), np.full(n, 2.0)], -1).astype(np.float32)
    tgt = np.stack([rng.uniform(0.05, 0.95, n), rng.uniform(0.1, 0.9, n), np.full(n, 0.9)], -1).astype(np.float32)
    d = tgt - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    from pathtracing_b200 import capi
    rays = np.zeros(n, dtype=capi.RAY_DTYPE)
    rays["o"], rays["d"] = o, d
    rays["tmax"] = np.inf
    return rays


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
@pytest.mark.parametrize("four_channel", [False, True])
def test_alpha_mask_inside_traversal_bit_exact(width, four_channel):
    """AlphaMode::Mask is deterministic: rejected candidates let the ray continue to exactly the hit the
    reference finds (Primitive.cpp:15-18), for closest-hit and for the any-hit path (Primitive.cpp:6-11)."""
    from oracle.refapi import RefScene
    from pathtracing_b200 import capi
    desc, _ = synth.alpha_scene(capi.ALPHA_MASK, four_channel)
    ref, gpu = RefScene(desc, width), GpuScene(desc, width)
    rays = _alpha_rays(60000)
    rh, _ = ref.trace_closest(rays)
    gh, gi = gpu.trace_closest(rays)
    assert np.array_equal(rh["t"].view(np.uint32), gh["t"].view(np.uint32))
    hit = gh["prim"] != 0xFFFFFFFF
    # the cards are instances 1 and 2: some rays stop on them, some pass through the cut-outs
    assert 0.05 < (gi[hit] == 1).mean() < 0.6 and (gi[hit] == 2).any() and (gi[hit] == 0).mean() > 0.2
    rays["tmax"] = 1.6
    assert np.array_equal(ref.trace_any(rays), gpu.trace_any(rays))
    gpu.close(); ref.close()


@needs_ref
def test_alpha_blend_inside_traversal_statistics():
    """AlphaMode::Blend accepts a candidate with probability alpha(uv): per-ray outcomes differ from the
    reference's (its draw is an unseeded thread-local PCG), the acceptance rate per opacity bin may not."""
    from oracle.refapi import RefScene
    from pathtracing_b200 import capi
    desc, _ = synth.alpha_scene(capi.ALPHA_BLEND)
    ref, gpu = RefScene(desc, 4), GpuScene(desc, 4)
    rays = _alpha_rays(400000)
    rh, ri = ref.trace_closest(rays, want_inst=True)
    gh, gi = gpu.trace_closest(rays)
    # card 1 is the nearest surface wherever it is hit: its hits have t below any other surface's
    for k in (1, 2):
        pr, pg = float((ri == k).mean()), float((gi == k).mean())
        se = np.sqrt(pr * (1 - pr) / len(rays) * 2)
        assert abs(pr - pg) < 5 * se + 1e-4, (k, pr, pg)
    t_r, t_g = rh["t"][np.isfinite(rh["t"])], gh["t"][np.isfinite(gh["t"])]
    assert abs(t_r.mean() - t_g.mean()) < 5 * np.sqrt((t_r.var() + t_g.var()) / len(t_r))
    rays["tmax"] = 1.6
    o_r, o_g = ref.trace_any(rays).mean(), gpu.trace_any(rays).mean()
    assert abs(o_r - o_g) < 5 * np.sqrt(o_r * (1 - o_r) * 2 / len(rays)) + 1e-4, (o_r, o_g)
    # two launches draw differently (the salt advances)
    gh2, _ = gpu.trace_closest(rays)
    gh3, _ = gpu.trace_closest(rays)
    assert (gh2["t"] != gh3["t"]).mean() > 0.01
    gpu.close(); ref.close()


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
@pytest.mark.parametrize("mesh", [True, False])
def test_sphere_and_quad_shapes_bit_exact(width, mesh):
    """SphereShape / QuadShape GeometricPrimitives as TLAS leaf items (Shape.cpp:3-56, 316-359), alone and
    next to BLAS instances: ids, t and the quad's (alpha, beta) bit-exact, any-hit equal."""
    from oracle.refapi import RefScene
    d, _ = synth.shapes_scene(mesh=mesh)
    ref, gpu = RefScene(d, width, id_mode=True), GpuScene(d, width)
    lo, hi = ref.bounds()
    glo, ghi = gpu.bounds()
    assert np.array_equal(lo, glo) and np.array_equal(hi, ghi)
    shapes = [k for k, (b, _) in enumerate(d.instances) if b < 0]
    for seed, tmax, box in ((7, np.inf, ((-3, -0.3, -2.5), (3, 2.2, 1.5))), (9, 2.0, ((-3, -0.3, -2.5), (3, 2.2, 1.5))), (11, np.inf, (lo, hi))):
        rays = synth.random_rays(100000, box[0], box[1], seed=seed, tmax=tmax)
        rh, ri = ref.trace_closest(rays, want_inst=True)
        gh, gi = gpu.trace_closest(rays)
        assert np.array_equal(ri, gi) and np.array_equal(rh["prim"], gh["prim"])
        assert np.array_equal(rh["t"].view(np.uint32), gh["t"].view(np.uint32))
        sh = np.isin(ri, shapes)
        assert np.array_equal(rh["u"][sh].view(np.uint32), gh["u"][sh].view(np.uint32))
        assert np.array_equal(rh["v"][sh].view(np.uint32), gh["v"][sh].view(np.uint32))
        assert np.array_equal(ref.trace_any(rays), gpu.trace_any(rays))
    gpu.close()


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
def test_wrapped_primitives_bit_exact(width):
    """TransformedPrimitive over sphere / quad GeometricPrimitives, AnimatedPrimitive over shapes, transformed shapes,
    BLASes and transformed BLASes (Primitive.cpp:32-96), rays carrying a time: ids, t, (alpha, beta), any-hit."""
    from oracle.refapi import RefScene
    d, _ = synth.wrapped_scene()
    ref, gpu = RefScene(d, width, id_mode=True), GpuScene(d, width)
    lo, hi = ref.bounds()
    glo, ghi = gpu.bounds()
    assert np.array_equal(lo, glo) and np.array_equal(hi, ghi)
    shapes = [k for k, (b, _) in enumerate(d.instances) if b < 0]
    wrapped = sorted(set(d.shape_xf) | set(d.animated))
    seen = set()
    for seed, tmax, box in ((7, np.inf, ((-3, -0.3, -2.5), (3, 2.2, 1.5))), (9, 2.0, ((-3, -0.3, -2.5), (3, 2.2, 1.5))), (11, np.inf, (lo, hi))):
        rays = synth.with_times(synth.random_rays(100000, box[0], box[1], seed=seed, tmax=tmax), seed=seed)
        rh, ri = ref.trace_closest(rays, want_inst=True)
        gh, gi = gpu.trace_closest(rays)
        assert np.array_equal(ri, gi) and np.array_equal(rh["prim"], gh["prim"])
        assert np.array_equal(rh["t"].view(np.uint32), gh["t"].view(np.uint32))
        sh = np.isin(ri, shapes)
        assert np.array_equal(rh["u"][sh].view(np.uint32), gh["u"][sh].view(np.uint32))
        assert np.array_equal(rh["v"][sh].view(np.uint32), gh["v"][sh].view(np.uint32))
        assert np.array_equal(ref.trace_any(rays), gpu.trace_any(rays))
        seen |= set(np.unique(ri[np.isin(ri, wrapped)]).tolist())
    assert seen == set(wrapped)
    gpu.close()


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
def test_alpha_mask_on_shapes(width):
    """Alpha-tested materials on sphere / quad GeometricPrimitives (Primitive.cpp:6-26), plain and inside a
    TransformedPrimitive.  Quads are bit-exact (their uv is the hit's (alpha, beta)); a sphere's uv goes through
    acos / atan2, whose last bit differs between libm and CUDA, so a candidate within an ulp of a checker edge may
    fall on the other side: a handful of rays in 10^5."""
    from oracle.refapi import RefScene
    from pathtracing_b200 import capi
    desc, _ = synth.alpha_shapes_scene(capi.ALPHA_MASK)
    ref, gpu = RefScene(desc, width), GpuScene(desc, width)      # not id mode: the cards' alpha needs their real texcoords
    rays = _alpha_rays(100000)
    rays["o"][:, 2] = 2.4
    rh, ri = ref.trace_closest(rays, want_inst=True)
    gh, gi = gpu.trace_closest(rays)
    same = (rh["t"].view(np.uint32) == gh["t"].view(np.uint32)) & (ri == gi)
    shapes = [k for k, (b, _) in enumerate(desc.instances) if b < 0]
    spheres = [k for k in shapes if desc.instances[k][1].kind == capi.SHAPE_SPHERE]
    assert (~same).mean() < 2e-4, (~same).mean()
    for k in shapes:
        hit_k = (gi == k).mean()
        assert 0.005 < hit_k, (k, hit_k)                           # every alpha-tested shape stops some rays ...
    # ... and lets others through: rays aimed at the sphere's disc that end up behind it
    c, r = np.array([0.3, 0.6, 1.25]), 0.18
    oc = rays["o"] - c
    b = (oc * rays["d"]).sum(1)
    disc = b * b - ((oc * oc).sum(1) - r * r) > 0
    assert 0.2 < (gi[disc] == spheres[0]).mean() < 0.95
    rays["tmax"] = 1.6
    ra, ga = ref.trace_any(rays), gpu.trace_any(rays)
    assert (ra != ga).mean() < 2e-4
    gpu.close(); ref.close()


def test_wrapper_entry_points_reject_bad_arguments():
    """pt_scene_animate / pt_scene_add_shape_xf: error codes + pt_last_error instead of undefined behaviour"""
    import ctypes as C
    from pathtracing_b200 import capi
    lib = capi.load_library()
    h = lib.pt_scene_create(0)
    assert h
    d = SceneDesc()
    d.sphere((0, 0, 0), 1.0, material=d.diffuse(d.solid((1, 1, 1))))
    for t in d.textures:
        assert lib.pt_scene_add_texture(h, C.byref(t)) >= 0
    for m in d.materials:
        assert lib.pt_scene_add_material(h, C.byref(m)) >= 0
    shape = d.instances[0][1]
    assert lib.pt_scene_add_shape_xf(h, C.byref(shape), None) < 0 and b"transform" in lib.pt_last_error()
    pid = lib.pt_scene_add_shape(h, C.byref(shape))
    assert pid == 0
    v, tb = (C.c_float * 3)(0, 1, 0), (C.c_float * 2)(0, 1)
    assert lib.pt_scene_animate(h, 5, v, tb) < 0 and b"out of range" in lib.pt_last_error()
    assert lib.pt_scene_animate(h, pid, None, tb) < 0
    assert lib.pt_scene_animate(h, pid, v, tb) == 0
    assert lib.pt_scene_animate(h, pid, v, tb) < 0 and b"nested" in lib.pt_last_error()
    assert lib.pt_scene_commit(h, 4, 1) == 0
    assert lib.pt_scene_animate(h, pid, v, tb) < 0 and b"committed" in lib.pt_last_error()
    lib.pt_scene_destroy(h)


@pytest.mark.parametrize("width", [2, 4, 8])
def test_streamed_host_buffer_ray_cast_equals_device_resident(width):
    """pt_trace_closest / pt_trace_any with host buffers of >= 1 Mi rays run ONE persistent kernel that is fed by the H2D
    stream through a watermark and hands results back per finished granule (capi.cu trace_streamed): same bits as the
    device-resident launch of the same rays (which the tests above pin to the reference), odd batch size included."""
    import torch
    from pathtracing_b200 import capi
    desc = synth.soup_scene(100000, opaque=True)
    scene = GpuScene(desc, width)
    lo, hi = scene.bounds()
    n = (1 << 20) + (1 << 18) + 4321
    rays = synth.random_rays(n, lo, hi, seed=5)
    for rep in range(2):                      # second call reuses the staging and control buffers
        h, i = scene.trace_closest(rays)
        d = torch.from_numpy(rays.view(np.uint8).reshape(-1)).cuda()
        dh = torch.empty(n * 16, dtype=torch.uint8, device="cuda")
        di = torch.empty(n * 4, dtype=torch.uint8, device="cuda")
        scene.trace_closest_dev(d.data_ptr(), n, dh.data_ptr(), di.data_ptr(), torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert np.array_equal(h.view(np.uint8), dh.cpu().numpy()) and np.array_equal(i, di.cpu().numpy().view(np.uint32))
        assert (h["prim"] != 0xFFFFFFFF).mean() > 0.1
    short = rays.copy()
    short["tmax"] = np.float32(0.1)
    occ = scene.trace_any(short)
    d = torch.from_numpy(short.view(np.uint8).reshape(-1)).cuda()
    do = torch.empty(n, dtype=torch.uint8, device="cuda")
    scene.trace_any_dev(d.data_ptr(), n, do.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(occ, do.cpu().numpy()) and 0.01 < occ.mean() < 0.9
    scene.close()
