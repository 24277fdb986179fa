"""CPU tier: the product's shading code (pathtracing_b200/csrc/shade*.cuh, camera.cuh, compiled for the host
by tests/emul) against the reference's own routines at the same, caller-chosen inputs (oracle/ref_driver.cpp
ref_unit_*).  Unlike the image tests this is deterministic: the reference's materials, lights, camera and
interaction code take their random numbers as arguments.  Tolerances are float round-off (both sides are
IEEE here; the GPU build of the same headers uses fast math and is covered by the statistical tests)."""
import ctypes as C

import numpy as np
import pytest

from conftest import needs_ref
from pathtracing_b200 import capi, synth

STRIDE = 44
vp, sz, i32, f32 = C.c_void_p, C.c_size_t, C.c_int, C.c_float


def _sig(lib, prefix):
    for name, args in (("unit_shade", [vp, vp, sz, vp, vp, i32]), ("unit_light_count", [vp]), ("unit_lights", [vp, vp]),
                       ("unit_light_sample", [vp, i32, sz, vp, vp, i32, f32]), ("unit_escape", [vp, sz, vp, vp]),
                       ("unit_camera", [vp, i32, i32, sz, vp, vp]), ("unit_hg", [f32, sz, vp, vp])):
        f = getattr(lib, f"{prefix}_{name}")
        f.restype, f.argtypes = i32, args
    pe = getattr(lib, f"{prefix}_unit_permutation_element")
    pe.restype, pe.argtypes = C.c_uint32, [C.c_uint32, C.c_uint32, C.c_uint32]
    h = getattr(lib, f"{prefix}_unit_hash")
    h.restype, h.argtypes = C.c_uint64, [C.c_uint32, C.c_uint32, C.c_uint64]


@pytest.fixture(scope="module")
def libs():
    from oracle import refapi
    import emulapi
    ref, emu = refapi.load("strict"), emulapi.load()
    _sig(ref, "ref")
    _sig(emu, "emul")
    emu.emul_shade_tables.restype, emu.emul_shade_tables.argtypes = i32, [vp, i32]
    return ref, emu


def _scenes(desc, width=4, light_sampler=1):
    from oracle.refapi import RefScene
    from emulapi import EmulScene
    r, e = RefScene(desc, width, light_sampler=light_sampler), EmulScene(desc, width)
    assert e.lib.emul_shade_tables(e.h, light_sampler) == 0, e.lib.emul_last_error()
    return r, e


def _camera_rays(cam, W, H, n, seed):
    """Rays through the scene as a camera would send them, plus second-bounce-like random ones."""
    rng = np.random.default_rng(seed)
    fr, at = np.array(cam.look_from[:], np.float32), np.array(cam.look_at[:], np.float32)
    w = fr - at
    w /= np.linalg.norm(w)
    u = np.cross([0, 1, 0], w)
    u /= np.linalg.norm(u)
    v = np.cross(w, u)
    hw = np.tan(cam.fov / 2)
    s, t = rng.uniform(-1, 1, n), rng.uniform(-1, 1, n)
    d = -w[None] + (s * hw)[:, None] * u[None] + (t * hw * H / W)[:, None] * v[None]
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros(n, capi.RAY_DTYPE)
    rays["o"], rays["d"], rays["tmax"] = fr, d.astype(np.float32), np.inf
    return rays


def _shade(libs, r, e, rays, seed=1):
    ref, emu = libs
    rng = np.random.default_rng(seed)
    n = len(rays)
    inp = np.concatenate([rng.random((n, 3), dtype=np.float32), rng.normal(size=(n, 3)).astype(np.float32)], 1)
    inp = np.ascontiguousarray(inp)
    a, b = np.zeros((n, STRIDE), np.float32), np.zeros((n, STRIDE), np.float32)
    assert ref.ref_unit_shade(r.h, rays.ctypes.data, n, inp.ctypes.data, a.ctypes.data, STRIDE) == 0
    assert emu.emul_unit_shade(e.h, rays.ctypes.data, n, inp.ctypes.data, b.ctypes.data, STRIDE) == 0
    return a, b


def _close(x, y, rtol, atol, what, frac=0.0):
    bad = ~np.isclose(x, y, rtol=rtol, atol=atol)
    if bad.ndim > 1:
        bad = bad.any(axis=tuple(range(1, bad.ndim)))
    assert bad.mean() <= frac, (what, float(bad.mean()), x[bad][:3], y[bad][:3])


@needs_ref
@pytest.mark.parametrize("width", [2, 4, 8])
def test_surface_interaction_and_materials(libs, width):
    """Shape.cpp:205-243 + Primitive.cpp:20-24,54-63 (p, n, ns, tangent, uv incl. normal maps and instance
    transforms) and Material::{scatter, calc_attenuation, PDF} of all four material classes."""
    desc, cam = synth.materials_scene(subdiv=3)
    r, e = _scenes(desc, width)
    rays = _camera_rays(cam, 64, 48, 20000, 3)
    # second generation: leave from the first hits in random directions (inside glass, grazing, from below)
    a, _ = _shade(libs, r, e, rays)
    hit = a[:, 0] == 1
    rng = np.random.default_rng(9)
    d2 = rng.normal(size=(hit.sum(), 3)).astype(np.float32)
    d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    r2 = np.zeros(hit.sum(), capi.RAY_DTYPE)
    r2["o"], r2["d"], r2["tmax"] = a[hit, 1:4] + 1e-3 * d2, d2, np.inf
    for rr in (rays, r2):
        a, b = _shade(libs, r, e, rr)
        assert np.array_equal(a[:, 0], b[:, 0]) and a[:, 0].mean() > 0.3
        h = a[:, 0] == 1
        assert np.array_equal(a[h, 16], b[h, 16]) and np.array_equal(a[h, 17], b[h, 17])          # material id, area light flag
        assert np.array_equal(a[h, 15].view(np.uint32), b[h, 15].view(np.uint32))                  # t: bit exact
        _close(a[h, 1:16], b[h, 1:16], 2e-5, 2e-6, "interaction")
        assert len(set(a[h, 16])) >= 7                                                           # every material was hit
        ok = h & (a[:, 16] >= 0)
        assert np.array_equal(a[ok, 18], b[ok, 18]), "scatter accepted / absorbed"
        s = ok & (a[:, 18] == 1)
        assert np.array_equal(a[s, 23], b[s, 23]), "BxDF flags"
        # sampled direction, f, pdf, then eval / pdf at that direction and at an unrelated one; a handful of
        # samples sit on a branch edge (total internal reflection, wi.z == 0) where round-off picks a side
        _close(a[s, 24:30], b[s, 24:30], 1e-4, 1e-5, "scattered ray", frac=2e-3)
        _close(a[s, 19:23], b[s, 19:23], 2e-3, 1e-6, "f, pdf", frac=5e-3)
        _close(a[s, 30:34], b[s, 30:34], 2e-3, 1e-6, "eval / pdf at the sampled direction", frac=5e-3)
        _close(a[ok, 34:38], b[ok, 34:38], 2e-3, 1e-6, "eval / pdf at w2", frac=2e-3)
        em = h & (a[:, 17] == 1)
        if em.any():
            _close(a[em, 38:43], b[em, 38:43], 1e-4, 1e-7, "emission, PMF, PDF of the hit light")
    r.close(); e.close()


@needs_ref
@pytest.mark.parametrize("sampler", [0, 1])
def test_lights_power_pmf_sample_pdf(libs, sampler):
    """Light::{sample, PDF, L, Power} + LightSampler::PMF for area lights (plain and inside a transformed
    instance), gradient sky, point and distant lights; Uniform and Power light samplers."""
    ref, emu = libs
    desc, cam = synth.materials_scene(subdiv=2, sky=True, point_light=True, distant_light=True)
    r, e = _scenes(desc, 4, light_sampler=sampler)
    n = ref.ref_unit_light_count(r.h)
    assert n == emu.emul_unit_light_count(e.h) and n >= 5
    ra, eb = np.zeros((n, 3), np.float32), np.zeros((n, 4), np.float32)
    ref.ref_unit_lights(r.h, ra.ctypes.data)
    emu.emul_unit_lights(e.h, eb.ctypes.data)
    assert np.array_equal(ra[:, 2], eb[:, 2])                                    # same light kinds in the same order
    _close(ra[:, 1], eb[:, 0], 1e-4, 1e-7, "sampler PMF per light")
    assert abs(ra[:, 1].sum() - 1.0) < 1e-4
    rng = np.random.default_rng(4)
    m = 2000
    for k in range(n):
        inp = np.ascontiguousarray(np.concatenate([rng.random((m, 2)), rng.uniform(-2, 2, (m, 3))], 1).astype(np.float32))
        a, b = np.zeros((m, 22), np.float32), np.zeros((m, 22), np.float32)
        assert ref.ref_unit_light_sample(r.h, k, m, inp.ctypes.data, a.ctypes.data, 22, 0.0) == 0
        assert emu.emul_unit_light_sample(e.h, k, m, inp.ctypes.data, b.ctypes.data, 22, 0.0) == 0
        if not a[:, 6:9].any():
            a[:, 9:11] = b[:, 9:11] = 0     # a direction sample's uv (GetSphereUV(dir) in the reference) is never read
        _close(a[:, :18], b[:, :18], 1e-4, 1e-5, f"light {k}: sample, shadow ray")
        _close(a[:, 18:22], b[:, 18:22], 2e-3, 1e-6, f"light {k}: PDF and L along the shadow ray", frac=1e-3)
    dirs = np.ascontiguousarray(rng.normal(size=(m, 3)).astype(np.float32))
    a, b = np.zeros((m, 4), np.float32), np.zeros((m, 4), np.float32)
    ref.ref_unit_escape(r.h, m, dirs.ctypes.data, a.ctypes.data)
    emu.emul_unit_escape(e.h, m, dirs.ctypes.data, b.ctypes.data)
    _close(a, b, 1e-5, 1e-7, "Le and PMF*PDF of the infinite lights for escaping rays")
    r.close(); e.close()


@needs_ref
def test_hdr_environment_le_and_pdf(libs):
    """TextureInfiniteLight::{Le, PDF} (Light.cpp:110-150): the environment lookup and the solid-angle pdf
    the reference derives from its 1920x1080 table total (the table itself is jitter-sampled there)."""
    ref, emu = libs
    desc, cam = synth.c5_scene(grid=24, n_textures=4, tex_size=8, n_lights=0, n_objects=2)
    r, e = _scenes(desc, 4)
    rng = np.random.default_rng(6)
    m = 4000
    dirs = np.ascontiguousarray(rng.normal(size=(m, 3)).astype(np.float32))
    a, b = np.zeros((m, 4), np.float32), np.zeros((m, 4), np.float32)
    ref.ref_unit_escape(r.h, m, dirs.ctypes.data, a.ctypes.data)
    emu.emul_unit_escape(e.h, m, dirs.ctypes.data, b.ctypes.data)
    _close(a[:, :3], b[:, :3], 1e-5, 1e-6, "Le of the environment map")
    # pdf = lum(Le) / totalWeight / cellOmega: totalWeight is a 64-sample jittered estimate per cell in the
    # reference and 4 strata centres here -> equal to well under a percent
    ratio = b[:, 3] / a[:, 3]
    assert abs(np.median(ratio) - 1.0) < 5e-3 and ratio.std() < 1e-4
    r.close(); e.close()


@needs_ref
def test_camera_generate_ray(libs):
    ref, emu = libs
    rng = np.random.default_rng(2)
    n = 5000
    for focus in ((0.0, 0.0), (0.15, 2.5)):
        from pathtracing_b200.scene import camera_desc
        cam = camera_desc((0.3, 1.2, 3.0), (-0.2, 0.4, 0.1), 1.1, focus_angle=focus[0], focus_distance=focus[1], shutter=(0.2, 0.7))
        inp = np.ascontiguousarray(np.concatenate([rng.uniform(0, 640, (n, 1)), rng.uniform(0, 360, (n, 1)), rng.random((n, 3))], 1))
        a, b = np.zeros((n, 7), np.float32), np.zeros((n, 7), np.float32)
        assert ref.ref_unit_camera(C.byref(cam), 640, 360, n, inp.ctypes.data, a.ctypes.data) == 0
        assert emu.emul_unit_camera(C.byref(cam), 640, 360, n, inp.ctypes.data, b.ctypes.data) == 0
        _close(a, b, 1e-5, 2e-6, f"Camera::GenerateRay, focus {focus}")


@needs_ref
def test_stratified_sampler_building_blocks_bit_exact(libs):
    """MurmurHash64A-based Hash(px, py, dimension) and PermutationElement (Util.hpp:45-168)."""
    ref, emu = libs
    rng = np.random.default_rng(5)
    for _ in range(3000):
        px, py, dim = int(rng.integers(0, 4096)), int(rng.integers(0, 4096)), int(rng.integers(0, 64))
        h = ref.ref_unit_hash(px, py, dim)
        assert h == emu.emul_unit_hash(px, py, dim)
        l = int(rng.integers(1, 2048))
        i = int(rng.integers(0, l))
        assert ref.ref_unit_permutation_element(i, l, h & 0xFFFFFFFF) == emu.emul_unit_permutation_element(i, l, h & 0xFFFFFFFF)
    l, p = 64, 0x9E3779B9
    assert sorted(emu.emul_unit_permutation_element(i, l, p) for i in range(l)) == list(range(l))


@needs_ref
@pytest.mark.parametrize("g", [0.0, 0.75, -0.6, 0.999])
def test_henyey_greenstein(libs, g):
    ref, emu = libs
    rng = np.random.default_rng(8)
    n = 4000
    inp = np.ascontiguousarray(np.concatenate([rng.normal(size=(n, 3)), rng.random((n, 2))], 1).astype(np.float32))
    a, b = np.zeros((n, 5), np.float32), np.zeros((n, 5), np.float32)
    ref.ref_unit_hg(g, n, inp.ctypes.data, a.ctypes.data)
    emu.emul_unit_hg(g, n, inp.ctypes.data, b.ctypes.data)
    _close(a[:, :3], b[:, :3], 1e-4, 2e-5, "HenyeyGreenstein::Sample direction", frac=1e-3)
    _close(a[:, 3:], b[:, 3:], 1e-3, 1e-6, "HenyeyGreenstein pdf", frac=1e-3)


@needs_ref
@pytest.mark.parametrize("one_sided", [False, True])
def test_sphere_and_quad_shapes(libs, one_sided):
    """SphereShape / QuadShape interactions (Shape.cpp:3-37, 316-341), their AreaLights (one- and two-sided:
    Light.cpp:257-287) incl. Sample / PDF / L / Power, next to a triangle mesh in the same TLAS."""
    ref, emu = libs
    desc, cam = synth.shapes_scene(one_sided=one_sided)
    r, e = _scenes(desc, 4)
    rays = _camera_rays(cam, 64, 48, 20000, 5)
    a, _ = _shade(libs, r, e, rays)
    hit = a[:, 0] == 1
    rng = np.random.default_rng(2)
    d2 = rng.normal(size=(hit.sum(), 3)).astype(np.float32)
    d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    r2 = np.zeros(hit.sum(), capi.RAY_DTYPE)
    r2["o"], r2["d"], r2["tmax"] = a[hit, 1:4] + 1e-3 * d2, d2, np.inf
    # rays aimed at the two lights from all around (front and back of the one-sided quad)
    tgt = np.where(rng.random((6000, 1)) < 0.5, np.array([[0.0, 2.05, -0.1]]), np.array([[-1.6, 1.0, -1.0]])) + rng.normal(size=(6000, 3)) * 0.25
    o3 = rng.uniform((-3, 0, -3), (3, 4, 3), (6000, 3))
    d3 = tgt - o3
    d3 /= np.linalg.norm(d3, axis=1, keepdims=True)
    r3 = np.zeros(6000, capi.RAY_DTYPE)
    r3["o"], r3["d"], r3["tmax"] = o3.astype(np.float32), d3.astype(np.float32), np.inf
    seen_lights = 0
    for rr in (rays, r2, r3):
        a, b = _shade(libs, r, e, rr)
        assert np.array_equal(a[:, 0], b[:, 0])
        h = a[:, 0] == 1
        assert np.array_equal(a[h, 16], b[h, 16]) and np.array_equal(a[h, 17], b[h, 17])
        assert np.array_equal(a[h, 15].view(np.uint32), b[h, 15].view(np.uint32))
        _close(a[h, 1:16], b[h, 1:16], 2e-5, 2e-6, "interaction", frac=1e-4)     # atan2 / acos round-off in GetSphereUV
        ok = h & (a[:, 16] >= 0)
        assert np.array_equal(a[ok, 18], b[ok, 18])
        s = ok & (a[:, 18] == 1)
        _close(a[s, 24:30], b[s, 24:30], 1e-4, 1e-5, "scattered ray", frac=2e-3)
        _close(a[s, 19:23], b[s, 19:23], 2e-3, 1e-6, "f, pdf", frac=5e-3)
        em = h & (a[:, 17] == 1)
        seen_lights += int(em.sum())
        if em.any():
            _close(a[em, 38:43], b[em, 38:43], 1e-4, 1e-7, "emission, PMF, PDF of the hit light")
    assert seen_lights > 500
    n = ref.ref_unit_light_count(r.h)
    assert n == emu.emul_unit_light_count(e.h) == 3
    ra, eb = np.zeros((n, 3), np.float32), np.zeros((n, 4), np.float32)
    ref.ref_unit_lights(r.h, ra.ctypes.data)
    emu.emul_unit_lights(e.h, eb.ctypes.data)
    _close(ra[:, 1], eb[:, 0], 2e-3, 1e-6, "sampler PMF per light (the reference jitters its 256 power samples)")
    m = 3000
    for k in range(n):
        inp = np.ascontiguousarray(np.concatenate([rng.random((m, 2)), rng.uniform(-2, 3, (m, 3))], 1).astype(np.float32))
        a, b = np.zeros((m, 22), np.float32), np.zeros((m, 22), np.float32)
        assert ref.ref_unit_light_sample(r.h, k, m, inp.ctypes.data, a.ctypes.data, 22, 0.0) == 0
        assert emu.emul_unit_light_sample(e.h, k, m, inp.ctypes.data, b.ctypes.data, 22, 0.0) == 0
        if not a[:, 6:9].any():
            a[:, 9:11] = b[:, 9:11] = 0
        _close(a[:, :18], b[:, :18], 1e-4, 1e-5, f"light {k}: sample, shadow ray", frac=1e-3)
        _close(a[:, 18:22], b[:, 18:22], 2e-3, 1e-6, f"light {k}: PDF and L along the shadow ray", frac=2e-3)
    r.close(); e.close()


@needs_ref
def test_wrapped_primitives_shading(libs):
    """TransformedPrimitive over sphere / quad GeometricPrimitives and AnimatedPrimitive wrappers (Primitive.cpp:32-96):
    the interaction that comes back through the wrappers at the ray's time, the inner AreaLight seen by a hit, and
    TransformedLight / AnimatedLight::{sample, PDF, L, Power} (Light.cpp:296-365) at several shutter times."""
    ref, emu = libs
    desc, cam = synth.wrapped_scene()
    r, e = _scenes(desc, 4)
    rng = np.random.default_rng(6)
    rays = synth.with_times(_camera_rays(cam, 64, 48, 20000, 5), seed=1)
    lo, hi = r.bounds()
    r2 = synth.with_times(synth.random_rays(30000, (-3, -0.3, -2.5), (3, 2.2, 1.5), seed=8), seed=2)
    # rays aimed at the two wrapped lights (a transformed two-sided quad, an animated sphere)
    tgt = np.where(rng.random((8000, 1)) < 0.5, np.array([[1.2, 1.7, -0.3]]), np.array([[-0.7, 1.5, 0.5]])) + rng.normal(size=(8000, 3)) * 0.2
    o3 = rng.uniform((-3, 0, -3), (3, 4, 3), (8000, 3))
    d3 = tgt - o3
    d3 /= np.linalg.norm(d3, axis=1, keepdims=True)
    r3 = np.zeros(8000, capi.RAY_DTYPE)
    r3["o"], r3["d"], r3["tmax"] = o3.astype(np.float32), d3.astype(np.float32), np.inf
    r3 = synth.with_times(r3, seed=3)
    wrapped = sorted(set(desc.shape_xf) | set(desc.animated))
    seen_lights = 0
    for rr in (rays, r2, r3):
        a, b = _shade(libs, r, e, rr)
        assert np.array_equal(a[:, 0], b[:, 0])
        h = a[:, 0] == 1
        assert np.array_equal(a[h, 16], b[h, 16]) and np.array_equal(a[h, 17], b[h, 17])
        assert np.array_equal(a[h, 15].view(np.uint32), b[h, 15].view(np.uint32))
        _close(a[h, 1:16], b[h, 1:16], 3e-5, 3e-6, "interaction", frac=2e-4)
        ok = h & (a[:, 16] >= 0)
        assert np.array_equal(a[ok, 18], b[ok, 18])
        s = ok & (a[:, 18] == 1)
        _close(a[s, 24:30], b[s, 24:30], 1e-4, 1e-5, "scattered ray", frac=2e-3)
        _close(a[s, 19:23], b[s, 19:23], 2e-3, 1e-6, "f, pdf", frac=5e-3)
        em = h & (a[:, 17] == 1)
        seen_lights += int(em.sum())
        if em.any():
            _close(a[em, 38:43], b[em, 38:43], 1e-4, 1e-7, "emission, PMF, PDF of the hit light")
    assert seen_lights > 500
    # the hits above did land on the wrapped primitives
    _, inst = e.trace_closest(r2)
    assert set(wrapped) <= set(np.unique(inst).tolist())
    n = ref.ref_unit_light_count(r.h)
    assert n == emu.emul_unit_light_count(e.h) == 5      # quad, sphere, transformed quad, animated sphere, sky
    ra, eb = np.zeros((n, 3), np.float32), np.zeros((n, 4), np.float32)
    ref.ref_unit_lights(r.h, ra.ctypes.data)
    emu.emul_unit_lights(e.h, eb.ctypes.data)
    assert np.array_equal(ra[:, 2], eb[:, 2])
    try:
        _close(ra[:, 1], eb[:, 0], 2e-3, 1e-6, "sampler PMF per light (the reference jitters its 256 power samples)")
    except AssertionError as err:
        # the only comparison here that depends on the reference's unseeded generator (AreaLight::PreProcess draws its 256
        # power samples from random_device-seeded PCG): repeat THIS comparison once against a fresh reference scene and say so.
        # Every other assert of this test is seeded and single-shot.
        import warnings
        warnings.warn(f"sampler PMF comparison failed once, repeating against a fresh reference scene: {err}")
        r_b, _e_b = _scenes(desc, 4)
        ra2 = np.zeros((n, 3), np.float32)
        ref.ref_unit_lights(r_b.h, ra2.ctypes.data)
        r_b.close(); _e_b.close()
        _close(ra2[:, 1], eb[:, 0], 2e-3, 1e-6, "sampler PMF per light, second reference scene")
    m = 3000
    for time in (0.0, 0.37, 0.9, 1.4):
        for k in range(n):
            inp = np.ascontiguousarray(np.concatenate([rng.random((m, 2)), rng.uniform(-2, 3, (m, 3))], 1).astype(np.float32))
            a, b = np.zeros((m, 22), np.float32), np.zeros((m, 22), np.float32)
            assert ref.ref_unit_light_sample(r.h, k, m, inp.ctypes.data, a.ctypes.data, 22, time) == 0
            assert emu.emul_unit_light_sample(e.h, k, m, inp.ctypes.data, b.ctypes.data, 22, time) == 0
            if not a[:, 6:9].any():
                a[:, 9:11] = b[:, 9:11] = 0
            _close(a[:, :18], b[:, :18], 1e-4, 1e-5, f"light {k} at time {time}: sample, shadow ray", frac=1e-3)
            _close(a[:, 18:22], b[:, 18:22], 2e-3, 1e-6, f"light {k} at time {time}: PDF and L along the shadow ray", frac=2e-3)
    r.close(); e.close()
