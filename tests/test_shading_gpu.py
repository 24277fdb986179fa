"""GPU tier: the DEVICE build of the shading code (shade*.cuh as render.cu compiles it: sm_100a, -use_fast_math),
probed one record per thread through the C-ABI (pt_unit_*, csrc/unit.cu), against the reference's own functions at the
same seeded inputs (oracle/ref_driver.cpp ref_unit_*: Material.hpp:200-673, Light.cpp:9-365, PhaseFunction.cpp:4-26,
Shape.cpp:205-243, Primitive.cpp:20-24,54-63).  Deterministic -- no Monte-Carlo noise in between, unlike the image
tests.  The host (IEEE) build of the same headers is checked in the CPU tier (test_shading_units.py) at round-off
tolerances; here the tolerances are those of fast math (approximate division / sqrt / rsqrt: 2 ulp; __sinf / __cosf /
__expf / __logf: absolute error about 2^-21) and are written next to every comparison:

  geometry of the hit (p, n, ns, tangent, uv)     rtol 1e-4, atol 1e-5
  t, material id, light flag, scatter accepted     exact (traversal is IEEE in every build)
  sampled direction and origin                     rtol 2e-3, atol 2e-4   on all but 0.5 % (branch edges: TIR, wi.z == 0)
  f, pdf, eval                                     rtol 1e-2, atol 1e-5   on all but 1 %   (GGX D near grazing amplifies round-off)
  emission, PMF, light PDF                         rtol 2e-3, atol 1e-6
"""
import ctypes as C

import numpy as np
import pytest

from conftest import needs_ref
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene
from test_shading_units import STRIDE, _camera_rays, _close, _sig

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref():
    from oracle import refapi
    lib = refapi.load("strict")
    _sig(lib, "ref")
    return lib


def _shade(ref, r, g, rays, seed=1):
    rng = np.random.default_rng(seed)
    n = len(rays)
    inp = np.ascontiguousarray(np.concatenate([rng.random((n, 3), dtype=np.float32), rng.normal(size=(n, 3)).astype(np.float32)], 1))
    a, b = np.zeros((n, STRIDE), np.float32), np.zeros((n, STRIDE), np.float32)
    assert ref.ref_unit_shade(r.h, rays.ctypes.data, n, inp.ctypes.data, a.ctypes.data, STRIDE) == 0
    assert g.lib.pt_unit_shade(g.h, rays.ctypes.data, n, inp.ctypes.data, b.ctypes.data, STRIDE) == 0, g._err()
    return a, b


def _check_shade(a, b, min_hit=0.3, min_mats=1):
    assert np.array_equal(a[:, 0], b[:, 0]) and a[:, 0].mean() > min_hit
    h = a[:, 0] == 1
    assert np.array_equal(a[h, 16], b[h, 16]) and np.array_equal(a[h, 17], b[h, 17])          # material id, area light flag
    assert np.array_equal(a[h, 15].view(np.uint32), b[h, 15].view(np.uint32))                  # t: bit exact
    _close(a[h, 1:15], b[h, 1:15], 1e-4, 1e-5, "interaction: p, n, ns, tangent, uv", frac=1e-3)
    assert len(set(a[h, 16])) >= min_mats
    ok = h & (a[:, 16] >= 0)
    flip = a[ok, 18] != b[ok, 18]
    assert flip.mean() <= 2e-3, ("scatter accepted / absorbed", float(flip.mean()))            # Fresnel / lobe choice right at u == threshold
    s = ok & (a[:, 18] == 1) & (b[:, 18] == 1)
    fl = a[s, 23] != b[s, 23]
    assert fl.mean() <= 2e-3, ("BxDF flags", float(fl.mean()))
    s[s] &= ~fl
    _close(a[s, 24:30], b[s, 24:30], 2e-3, 2e-4, "scattered ray", frac=5e-3)
    _close(a[s, 19:23], b[s, 19:23], 1e-2, 1e-5, "f, pdf", frac=1e-2)
    _close(a[s, 30:34], b[s, 30:34], 1e-2, 1e-5, "eval / pdf at the sampled direction", frac=1e-2)
    _close(a[ok, 34:38], b[ok, 34:38], 1e-2, 1e-5, "eval / pdf at w2", frac=5e-3)
    em = h & (a[:, 17] == 1)
    if em.any():
        _close(a[em, 38:43], b[em, 38:43], 2e-3, 1e-6, "emission, PMF, PDF of the hit light", frac=1e-3)
    return int(em.sum())


@needs_ref
@pytest.mark.parametrize("width,layout", [(4, 0), (4, 1), (2, 0), (8, 0)])
def test_device_interaction_and_materials(ref, width, layout):
    from oracle.refapi import RefScene
    desc, cam = synth.materials_scene(subdiv=3)
    r, g = RefScene(desc, width), GpuScene(desc, width, layout=layout)
    rays = _camera_rays(cam, 64, 48, 20000, 3)
    a, _ = _shade(ref, r, g, rays)
    hit = a[:, 0] == 1
    rng = np.random.default_rng(9)
    d2 = rng.normal(size=(hit.sum(), 3)).astype(np.float32)
    d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    r2 = np.zeros(hit.sum(), capi.RAY_DTYPE)
    r2["o"], r2["d"], r2["tmax"] = a[hit, 1:4] + 1e-3 * d2, d2, np.inf
    for rr in (rays, r2):
        _check_shade(*_shade(ref, r, g, rr), min_mats=7)
    r.close(); g.close()


@needs_ref
def test_device_shapes_wrappers_and_their_lights(ref):
    from oracle.refapi import RefScene
    rng = np.random.default_rng(2)
    seen = 0
    for desc, cam, timed in ((*synth.shapes_scene(one_sided=False), False), (*synth.shapes_scene(one_sided=True), False), (*synth.wrapped_scene(), True)):
        r, g = RefScene(desc, 4), GpuScene(desc, 4)
        rays = _camera_rays(cam, 64, 48, 20000, 5)
        tgt = np.where(rng.random((6000, 1)) < 0.5, np.array([[0.0, 2.05, -0.1]]), np.array([[-1.6, 1.0, -1.0]])) + rng.normal(size=(6000, 3)) * 0.25
        o3 = rng.uniform((-3, 0, -3), (3, 4, 3), (6000, 3))
        d3 = tgt - o3
        d3 /= np.linalg.norm(d3, axis=1, keepdims=True)
        r3 = np.zeros(6000, capi.RAY_DTYPE)
        r3["o"], r3["d"], r3["tmax"] = o3.astype(np.float32), d3.astype(np.float32), np.inf
        for k, rr in enumerate((rays, r3)):
            if timed:
                rr = synth.with_times(rr, seed=k + 1)
            seen += _check_shade(*_shade(ref, r, g, rr), min_hit=0.2)
        r.close(); g.close()
    assert seen > 500


@needs_ref
@pytest.mark.parametrize("sampler", [0, 1])
def test_device_light_sampling(ref, sampler):
    """Light::{sample, PDF, L} for area lights (plain and inside a transformed instance), gradient sky, point and
    distant lights, the wrapped lights at several shutter times; rtol 2e-3 on sample / shadow ray, 5e-3 on PDF and L."""
    from oracle.refapi import RefScene
    rng = np.random.default_rng(6)
    m = 3000
    for desc, times in ((synth.materials_scene(subdiv=2, sky=True, point_light=True, distant_light=True)[0], (0.0,)), (synth.wrapped_scene()[0], (0.0, 0.37, 1.4))):
        r, g = RefScene(desc, 4, light_sampler=sampler), GpuScene(desc, 4, light_sampler=sampler)
        n = ref.ref_unit_light_count(r.h)
        assert n == g.lib.pt_unit_light_count(g.h) and n >= 5
        for time in times:
            for k in range(n):
                inp = np.ascontiguousarray(np.concatenate([rng.random((m, 2)), rng.uniform(-2, 3, (m, 3))], 1).astype(np.float32))
                a, b = np.zeros((m, 22), np.float32), np.zeros((m, 22), np.float32)
                assert ref.ref_unit_light_sample(r.h, k, m, inp.ctypes.data, a.ctypes.data, 22, time) == 0
                assert g.lib.pt_unit_light_sample(g.h, k, m, inp.ctypes.data, b.ctypes.data, 22, C.c_float(time)) == 0, g._err()
                if not a[:, 6:9].any():
                    a[:, 9:11] = b[:, 9:11] = 0
                _close(a[:, :18], b[:, :18], 2e-3, 2e-4, f"light {k} at time {time}: sample, shadow ray", frac=2e-3)
                _close(a[:, 18:22], b[:, 18:22], 5e-3, 1e-6, f"light {k} at time {time}: PDF and L along the shadow ray", frac=5e-3)
        r.close(); g.close()


@needs_ref
@pytest.mark.parametrize("g", [0.0, 0.75, -0.6, 0.999])
def test_device_henyey_greenstein(ref, g):
    """HenyeyGreenstein::Sample / PDF: direction atol 2e-4 (fast sin / cos / sqrt), pdf rtol 2e-3."""
    lib = capi.load_library()
    rng = np.random.default_rng(8)
    n = 4000
    inp = np.ascontiguousarray(np.concatenate([rng.normal(size=(n, 3)), rng.random((n, 2))], 1).astype(np.float32))
    a, b = np.zeros((n, 5), np.float32), np.zeros((n, 5), np.float32)
    ref.ref_unit_hg(g, n, inp.ctypes.data, a.ctypes.data)
    assert lib.pt_unit_hg(0, C.c_float(g), n, inp.ctypes.data, b.ctypes.data) == 0
    _close(a[:, :3], b[:, :3], 1e-3, 2e-4, "HenyeyGreenstein::Sample direction", frac=2e-3)
    _close(a[:, 3:], b[:, 3:], 2e-3, 1e-6, "HenyeyGreenstein pdf", frac=2e-3)
