"""GPU tier, collected last on purpose: pt_trace_closest_static, the
host-buffer closest-hit ray cast that allocates nothing (its staging buffers and ray-dispenser cursor live in a
`static __device__` array inside libptb200.so), against the existing entry point pt_trace_closest.

Both run the same kernel, so the comparison is bit for bit (integer ids and the bits of t, u, v) -- the tolerance
every ray-cast test of this repo uses.  Two calls in a row: the scratch keeps the previous call's contents, so the
second call (a different, shorter batch) would show anything that was not cleared or fully overwritten.
No counterpart in the reference (include/pt_b200.h)."""
import ctypes as C

import numpy as np
import pytest

from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene

pytestmark = pytest.mark.gpu


def _static(gpu, rays, want_inst=True):
    rays = np.ascontiguousarray(rays)
    n = len(rays)
    hits = np.full(n, 0x7F, np.uint8).repeat(16).view(capi.HIT_DTYPE).copy()      # poisoned: every field must be written
    inst = np.full(n, 0x12345678, np.uint32)
    rc = gpu.lib.pt_trace_closest_static(gpu.h, rays.ctypes.data, n, hits.ctypes.data, inst.ctypes.data if want_inst else None)
    assert rc == 0, gpu.lib.pt_last_error().decode()
    return hits, inst


def _same(a, b):
    for f in ("t", "u", "v"):
        assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), f
    assert np.array_equal(a["prim"], b["prim"])


@pytest.mark.parametrize("width", [4, 2, 8])
def test_static_entry_matches_pt_trace_closest_twice_in_a_row(width):
    desc = synth.instanced_scene(3, 24)          # TLAS + transformed BLAS instances: instance ids matter too
    gpu = GpuScene(desc, width)
    lo, hi = gpu.bounds()
    first = synth.random_rays(capi.STATIC_MAX_RAYS, lo, hi, seed=21)              # exactly the maximum
    second = synth.random_rays(40001, lo, hi, seed=22, tmax=2.5)                  # shorter, other rays: stale scratch would show
    for rays in (first, second):
        want_h, want_i = gpu.trace_closest(rays)
        got_h, got_i = _static(gpu, rays)
        _same(got_h, want_h)
        assert np.array_equal(got_i, want_i)
        assert (want_h["prim"] != capi.PT_MISS).mean() > 0.05
    # inst == NULL is allowed, like in pt_trace_closest
    got_h, untouched = _static(gpu, second, want_inst=False)
    _same(got_h, gpu.trace_closest(second)[0])
    assert (untouched == 0x12345678).all()
    gpu.close()


def test_static_entry_rejects_sizes_beyond_its_scratch():
    gpu = GpuScene(synth.soup_scene(2000, seed=3), 4)
    lo, hi = gpu.bounds()
    rays = synth.random_rays(capi.STATIC_MAX_RAYS + 1, lo, hi, seed=5)
    hits = np.zeros(len(rays), capi.HIT_DTYPE)
    rc = gpu.lib.pt_trace_closest_static(gpu.h, rays.ctypes.data, len(rays), hits.ctypes.data, None)
    assert rc == -1                                                                # PT_ERR_INVALID
    msg = gpu.lib.pt_last_error().decode()
    assert "PT_STATIC_MAX_RAYS" in msg and str(capi.STATIC_MAX_RAYS) in msg, msg
    assert not hits.view(np.uint8).any()                                           # nothing was written
    # n == 0 is a no-op, and the scene still works afterwards
    assert gpu.lib.pt_trace_closest_static(gpu.h, None, 0, None, None) == 0
    got, _ = _static(gpu, rays[:1000])
    _same(got, gpu.trace_closest(rays[:1000])[0])
    gpu.close()


def test_static_entry_on_a_second_scene_shares_the_scratch():
    """Two scenes on one device use the same array one after the other (not overlapping): results stay right."""
    a = GpuScene(synth.soup_scene(5000, seed=1), 4)
    b = GpuScene(synth.terrain_scene(41, 31), 2)
    for gpu, seed in ((a, 1), (b, 2), (a, 3)):
        lo, hi = gpu.bounds()
        rays = synth.random_rays(30000, lo, hi, seed=seed)
        got, gi = _static(gpu, rays)
        want, wi = gpu.trace_closest(rays)
        _same(got, want)
        assert np.array_equal(gi, wi)
    a.close(); b.close()
