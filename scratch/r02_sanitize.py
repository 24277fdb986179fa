"""A small tour of every kernel family for compute-sanitizer (one tool per run: memcheck, then racecheck):
ray casts in every width and layout (host and device buffers, the static entry), GPU builder, the three integrators incl. a medium
boundary scene (k_shadow_tr), the shade-order sort (VolPath), adaptive sampling, virtual multi-GPU ranks, the unit probes.
Sizes are tiny: the tools slow kernels down 10-100x."""
import ctypes as C
import sys
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene, MultiGpu, render_params

def log(*a):
    print(*a, flush=True)

d = synth.instanced_scene(2, 8)
for width, layout in ((2, 0), (4, 0), (8, 0), (4, 1)):
    g = GpuScene(d, width, layout=layout)
    lo, hi = g.bounds()
    rays = synth.random_rays(6000, lo, hi, seed=width)
    h, i = g.trace_closest(rays)
    occ = g.trace_any(rays)
    hit = h["prim"] != 0xFFFFFFFF
    assert (occ.astype(bool) != hit).mean() < 1e-2
    if width == 4 and layout == 0:
        hs = np.empty(len(rays), capi.HIT_DTYPE); ins = np.empty(len(rays), np.uint32)
        assert g.lib.pt_trace_closest_static(g.h, rays.ctypes.data, len(rays), hs.ctypes.data, ins.ctypes.data) == 0
        assert np.array_equal(hs.view(np.uint8), h.view(np.uint8)) and np.array_equal(ins, i)
    log("ray cast", width, layout, "hit rate", float(hit.mean()), "overflows", g.stack_overflows())
    g.close()

for name, (desc, cam) in (("shapes", synth.shapes_scene()), ("wrapped", synth.wrapped_scene())):
    g = GpuScene(desc, 4)
    lo, hi = g.bounds()
    rays = synth.with_times(synth.random_rays(4000, lo, hi, seed=2), seed=1)
    g.trace_closest(rays); g.trace_any(rays)
    for integ in (0, 1, 2):
        st = g.render(cam, render_params(24, 16, 3, 5, seed=1, integrator=integ))
        assert st["samples"] == 24 * 16 * 3
    log(name, "ok")
    g.close()

desc, cam = synth.cornell_scene()
g = GpuScene(desc, 4)
st = g.render(cam, render_params(32, 24, 4, 6, seed=2, adaptive=1))
log("adaptive samples", st["samples"])
m = MultiGpu(g, devices=[0, 0, 0])
st = m.render(cam, render_params(40, 40, 2, 4, seed=3))          # 2 passes over 3 ranks: tile sharding
st = m.render(cam, render_params(32, 24, 7, 4, seed=3))          # pass sharding
m.close(); g.close()

desc, cam = synth.materials_scene(subdiv=1, sky=True, point_light=True, distant_light=True)
g = GpuScene(desc, 4)
for integ in (0, 1, 2):
    g.render(cam, render_params(24, 16, 2, 6, seed=4, integrator=integ))
n = 300
rays = synth.random_rays(n, *g.bounds(), seed=5)
inp = np.random.default_rng(1).random((n, 6), dtype=np.float32)
out = np.zeros((n, 48), np.float32)
assert g.lib.pt_unit_shade(g.h, rays.ctypes.data, n, inp.ctypes.data, out.ctypes.data, 48) == 0
g.close()
log("sanitize tour ok")
