import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene, render_params
from test_bvh_build import _boxes
mode = sys.argv[1]
if mode == "build":
    for kind, n in (("soup", 20000), ("lattice", 9000), ("line", 7000), ("soup", 1500)):
        b = _boxes(kind, n, seed=3)
        hn, ho, _ = capi.build_bvh2(b, capi.BUILDER_HOST)
        gn, go, _ = capi.build_bvh2(b, capi.BUILDER_GPU)
        print(kind, n, "identical", np.array_equal(ho, go) and all(np.array_equal(hn[f], gn[f]) for f in ("left", "right", "count", "axis")), flush=True)
else:
    desc, cam = synth.cornell_scene()
    s = GpuScene(desc, 4)
    st = s.render(cam, render_params(96, 96, 4, 6, seed=3, adaptive=1))
    print("render samples", st["samples"], flush=True)
    d2 = synth.instanced_scene(3, 8)
    s2 = GpuScene(d2, 4, builder=capi.BUILDER_GPU)
    lo, hi = s2.bounds()
    rays = synth.random_rays(50000, lo, hi, seed=2)
    h, i = s2.trace_closest(rays)
    print("hits", int((h["prim"] != 0xFFFFFFFF).sum()), flush=True)
