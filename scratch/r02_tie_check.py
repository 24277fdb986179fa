"""Fast layout (distance-sorted pushes) against the exact layout (the reference's visit order) on the C2 rays: every difference must be a tie (bit-equal t)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene
n = 1 << 24
for name, desc in (("soup", synth.soup_scene(1_000_000)), ("surface", synth.terrain_scene(708, 708))):
    a = GpuScene(desc, 4, layout=capi.LAYOUT_EXACT)
    b = GpuScene(desc, 4, layout=capi.LAYOUT_FAST)
    lo, hi = a.bounds()
    rays = synth.random_rays(n, lo, hi, seed=1)
    ha, ia = a.trace_closest(rays)
    hb, ib = b.trace_closest(rays)
    d = ha["prim"] != hb["prim"]
    td = ha["t"].view(np.uint32) != hb["t"].view(np.uint32)
    print(name, "rays", n, "prim differs", int(d.sum()), "t differs", int(td.sum()), "prim differs with equal t", int((d & ~td).sum()), flush=True)
    if td.any():
        k = np.nonzero(td)[0]
        ulp = np.abs(ha["t"][k].view(np.int32).astype(np.int64) - hb["t"][k].view(np.int32).astype(np.int64))
        print("  t differs by (ulp):", ulp.tolist(), "t", ha["t"][k].tolist(), hb["t"][k].tolist(), "prims", ha["prim"][k].tolist(), hb["prim"][k].tolist())
        # which one does the reference-order kernel on the OTHER width give?
    if d.any():
        k = np.nonzero(d)[0][:5]
        print("  e.g.", [(int(i), float(ha["t"][i]), int(ha["prim"][i]), int(hb["prim"][i])) for i in k])
    a.close(); b.close()
