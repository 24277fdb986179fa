"""Scene commit times: host vs GPU builder, C2 soup 1M and C5 10M."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene
def run(name, desc):
    res = {}
    lo = hi = rays = None
    for bname, b in (("host", capi.BUILDER_HOST), ("gpu", capi.BUILDER_GPU)):
        t0 = time.perf_counter()
        g = GpuScene(desc, 4, builder=b)
        wall = time.perf_counter() - t0
        st = g.build_stats()
        if rays is None:
            lo, hi = g.bounds()
            rays = synth.random_rays(1 << 20, lo, hi, seed=3)
        h, i = g.trace_closest(rays)
        res[bname] = (h, i)
        print(json.dumps({"scene": name, "builder": bname, "ctor_wall_s": round(wall, 3), **{k: (round(v, 4) if isinstance(v, float) else v) for k, v in st.items()}}), flush=True)
        g.close()
    same = np.array_equal(res["host"][0].view(np.uint8), res["gpu"][0].view(np.uint8)) and np.array_equal(res["host"][1], res["gpu"][1])
    print(json.dumps({"scene": name, "gpu_builder_hits_identical": bool(same)}), flush=True)
run("soup 1M", synth.soup_scene(1_000_000))
run("c3", synth.c3_scene()[0])
run("c5", synth.c5_scene()[0])
