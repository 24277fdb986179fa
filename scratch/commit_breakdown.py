import sys, time, numpy as np
sys.path.insert(0, '.')
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene
descs = {"soup1M": synth.soup_scene(1000000), "c3": synth.c3_scene()}
if len(sys.argv) > 1: descs["c5"] = synth.c5_scene()
for name, desc in descs.items():
    for w in (4,):
        for b in (capi.BUILDER_HOST, capi.BUILDER_GPU, capi.BUILDER_GPU):
            t0 = time.perf_counter(); s = GpuScene(desc, w, builder=b); t1 = time.perf_counter()
            st = s.build_stats(); s.close()
            print(name, "bvh", w, "builder", b, f"ctor {t1-t0:.3f}s", {k: (round(v, 4) if isinstance(v, float) else v) for k, v in st.items()}, flush=True)
