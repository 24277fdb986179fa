import sys, time, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from pathtracing_b200 import capi, synth
from test_bvh_build import _boxes
for kind, n in (("soup", 1000000), ("terrain", 1000000), ("soup", 10000000)):
    boxes = _boxes(kind, n, seed=3)
    capi.build_bvh2(boxes[:5000], capi.BUILDER_GPU)
    t0 = time.perf_counter(); hn, ho, _ = capi.build_bvh2(boxes, capi.BUILDER_HOST); t1 = time.perf_counter()
    gn, go, ms = capi.build_bvh2(boxes, capi.BUILDER_GPU); t2 = time.perf_counter()
    gn, go, ms = capi.build_bvh2(boxes, capi.BUILDER_GPU); t3 = time.perf_counter()
    same = len(hn) == len(gn) and np.array_equal(ho, go) and all(np.array_equal(hn[f], gn[f]) for f in ("left", "right", "count", "axis", "lo", "hi"))
    print(f"{kind} n={len(boxes)} nodes={len(hn)} host {t1-t0:.3f}s gpu call {t3-t2:.3f}s (first {t2-t1:.3f}s) kernels {ms:.2f} ms identical={same}", flush=True)
