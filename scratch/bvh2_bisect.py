"""BVH2 / BVH4 closest-hit rate of an older libptb200.so (PTB_LIB=path) on the C2 soup; missing newer symbols are tolerated."""
import os, sys, ctypes as C, numpy as np
sys.path.insert(0, '.')
import torch
from pathtracing_b200 import capi
path = os.environ.get("PTB_LIB")
if path:
    real = C.CDLL
    class Dummy:
        restype = None; argtypes = None
        def __call__(self, *a): return 0
    class Tolerant:
        def __init__(self, lib): object.__setattr__(self, "_l", lib)
        def __getattr__(self, n):
            try: return getattr(self._l, n)
            except AttributeError: return Dummy()
    capi.C.CDLL = lambda p, *a, **k: Tolerant(real(path, *a, **k))
from pathtracing_b200 import synth
from pathtracing_b200.scene import GpuScene
desc = synth.soup_scene(1000000)
n = 1 << 24
for width in tuple(int(x) for x in os.environ.get("PTB_WIDTHS", "2,4").split(",")):
    scene = GpuScene(desc, width)
    lo, hi = scene.bounds()
    rays = torch.from_numpy(synth.random_rays(n, lo, hi, seed=7).view(np.uint8).reshape(-1)).cuda()
    hits = torch.empty(n * 16, dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    f = lambda: scene.lib.pt_trace_closest_dev(scene.h, rays.data_ptr(), n, hits.data_ptr(), None, st)
    for _ in range(3): f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5): f()
    b.record(); torch.cuda.synchronize()
    print(os.path.basename(path or "HEAD"), f"BVH{width} {n * 5 / a.elapsed_time(b) / 1e3:.1f} Mrays/s", flush=True)
    scene.close()
