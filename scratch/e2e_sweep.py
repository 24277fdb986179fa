import os, sys, time, numpy as np, ctypes as C
sys.path.insert(0, '.')
import torch
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene
desc = synth.soup_scene(1000000)
scene = GpuScene(desc, 4, builder=capi.BUILDER_GPU)
lo, hi = scene.bounds()
n = 1 << 24
rays = synth.random_rays(n, lo, hi, seed=7)
host_rays = torch.from_numpy(rays.view(np.uint8).reshape(-1)).pin_memory()
host_hits = torch.empty(n * 16, dtype=torch.uint8).pin_memory()
lib = scene.lib
for _ in range(2): lib.pt_trace_closest(scene.h, host_rays.data_ptr(), n, host_hits.data_ptr(), None)
t0 = time.perf_counter()
for _ in range(5): rc = lib.pt_trace_closest(scene.h, host_rays.data_ptr(), n, host_hits.data_ptr(), None)
ms = (time.perf_counter() - t0) * 1e3 / 5
print("chunk", os.environ.get("PTB_CHUNK"), f"e2e {n/ms/1e3:.1f} Mrays/s {ms:.2f} ms", flush=True)
