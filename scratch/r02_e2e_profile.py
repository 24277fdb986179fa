"""Where the host-buffer ray cast (pt_trace_closest, streamed form) spends its time on the C2 workload:
PTB_PROFILE prints, per call, when the last H2D chunk landed, the kernel's duration and when the last D2H ended."""
import os, sys, time
os.environ["PTB_PROFILE"] = "1"
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene

n = 1 << 24
for layout in (capi.LAYOUT_FAST, capi.LAYOUT_EXACT):
    desc = synth.soup_scene(1_000_000)
    gpu = GpuScene(desc, 4, layout=layout)
    lo, hi = gpu.bounds()
    rays = synth.random_rays(n, lo, hi, seed=1)
    host_rays = torch.from_numpy(rays.view(np.uint8).reshape(-1)).pin_memory()
    host_hits = torch.empty(n * 16, dtype=torch.uint8).pin_memory()
    d_rays = host_rays.cuda()
    d_hits = torch.empty(n * 16, dtype=torch.uint8, device="cuda")
    print("layout", layout, flush=True)
    for it in range(5):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rc = gpu.lib.pt_trace_closest(gpu.h, host_rays.data_ptr(), n, host_hits.data_ptr(), None)
        assert rc == 0
        print("  host buffers %.3f ms" % ((time.perf_counter() - t0) * 1e3), file=sys.stderr, flush=True)
    for it in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        gpu.trace_closest_dev(d_rays.data_ptr(), n, d_hits.data_ptr())
        torch.cuda.synchronize()
        print("  device buffers %.3f ms" % ((time.perf_counter() - t0) * 1e3), file=sys.stderr, flush=True)
    gpu.close()

# compact transfer form on the same workload
for layout in (capi.LAYOUT_FAST,):
    gpu = GpuScene(synth.soup_scene(1_000_000), 4, layout=layout)
    lo, hi = gpu.bounds()
    rays = synth.random_rays(n, lo, hi, seed=1)
    r24 = np.zeros(n, capi.RAY24_DTYPE); r24["o"], r24["d"] = rays["o"], rays["d"]
    h24 = torch.from_numpy(r24.view(np.uint8).reshape(-1)).pin_memory()
    hh8 = torch.empty(n * 8, dtype=torch.uint8).pin_memory()
    for it in range(5):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rc = gpu.lib.pt_trace_closest_compact(gpu.h, h24.data_ptr(), n, hh8.data_ptr(), None)
        assert rc == 0
        print("  compact host buffers %.3f ms" % ((time.perf_counter() - t0) * 1e3), file=sys.stderr, flush=True)
    gpu.close()
