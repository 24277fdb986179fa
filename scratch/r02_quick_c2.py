"""C2 device-resident timing only (closest + any, fast and exact), for kernel experiments."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene
n = 1 << 24
layouts = [capi.LAYOUT_FAST] if "fast" in sys.argv else [capi.LAYOUT_FAST, capi.LAYOUT_EXACT]
for layout in layouts:
    gpu = GpuScene(synth.soup_scene(1_000_000), 4, layout=layout)
    lo, hi = gpu.bounds()
    rays = synth.random_rays(n, lo, hi, seed=1)
    d_rays = torch.from_numpy(rays.view(np.uint8).reshape(-1)).cuda()
    d_hits = torch.empty(n * 16, dtype=torch.uint8, device="cuda")
    d_occ = torch.empty(n, dtype=torch.uint8, device="cuda")
    for name, fn in (("closest", lambda: gpu.trace_closest_dev(d_rays.data_ptr(), n, d_hits.data_ptr())), ("any", lambda: gpu.trace_any_dev(d_rays.data_ptr(), n, d_occ.data_ptr()))):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ts = []
        for _ in range(5):
            torch.cuda.synchronize(); e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = float(np.median(ts))
        print(f"layout {layout} {name}: {ms:.3f} ms  {n / ms / 1e3:.1f} Mrays/s  checksum {int(d_hits.view(torch.int32).sum().item()) if name == 'closest' else int(d_occ.sum().item())}", flush=True)
    gpu.close()
