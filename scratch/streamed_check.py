"""host-buffer ray cast (streamed single kernel for >= 1 Mi rays) against the device-resident launch: identical results; rate"""
import os, sys, time, numpy as np
sys.path.insert(0, '.')
import torch
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene
desc = synth.soup_scene(1000000)
for width in (4, 2, 8):
    scene = GpuScene(desc, width, builder=capi.BUILDER_GPU)
    lo, hi = scene.bounds()
    n = (1 << 22) + 12345
    rays = synth.random_rays(n, lo, hi, seed=7)
    h, i = scene.trace_closest(rays)               # host buffers
    d = torch.from_numpy(rays.view(np.uint8).reshape(-1)).cuda()
    dh = torch.empty(n * 16, dtype=torch.uint8, device="cuda"); di = torch.empty(n * 4, dtype=torch.uint8, device="cuda")
    scene.trace_closest_dev(d.data_ptr(), n, dh.data_ptr(), di.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ref_h = dh.cpu().numpy().view(capi.HIT_DTYPE); ref_i = di.cpu().numpy().view(np.uint32)
    ok = np.array_equal(h.view(np.uint8), ref_h.view(np.uint8)) and np.array_equal(i, ref_i)
    rays["tmax"] = 0.05
    occ = scene.trace_any(rays)
    d2 = torch.from_numpy(rays.view(np.uint8).reshape(-1)).cuda(); do = torch.empty(n, dtype=torch.uint8, device="cuda")
    scene.trace_any_dev(d2.data_ptr(), n, do.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ok2 = np.array_equal(occ, do.cpu().numpy())
    print("width", width, "closest identical", ok, "any identical", ok2, "occluded", float(occ.mean()), flush=True)
    assert ok and ok2
    scene.close()
print("streamed check ok")
