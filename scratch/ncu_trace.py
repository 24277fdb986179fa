"""One closest-hit launch series for ncu: soup 1M, BVH4, layout from argv[1] (exact|fast), rays from argv[2] (default 4 Mi)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene
layout = 1 if (len(sys.argv) > 1 and sys.argv[1] == "fast") else 0
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 22
mesh = sys.argv[3] if len(sys.argv) > 3 else "soup"
if mesh == "c5":
    desc, _ = synth.c5_scene()
else:
    desc = synth.soup_scene(1_000_000)
gpu = GpuScene(desc, 4, layout=layout)
lo, hi = gpu.bounds()
rays = torch.from_numpy(synth.random_rays(n, lo, hi, seed=7).view(np.uint8)).cuda()
hits = torch.empty(n * 16, dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(4):
    gpu.trace_closest_dev(rays.data_ptr(), n, hits.data_ptr(), None, st)
torch.cuda.synchronize()
print("ok", layout, n)
gpu.close()
