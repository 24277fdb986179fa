"""Host-buffer ray cast (streamed form) on the C2 workload under the PTB_STREAM_* knobs of csrc/capi.cu: median ms of the full and the compact form."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene

n = 1 << 24
gpu = GpuScene(synth.soup_scene(1_000_000), 4, layout=capi.LAYOUT_FAST)
lo, hi = gpu.bounds()
rays = synth.random_rays(n, lo, hi, seed=1)
host_rays = torch.from_numpy(rays.view(np.uint8).reshape(-1)).pin_memory()
host_hits = torch.empty(n * 16, dtype=torch.uint8).pin_memory()
r24 = np.zeros(n, capi.RAY24_DTYPE); r24["o"], r24["d"] = rays["o"], rays["d"]
h24 = torch.from_numpy(r24.view(np.uint8).reshape(-1)).pin_memory()
hh8 = torch.empty(n * 8, dtype=torch.uint8).pin_memory()
def run(f):
    ts = []
    for it in range(9):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        assert f() == 0
        ts.append((time.perf_counter() - t0) * 1e3)
    return sorted(ts[2:])[len(ts[2:]) // 2]
full = run(lambda: gpu.lib.pt_trace_closest(gpu.h, host_rays.data_ptr(), n, host_hits.data_ptr(), None))
comp = run(lambda: gpu.lib.pt_trace_closest_compact(gpu.h, h24.data_ptr(), n, hh8.data_ptr(), None))
print("chunk %s taper %s: full %.3f ms (%.0f Mrays/s)  compact %.3f ms (%.0f Mrays/s)  hits checksum %d" % (
    os.environ.get("PTB_STREAM_CHUNK_LOG2", "-"), os.environ.get("PTB_STREAM_TAPER_LOG2", "-"), full, n / full / 1e3, comp, n / comp / 1e3,
    int(host_hits.view(torch.int32).to(torch.int64).sum())), flush=True)
gpu.close()
