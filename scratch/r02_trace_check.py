"""C2-style ray cast, device resident: exact vs fast layout, closest + any, with visit counts."""
import os, sys, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene

def timeit(fn, reps=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

def run(name, desc, n=1 << 24, widths=((4, 0), (4, 1), (2, 0), (8, 0))):
    rays = None
    for width, layout in widths:
        t0 = time.perf_counter()
        gpu = GpuScene(desc, width, layout=layout)
        build = time.perf_counter() - t0
        if rays is None:
            lo, hi = gpu.bounds()
            rays = torch.from_numpy(synth.random_rays(n, lo, hi, seed=7).view(np.uint8)).cuda()
            hits = torch.empty(n * 16, dtype=torch.uint8, device="cuda")
            occ = torch.empty(n, dtype=torch.uint8, device="cuda")
        st = torch.cuda.current_stream().cuda_stream
        ms_c = timeit(lambda: gpu.trace_closest_dev(rays.data_ptr(), n, hits.data_ptr(), None, st))
        ms_a = timeit(lambda: gpu.trace_any_dev(rays.data_ptr(), n, occ.data_ptr(), st))
        stt = gpu.trace_stats(rays.data_ptr(), n)
        print(json.dumps({"scene": name, "width": width, "layout": "fast" if layout else "exact", "build_s": round(build, 2),
                          "closest_mrays": round(n / ms_c / 1e3, 1), "any_mrays": round(n / ms_a / 1e3, 1), "ms_closest": round(ms_c, 3),
                          "nodes_per_ray": round((stt["tlas_nodes"] + stt["blas_nodes"]) / n, 2), "tris_per_ray": round(stt["triangles"] / n, 2),
                          "hit_rate": round(stt["hits"] / n, 4), "overflows": gpu.stack_overflows()}), flush=True)
        gpu.close()

which = sys.argv[1:] or ["soup", "terrain"]
if "soup" in which: run("soup 1M", synth.soup_scene(1_000_000))
if "soup4" in which: run("soup 1M", synth.soup_scene(1_000_000), widths=((4, 0), (4, 1)))
if "terrain" in which: run("terrain 1M", synth.terrain_scene(1001, 501), widths=((4, 0), (4, 1)))
if "c5" in which:
    desc, cam = synth.c5_scene()
    run("c5 geometry 10M", desc, widths=((4, 0), (4, 1)))
