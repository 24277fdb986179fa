"""Quick GPU check of the device-driven wavefront loop: frame times of C1 / C3 / C4 (+ virtual ranks)."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene, MultiGpu, render_params

def run(name, builder, W, H, spp, depth, integ, reps=3):
    desc, cam = builder()
    t0 = time.perf_counter()
    gpu = GpuScene(desc, 4)
    build = time.perf_counter() - t0
    p = render_params(W, H, spp, depth, integrator=integ, seed=1)
    gpu.render(cam, p)
    best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        st = gpu.render(cam, p)
        wall = (time.perf_counter() - t0) * 1e3
        best = min(best, st["gpu_ms"])
    print(json.dumps({"cfg": name, "build_s": build, "gpu_ms": best, "wall_ms_last": wall, "msamples_s": W * H * spp / best / 1e3,
                      "rays_per_sample": (st["closest_rays"] + st["shadow_rays"]) / st["samples"], "launches": st["kernel_launches"]}), flush=True)
    gpu.close()

which = sys.argv[1:] or ["c1", "c3", "c4"]
if "c1" in which: run("c1", synth.cornell_scene, 512, 512, 16, 8, capi.INTEGRATOR_PATH, reps=10)
if "c1s" in which: run("c1 2spp (one of 8 shards)", synth.cornell_scene, 512, 512, 2, 8, capi.INTEGRATOR_PATH, reps=10)
if "c3" in which: run("c3", synth.c3_scene, 1920, 1080, 64, 8, capi.INTEGRATOR_PATH)
if "c4" in which: run("c4", synth.c4_scene, 1920, 1080, 128, 8, capi.INTEGRATOR_VOLPATH)
if "c5" in which: run("c5 64spp", synth.c5_scene, 1920, 1080, 64, 128, capi.INTEGRATOR_PATH, reps=1)
