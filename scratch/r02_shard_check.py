"""What one of 8 ranks renders of a C3 / C4 frame (spp / 8 passes), for different slot-pool sizes (PTB_BATCH): is the per-rank
inefficiency at N = 8 the missing slot regeneration of a share that fits the pool?"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene, render_params
which = sys.argv[1]
name, builder, spp, integ = {"c3": ("c3", synth.c3_scene, 64, capi.INTEGRATOR_PATH), "c4": ("c4", synth.c4_scene, 128, capi.INTEGRATOR_VOLPATH)}[which]
desc, cam = builder()
gpu = GpuScene(desc, 4, layout=capi.LAYOUT_FAST)
for share in (8, 1):
    p = render_params(1920, 1080, spp // share, 8, integrator=integ, seed=1)
    gpu.render(cam, p)
    best = 1e9
    for _ in range(3):
        st = gpu.render(cam, p)
        best = min(best, st["gpu_ms"])
    print(json.dumps({"cfg": name, "batch": os.environ.get("PTB_BATCH", "default"), "spp": spp // share, "gpu_ms": best, "x_share": best * share, "launches": st["kernel_launches"]}), flush=True)
    if share == 8 and os.environ.get("PTB_BATCH"):
        break
gpu.close()
