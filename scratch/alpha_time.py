import sys
sys.path.insert(0, ".")
from pathtracing_b200 import synth, capi
from pathtracing_b200.scene import GpuScene, render_params
for mode in (capi.ALPHA_MASK, None):
    desc, cam = synth.alpha_scene(mode) if mode is not None else synth.cornell_scene()
    g = GpuScene(desc, 4)
    for k in range(3):
        st = g.render(cam, render_params(1024, 1024, 16, 8, seed=1))
        print(mode, k, st["gpu_ms"], st["kernel_launches"])
    g.close()
