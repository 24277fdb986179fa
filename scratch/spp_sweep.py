import sys, time
sys.path.insert(0, ".")
from pathtracing_b200.scene import GpuScene, render_params
import bench
key = sys.argv[1]
name, builder, W, H, _, depth, integ = bench.render_configs()[key]
desc, cam = builder()
g = GpuScene(desc, 4)
for spp in [int(x) for x in sys.argv[2:]]:
    for rep in range(2):
        st = g.render(cam, render_params(W, H, spp, depth, integrator=integ, seed=1))
    print(key, spp, "spp", round(st["samples"] / st["gpu_ms"] / 1e3, 1), "Ms/s", round(st["gpu_ms"], 1), "ms", st["kernel_launches"], "launches", flush=True)
g.close()
