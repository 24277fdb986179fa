import sys
sys.path.insert(0, ".")
from pathtracing_b200 import synth, capi
from pathtracing_b200.scene import GpuScene, render_params
key = sys.argv[1] if len(sys.argv) > 1 else "c3"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
import bench
name, builder, W, H, _, depth, integ = bench.render_configs()[key]
desc, cam = builder()
g = GpuScene(desc, 4, layout=capi.LAYOUT_FAST if (len(sys.argv) > 3 and sys.argv[3] == "fast") else None)
st = g.render(cam, render_params(W, H, spp, depth, integrator=integ, seed=1))
print(key, st)
g.close()
