import sys
sys.path.insert(0, ".")
from pathtracing_b200 import synth, capi
from pathtracing_b200.scene import GpuScene, render_params
which = sys.argv[1] if len(sys.argv) > 1 else "cornell"
desc, cam = synth.cornell_scene() if which == "cornell" else synth.materials_scene(subdiv=5)
g = GpuScene(desc, 4)
st = g.render(cam, render_params(1024, 1024, 4, 8, seed=1))
print(st)
g.close()
