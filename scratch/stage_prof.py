import os, sys, time, numpy as np
sys.path.insert(0, '.')
os.environ["PTB_PROFILE"] = "1"
import bench
from pathtracing_b200 import capi, multi
from pathtracing_b200.scene import GpuScene, render_params
keys = sys.argv[1].split(",")
spp_override = int(sys.argv[2]) if len(sys.argv) > 2 else 0
for key in keys:
    name, builder, W, H, spp, depth, integ = bench.render_configs()[key]
    if spp_override: spp = spp_override
    t0 = time.perf_counter(); desc, cam = builder(); t1 = time.perf_counter()
    scene = GpuScene(desc, 4, builder=capi.BUILDER_GPU); t2 = time.perf_counter()
    print(key, f"synth {t1-t0:.2f}s commit {t2-t1:.2f}s", scene.build_stats(), flush=True)
    params = render_params(W, H, spp, depth, integrator=integ, seed=1)
    for _ in range(2):
        st, film = multi.render_sharded(scene, cam, params, None)
        sys.stderr.flush()
        print(key, "spp", spp, f"gpu_ms {st['gpu_ms']:.2f} Msamples/s {W*H*spp/st['gpu_ms']/1e3:.1f} rays/sample {(st['closest_rays']+st['shadow_rays'])/(W*H*spp):.2f} launches {st['kernel_launches']}", flush=True)
    scene.close()
