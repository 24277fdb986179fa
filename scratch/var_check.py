import sys, json
import numpy as np
sys.path.insert(0, ".")
from pathtracing_b200 import synth, capi
from pathtracing_b200.scene import GpuScene, render_params
from oracle.refapi import RefScene, available
fl = "fast" if available("fast") else "strict"
for integ, name in ((capi.INTEGRATOR_SIMPLE, "simple"), (capi.INTEGRATOR_PATH, "path")):
    desc, cam = synth.cornell_scene(light_scale=40.0, light_size=0.12)
    spp = 2048
    p = render_params(48, 48, spp, 5, integrator=integ, seed=3)
    r = RefScene(desc, 4, flavour=fl, count_rays=True); rr, rs, rst = r.render(cam, p); r.close()
    g = GpuScene(desc, 4); gst = g.render(cam, p); gr, gs = g.film(), g.film_sumsq(); g.close()
    rm, gm = rr[..., :3] / rr[..., 3:4], gr[..., :3] / gr[..., 3:4]
    print(name, "mean ratio", (gm.mean((0, 1)) / rm.mean((0, 1))).round(4), "sumsq ratio", (gs.sum((0, 1)) / rs.sum((0, 1))).round(4),
          "rays", gst["closest_rays"] / rst["closest_rays"])
    # per-pixel second moment ratio, block averaged
    blk = lambda a: a.reshape(6, 8, 6, 8, 3).mean(axis=(1, 3))
    print(" block sumsq ratio quantiles", np.quantile(blk(gs) / np.maximum(blk(rs), 1e-12), [0.05, 0.25, 0.5, 0.75, 0.95]).round(3))
