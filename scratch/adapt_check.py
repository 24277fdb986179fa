import sys
import numpy as np
sys.path.insert(0, ".")
from pathtracing_b200 import synth, capi
from pathtracing_b200.scene import GpuScene, render_params
from oracle.refapi import RefScene, available
fl = "fast" if available("fast") else "strict"
desc, cam = synth.cornell_scene(light_scale=40.0, light_size=0.12)
for integ, spp in ((capi.INTEGRATOR_SIMPLE, 4), (capi.INTEGRATOR_PATH, 2)):
    for seed in (1, 2):
        p = render_params(48, 48, spp, 5, integrator=integ, seed=seed, adaptive=1)
        r = RefScene(desc, 4, flavour=fl); rr, rs, rst = r.render(cam, p); r.close()
        g = GpuScene(desc, 4); gst = g.render(cam, p); gr = g.film(); g.close()
        rn, gn = rr[..., 3] / spp, gr[..., 3] / spp
        print(integ, seed, "samples", gst["samples"], rst["samples"])
        print("  ref hist", np.bincount(np.minimum(rn.astype(int), 12).ravel(), minlength=13))
        print("  gpu hist", np.bincount(np.minimum(gn.astype(int), 12).ravel(), minlength=13))
        print("  ref max", rn.max(), "gpu max", gn.max(), " >=20:", (rn >= 20).sum(), (gn >= 20).sum())
