"""Null calibration of the per-pixel z-test: ref vs ref, gpu vs gpu, gpu vs ref on one scene."""
import sys, json
import numpy as np
sys.path.insert(0, ".")
from pathtracing_b200 import synth, capi
from pathtracing_b200.scene import GpuScene, render_params
from oracle.refapi import RefScene, available

def mean(rgbw): return rgbw[..., :3] / np.maximum(rgbw[..., 3:4], 1e-300)

def z(a, asq, b, bsq, spp):
    var = (np.maximum(asq / spp - a ** 2, 0) + np.maximum(bsq / spp - b ** 2, 0)) * spp / (spp - 1)
    s = np.sqrt(var / spp); ok = s > 1e-12
    zz = (a - b)[ok] / s[ok]
    return {"z_frac": float((np.abs(zz) > 3).mean()), "z_bias": float(zz.mean()), "z_std": float(zz.std()),
            "dmean": ((a.mean((0, 1)) - b.mean((0, 1))) / b.mean((0, 1))).round(5).tolist()}

which = sys.argv[1] if len(sys.argv) > 1 else "materials"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
if which == "materials": (desc, cam), W, H, depth, integ = synth.materials_scene(), 64, 48, 12, capi.INTEGRATOR_PATH
elif which == "medium": (desc, cam), W, H, depth, integ = synth.medium_scene(), 48, 36, 12, capi.INTEGRATOR_VOLPATH
else: (desc, cam), W, H, depth, integ = synth.cornell_scene(), 64, 64, 8, capi.INTEGRATOR_PATH
fl = "fast" if available("fast") else "strict"
res = []
for seed in (1, 2):
    p = render_params(W, H, spp, depth, integrator=integ, seed=seed)
    r = RefScene(desc, 4, flavour=fl); rr, rs, _ = r.render(cam, p); r.close()
    g = GpuScene(desc, 4); g.render(cam, p); gr, gs = g.film(), g.film_sumsq(); g.close()
    res.append((mean(rr), rs, mean(gr), gs))
(r1, r1s, g1, g1s), (r2, r2s, g2, g2s) = res
print(json.dumps({"ref_ref": z(r1, r1s, r2, r2s, spp), "gpu_gpu": z(g1, g1s, g2, g2s, spp),
                  "gpu1_ref1": z(g1, g1s, r1, r1s, spp), "gpu2_ref2": z(g2, g2s, r2, r2s, spp),
                  "gpu1_ref2": z(g1, g1s, r2, r2s, spp), "gpu2_ref1": z(g2, g2s, r1, r1s, spp)}))
np.savez_compressed(f"gpurun_out/calib_{which}_{spp}.npz", r1=r1, r2=r2, g1=g1, g2=g2, r1s=r1s, r2s=r2s, g1s=g1s, g2s=g2s)
