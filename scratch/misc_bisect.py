import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from pathtracing_b200 import synth
import test_render_gpu as T
for name, kw in (("full", {}), ("no_uv", dict(uv_tex=False)), ("no_normal", dict(normal_tex=False)), ("no_thin", dict(thin=False)),
                 ("no_lens", dict(lens=False)), ("gradient_sky", dict(uniform_sky=False)), ("no_metal", dict(metal_half=False)),
                 ("plain", dict(uv_tex=False, normal_tex=False, thin=False, lens=False, metal_half=False))):
    desc, cam = synth.misc_scene(**kw)
    try:
        out = T.compare(desc, cam, 64, 48, 512, 8)
        print(name, "PASS", {k: round(out[k], 4) for k in ("z_frac", "z_bias", "z_std", "rel_mse")})
    except AssertionError as e:
        print(name, "FAIL", str(e)[:330].replace("\n", " "))
