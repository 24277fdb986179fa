import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from pathtracing_b200 import synth, capi
import test_render_gpu as T
for name, kw in (("full", {}), ("no_env", dict(env=False)), ("env_only", dict(n_lights=0, n_objects=0)), ("no_lights", dict(n_lights=0)),
                 ("one_light_no_env", dict(env=False, n_lights=1)), ("no_objects", dict(n_objects=0))):
    desc, cam = synth.c5_scene(grid=48, n_textures=4, tex_size=16, **kw)
    try:
        out = T.compare(desc, cam, 48, 27, 1024, 8)
        print(name, "PASS", {k: out[k] for k in ("z_frac", "z_bias", "z_std", "rel_mse")})
    except AssertionError as e:
        msg = str(e)
        print(name, "FAIL", msg[:400].replace("\n", " "))
