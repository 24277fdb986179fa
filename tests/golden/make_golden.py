"""Regenerates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref, built from /root/reference by
oracle/Makefile).  Run where the reference tree is mounted:

    python tests/golden/make_golden.py

Ray casts: per (scene, BVH width) the rays and the reference's answers (hit distance bits, primitive id, instance id,
the quad's (alpha, beta) bits, any-hit flag at a finite tmax).  Render: the reference's converged Cornell image
(per-pixel mean and sum of squares).  Every file carries a digest of the scene arrays it was made from, so that a
box whose numpy builds a different scene says so instead of failing the comparison."""
import hashlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
from pathtracing_b200 import capi, synth  # noqa: E402
from pathtracing_b200.scene import render_params  # noqa: E402

OUT = Path(__file__).resolve().parent
N_RAYS = 6000


def scene_digest(desc) -> str:
    """sha256 over everything that shapes the scene: mesh arrays, instance transforms, shape descriptors, wrappers"""
    import ctypes as C
    h = hashlib.sha256()
    for a in desc._keep:
        h.update(np.ascontiguousarray(a).tobytes())
    for k, (b, x) in enumerate(desc.instances):
        h.update(np.int32(b).tobytes())
        if b >= 0:
            h.update(b"-" if x is None else np.ascontiguousarray(x, np.float32).tobytes())
        else:
            h.update(bytes(C.string_at(C.addressof(x), C.sizeof(x))))
        if k in desc.shape_xf:
            h.update(np.ascontiguousarray(desc.shape_xf[k], np.float32).tobytes())
        if k in desc.animated:
            h.update(np.asarray(desc.animated[k][0] + desc.animated[k][1], np.float32).tobytes())
    return h.hexdigest()


def raycast_cases():
    """name -> (scene description, ray box or None for the scene bounds, rays carry a time)"""
    yield "soup", synth.soup_scene(100000, opaque=True), None, False
    yield "instanced", synth.instanced_scene(2, 8), None, False
    yield "shapes", synth.shapes_scene()[0], ((-3, -0.3, -2.5), (3, 2.2, 1.5)), False
    yield "wrapped", synth.wrapped_scene()[0], ((-3, -0.3, -2.5), (3, 2.2, 1.5)), True


def render_case():
    desc, cam = synth.cornell_scene()
    return desc, cam, render_params(24, 24, 4096, 6, seed=11)


def c1_case():
    """BASELINE.json configs[0] as it stands: Cornell box 512x512, depth 8, NEE + MIS (PathIntegrator, PowerLightSampler);
    the test renders it at the config's 16 spp, the stored reference film is converged further (C1_REF_SPP)."""
    desc, cam = synth.cornell_scene()
    return desc, cam, render_params(512, 512, 16, 8, seed=21)


C1_REF_SPP = 256


def make_c1():
    """The reference's film of C1 at 256 spp, stored as per-pixel mean and per-sample variance in float16 (3 MB: the rounding,
    5e-4 relative, is far below the Monte-Carlo noise of the 16-spp frame it is compared with)."""
    from oracle.refapi import RefScene
    desc, cam, p = c1_case()
    q = render_params(p.width, p.height, C1_REF_SPP, p.max_depth, seed=5)
    ref = RefScene(desc, 4, flavour="fast" if __import__("oracle.refapi").refapi.available("fast") else "strict")
    rgbw, sq, stats = ref.render(cam, q)
    ref.close()
    mean = rgbw[..., :3] / rgbw[..., 3:4]
    var = np.maximum(sq / C1_REF_SPP - mean ** 2, 0.0) * C1_REF_SPP / (C1_REF_SPP - 1)
    np.savez_compressed(OUT / "render_c1_512_depth8.npz", mean=mean.astype(np.float16), var=var.astype(np.float16), spp=C1_REF_SPP,
                        digest=scene_digest(desc))
    print("c1", stats, "mean radiance", float(mean.mean()))


def main():
    from oracle.refapi import RefScene
    for name, desc, box, timed in raycast_cases():
        for width in (2, 4, 8):
            ref = RefScene(desc, width, id_mode=True)
            lo, hi = ref.bounds()
            b = box if box is not None else (lo, hi)
            rays = synth.random_rays(N_RAYS, b[0], b[1], seed=17 + width)
            if timed:
                rays = synth.with_times(rays, seed=width)
            h, inst = ref.trace_closest(rays, want_inst=True)
            short = rays.copy()
            short["tmax"] = np.float32(0.6 * float(np.median(h["t"][np.isfinite(h["t"])])))
            occ = ref.trace_any(short)
            ref.close()
            np.savez_compressed(OUT / f"raycast_{name}_bvh{width}.npz", rays=rays, t_bits=h["t"].view(np.uint32), prim=h["prim"], inst=inst,
                                u_bits=np.nan_to_num(h["u"]).view(np.uint32), v_bits=np.nan_to_num(h["v"]).view(np.uint32),
                                any_tmax=short["tmax"][0], occluded=occ, bounds=np.stack([lo, hi]), digest=scene_digest(desc))
            print(name, width, "hit rate", float((h["prim"] != 0xFFFFFFFF).mean()), "occluded", float(occ.mean()))
    desc, cam, p = render_case()
    ref = RefScene(desc, 4, flavour="fast" if __import__("oracle.refapi").refapi.available("fast") else "strict")
    rgbw, sq, stats = ref.render(cam, p)
    ref.close()
    np.savez_compressed(OUT / "render_cornell_path.npz", rgbw=rgbw, sumsq=sq, spp=p.spp, digest=scene_digest(desc))
    print("render", stats)
    make_c1()


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "c1":
        make_c1()           # only the C1 film (the other fixtures stay as they are)
    else:
        main()
