"""Parity against committed golden vectors (tests/golden/*.npz, written by tests/golden/make_golden.py from the
UNMODIFIED reference): the checks that still run where neither /root/reference nor oracle/_ref exists.

  CPU tier : the host emulation of the product's traversal (tests/emul) against the vectors; where oracle/_ref is
             present, the reference against the vectors too (the fixtures are what the reference says today).
  GPU tier : the CUDA kernels through the C-ABI against the vectors, bit for bit; the Cornell render against the
             reference's stored converged image, statistically."""
import sys
from pathlib import Path

import numpy as np
import pytest

from conftest import needs_ref
from pathtracing_b200 import synth

GOLDEN = Path(__file__).resolve().parent / "golden"
sys.path.insert(0, str(GOLDEN))
import make_golden  # noqa: E402

CASES = {name: (desc, timed) for name, desc, _, timed in make_golden.raycast_cases()}
PARAMS = [(n, w) for n in CASES for w in (2, 4, 8)]


def _load(name, width):
    g = np.load(GOLDEN / f"raycast_{name}_bvh{width}.npz")
    desc = CASES[name][0]
    if str(g["digest"]) != make_golden.scene_digest(desc):
        pytest.fail(f"this box's numpy builds a different '{name}' scene than the one the golden vectors were made from: "
                    "regenerate them with tests/golden/make_golden.py where the reference is mounted")
    return g, desc


def _check(g, desc, hits, inst, trace_any):
    assert np.array_equal(hits["t"].view(np.uint32), g["t_bits"])
    assert np.array_equal(hits["prim"], g["prim"])
    assert np.array_equal(inst, g["inst"])
    shapes = [k for k, (b, _) in enumerate(desc.instances) if b < 0]
    sh = np.isin(inst, shapes)
    assert np.array_equal(hits["u"][sh].view(np.uint32), g["u_bits"][sh]) and np.array_equal(hits["v"][sh].view(np.uint32), g["v_bits"][sh])
    short = g["rays"].copy()
    short["tmax"] = g["any_tmax"]
    assert np.array_equal(trace_any(short), g["occluded"])
    assert 0.01 < g["occluded"].mean() < 0.9 and (g["prim"] != 0xFFFFFFFF).mean() > 0.05


@pytest.mark.parametrize("name,width", PARAMS)
def test_emulation_matches_golden(name, width):
    from emulapi import EmulScene
    g, desc = _load(name, width)
    em = EmulScene(desc, width)
    lo, hi = em.bounds()
    assert np.array_equal(np.stack([lo, hi]), g["bounds"])
    h, i = em.trace_closest(g["rays"])
    _check(g, desc, h, i, em.trace_any)


@needs_ref
@pytest.mark.parametrize("name,width", PARAMS)
def test_reference_matches_golden(name, width):
    """the fixtures are what the reference computes today (the oracle is pinned to them and they to it)"""
    from oracle.refapi import RefScene
    g, desc = _load(name, width)
    ref = RefScene(desc, width, id_mode=True)
    h, i = ref.trace_closest(g["rays"], want_inst=True)
    h = h.copy()
    h["u"], h["v"] = np.nan_to_num(h["u"]), np.nan_to_num(h["v"])
    _check(g, desc, h, i, ref.trace_any)
    ref.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name,width", PARAMS)
def test_gpu_matches_golden(name, width):
    from pathtracing_b200.scene import GpuScene
    g, desc = _load(name, width)
    gpu = GpuScene(desc, width)
    lo, hi = gpu.bounds()
    assert np.array_equal(np.stack([lo, hi]), g["bounds"])
    h, i = gpu.trace_closest(g["rays"])
    _check(g, desc, h, i, gpu.trace_any)
    gpu.close()


@pytest.mark.gpu
def test_gpu_render_matches_golden_image():
    """PathIntegrator on the Cornell box against the reference's stored 4096-spp film: relMSE above the noise floor
    <= 1e-3, per-pixel z-scores N(0,1)-like, channel means within 1 % (the criteria of tests/test_render_gpu.py)"""
    from pathtracing_b200.scene import GpuScene
    g = np.load(GOLDEN / "render_cornell_path.npz")
    desc, cam, p = make_golden.render_case()
    if str(g["digest"]) != make_golden.scene_digest(desc):
        pytest.fail("golden render was made from a different Cornell scene: regenerate it (tests/golden/make_golden.py)")
    spp = int(g["spp"])
    assert spp == p.spp
    try:
        _render_vs_golden(g, desc, cam, p, spp)
    except AssertionError:
        p.seed += 1          # a chance failure does not reproduce with independent samples, a bias does
        _render_vs_golden(g, desc, cam, p, spp)


def _render_vs_golden(g, desc, cam, p, spp):
    from pathtracing_b200.scene import GpuScene
    gpu = GpuScene(desc, 4)
    st = gpu.render(cam, p)
    rgbw, sq = gpu.film(), gpu.film_sumsq()
    gpu.close()
    assert st["samples"] == p.width * p.height * spp
    a, b = rgbw[..., :3] / rgbw[..., 3:4], g["rgbw"][..., :3] / g["rgbw"][..., 3:4]
    var = (np.maximum(g["sumsq"] / spp - b ** 2, 0.0) + np.maximum(sq / spp - a ** 2, 0.0)) * spp / (spp - 1)
    sigma = np.sqrt(var / spp)
    ok = sigma > 1e-12
    z = (a - b)[ok] / sigma[ok]
    floor = float((sigma ** 2).sum() / (b ** 2).sum())
    mse = float(((a - b) ** 2).sum() / (b ** 2).sum())
    assert mse <= 1e-3 + 2.0 * floor, (mse, floor)
    assert (np.abs(z) > 3).mean() <= 0.01 and abs(z.mean()) <= 0.08 and z.std() <= 1.12, ((np.abs(z) > 3).mean(), z.mean(), z.std())
    assert np.allclose(a.mean(axis=(0, 1)), b.mean(axis=(0, 1)), rtol=0.01)


# ---- BASELINE.json configs[0] at its real size: Cornell 512 x 512, 16 spp, depth 8, NEE + MIS ------------------------------
# A 16-spp estimate of this integrand is heavy-tailed (rare bright paths), and a pixel's sample variance is correlated with
# its own mean error, so a z-score built from the frame's OWN 16-sample variance is neither normal nor centred: the reference
# rendering the config against its own stored 256-spp film gives 3.3 % of pixels beyond 3 sigma and a mean z of -0.33 that way.
# The statistics below were therefore calibrated on exactly that experiment (reference at 16 spp vs the stored film, three
# draws; numbers in brackets), and the thresholds leave the usual margin:
#   pixel level, variance = mean of the two films' per-sample variances:  beyond 3 sigma [0.21-0.23 %] <= 0.6 %, std(z) [0.966-0.968] <= 1.06
#   16 x 16 blocks (256 px x 16 spp: near normal), variance from the stored film:  mean z [-0.012..0.000] within 0.05, std(z) [1.02-1.07] <= 1.15,
#                                                                                 beyond 3 sigma [0.8-1.0 %] <= 2 %
#   relative MSE [0.0014-0.0015, all of it Monte-Carlo noise: floor 0.0015] <= 1e-3 + 2 x floor;  channel means within 1 %
def _blocks(x, k=16):
    h, w, c = x.shape
    return x.reshape(h // k, k, w // k, k, c).mean(axis=(1, 3))


def _c1_compare(a, va, na, b, vb, nb):
    """GPU film (mean a, per-sample variance va, na spp) against a reference film (b, vb, nb spp) of C1"""
    s2 = 0.5 * (va + vb) / na + vb / nb
    ok = s2 > 1e-24
    z = (a - b)[ok] / np.sqrt(s2[ok])
    floor = float((vb * (1.0 / na + 1.0 / nb)).sum() / (b ** 2).sum())
    mse = float(((a - b) ** 2).sum() / (b ** 2).sum())
    k = 16
    sb = _blocks(vb, k) / (k * k) * (1.0 / na + 1.0 / nb)
    zb = ((_blocks(a, k) - _blocks(b, k)) / np.sqrt(np.maximum(sb, 1e-30)))[sb > 1e-24]
    out = {"rel_mse": mse, "noise_floor": floor, "pixel_z_frac": float((np.abs(z) > 3).mean()), "pixel_z_std": float(z.std()),
           "block_z_bias": float(zb.mean()), "block_z_std": float(zb.std()), "block_z_frac": float((np.abs(zb) > 3).mean()),
           "gpu_mean": a.mean(axis=(0, 1)), "ref_mean": b.mean(axis=(0, 1))}
    assert mse <= 1e-3 + 2.0 * floor, out
    assert out["pixel_z_frac"] <= 0.006 and out["pixel_z_std"] <= 1.06, out
    assert abs(out["block_z_bias"]) <= 0.05 and out["block_z_std"] <= 1.15 and out["block_z_frac"] <= 0.02, out
    assert np.allclose(out["gpu_mean"], out["ref_mean"], rtol=0.01), out
    return out


def _c1_gpu(p):
    from pathtracing_b200.scene import GpuScene
    desc, cam, _ = make_golden.c1_case()
    gpu = GpuScene(desc, 4)
    st = gpu.render(cam, p)
    rgbw, sq = gpu.film(), gpu.film_sumsq()
    gpu.close()
    assert st["samples"] == p.width * p.height * p.spp
    a = rgbw[..., :3] / rgbw[..., 3:4]
    va = np.maximum(sq / p.spp - a ** 2, 0.0) * p.spp / (p.spp - 1)
    return a, va, st


def _twice(fn, what, seeds):
    """a failed statistical comparison is REPORTED and has to reproduce with independent samples before it counts"""
    try:
        return fn(seeds[0])
    except AssertionError as first:
        import warnings
        warnings.warn(f"{what} failed once, repeating with independent samples: {str(first)[:600]}")
        return fn(seeds[1])


@pytest.mark.gpu
def test_c1_full_config_against_the_stored_reference_film():
    """The config as BASELINE.json states it (512 x 512, 16 spp, depth 8) against the reference's stored 256-spp film of the
    same scene (tests/golden/render_c1_512_depth8.npz, made by make_golden.py c1)."""
    g = np.load(GOLDEN / "render_c1_512_depth8.npz")
    desc, cam, p = make_golden.c1_case()
    if str(g["digest"]) != make_golden.scene_digest(desc):
        pytest.fail("golden C1 film was made from a different Cornell scene: regenerate it (python tests/golden/make_golden.py c1)")
    b, vb, nb = g["mean"].astype(np.float64), g["var"].astype(np.float64), int(g["spp"])

    def once(seed):
        p.seed = seed
        a, va, st = _c1_gpu(p)
        return _c1_compare(a, va, p.spp, b, vb, nb)

    _twice(once, "C1 against the stored film", (21, 22))


@needs_ref
@pytest.mark.gpu
def test_c1_full_config_against_the_live_reference():
    """The same frame against the reference itself rendering the config now (oracle/_ref, all host threads, 16 spp): the criteria
    of tests/test_render_gpu.py (symmetric: both films carry 16 samples per pixel), plus the rays traced per sample."""
    from oracle.refapi import RefScene
    from oracle import refapi
    desc, cam, p = make_golden.c1_case()
    ref = RefScene(desc, 4, count_rays=True, flavour="fast" if refapi.available("fast") else "strict")
    r_rgbw, r_sq, r_st = ref.render(cam, p)
    ref.close()
    b = r_rgbw[..., :3] / r_rgbw[..., 3:4]
    vb = np.maximum(r_sq / p.spp - b ** 2, 0.0) * p.spp / (p.spp - 1)

    def once(seed):
        p.seed = seed
        a, va, st = _c1_gpu(p)
        s2 = (va + vb) / p.spp
        ok = s2 > 1e-24
        z = (a - b)[ok] / np.sqrt(s2[ok])
        floor = float(s2.sum() / (b ** 2).sum())
        mse = float(((a - b) ** 2).sum() / (b ** 2).sum())
        out = {"rel_mse": mse, "noise_floor": floor, "z_frac": float((np.abs(z) > 3).mean()), "z_bias": float(z.mean()), "z_std": float(z.std())}
        assert r_st["samples"] == st["samples"]
        assert mse <= 1e-3 + 2.0 * floor and out["z_frac"] <= 0.01 and abs(out["z_bias"]) <= 0.08 and out["z_std"] <= 1.12, out
        assert np.allclose(a.mean(axis=(0, 1)), b.mean(axis=(0, 1)), rtol=0.01), out
        # and the two trace the same number of rays per sample (path lengths, NEE decisions)
        assert abs((st["closest_rays"] + st["shadow_rays"]) / st["samples"] - (r_st["closest_rays"] + r_st["shadow_rays"]) / r_st["samples"]) < 0.02

    _twice(once, "C1 against the live reference", (31, 32))
