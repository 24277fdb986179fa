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
def _c1_stats(a, va, na, b, vb, nb):
    """per-pixel z between two films with per-sample variances va / vb at na / nb samples per pixel"""
    sigma = np.sqrt(va / na + vb / nb)
    ok = sigma > 1e-12
    z = (a - b)[ok] / sigma[ok]
    floor = float((sigma ** 2).sum() / (b ** 2).sum())
    mse = float(((a - b) ** 2).sum() / (b ** 2).sum())
    return z, mse, floor


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


@pytest.mark.gpu
def test_c1_full_config_against_the_stored_reference_film():
    """The config as BASELINE.json states it (512 x 512, 16 spp, depth 8) against the reference's stored 256-spp film of the
    same scene: relMSE above the noise floor <= 1e-3, per-pixel means within 3 sigma as a normal law allows, no bias, channel
    means within 1 %.  A first failure is reported (not swallowed) and must reproduce with independent samples to count."""
    g = np.load(GOLDEN / "render_c1_512_depth8.npz")
    desc, cam, p = make_golden.c1_case()
    if str(g["digest"]) != make_golden.scene_digest(desc):
        pytest.fail("golden C1 film was made from a different Cornell scene: regenerate it (python tests/golden/make_golden.py c1)")
    b, vb, nb = g["mean"].astype(np.float64), g["var"].astype(np.float64), int(g["spp"])

    def once(seed):
        p.seed = seed
        a, va, st = _c1_gpu(p)
        z, mse, floor = _c1_stats(a, va, p.spp, b, vb, nb)
        out = {"rel_mse": mse, "noise_floor": floor, "z_frac": float((np.abs(z) > 3).mean()), "z_bias": float(z.mean()), "z_std": float(z.std()),
               "gpu_mean": a.mean(axis=(0, 1)), "ref_mean": b.mean(axis=(0, 1)), "rays_per_sample": (st["closest_rays"] + st["shadow_rays"]) / st["samples"]}
        assert mse <= 1e-3 + 2.0 * floor, out
        # a 16-spp estimate of a skewed integrand is not normal: 3-sigma outliers run a little above 0.27 % (calibrated on
        # reference-vs-reference renders of this scene, scratch/calib.py); a lobe or stream error moves all three numbers
        assert out["z_frac"] <= 0.012 and abs(out["z_bias"]) <= 0.05 and out["z_std"] <= 1.12, out
        assert np.allclose(out["gpu_mean"], out["ref_mean"], rtol=0.01), out
        return out

    try:
        once(21)
    except AssertionError as first:
        import warnings
        warnings.warn(f"C1 comparison failed once, repeating with independent samples: {str(first)[:400]}")
        once(22)


@needs_ref
@pytest.mark.gpu
def test_c1_full_config_against_the_live_reference():
    """The same frame against the reference itself rendering it now (oracle/_ref, all host threads, 16 spp like the config)."""
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
        z, mse, floor = _c1_stats(a, va, p.spp, b, vb, p.spp)
        out = {"rel_mse": mse, "noise_floor": floor, "z_frac": float((np.abs(z) > 3).mean()), "z_bias": float(z.mean()), "z_std": float(z.std())}
        assert r_st["samples"] == st["samples"]
        assert mse <= 1e-3 + 2.0 * floor and out["z_frac"] <= 0.012 and abs(out["z_bias"]) <= 0.05 and out["z_std"] <= 1.12, out
        assert np.allclose(a.mean(axis=(0, 1)), b.mean(axis=(0, 1)), rtol=0.01), out
        # and the two trace the same number of rays per sample (path lengths, NEE decisions)
        assert abs((st["closest_rays"] + st["shadow_rays"]) / st["samples"] - (r_st["closest_rays"] + r_st["shadow_rays"]) / r_st["samples"]) < 0.02

    try:
        once(31)
    except AssertionError as first:
        import warnings
        warnings.warn(f"C1 live comparison failed once, repeating with independent samples: {str(first)[:400]}")
        once(32)
