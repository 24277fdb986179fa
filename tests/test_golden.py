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
