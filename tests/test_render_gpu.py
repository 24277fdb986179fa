"""GPU tier: the wavefront integrator (pt_render through the C-ABI) against the UNMODIFIED
reference integrators (oracle/_ref: {Simple,,Vol}PathIntegrator::Li driven by the fixed-spp
mirror of TileIntegrator::Render in oracle/ref_driver.cpp).

The reference seeds its RNG from std::random_device, so parity is statistical (north_star):
  * converged images agree to relative MSE <= 1e-3,
  * per-pixel means agree within 3 sigma (sigma from the per-pixel sample variance of both
    estimators), checked as: the fraction of pixels beyond 3 sigma is
    what a normal law allows (plus slack for heavy tails) and the mean z-score shows no bias.
"""
import math

import numpy as np
import pytest

from conftest import needs_ref
from pathtracing_b200 import capi, synth
from pathtracing_b200.scene import GpuScene, render_params

pytestmark = pytest.mark.gpu


def film_mean(rgbw):
    w = np.maximum(rgbw[..., 3:4], 1e-300)
    return rgbw[..., :3] / w


def compare(*args, **kw):
    """Statistical comparison of one GPU render with one reference render.  Each of its thresholds is calibrated to
    fail an unbiased renderer about once in a few hundred draws, and the GPU tier makes some forty such comparisons;
    so a failure has to REPRODUCE with fresh, independent samples on both sides (another GPU seed; the reference
    seeds itself from random_device) before it counts: a bias fails both draws, chance almost never does."""
    try:
        return _compare_once(1234, *args, **kw)
    except AssertionError as first:
        import warnings
        warnings.warn(f"statistical comparison failed once, repeating with independent samples: {str(first)[:300]}")
        return _compare_once(98765, *args, **kw)


def _compare_once(seed, desc, cam, width, height, spp, depth, integrator=capi.INTEGRATOR_PATH, bvh=4, light_sampler=1,
                  rel_mse=1e-3, z_frac=0.01, z_bias=0.08, filt=capi.FILTER_BOX, radius=(0.5, 0.5), fparam=(0.0, 0.0),
                  sampler=0, strata=(1, 1)):
    from oracle.refapi import RefScene
    params = render_params(width, height, spp, depth, integrator=integrator, filter=filt, filter_radius=radius,
                           filter_param=fparam, seed=seed, sampler=sampler, strata=strata)
    ref = RefScene(desc, bvh, light_sampler=light_sampler, count_rays=True, flavour="fast" if __import__("oracle.refapi").refapi.available("fast") else "strict")
    r_rgbw, r_sq, r_stats = ref.render(cam, params)
    gpu = GpuScene(desc, bvh, light_sampler=light_sampler)
    g_stats = gpu.render(cam, params)
    g_rgbw = gpu.film()
    g_sq = gpu.film_sumsq()
    gpu.close()
    ref.close()
    assert g_stats["samples"] == width * height * spp == r_stats["samples"]
    a, b = film_mean(g_rgbw), film_mean(r_rgbw)
    assert np.isfinite(a).all()
    mse = float(((a - b) ** 2).sum() / max((b ** 2).sum(), 1e-30))
    floor = 0.0
    out = {"rel_mse": mse, "gpu_mean": a.mean(axis=(0, 1)), "ref_mean": b.mean(axis=(0, 1)),
           "gpu_rays": (g_stats["closest_rays"], g_stats["shadow_rays"]), "ref_rays": (r_stats["closest_rays"], r_stats["shadow_rays"])}
    if filt == capi.FILTER_BOX:
        # per-pixel z-score with the sample variance of BOTH estimators (symmetric, so that the
        # correlation between a pixel's mean and its own variance estimate cancels under H0)
        var = (np.maximum(r_sq / spp - b ** 2, 0.0) + np.maximum(g_sq / spp - a ** 2, 0.0)) * spp / max(spp - 1, 1)
        sigma = np.sqrt(var / spp)
        ok = sigma > 1e-12
        z = (a - b)[ok] / sigma[ok]
        out["z_frac"] = float((np.abs(z) > 3).mean())
        out["z_bias"] = float(z.mean())
        out["z_std"] = float(z.std())
        # pixels with zero reference variance must agree almost exactly (e.g. directly seen light, black)
        flat = ~ok
        if flat.any():
            assert np.allclose(a[flat], b[flat], rtol=2e-2, atol=2e-3), out
        assert out["z_frac"] <= z_frac, out
        assert abs(out["z_bias"]) <= z_bias, out
        # calibrated on reference-vs-reference renders of these scenes: z is N(0,1) to within a few percent
        # (scratch/calib.py); correlated sample streams or a biased lobe push the spread well above 1
        assert out["z_std"] <= 1.12, out
        # what two unbiased estimators of the same image differ by at this sample count
        out["noise_floor"] = floor = float((sigma ** 2).sum() / max((b ** 2).sum(), 1e-30))
    # relMSE <= 1e-3 once converged: the excess over the Monte-Carlo noise floor of this spp must stay below it
    assert mse <= rel_mse + 2.0 * floor, out
    # global energy: channel means agree to 1 % (plus the standard error of the two estimates, which
    # rare bright paths -- caustics seen through glass -- dominate)
    stderr = float(np.sqrt((sigma ** 2).sum() / 3.0) / (width * height)) if filt == capi.FILTER_BOX else 0.0
    out["mean_stderr"] = stderr
    assert np.allclose(out["gpu_mean"], out["ref_mean"], rtol=1e-2, atol=1e-4 + 4.0 * stderr), out
    return out


@needs_ref
def test_cornell_path_nee_mis():
    """C1: Cornell box, Lambertian + area light, NEE + MIS, depth 8."""
    desc, cam = synth.cornell_scene()
    out = compare(desc, cam, 64, 64, 2048, 8)
    # rays per sample in the same ballpark (same termination statistics)
    g = out["gpu_rays"][0] / out["ref_rays"][0]
    assert 0.97 < g < 1.03, out


@needs_ref
@pytest.mark.parametrize("integrator", [capi.INTEGRATOR_PATH, capi.INTEGRATOR_VOLPATH])
def test_small_slot_pool_regenerates_slots(monkeypatch, integrator):
    """A pool much smaller than the call (PTB_BATCH): finished slots are splatted and handed to new camera samples
    mid-flight, path state lives in 64 B records, the queue order becomes a random order over the pool
    (render.cu: PathState layout, done list, k_generate into free slots).  The default pool holds these test frames whole."""
    monkeypatch.setenv("PTB_BATCH", "98304")
    desc, cam = synth.materials_scene(subdiv=2, sky=True)
    out = compare(desc, cam, 64, 48, 512, 8, integrator=integrator)
    assert out["gpu_rays"][0] > 64 * 48 * 512


@needs_ref
def test_cornell_bvh2_uniform_light_sampler():
    desc, cam = synth.cornell_scene()
    compare(desc, cam, 48, 48, 1024, 5, bvh=2, light_sampler=0)


@needs_ref
def test_cornell_stratified_sampler():
    """StratifiedSampler (Sampler.hpp:66-128): jittered, permuted strata for the camera dimensions.
    The per-pixel sample variance over-estimates a stratified estimator's, so the z-test is conservative;
    the edge pixels (where stratification of the pixel area matters) carry the check."""
    desc, cam = synth.cornell_scene()
    compare(desc, cam, 48, 48, 1024, 5, sampler=1, strata=(32, 32))


def test_stratified_sampler_covers_every_stratum():
    """The film weight of a pixel is the sum of filter values at its sample offsets (radius 0.5: own pixel
    only).  With n x n strata that sum spreads far less across pixels than with n*n independent offsets --
    a property of the sampler alone, whatever the scene."""
    desc, cam = synth.cornell_scene()
    spread, mean = {}, {}
    for sampler, strata in ((0, (1, 1)), (1, (8, 8))):
        gpu = GpuScene(desc, 4)
        gpu.render(cam, render_params(32, 32, 64, 1, sampler=sampler, strata=strata, seed=7, filter=capi.FILTER_GAUSSIAN,
                                      filter_radius=(0.5, 0.5), filter_param=(0.25, 0.0)))
        w = gpu.film()[..., 3]
        gpu.close()
        mean[sampler] = float(w.mean())
        spread[sampler] = float(w.std() / w.mean())
    assert abs(mean[0] / mean[1] - 1.0) < 0.02, mean
    assert spread[1] < 0.35 * spread[0], spread


@needs_ref
@pytest.mark.parametrize("mode", [capi.ALPHA_MASK, capi.ALPHA_BLEND])
def test_alpha_tested_cards(mode):
    """Cut-out cards: Material::Alpha rejects hits inside traversal for camera, bounce and shadow rays."""
    desc, cam = synth.alpha_scene(mode)
    compare(desc, cam, 48, 48, 1024, 5)


@needs_ref
def test_cornell_simple_integrator():
    """SimplePathIntegrator: BSDF sampling only (a19)."""
    desc, cam = synth.cornell_scene(light_scale=4.0, light_size=0.7)
    compare(desc, cam, 32, 32, 4096, 6, integrator=capi.INTEGRATOR_SIMPLE)


@needs_ref
@pytest.mark.parametrize("filt,radius,fparam", [(capi.FILTER_GAUSSIAN, (1.5, 1.5), (0.5, 0.0)),
                                               (capi.FILTER_MITCHELL, (1.5, 1.5), (1 / 3, 1 / 3)),
                                               (capi.FILTER_LANCZOS, (1.5, 1.5), (3.0, 0.0))])
def test_film_filters(filt, radius, fparam):
    desc, cam = synth.cornell_scene()
    compare(desc, cam, 48, 48, 1024, 4, filt=filt, radius=radius, fparam=fparam, rel_mse=2e-3)


def test_pass_sharding_is_independent_of_rank_count():
    """sample passes {r, r+G, ...} over G 'ranks' accumulate to the 1-rank film (same sample set)."""
    desc, cam = synth.cornell_scene()
    gpu = GpuScene(desc, 4)
    p = render_params(40, 40, 12, 4, seed=9)
    gpu.render(cam, p)
    one = gpu.film().copy()
    for r in range(3):
        gpu.render(cam, render_params(40, 40, 12, 4, seed=9, pass_begin=r, pass_stride=3), clear=(r == 0))
    three = gpu.film()
    assert np.array_equal(one[..., 3], three[..., 3])
    assert np.allclose(one[..., :3], three[..., :3], rtol=1e-9, atol=1e-12)
    # and of the batch size
    import os
    os.environ["PTB_BATCH"] = "5000"
    gpu.render(cam, p)
    small = gpu.film()
    del os.environ["PTB_BATCH"]
    assert np.allclose(one[..., :3], small[..., :3], rtol=1e-9, atol=1e-12)
    gpu.close()


def test_render_error_paths():
    desc, cam = synth.cornell_scene()
    gpu = GpuScene(desc, 4)
    with pytest.raises(capi.PtError):
        gpu.render(cam, render_params(0, 10, 1, 1))
    with pytest.raises(capi.PtError):
        gpu.render(cam, render_params(8, 8, 1, 1, integrator=7))
    with pytest.raises(capi.PtError):
        gpu.film_device_ptr() if False else gpu.render(cam, render_params(8, 8, 1, 1, filter=9))
    st = gpu.render(cam, render_params(8, 8, 2, 0))      # max depth 0: black, but counted
    assert st["samples"] == 128 and st["closest_rays"] == 0
    assert gpu.film()[..., :3].sum() == 0
    gpu.close()


@needs_ref
def test_materials_instances_sky():
    """C3 family: TLAS/BLAS instances, microfacet dielectric / diffuse / conductor / thin dielectric,
    image + checker + normal-map textures, TransformedLight, gradient sky."""
    desc, cam = synth.materials_scene()
    compare(desc, cam, 64, 48, 1024, 12)


@needs_ref
def test_function_infinite_light_baked_into_a_table():
    """Any sky function other than the GradientSky functor reaches the device as a baked lat-long table (FunctionInfiniteLight,
    Light.hpp:74; host/PathTracing.hpp).  Here the table holds the gradient sky itself, so the reference (closed form) renders
    the same scene: same image within the usual criteria, uniform-sphere sampling and its pdf unchanged."""
    desc, cam = synth.materials_scene(baked_sky=True)
    compare(desc, cam, 64, 48, 1024, 12)


@needs_ref
def test_materials_point_and_distant_lights():
    desc, cam = synth.materials_scene(sky=False, point_light=True, distant_light=True)
    compare(desc, cam, 48, 36, 1024, 8)


@needs_ref
def test_volpath_homogeneous_medium():
    """C4 family: global homogeneous medium, Henyey-Greenstein phase function, VolPathIntegrator."""
    desc, cam = synth.medium_scene()
    compare(desc, cam, 48, 36, 1024, 12, integrator=capi.INTEGRATOR_VOLPATH)


@needs_ref
def test_c3_instanced_materials_small():
    """C3 at test size: the same builder as the bench scene (instanced displaced spheres, four materials,
    checkered ground, transformed quad light, gradient sky)."""
    desc, cam = synth.c3_scene(subdiv=2, n_instances=9)
    compare(desc, cam, 64, 36, 1024, 10)


@needs_ref
def test_c4_medium_textures_small():
    desc, cam = synth.c4_scene(tex_size=64, subdiv=2)
    compare(desc, cam, 48, 27, 1024, 12, integrator=capi.INTEGRATOR_VOLPATH)


@needs_ref
def test_c5_terrain_hdr_environment_small():
    """C5 at test size: multi-mesh terrain BLAS with per-patch image textures, glass / metal instances,
    emissive quads and an HDR environment through TextureInfiniteLight (importance-sampled sun)."""
    desc, cam = synth.c5_scene(grid=48, n_textures=4, tex_size=16)
    compare(desc, cam, 48, 27, 2048, 8)


@needs_ref
def test_misc_uniform_sky_uv_normal_textures_thin_lens():
    desc, cam = synth.misc_scene()
    compare(desc, cam, 64, 48, 1024, 8)


@needs_ref
@pytest.mark.parametrize("one_sided", [True, False])
def test_sphere_and_quad_shapes_path(one_sided):
    """main.cpp-style scene of analytic GeometricPrimitives: checker ground QuadShape, SphereShapes with every
    material, a QuadShape AreaLight (one- / two-sided), an emissive textured SphereShape, a mesh instance."""
    desc, cam = synth.shapes_scene(one_sided=one_sided)
    compare(desc, cam, 64, 48, 1024, 10)


@needs_ref
def test_sphere_and_quad_shapes_simple_integrator_bvh8_uniform_lights():
    desc, cam = synth.shapes_scene()
    compare(desc, cam, 48, 36, 1024, 6, integrator=capi.INTEGRATOR_SIMPLE, bvh=8)
    compare(desc, cam, 48, 36, 512, 6, bvh=2, light_sampler=0)


@needs_ref
def test_medium_boundary_sphere_volpath_intersect_tr():
    """A material-less SphereShape holding a HomogeneusMedium (main.cpp:208): the path passes through the
    boundary and switches medium (Integrators.cpp:378-382), shadow rays walk through it collecting
    transmittance (Scene::IntersectTr, Scene.cpp:8-29)."""
    desc, cam = synth.shapes_scene(boundary=True)
    compare(desc, cam, 64, 48, 1024, 12, integrator=capi.INTEGRATOR_VOLPATH)


@needs_ref
def test_medium_boundary_sphere_path_integrator():
    """PathIntegrator with the same boundary: camera paths continue through it (Integrators.cpp:228-232),
    shadow rays are blocked by it (Unoccluded -> IntersectPred)."""
    desc, cam = synth.shapes_scene(boundary=True)
    compare(desc, cam, 48, 36, 1024, 8)


@needs_ref
def test_wrapped_primitives_path_motion_blur():
    """TransformedPrimitive over sphere / quad GeometricPrimitives and AnimatedPrimitives (motion blur over the
    camera's shutter sample, main.cpp:187-190) incl. a transformed quad light and an animated sphere light
    (TransformedLight / AnimatedLight, Light.cpp:296-365)."""
    desc, cam = synth.wrapped_scene()
    compare(desc, cam, 64, 48, 1024, 10)


@needs_ref
def test_wrapped_primitives_volpath_boundary_bvh2():
    desc, cam = synth.wrapped_scene(boundary=True)
    compare(desc, cam, 48, 36, 1024, 8, integrator=capi.INTEGRATOR_VOLPATH, bvh=2)


@needs_ref
@pytest.mark.parametrize("mode", [capi.ALPHA_MASK, capi.ALPHA_BLEND])
def test_alpha_tested_shapes(mode):
    """alpha-tested sphere / quad GeometricPrimitives (plain and transformed) next to the alpha-tested cards"""
    desc, cam = synth.alpha_shapes_scene(mode)
    compare(desc, cam, 48, 48, 1024, 6)


@needs_ref
def test_volpath_without_medium_equals_reference():
    desc, cam = synth.cornell_scene()
    compare(desc, cam, 40, 40, 1024, 6, integrator=capi.INTEGRATOR_VOLPATH)


@needs_ref
@pytest.mark.parametrize("integrator,spp", [(capi.INTEGRATOR_SIMPLE, 4), (capi.INTEGRATOR_PATH, 2)])
def test_adaptive_sampling_loop(integrator, spp):
    try:
        _adaptive_sampling_loop(integrator, spp, 9)
    except AssertionError:       # statistical: has to reproduce with independent samples
        _adaptive_sampling_loop(integrator, spp, 10)


def _adaptive_sampling_loop(integrator, spp, seed):
    """TileIntegrator::Render's per-pixel stopping rule (Integrators.cpp:59-86): rounds of spp samples until
    every channel's 95 % interval is within 1.5 x its mean, at most 128 rounds.  With a box filter the film
    weight of a pixel IS its sample count, so the rule itself can be compared: total samples, the
    distribution of rounds per pixel, and where in the image the extra rounds go."""
    from oracle.refapi import RefScene
    desc, cam = synth.cornell_scene(light_scale=40.0, light_size=0.12)
    W = H = 48
    p = render_params(W, H, spp, 5, integrator=integrator, seed=seed, adaptive=1)
    ref = RefScene(desc, 4, flavour="fast" if __import__("oracle.refapi").refapi.available("fast") else "strict")
    r_rgbw, _, r_stats = ref.render(cam, p)
    ref.close()
    gpu = GpuScene(desc, 4)
    g_stats = gpu.render(cam, p)
    g_rgbw = gpu.film()
    gpu.close()
    rn, gn = r_rgbw[..., 3] / spp, g_rgbw[..., 3] / spp                 # rounds per pixel
    assert np.allclose(gn, np.round(gn)) and gn.min() >= 1 and gn.max() <= 128
    assert g_stats["samples"] == int(round(g_rgbw[..., 3].sum())) and r_stats["samples"] == int(round(r_rgbw[..., 3].sum()))
    assert rn.max() > 4, "scene too easy: the reference never iterates"
    # the raw total is dominated by the two or three dozen pixels that run to the 128-round cap (512 samples each;
    # their count is Poisson-like, and the reference is seeded from random_device: 25144 vs 16704 was seen once),
    # so the totals are compared with the rounds per pixel clipped at 16; the rule itself is compared through
    # the distribution of rounds per pixel
    cr, cg = float(np.minimum(rn, 16).sum()), float(np.minimum(gn, 16).sum())
    assert abs(cg / cr - 1.0) < 0.15, (cg, cr, g_stats["samples"], r_stats["samples"])
    assert abs(g_stats["samples"] / r_stats["samples"] - 1.0) < 0.6, (g_stats["samples"], r_stats["samples"])
    npx = W * H
    for lo, hi in ((1, 1), (2, 3), (4, 11), (12, 128)):
        cr, cg = int(((rn >= lo) & (rn <= hi)).sum()), int(((gn >= lo) & (gn <= hi)).sum())
        pr = max(cr, cg, 1) / npx
        assert abs(cr - cg) <= 5.0 * np.sqrt(2 * npx * pr * (1 - pr)) + 4, ((lo, hi), cr, cg)
    if integrator == capi.INTEGRATOR_PATH:
        # and it iterates in the same places (penumbra, ceiling): block-averaged "needed more than one round" maps
        blk = lambda a: (a > 1).reshape(H // 8, 8, W // 8, 8).mean(axis=(1, 3)).ravel()
        assert np.corrcoef(blk(gn), blk(rn))[0, 1] > 0.7
    a, b = film_mean(g_rgbw), film_mean(r_rgbw)
    assert np.allclose(a.mean(axis=(0, 1)), b.mean(axis=(0, 1)), rtol=0.25)     # a few samples per pixel: sanity only


def test_adaptive_rejects_pass_sharding():
    desc, cam = synth.cornell_scene()
    gpu = GpuScene(desc, 4)
    with pytest.raises(capi.PtError, match="pass sharding"):
        gpu.render(cam, render_params(16, 16, 4, 3, adaptive=1, pass_begin=1, pass_stride=2))
    gpu.close()


def test_adaptive_with_pass_sharding_equals_single_rank():
    """Two ranks (two scenes on this GPU, one host thread each) shard the passes of an adaptive render and sum
    their per-pixel statistics every round through pt_scene_set_allreduce: every pixel must stop after the
    same round as in the single-rank render, and the summed films must agree (sample streams are keyed by
    (pixel, sample index), so the two runs draw identical samples)."""
    import threading

    import torch

    from pathtracing_b200 import multi
    desc, cam = synth.cornell_scene(light_scale=40.0, light_size=0.12)
    W = H = 40
    spp = 4
    one = GpuScene(desc, 4)
    one.render(cam, render_params(W, H, spp, 5, seed=5, adaptive=1))
    ref = one.film()
    one.close()
    assert (ref[..., 3] / spp).max() > 3, "scene too easy: no pixel iterates"

    G = 2
    scenes = [GpuScene(desc, 4) for _ in range(G)]
    barrier = threading.Barrier(G)
    bufs = [None] * G
    calls = [0] * G
    keep = []

    def make_cb(r):
        def cb(user, ptr, n, stream):
            bufs[r] = torch.as_tensor(multi._DevArray(ptr, n), device="cuda:0")
            calls[r] += 1
            barrier.wait()
            if r == 0:
                total = bufs[0] + bufs[1]
                bufs[0].copy_(total); bufs[1].copy_(total)
                torch.cuda.synchronize()
            barrier.wait()
            return 0
        return capi.ALLREDUCE_FN(cb)

    errors = []

    def work(r):
        try:
            keep.append(make_cb(r))
            scenes[r]._ck(scenes[r].lib.pt_scene_set_allreduce(scenes[r].h, keep[-1], None))
            scenes[r].render(cam, render_params(W, H, spp, 5, seed=5, adaptive=1, pass_begin=r, pass_stride=G))
        except Exception as e:
            errors.append(e)
            barrier.abort()

    threads = [threading.Thread(target=work, args=(r,)) for r in range(G)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=300)
    assert not errors, errors
    assert calls[0] == calls[1] and calls[0] >= 3
    films = [sc.film() for sc in scenes]
    for sc in scenes:
        sc.close()
    total = films[0] + films[1]
    assert np.array_equal(total[..., 3], ref[..., 3])                      # same number of rounds for every pixel
    assert np.allclose(total[..., :3], ref[..., :3], rtol=1e-9, atol=1e-12)
    assert films[0][..., 3].sum() > 0 and films[1][..., 3].sum() > 0


def test_sample_streams_are_independent():
    """Two renders with different seeds must differ per pixel by what their own sample variance predicts
    (z ~ N(0,1)).  Guards the per-sample generator seeding: PCG streams that differ only in a small
    increment are correlated and gave z_std 1.7 here."""
    desc, cam = synth.materials_scene()
    spp, img = 1024, []
    for seed in (11, 12):
        gpu = GpuScene(desc, 4)
        gpu.render(cam, render_params(64, 48, spp, 8, seed=seed))
        img.append((film_mean(gpu.film()), gpu.film_sumsq()))
        gpu.close()
    (a, asq), (b, bsq) = img
    var = (np.maximum(asq / spp - a ** 2, 0.0) + np.maximum(bsq / spp - b ** 2, 0.0)) * spp / (spp - 1)
    sigma = np.sqrt(var / spp)
    ok = sigma > 1e-12
    z = (a - b)[ok] / sigma[ok]
    assert 0.85 < z.std() < 1.12 and (np.abs(z) > 3).mean() < 0.01 and abs(z.mean()) < 0.08, (z.std(), (np.abs(z) > 3).mean(), z.mean())


def test_film_tensor_view_and_single_rank_reduce():
    """multi.render_sharded on one rank: the torch view aliases the device film (no copy)."""
    import torch
    from pathtracing_b200 import multi
    desc, cam = synth.cornell_scene()
    gpu = GpuScene(desc, 4)
    p = render_params(32, 32, 4, 3, seed=3)
    st, film = multi.render_sharded(gpu, cam, p, None)
    host = gpu.film()
    assert torch.equal(film.cpu(), torch.from_numpy(host.reshape(-1)))
    assert float(host[..., 3].min()) == 4.0 == float(host[..., 3].max())
    gpu.close()


@pytest.mark.parametrize("case", ["c3", "c4", "materials", "wrapped", "boundary"])
def test_render_on_the_fast_layout_gives_the_same_film(case):
    """PT_LAYOUT_FAST under pt_render (what bench.py uses for C3 / C4 / C5): the quantised clusters only decide which triangles get
    tested, the sample streams are keyed by sample id, so the film must be the exact layout's film up to the order of the fp64 sums --
    except for the rare paths that meet a tie (two triangles at bit-equal t; the fast closest-hit kernel visits children by entry
    distance).  Every integrator, textures, a medium, wrappers with motion blur, a medium boundary (k_shadow_tr on the fast layout)."""
    desc, cam, integ = {
        "c3": lambda: (*synth.c3_scene(subdiv=3, n_instances=20), capi.INTEGRATOR_PATH),
        "c4": lambda: (*synth.c4_scene(tex_size=64, subdiv=3), capi.INTEGRATOR_VOLPATH),
        "materials": lambda: (*synth.materials_scene(subdiv=3), capi.INTEGRATOR_SIMPLE),
        "wrapped": lambda: (*synth.wrapped_scene(), capi.INTEGRATOR_PATH),
        "boundary": lambda: (*synth.shapes_scene(boundary=True), capi.INTEGRATOR_VOLPATH),
    }[case]()
    p = render_params(96, 64, 8, 6, integrator=integ, seed=3)
    films = []
    for layout in (capi.LAYOUT_EXACT, capi.LAYOUT_FAST):
        g = GpuScene(desc, 4, layout=layout)
        st = g.render(cam, p)
        films.append((st, g.film().copy()))
        g.close()
    (sa, fa), (sb, fb) = films
    assert np.array_equal(fa[..., 3], fb[..., 3]) and sa["samples"] == sb["samples"]
    same = np.isclose(fa[..., :3], fb[..., :3], rtol=1e-9, atol=1e-12).all(axis=-1)
    assert same.mean() >= 0.999, float(same.mean())
    assert abs(sa["closest_rays"] - sb["closest_rays"]) <= 1e-4 * sa["closest_rays"]
