#!/usr/bin/env python
"""bench.py — BASELINE.json's metric on its C2 config (ray-cast-only: 16 Mi random
incoherent rays vs a synthetic 1M-triangle mesh, BVH4 closest-hit, BVH2 beside it).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N ...            # the reference's own CPU code

One JSON line on stdout (rank 0).  A "step" is one pass of the hot path over one batch
of rays.  `value` is device-resident throughput (CUDA events, max over ranks); `e2e` is
the same metric through the C-ABI's host-buffer entry point (pinned host rays in, host
hits out, copies inside the timed region).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "incoherent closest-hit ray throughput"
UNIT = "Mrays/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--tris", type=int, default=1_000_000)
    ap.add_argument("--rays", type=int, default=1 << 24)
    ap.add_argument("--mesh", default="soup", choices=["soup", "terrain"])
    ap.add_argument("--bvh", type=int, default=4, choices=[2, 4, 8])
    ap.add_argument("--render-layout", default="auto", choices=["auto", "fast", "exact"],
                    help="node layout of the render configs: auto = RENDER_LAYOUT per config (fast for C3 / C4 / C5, exact for C1)")
    ap.add_argument("--layout", default="fast", choices=["fast", "exact"],
                    help="BVH4 node layout of the headline: 'fast' = 64 B quantised clusters of the reference's tree (same hits, "
                         "tests/test_raycast_gpu.py), 'exact' = the reference's 128 B cluster content; the other one is reported in extras")
    ap.add_argument("--no-extras", action="store_true", help="skip the BVH2 / any-hit / second-mesh side measurements")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--renders", default="c1,c3,c4,c5", help="render configs measured as extras: subset of c1,c3,c4,c5 (c5: 10 M triangles, 1024 spp, 128 bounces)")
    ap.add_argument("--no-dram-cast", action="store_true", help="skip the DRAM-bound ray cast (16 Mi rays vs the 10 M-triangle C5 geometry)")
    return ap.parse_args()


def make_scene(mesh, tris):
    from pathtracing_b200 import synth
    if mesh == "soup":
        return synth.soup_scene(tris)
    nx = int(round(math.sqrt(tris / 2.0 * 2.0))) + 1
    nz = tris // (2 * (nx - 1)) + 1
    return synth.terrain_scene(nx, nz)


def workload_name(a):
    return (f"C2 ray-cast: {a.rays} random incoherent rays (origin uniform in scene box, direction uniform on S2, "
            f"tmax=inf) vs {a.tris}-triangle synthetic {a.mesh}, BVH{a.bvh} closest-hit")


def config_of(a):
    """The SAME dict in both arms (the driver compares them): what is measured, nothing about how."""
    return {"workload": workload_name(a), "bvh": a.bvh}


def tracked_profile(name):
    """Numbers that come from an ncu capture are read from a tracked file under profiles/, never typed in here."""
    f = ROOT / "profiles" / name
    return json.loads(f.read_text()) if f.exists() else None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.rows = []
        self.proc = None
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# ---------------------------------------------------------------------------------------------
# reference arm: the reference's own CPU implementation (oracle/_ref, Release flags) on host cores
# ---------------------------------------------------------------------------------------------

def cpu_reference_rate(a, desc, rays, budget_s=12.0, steps=1, warmup=0):
    from oracle import refapi
    flavour = "fast" if refapi.available("fast") else "strict"
    kind = "reference"
    ref = refapi.RefScene(desc, a.bvh, id_mode=False, flavour=flavour)
    cores = ref.threads
    pilot = min(len(rays), 100_000)
    ref.trace_closest(rays[:pilot])
    rate = pilot / max(ref.last_seconds, 1e-6)
    n = int(min(len(rays), max(200_000, rate * budget_s)))
    for _ in range(warmup):
        ref.trace_closest(rays[:min(n, 200_000)])
    secs = []
    for _ in range(max(steps, 1)):
        ref.trace_closest(rays[:n])
        secs.append(ref.last_seconds)
    mr = n / np.mean(secs) / 1e6
    ref.close()
    return {"value": mr, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"first {n} rays of the {len(rays)}-ray set, all {cores} host threads, reference compiled -O3 ({flavour})"}, float(np.mean(secs)) * 1e3


def run_reference(a):
    rank, world, local = dist_env()
    if rank != 0:
        return
    from pathtracing_b200 import synth
    desc = make_scene(a.mesh, a.tris)
    pos_lo, pos_hi = bounds_of(desc)
    rays = synth.random_rays(min(a.rays, 1 << 22), pos_lo, pos_hi)
    budget = max(4.0, min(20.0, 150.0 / max(a.steps + a.warmup, 1)))
    base, ms = cpu_reference_rate(a, desc, rays, budget_s=budget, steps=a.steps, warmup=min(a.warmup, 1))
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config_of(a),
            "sample_rays": len(rays),
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def bounds_of(desc):
    lo = np.full(3, np.inf, np.float32)
    hi = np.full(3, -np.inf, np.float32)
    for arr in desc._keep:
        if arr.dtype == np.float32 and arr.ndim == 2 and arr.shape[1] == 3:
            lo = np.minimum(lo, arr.min(0))
            hi = np.maximum(hi, arr.max(0))
            break   # first float3 array of the single mesh = positions
    return lo, hi


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------

def time_dev(fn, steps, warmup, torch, dist, world):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms / steps


def render_configs():
    """BASELINE.json configs[0], [2], [3], [4] as (key, workload text, scene builder, W, H, spp, depth, integrator)."""
    from pathtracing_b200 import capi, synth
    return {
        "c1": ("C1 Cornell 512x512, 16 spp, depth 8, NEE+MIS, PathIntegrator, BVH4", synth.cornell_scene, 512, 512, 16, 8, capi.INTEGRATOR_PATH),
        "c3": ("C3 instanced 10 M triangles (122 x 81 920-triangle BLAS instances via TLAS4/BLAS4), rough dielectric / rough metal / "
               "Lambertian / specular conductor, 1920x1080, 64 spp, depth 8, PathIntegrator", synth.c3_scene, 1920, 1080, 64, 8, capi.INTEGRATOR_PATH),
        "c4": ("C4 homogeneous medium (HG g=0.75) + generated 1024^2 albedo/roughness/normal textures, NEE+MIS through media, "
               "1920x1080, 128 spp, depth 8, VolPathIntegrator", synth.c4_scene, 1920, 1080, 128, 8, capi.INTEGRATOR_VOLPATH),
        "c5": ("C5 10 M unique triangles, 64 generated texture sets + HDR environment (TextureInfiniteLight), 1920x1080, 1024 spp, "
               "128 bounces, PathIntegrator", synth.c5_scene, 1920, 1080, 1024, 128, capi.INTEGRATOR_PATH),
    }


# per-bounce state traffic of the wavefront loop (bytes): k_shade reads and writes the 64 B path state and
# reads the 32 B ray + 4 B slot + 16 B hit + 4 B instance id that k_trace wrote; a continuing path writes the next ray + slot;
# shading fetches 3 indices + 3 x (position 12 + normal 12 + uv 8 + tangent 12) = 144 B
B_STATE_PER_BOUNCE = 2 * 64 + (32 + 4 + 16 + 4) + (32 + 4) + 144
B_SHADOW_QUEUE = 2 * (32 + 4 + 16)        # a shadow ray, its slot and contribution: written by k_shade, read by the any-hit kernel
B_SAMPLE_FIXED = 64 + 32 + 4 + 7 * 8      # generate writes state + first ray + slot; the film splat is a read-modify-write of 7 doubles (box filter)
RENDER_LAYOUT = {"c1": "exact", "c3": "fast", "c4": "fast", "c5": "fast"}     # --render-layout auto
TEX_LOOKUPS = {"c1": 0, "c3": 0, "c4": 3, "c5": 2}   # image textures evaluated per surface bounce (4 texels x 4 B each, bilinear)


def sample_roofline(key, counted, ms, W, H, spp, peak, peak_src, fast=False):
    """B_sample = sum over bounces [B_ray(closest) + B_ray(shadow) + path-state traffic + shading fetches] + film,
    with the visit counts MEASURED by the counting build of the same kernels on this scene (a counted frame at low spp)."""
    st, c, sh = counted
    n_c, n_s, n = max(st["closest_rays"], 1), max(st["shadow_rays"], 1), max(st["samples"], 1)
    s_node, s_tri = (64, 64) if fast else (128, 48)       # bytes of a cluster / a triangle record in the scene's node layout
    b_closest = 32 + 16 + 4 + ((c["tlas_nodes"] + c["blas_nodes"]) * s_node + c["triangles"] * s_tri + c["instances"] * 96) / n_c
    b_shadow = 32 + ((sh["tlas_nodes"] + sh["blas_nodes"]) * s_node + sh["triangles"] * s_tri + sh["instances"] * 96) / n_s
    bounces, shadows = n_c / n, st["shadow_rays"] / n
    b_sample = bounces * (b_closest + B_STATE_PER_BOUNCE + TEX_LOOKUPS[key] * 16) + shadows * (b_shadow + B_SHADOW_QUEUE) + B_SAMPLE_FIXED
    achieved = W * H * spp * b_sample / (ms * 1e-3) / 1e9
    return {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
            "peak_source": peak_src, "bytes_per_sample": b_sample, "bounces_per_sample": bounces, "shadow_rays_per_sample": shadows,
            "closest": {"nodes_per_ray": (c["tlas_nodes"] + c["blas_nodes"]) / n_c, "tris_per_ray": c["triangles"] / n_c, "instances_per_ray": c["instances"] / n_c, "bytes_per_ray": b_closest},
            "shadow": {"nodes_per_ray": (sh["tlas_nodes"] + sh["blas_nodes"]) / n_s, "tris_per_ray": sh["triangles"] / n_s, "instances_per_ray": sh["instances"] / n_s, "bytes_per_ray": b_shadow},
            "counted_on": f"{st['samples']} samples of the same frame with the counting build of the kernels",
            "note": "algorithmic bytes as SURVEY.md 8d defines them (no cache reuse assumed): the frames are traversal-bound (75-80 % of the time in k_trace), "
                    "and k_trace is bound by L1 wavefronts / latency, not DRAM -- a fraction near or above 1 means the caches serve most node reads, not that HBM is saturated"}


def bench_render(a, key, torch, dist, world, rank, local, peak, peak_src):
    """One render config: sample passes sharded over the ranks, one NCCL reduce of the film per frame, all inside
    libptb200.so (pt_dist_render; strong scaling: total work fixed).  Frame time = CUDA events around render + reduce,
    max over ranks."""
    from pathtracing_b200 import capi, multi
    from pathtracing_b200.scene import GpuScene, render_params
    name, builder, W, H, spp, depth, integ = render_configs()[key]
    t0 = time.perf_counter()
    desc, cam = builder()
    synth_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    # node layout per config: the quantised 64 B clusters win wherever rays diverge (C3 +8 %, C4 +7 %, measured with this build); the
    # 512 x 512 Cornell frame is all coherent rays over 36 triangles and is a few % faster on the exact clusters
    fast = RENDER_LAYOUT.get(key, "exact") == "fast" if a.render_layout == "auto" else a.render_layout == "fast"
    scene = GpuScene(desc, 4, device=local, layout=capi.LAYOUT_FAST if fast else capi.LAYOUT_EXACT)
    build_s = time.perf_counter() - t0
    scene.dist_init(dist if world > 1 else None)
    params = render_params(W, H, spp, depth, integrator=integ, seed=1)
    total = W * H * spp
    heavy = total > (1 << 25)
    very_heavy = total > (1 << 30)
    # a light frame (C1: 5 ms) is timed 20 times: one host hiccup of 10 ms in five frames moved the mean by a third (r02_bench_v10.json)
    steps = 1 if very_heavy else 3 if heavy else max(2, min(2 * a.steps, 20))
    # warm-up: whole frames, except for the very heavy config (C5: 2 G samples), which warms up on a 64-spp frame of the same
    # scene -- enough samples per rank to size the wavefront arena as the full frame does (64 Mi slots), so that the timed
    # frame does not pay a 16 GB cudaMalloc
    warm = params if not very_heavy else render_params(W, H, min(spp, 64 * world), depth, integrator=integ, seed=2)
    # the barrier comes BEFORE the warm-up frames: every frame ends in a collective of its own (the film reduce), which keeps the ranks
    # together, and the first frame behind a torch NCCL barrier took 6.8 ms instead of 1.5 (C1 on 8 GPUs, r02_bench_n8_v3.json)
    import gc
    gc.collect()
    gc.disable()            # a generation-2 collection inside a 5 ms frame showed up as one 30-100 ms frame in five (frame_ms_rank0)
    if world > 1:
        dist.barrier()
    warm_ms = []
    for _ in range(2 if heavy else 3):
        tw = time.perf_counter()
        multi.render_sharded(scene, cam, warm, dist if world > 1 else None)
        warm_ms.append((time.perf_counter() - tw) * 1e3)
    # a light frame (C1: 5 ms) keeps warming up until the host side has settled: on some boxes the first frames behind the ray-cast part of this
    # script took 15-40 ms each against 5.3 ms on the device (r02_bench_v11.json frame_ms_rank0).  Up to 60 more frames, until three in a row are
    # within 10 % of the fastest one seen; every rank runs the same number (the decision is rank 0's, broadcast), a frame is a collective.
    # The heavy frames (C3, C4: 0.2-0.6 s) get up to three more, until two in a row are within 5 % (r02_bench_n4_v2.json: C3's first timed frame 83 ms, then 69).
    extra_warm = 0
    max_extra, need, tol = (60, 3, 1.1) if not heavy else (3, 2, 1.05) if not very_heavy else (0, 0, 0.0)
    while extra_warm < max_extra:
        settled = len(warm_ms) >= need and max(warm_ms[-need:]) <= tol * min(warm_ms)
        if world > 1:
            flag = torch.tensor([1 if settled else 0], device="cuda", dtype=torch.int32)
            dist.broadcast(flag, 0)
            settled = bool(flag.item())
        if settled:
            break
        tw = time.perf_counter()
        multi.render_sharded(scene, cam, warm, dist if world > 1 else None)
        warm_ms.append((time.perf_counter() - tw) * 1e3)
        extra_warm += 1
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_wall = time.perf_counter()
    e0.record()
    rays = 0
    launches = 0
    gpu_ms = 0.0
    frame_ms = []
    for _ in range(steps):
        tf = time.perf_counter()
        st, film = multi.render_sharded(scene, cam, params, dist if world > 1 else None)
        frame_ms.append((time.perf_counter() - tf) * 1e3)
        rays += st["closest_rays"] + st["shadow_rays"]
        launches += st["kernel_launches"]
        gpu_ms += st["gpu_ms"]
    e1.record()
    torch.cuda.synchronize()
    # pt_dist_render is synchronous and runs on the scene's own stream: the wall clock around it is the frame time
    # (torch's events bracket the same region on torch's stream and agree when the region is long)
    ms = (time.perf_counter() - t_wall) * 1e3 / steps
    gc.enable()
    if world > 1:
        t = torch.tensor([ms, float(rays)], device="cuda", dtype=torch.float64)
        dist.all_reduce(t[:1], op=dist.ReduceOp.MAX)
        dist.all_reduce(t[1:], op=dist.ReduceOp.SUM)
        ms, rays = float(t[0].item()), float(t[1].item())
    out = None
    if rank == 0:
        img = film.cpu().numpy().reshape(H, W, 4)
        out = {"workload": name, "msamples_per_s": total / (ms * 1e-3) / 1e6,
               "mrays_per_s": rays / steps / (ms * 1e-3) / 1e6, "rays_per_sample": rays / steps / total, "ms_per_frame": ms,
               "ms_render_rank0": gpu_ms / steps,
               "msamples_per_s_median_frame": total / (sorted(frame_ms)[len(frame_ms) // 2] * 1e-3) / 1e6,    # rank 0's median frame: the mean above carries every host hiccup of a 1-5 ms frame
               "n_gpus": world, "scaling": "strong", "frames_timed": steps, "frames_warmup": len(warm_ms), "frame_ms_rank0": frame_ms, "node_layout": "fast" if fast else "exact", "scene_build_s": build_s, "scene_synth_s": synth_s,
               "collective": "one ncclReduce(sum) of the fp64 film (4*W*H) and sum-of-squares (3*W*H) accumulators inside libptb200.so" if world > 1 else "none",
               "kernel_launches_per_frame": int(launches // steps), "mean_weight": float(img[..., 3].mean()),
               "mean_radiance": float((img[..., :3] / np.maximum(img[..., 3:4], 1e-300)).mean())}
        if world == 1:
            cspp = max(1, min(spp, 4))
            counted = scene.counted_render(cam, render_params(W, H, cspp, depth, integrator=integ, seed=3))
            out["roofline"] = sample_roofline(key, counted, ms, W, H, spp, peak, peak_src, fast=fast)
    scene.close()
    return out


def cpu_render_rate(a, key="c1"):
    """The reference's own integrator (oracle/_ref, all host threads) on the SAME scene as render config `key`,
    on a bounded sample: a reduced resolution (same camera, so the same mix of paths) and few spp."""
    from oracle import refapi
    from pathtracing_b200.scene import render_params
    name, builder, W, H, spp, depth, integ = render_configs()[key]
    w, h, s = {"c1": (256, 256, 16), "c3": (480, 270, 4), "c4": (480, 270, 4), "c5": (240, 135, 2)}[key]
    t0 = time.perf_counter()
    desc, cam = builder()
    flavour = "fast" if refapi.available("fast") else "strict"
    ref = refapi.RefScene(desc, 4, count_rays=True, flavour=flavour)
    build_s = time.perf_counter() - t0
    params = render_params(w, h, s, depth, integrator=integ)
    _, _, st = ref.render(cam, params, want_sumsq=False)
    cores = ref.threads
    ref.close()
    return {"msamples_per_s": st["samples"] / st["seconds"] / 1e6, "mrays_per_s": (st["closest_rays"] + st["shadow_rays"]) / st["seconds"] / 1e6,
            "cores": cores, "kind": "reference", "scene_build_s": build_s,
            "sample": f"{w}x{h} at {s} spp of the same scene and camera ({st['samples']} samples), fixed-spp mirror of TileIntegrator::Render, reference compiled -O3 ({flavour})"}


def dram_bound_cast(a, torch, dist, local, peak, peak_src):
    """The DRAM-bound ray cast SURVEY.md section 8d describes: 16 Mi incoherent rays against the 10 M-triangle C5 geometry, whose
    nodes + triangles (about 1 GB) cannot live in the 126 MB L2.  Algorithmic bytes from the counting build; DRAM bytes per
    ray from the tracked ncu capture of the same launch."""
    from pathtracing_b200 import capi, synth
    from pathtracing_b200.scene import GpuScene
    desc, _ = synth.c5_scene()
    n = a.rays
    out = {"workload": f"{n} random incoherent rays vs the C5 geometry (10 M unique triangles + instanced objects), BVH4 closest-hit"}
    rays = None
    for tag, layout in (("exact", capi.LAYOUT_EXACT), ("fast", capi.LAYOUT_FAST)):
        t0 = time.perf_counter()
        sc = GpuScene(desc, 4, device=local, layout=layout)
        build_s = time.perf_counter() - t0
        if rays is None:
            lo, hi = sc.bounds()
            rays = torch.from_numpy(synth.random_rays(n, lo, hi, seed=7).view(np.uint8)).cuda()
            hits = torch.empty(n * 16, dtype=torch.uint8, device="cuda")
        stream = torch.cuda.current_stream().cuda_stream
        ms = time_dev(lambda: sc.trace_closest_dev(rays.data_ptr(), n, hits.data_ptr(), None, stream), max(3, a.steps // 2), 3, torch, dist, 1)
        st = sc.trace_stats(rays.data_ptr(), n)
        s_node, s_tri = (64, 64) if layout == capi.LAYOUT_FAST else (128, 48)
        nodes = (st["tlas_nodes"] + st["blas_nodes"]) / n
        b_ray = 32 + 16 + nodes * s_node + st["triangles"] / n * s_tri + st["instances"] / n * 96
        kernel = f"k_trace<{'44' if layout == capi.LAYOUT_FAST else 4},false,false,16,NoAlpha>"
        prof = (tracked_profile("r02_kernels.json") or {}).get(f"{kernel}|c5|10M")
        traffic = prof["dram_bytes_per_ray"] * n if prof else None
        out[tag] = {"mrays_per_s": n / (ms * 1e-3) / 1e6, "ms": ms, "scene_build_s": build_s, "nodes_per_ray": nodes, "tris_per_ray": st["triangles"] / n,
                    "instances_per_ray": st["instances"] / n, "hit_rate": st["hits"] / n,
                    "roofline": {"bound": "hbm", "achieved": n * b_ray / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": n * b_ray / (ms * 1e-3) / 1e9 / peak,
                                 "traffic": traffic, "traffic_source": prof["source"] if prof else None, "bytes_per_ray": b_ray, "peak_source": peak_src,
                                 "dram_frac": (traffic / (ms * 1e-3) / 1e9 / peak) if traffic else None, "limiter": prof["limiter"] if prof else None}}
        sc.close()
    return out


def run_ours(a):
    import torch
    import torch.distributed as dist
    from pathtracing_b200 import synth
    from pathtracing_b200.scene import GpuScene

    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — pathtracing_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from pathtracing_b200 import capi
    layout = capi.LAYOUT_FAST if (a.layout == "fast" and a.bvh == 4) else capi.LAYOUT_EXACT
    peaks_file = ROOT / "MEASURED_PEAKS.json"
    if peaks_file.exists():
        peak, peak_src = json.loads(peaks_file.read_text())["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    desc = make_scene(a.mesh, a.tris)
    torch.zeros(1, device="cuda")       # the CUDA context exists before the commit is timed (0.3-0.5 s otherwise land in scene_build_s)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    scene = GpuScene(desc, a.bvh, device=local, layout=layout)
    build_s = time.perf_counter() - t0
    lo, hi = scene.bounds()
    n = a.rays
    rays_np = synth.random_rays(n, lo, hi, seed=7 + rank)   # weak scaling: every rank traces its own n rays
    host_rays = torch.empty(n * 32, dtype=torch.uint8).pin_memory()
    host_rays.numpy()[:] = rays_np.view(np.uint8)
    host_hits = torch.empty(n * 16, dtype=torch.uint8).pin_memory()
    d_rays = host_rays.cuda(non_blocking=True)
    d_hits = torch.empty(n * 16, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream

    def step():
        scene.trace_closest_dev(d_rays.data_ptr(), n, d_hits.data_ptr(), None, stream)

    sampler = ClockSampler(local) if rank == 0 else None
    ms = time_dev(step, a.steps, a.warmup, torch, dist, world)
    clocks = sampler.stop() if sampler else None
    value = world * n / (ms * 1e-3) / 1e6

    # end to end through the host-buffer C-ABI call (H2D + kernel + D2H inside the timed region)
    lib = scene.lib
    hp, hh = host_rays.data_ptr(), host_hits.data_ptr()

    def e2e_step():
        rc = lib.pt_trace_closest(scene.h, hp, n, hh, None)
        assert rc == 0, lib.pt_last_error()

    for _ in range(max(1, min(a.warmup, 2))):
        e2e_step()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    e2e_steps = max(2, min(a.steps, 5))
    for _ in range(e2e_steps):
        e2e_step()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
    if world > 1:
        t = torch.tensor([e2e_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    e2e_value = world * n / (e2e_ms * 1e-3) / 1e6
    # the same rays through the compact transfer form (pt_trace_closest_compact: 24 B in, 8 B out per ray; every C2 ray has
    # tmax = inf and time = 0).  Reported next to the full-format number, which stays the headline.
    host_rays24 = torch.empty(n * 24, dtype=torch.uint8).pin_memory()
    r24 = host_rays24.numpy().view(capi.RAY24_DTYPE)
    r24["o"], r24["d"] = rays_np["o"], rays_np["d"]
    host_hits8 = torch.empty(n * 8, dtype=torch.uint8).pin_memory()
    hp24, hh8 = host_rays24.data_ptr(), host_hits8.data_ptr()

    def compact_step():
        rc = lib.pt_trace_closest_compact(scene.h, hp24, n, hh8, None)
        assert rc == 0, lib.pt_last_error()

    for _ in range(max(1, min(a.warmup, 2))):
        compact_step()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        compact_step()
    compact_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
    if world > 1:
        t = torch.tensor([compact_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        compact_ms = float(t.item())
    compact_value = world * n / (compact_ms * 1e-3) / 1e6
    h8 = host_hits8.numpy().view(capi.HIT8_DTYPE)
    h16 = host_hits.numpy().view(capi.HIT_DTYPE)
    compact_equal = bool(np.array_equal(h8["t"].view(np.uint32), h16["t"].view(np.uint32)) and np.array_equal(h8["prim"], h16["prim"]))
    del host_rays24, host_hits8, r24, h8, h16
    # what the PCIe link of this box gives the same pinned buffers (plain copies, CUDA events): the floor of e2e
    pcie = {}
    for name, dst, src in (("h2d", d_rays, host_rays), ("d2h", host_hits, d_hits)):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(3):
            dst.copy_(src, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        pcie[name] = 3 * src.numel() * src.element_size() / (e0.elapsed_time(e1) * 1e-3) / 1e9
    e2e_floor_ms = max(n * 32 / pcie["h2d"], n * 16 / pcie["d2h"]) / 1e6
    # what the host memory system gives a rank while ALL ranks use it at once (one NUMA node feeds every GPU's copy engines):
    # a plain host-to-host copy of the ray buffer, every rank at the same time.  The e2e figure at N >= 4 is bound by this, not
    # by a kernel or a link (per-rank PCIe rates fall with N, SCALE_r01.json).
    scratch = np.empty(n * 32, np.uint8)
    src_np = host_rays.numpy()
    np.copyto(scratch, src_np)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(2):
        np.copyto(scratch, src_np)
    host_copy_gbs = 2 * 2 * n * 32 / (time.perf_counter() - t0) / 1e9       # bytes read + written
    del scratch
    if world > 1:
        t = torch.tensor([host_copy_gbs, pcie["h2d"], pcie["d2h"]], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        host_copy_gbs, pcie_min = float(t[0].item()), (float(t[1].item()), float(t[2].item()))
    else:
        pcie_min = (pcie["h2d"], pcie["d2h"])
    # the result really came back: hit count of the host buffer
    hits_np = host_hits.numpy().view(np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("prim", "<u4")]))
    hit_rate = float((hits_np["prim"] != 0xFFFFFFFF).mean())

    render_extras = {}
    if not a.no_extras:
        for key in [k for k in a.renders.split(",") if k]:
            render_extras[key] = bench_render(a, key, torch, dist, world, rank, local, peak, peak_src)   # collective: every rank takes part

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # roofline of the dominant (only) kernel: algorithmic bytes per ray from a counting build of the
    # same traversal: 32 B ray + 16 B hit + nodes*S_node + tris*S_tri + instances*96
    st = scene.trace_stats(d_rays.data_ptr(), n)
    fast = layout == capi.LAYOUT_FAST
    s_node = 64 if fast else {2: 64, 4: 128, 8: 256}[a.bvh]
    s_tri = 64 if fast else 48
    nodes_per_ray = (st["tlas_nodes"] + st["blas_nodes"]) / n
    tris_per_ray = st["triangles"] / n
    inst_per_ray = st["instances"] / n
    b_ray = 32 + 16 + nodes_per_ray * s_node + tris_per_ray * s_tri + inst_per_ray * 96
    achieved = n * b_ray / (ms * 1e-3) / 1e9
    # the soup and the surface mesh are ONE untransformed BLAS: the kernel variant without instance code (trace.cuh NoAlphaSingle)
    kernel = f"k_trace<{'44' if fast else a.bvh},false,false,16,NoAlphaSingle>"
    # what ncu measured for this kernel on this workload (dram__bytes per ray, L1 wavefront and issue utilisation): read from
    # the tracked summary of the capture, never typed in here; absent for workloads that were not captured
    prof = (tracked_profile("r02_kernels.json") or {}).get(f"{kernel}|{a.mesh}|{a.tris}")
    traffic = prof["dram_bytes_per_ray"] * n if prof else None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_source": prof["source"] if prof else None,
                "algorithmic_bytes": n * b_ray,
                "peak_source": peak_src, "kernel": kernel, "bytes_per_ray": b_ray, "node_bytes": s_node, "triangle_bytes": s_tri,
                "nodes_per_ray": nodes_per_ray, "tris_per_ray": tris_per_ray, "instances_per_ray": inst_per_ray,
                "limiter": prof["limiter"] if prof else None,
                "dram_frac": (traffic / (ms * 1e-3) / 1e9 / peak) if traffic else None,
                "l1_wavefront_frac": prof.get("l1_wavefront_frac") if prof else None, "issue_active_frac": prof.get("issue_active_frac") if prof else None,
                "note": "'bound' names the contract's roofline (algorithmic bytes / HBM peak).  What ncu shows binds the kernel is in 'limiter': the "
                        "87 MB scene is L2-resident at this size, DRAM traffic is far BELOW the algorithmic bytes (dram_frac), the kernel runs against "
                        "instruction issue and L1 wavefronts (DESIGN.md 4.1).  The DRAM-bound case is extras.dram_bound_cast."}

    extras = {}
    if not a.no_extras:
        d_occ = torch.empty(n, dtype=torch.uint8, device="cuda")
        ms_any = time_dev(lambda: scene.trace_any_dev(d_rays.data_ptr(), n, d_occ.data_ptr(), stream), max(2, a.steps // 2), 2, torch, dist, 1)
        extras[f"any_hit_bvh{a.bvh}_mrays"] = n / (ms_any * 1e-3) / 1e6
        if a.bvh == 4:
            # the same tree in the other node layout
            so = GpuScene(desc, 4, device=local, layout=capi.LAYOUT_EXACT if fast else capi.LAYOUT_FAST)
            tag = "exact" if fast else "fast"
            mso = time_dev(lambda: so.trace_closest_dev(d_rays.data_ptr(), n, d_hits.data_ptr(), None, stream), max(2, a.steps // 2), 2, torch, dist, 1)
            extras[f"closest_bvh4_{tag}_layout_mrays"] = n / (mso * 1e-3) / 1e6
            msa = time_dev(lambda: so.trace_any_dev(d_rays.data_ptr(), n, d_occ.data_ptr(), stream), max(2, a.steps // 2), 2, torch, dist, 1)
            extras[f"any_hit_bvh4_{tag}_layout_mrays"] = n / (msa * 1e-3) / 1e6
            sto = so.trace_stats(d_rays.data_ptr(), n)
            extras[f"bvh4_{tag}_layout_nodes_per_ray"] = (sto["tlas_nodes"] + sto["blas_nodes"]) / n
            so.close()
        other = 2 if a.bvh == 4 else 4
        s2 = GpuScene(desc, other, device=local)
        ms2 = time_dev(lambda: s2.trace_closest_dev(d_rays.data_ptr(), n, d_hits.data_ptr(), None, stream), max(2, a.steps // 2), 2, torch, dist, 1)
        extras[f"closest_bvh{other}_mrays"] = n / (ms2 * 1e-3) / 1e6
        s2.close()
        s8 = GpuScene(desc, 8, device=local)
        ms8 = time_dev(lambda: s8.trace_closest_dev(d_rays.data_ptr(), n, d_hits.data_ptr(), None, stream), max(2, a.steps // 2), 2, torch, dist, 1)
        extras["closest_bvh8_mrays"] = n / (ms8 * 1e-3) / 1e6
        st8 = s8.trace_stats(d_rays.data_ptr(), n)
        extras["bvh8_nodes_per_ray"] = (st8["tlas_nodes"] + st8["blas_nodes"]) / n
        s8.close()

    if not a.no_extras and not a.no_dram_cast:
        extras["dram_bound_cast"] = dram_bound_cast(a, torch, dist, local, peak, peak_src)
    for key, val in render_extras.items():
        extras[f"{key}_render"] = val
    for key in render_extras:
        if world == 1 and not a.no_cpu:
            try:
                extras[f"{key}_render"]["cpu_reference"] = cpu_render_rate(a, key)
            except Exception as e:
                extras[f"{key}_render"]["cpu_reference"] = {"unavailable": str(e)}

    cpu = None
    if world == 1 and not a.no_cpu:
        try:
            cpu, _ = cpu_reference_rate(a, desc, rays_np[: 1 << 22])
        except Exception as e:  # the oracle is a reported baseline, never a dependency of the product path
            cpu = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference", "sample": f"unavailable: {e}"}

    full = int(os.environ.get("PTB_CHUNK", 2 << 20))
    chunks, off = 0, 0
    while off < n:         # csrc/capi.cu trace_pipelined: 1/8, 1/2, then full chunks, the last one split so that a quarter chunk ends the step
        m = max(1024, full // 8 if chunks == 0 else full // 2 if chunks == 1 else full)
        left = n - off
        if m >= left:
            m = left - full // 4 if left > full // 2 else left
        off += m
        chunks += 1
    streamed = n >= (1 << 20) and os.environ.get("PTB_E2E_STREAMED", "1") != "0"
    e2e_api = ("pt_trace_closest (pinned host buffers; ONE persistent kernel fed by the H2D stream through a watermark, results copied out per finished 256 Ki-ray granule)"
               if streamed else f"pt_trace_closest (pinned host buffers, {chunks} chunks through an H2D / 2 x trace / D2H stream pipeline)")
    if streamed:
        chunks = 1
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": config_of(a),
            "details": {"rays_per_gpu": n, "triangles": a.tris, "mesh": a.mesh, "node_layout": a.layout if a.bvh == 4 else "exact",
                        "parallelism": f"scene replicated, rays sharded x{world}, no collective",
                        "l2": "ray+hit buffers (768 MiB per step) exceed the 126 MB L2; the scene (nodes + triangles) stays resident by design",
                        "hit_rate": hit_rate, "scene_build_s": build_s, "stack_overflows": scene.stack_overflows()},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n * 32, "d2h_bytes_per_step": n * 16,
                    "ms_per_step": e2e_ms, "api": e2e_api,
                    "pcie_h2d_gbs": pcie["h2d"], "pcie_d2h_gbs": pcie["d2h"], "pcie_floor_ms": e2e_floor_ms,
                    "pcie_h2d_gbs_min_over_ranks": pcie_min[0], "pcie_d2h_gbs_min_over_ranks": pcie_min[1],
                    "host_copy_gbs_per_rank_all_ranks_at_once": host_copy_gbs,
                    "limiter": ("the kernel (PCIe floor of the same buffers is below the step time)" if world == 1 else
                                "host memory: every rank moves 48 B per ray through one NUMA node's DRAM; per-rank pinned-copy and host-copy rates fall with the rank count"),
                    "note": "floor = the slower of the two plain pinned copies of one step's buffers on this box",
                    "compact": {"value": compact_value, "unit": UNIT, "ms_per_step": compact_ms, "h2d_bytes_per_step": n * 24, "d2h_bytes_per_step": n * 8,
                                "api": "pt_trace_closest_compact (pt_ray24 in, pt_hit8 out; same kernel family, same pipeline)",
                                "same_t_bits_and_ids_as_the_full_form": compact_equal}},
            "gpu_launches": a.steps, "gpu_launches_e2e": e2e_steps * chunks,
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "extras": extras}
    print(json.dumps(line), flush=True)
    scene.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
