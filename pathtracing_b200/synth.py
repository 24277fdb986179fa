"""Seeded synthetic inputs for BASELINE.json's configs.

Everything is generated with numpy's PCG64 from a fixed seed, in float32, so the
reference harness, the oracle and the GPU see byte-identical arrays.
"""
from __future__ import annotations

import math

import numpy as np

from . import capi
from .scene import SceneDesc, camera_desc, render_params


# ---- meshes -------------------------------------------------------------------------

def soup(n_tris: int, seed: int = 42, edge: float = 0.01):
    """C2 'soup': triangle centres uniform in [0,1]^3, vertices = centre + edge*U(-.5,.5)^3."""
    rng = np.random.default_rng(seed)
    c = rng.random((n_tris, 1, 3), dtype=np.float32)
    off = (rng.random((n_tris, 3, 3), dtype=np.float32) - np.float32(0.5)) * np.float32(edge)
    pos = (c + off).reshape(-1, 3).astype(np.float32)
    idx = np.arange(3 * n_tris, dtype=np.uint32).reshape(-1, 3)
    return pos, idx


def terrain(nx: int, nz: int, seed: int = 42, height: float = 0.15, octaves: int = 5):
    """C2 'surface': (nx x nz) displaced grid over [0,1]^2 -> 2*(nx-1)*(nz-1) triangles."""
    rng = np.random.default_rng(seed)
    x = np.linspace(0, 1, nx, dtype=np.float32)
    z = np.linspace(0, 1, nz, dtype=np.float32)
    X, Z = np.meshgrid(x, z, indexing="xy")
    Y = np.zeros_like(X)
    amp, freq = 1.0, 2.0
    for _ in range(octaves):
        ph = rng.random(4) * 2 * math.pi
        kx, kz = rng.normal(size=2) * freq, rng.normal(size=2) * freq
        Y += amp * (np.sin(2 * math.pi * (kx[0] * X + kz[0] * Z) + ph[0]) * np.cos(2 * math.pi * (kx[1] * X - kz[1] * Z) + ph[1]))
        amp *= 0.5
        freq *= 2.0
    Y = (0.5 + height * Y / 2.0).astype(np.float32)
    pos = np.stack([X, Y, Z], axis=-1).reshape(-1, 3).astype(np.float32)
    i = (np.arange(nz - 1)[:, None] * nx + np.arange(nx - 1)[None, :]).reshape(-1).astype(np.uint32)
    tri = np.stack([i, i + nx, i + 1, i + 1, i + nx, i + nx + 1], axis=-1).reshape(-1, 3)
    uv = np.stack([X, Z], axis=-1).reshape(-1, 2).astype(np.float32)
    return pos, tri.astype(np.uint32), uv


def icosphere(subdiv: int, radius: float = 1.0):
    t = (1.0 + 5 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], dtype=np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6],
                  [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10],
                  [8, 6, 7], [9, 8, 1]], dtype=np.int64)
    for _ in range(subdiv):
        edges = np.sort(np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]]), axis=1)
        uniq, inv = np.unique(edges, axis=0, return_inverse=True)
        mid = v[uniq[:, 0]] + v[uniq[:, 1]]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        base = len(v)
        v = np.concatenate([v, mid])
        n = len(f)
        a, b, c = base + inv[:n], base + inv[n:2 * n], base + inv[2 * n:]
        f = np.concatenate([np.stack([f[:, 0], a, c], 1), np.stack([f[:, 1], b, a], 1), np.stack([f[:, 2], c, b], 1),
                            np.stack([a, b, c], 1)])
    pos = (v * radius).astype(np.float32)
    nrm = v.astype(np.float32)
    uv = np.stack([np.arctan2(v[:, 2], v[:, 0]) / (2 * math.pi) + 0.5, np.arccos(np.clip(v[:, 1], -1, 1)) / math.pi], 1).astype(np.float32)
    return pos, f.astype(np.uint32), nrm, uv


def quad(origin, u, v):
    """Two triangles spanning origin + s*u + t*v (a parallelogram, so the reference's
    unfolded triangle sampling, Shape.cpp:277-297, stays unbiased for the pair)."""
    o, u, v = (np.asarray(a, np.float32) for a in (origin, u, v))
    pos = np.stack([o, o + u, o + u + v, o + v]).astype(np.float32)
    idx = np.array([[0, 1, 2], [0, 2, 3]], np.uint32)
    n = np.cross(u, v)
    n = n / np.linalg.norm(n)
    nrm = np.tile(n.astype(np.float32), (4, 1))
    uv = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32)
    return pos, idx, nrm, uv


def box(lo, hi, rot_y=0.0):
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    c = (lo + hi) / 2
    faces = []
    for axis in range(3):
        for s in (0, 1):
            a1, a2 = (axis + 1) % 3, (axis + 2) % 3
            o = lo.copy()
            o[axis] = hi[axis] if s else lo[axis]
            uu, vv = np.zeros(3, np.float32), np.zeros(3, np.float32)
            uu[a1] = hi[a1] - lo[a1]
            vv[a2] = hi[a2] - lo[a2]
            faces.append(quad(o, uu, vv) if s else quad(o, vv, uu))
    pos = np.concatenate([f[0] for f in faces])
    idx = np.concatenate([f[1] + 4 * k for k, f in enumerate(faces)]).astype(np.uint32)
    nrm = np.concatenate([f[2] for f in faces])
    uv = np.concatenate([f[3] for f in faces])
    if rot_y:
        cs, sn = math.cos(rot_y), math.sin(rot_y)
        R = np.array([[cs, 0, sn], [0, 1, 0], [-sn, 0, cs]], np.float32)
        pos = ((pos - c) @ R.T + c).astype(np.float32)
        nrm = (nrm @ R.T).astype(np.float32)
    return pos, idx, nrm, uv


# ---- rays ---------------------------------------------------------------------------

def random_rays(n: int, lo, hi, seed: int = 7, tmax: float = np.inf):
    """C2 rays: origin uniform in the box [lo,hi], direction uniform on S^2."""
    rng = np.random.default_rng(seed)
    rays = np.zeros(n, dtype=capi.RAY_DTYPE)
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    rays["o"] = lo + rng.random((n, 3), dtype=np.float32) * (hi - lo)
    z = np.float32(1) - np.float32(2) * rng.random(n, dtype=np.float32)
    phi = np.float32(2 * math.pi) * rng.random(n, dtype=np.float32)
    r = np.sqrt(np.maximum(np.float32(0), np.float32(1) - z * z))
    rays["d"] = np.stack([r * np.cos(phi), r * np.sin(phi), z], axis=-1).astype(np.float32)
    rays["tmax"] = np.float32(tmax)
    return rays


# ---- whole scenes ---------------------------------------------------------------------

def soup_scene(n_tris: int, seed: int = 42, opaque: bool = False):
    """C2: one BLAS, one untransformed instance, default MicrofacetDiffuse material."""
    d = SceneDesc()
    t = d.solid((0.7, 0.7, 0.7))
    m = d.diffuse(t, alpha_mode=capi.ALPHA_OPAQUE if opaque else capi.ALPHA_BLEND)
    pos, idx = soup(n_tris, seed)
    me = d.mesh(pos, idx, material=m)
    b = d.add_blas([me])
    d.instance(b)
    return d


def terrain_scene(nx: int, nz: int, seed: int = 42, opaque: bool = False):
    d = SceneDesc()
    t = d.solid((0.7, 0.7, 0.7))
    m = d.diffuse(t, alpha_mode=capi.ALPHA_OPAQUE if opaque else capi.ALPHA_BLEND)
    pos, idx, uv = terrain(nx, nz, seed)
    me = d.mesh(pos, idx, texcoords=uv, material=m)
    b = d.add_blas([me])
    d.instance(b)
    return d


def rot_trs(rng, scale_range=(0.5, 1.5), extent=4.0):
    """Random translate * rotate * uniform scale (row-major 4x4)."""
    axis = rng.normal(size=3)
    axis /= np.linalg.norm(axis)
    ang = rng.random() * 2 * math.pi
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    R = np.eye(3) + math.sin(ang) * K + (1 - math.cos(ang)) * (K @ K)
    s = scale_range[0] + rng.random() * (scale_range[1] - scale_range[0])
    M = np.eye(4)
    M[:3, :3] = R * s
    M[:3, 3] = (rng.random(3) - 0.5) * 2 * extent
    return M.astype(np.float32)


def instanced_scene(subdiv: int = 3, n_instances: int = 24, seed: int = 3, with_untransformed: bool = True):
    """Small C3-style scene: one icosphere BLAS + one box BLAS, many TransformedPrimitives."""
    rng = np.random.default_rng(seed)
    d = SceneDesc()
    t = d.solid((0.6, 0.6, 0.6))
    m = d.diffuse(t)
    pos, idx, nrm, uv = icosphere(subdiv)
    b0 = d.add_blas([d.mesh(pos, idx, nrm, uv, material=m)])
    bp, bi, bn, bu = box((-0.5, -0.5, -0.5), (0.5, 0.5, 0.5))
    qp, qi, qn, qu = quad((-0.8, 0.9, -0.8), (1.6, 0, 0), (0, 0, 1.6))
    b1 = d.add_blas([d.mesh(bp, bi, bn, bu, material=m), d.mesh(qp, qi, qn, qu, material=m)])
    if with_untransformed:
        d.instance(b0)
    for k in range(n_instances):
        d.instance(b0 if k % 2 == 0 else b1, rot_trs(rng))
    return d


def cornell_scene(light_scale: float = 15.0, light_size: float = 0.3):
    """C1: procedural Cornell box as triangle meshes, Lambertian (MicrofacetDiffuse,
    roughness 1) walls + a two-triangle parallelogram area light."""
    d = SceneDesc()
    white = d.diffuse(d.solid((0.73, 0.73, 0.73)))
    red = d.diffuse(d.solid((0.65, 0.05, 0.05)))
    green = d.diffuse(d.solid((0.12, 0.45, 0.15)))
    black = d.diffuse(d.solid((0.0, 0.0, 0.0)))
    emit = d.solid((1.0, 0.85, 0.6), scale=(light_scale,) * 3)
    meshes = []

    def add(q, mat, emissive=-1):
        meshes.append(d.mesh(q[0], q[1], q[2], q[3], material=mat, emissive_tex=emissive))

    add(quad((0, 0, 0), (0, 0, 1), (1, 0, 0)), white)        # floor (normal +y)
    add(quad((0, 1, 0), (1, 0, 0), (0, 0, 1)), white)        # ceiling (normal -y)
    add(quad((0, 0, 0), (1, 0, 0), (0, 1, 0)), white)        # back wall z=0
    add(quad((0, 0, 0), (0, 1, 0), (0, 0, 1)), red)          # left x=0
    add(quad((1, 0, 0), (0, 0, 1), (0, 1, 0)), green)        # right x=1
    add(box((0.53, 0.0, 0.09), (0.83, 0.3, 0.39), rot_y=-0.3), white)
    add(box((0.15, 0.0, 0.45), (0.45, 0.6, 0.75), rot_y=0.35), white)
    lo = 0.5 - light_size / 2
    add(quad((lo, 0.998, lo), (light_size, 0, 0), (0, 0, light_size)), black, emissive=emit)
    b = d.add_blas(meshes)
    d.instance(b)
    cam = camera_desc((0.5, 0.5, 2.4), (0.5, 0.5, 0.0), fov=2 * math.atan(0.5 / 1.4) * 1.0)
    return d, cam
