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


def deep_chain_scene(n: int = 230, ratio: float = 2.0 ** -0.25):
    """n self-similar triangles: triangle k lies in the plane x = ratio^k and spans about ratio^k in y and z.  Binned SAH peels the big outer
    ones off a shrinking inner group, so the tree is a chain tens of clusters deep, and a ray along +x through the origin finds every
    peeled child farther than the inner group it descends into: all of them wait on the traversal stack before the first triangle is
    tested.  n = 230 needs more than the 96 stack rows of k_trace (PT_ERR_OVERFLOW); n = 60 does not."""
    d = SceneDesc()
    m = d.diffuse(d.solid((0.7, 0.7, 0.7)), alpha_mode=capi.ALPHA_OPAQUE)
    sc = (ratio ** np.arange(n)).astype(np.float32)
    base = np.array([[1, -1, -1], [1, 2, -1], [1, -1, 2]], np.float32)
    pos = sc[:, None, None] * base[None]
    me = d.mesh(pos.reshape(-1, 3), np.arange(3 * n, dtype=np.uint32), material=m)
    d.instance(d.add_blas([me]))
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


def trs(translate=(0, 0, 0), scale=1.0, rot_axis=(0, 1, 0), rot_angle=0.0):
    axis = np.asarray(rot_axis, np.float64)
    axis = axis / np.linalg.norm(axis)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    R = np.eye(3) + math.sin(rot_angle) * K + (1 - math.cos(rot_angle)) * (K @ K)
    M = np.eye(4)
    M[:3, :3] = R * scale
    M[:3, 3] = translate
    return M.astype(np.float32)


def procedural_textures(size=64, seed=11):
    """Generated albedo / roughness-metallic / normal-map images (u8, 3 channels), the 'procedural
    textures' of C4/C5: the reference can only load image files, so its harness writes these very
    pixels to PNG and loads them back (oracle/ref_driver.cpp)."""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:size, 0:size].astype(np.float32) / size
    albedo = np.stack([0.55 + 0.4 * np.sin(6.28 * 3 * x) * np.cos(6.28 * 2 * y), 0.5 + 0.35 * np.cos(6.28 * 2 * x + 1.0),
                       0.45 + 0.3 * np.sin(6.28 * (x + y) * 2)], -1)
    albedo = np.clip(albedo + 0.05 * rng.standard_normal(albedo.shape), 0.05, 0.95)
    rough = np.clip(0.25 + 0.5 * (0.5 + 0.5 * np.sin(6.28 * 4 * x) * np.sin(6.28 * 4 * y)), 0.05, 1.0)
    metal = (np.sin(6.28 * 2 * (x - y)) > 0.3).astype(np.float32)
    orm = np.stack([np.ones_like(rough), rough, metal], -1)           # roughness in .g, metallic in .b
    nx, ny = 0.35 * np.sin(6.28 * 5 * x), 0.35 * np.cos(6.28 * 5 * y)
    nz = np.sqrt(np.maximum(1 - nx * nx - ny * ny, 0.05))
    nm = np.stack([nx, ny, nz], -1) * 0.5 + 0.5
    to8 = lambda a: np.ascontiguousarray(np.clip(np.round(a * 255), 0, 255).astype(np.uint8))
    return to8(albedo), to8(orm), to8(nm)


def baked_gradient_table(horizon, zenith, scale, w=1024, h=512):
    """The gradient sky evaluated the way the host API bakes ANY FunctionInfiniteLight sky function (host/PathTracing.hpp):
    a lat-long float table at the centres of SphereShape::GetSphereUV cells (u = phi / 2 pi, v = theta / pi, y up)."""
    v = (np.arange(h, dtype=np.float64) + 0.5) / h
    a = 0.5 * (np.cos(np.pi * v) + 1.0)
    col = scale * ((1.0 - a)[:, None] * np.asarray(horizon, np.float64)[None] + a[:, None] * np.asarray(zenith, np.float64)[None])
    return np.ascontiguousarray(np.broadcast_to(col[:, None, :], (h, w, 3)).astype(np.float32))


def materials_scene(subdiv=3, sky=True, point_light=False, distant_light=False, baked_sky=False):
    """C3-style: instanced icospheres (TransformedPrimitive) with the four README materials +
    ThinDielectric on a textured ground, a transformed emissive quad (TransformedLight) and a gradient sky."""
    d = SceneDesc()
    alb, orm, nm = procedural_textures()
    t_alb, t_orm, t_nm = d.image_u8(alb), d.image_u8(orm), d.image_u8(nm)
    white = d.solid((0.8, 0.8, 0.8))
    ground_tex = d.checker(d.solid((0.75, 0.75, 0.75)), d.solid((0.2, 0.25, 0.3)), (0.125, 0.125))
    mats = [
        d.dielectric(1.5, d.solid((0.95, 0.97, 1.0)), rough=d.solid((0.15, 0.15, 0.15))),     # rough glass
        d.diffuse(d.solid((0.95, 0.64, 0.54)), rough=d.solid((0.3, 0.3, 0.3)), metal=d.solid((1, 1, 1))),   # rough copper
        d.diffuse(d.solid((0.2, 0.5, 0.8))),                                                   # Lambertian
        d.conductor((0.9, 0.85, 0.6)),                                                         # mirror
        d.dielectric(1.5, white),                                                              # smooth glass
        d.thin_dielectric(1.5, d.solid((0.9, 1.0, 0.9))),
        d.diffuse(t_alb, norm=-1, rough=t_orm, metal=t_orm),                                   # textured metal/rough
    ]
    pos, idx, nrm, uv = icosphere(subdiv, 1.0)
    blas = [d.add_blas([d.mesh(pos, idx, nrm, uv, material=m)]) for m in mats]
    rng = np.random.default_rng(5)
    for k, b in enumerate(blas):
        ang = 2 * math.pi * k / len(blas)
        d.instance(b, trs((1.9 * math.cos(ang), 0.5 + 0.1 * (k % 2), 1.9 * math.sin(ang)), 0.5 + 0.05 * (k % 3),
                          rng.normal(size=3), rng.random() * 6.28))
    # ground with a normal map and explicit tangents, untransformed
    gp, gi, gn, gu = quad((-4, 0, -4), (0, 0, 8), (8, 0, 0))
    gt = np.tile(np.array([[0, 0, 1]], np.float32), (4, 1))
    gmat = d.diffuse(ground_tex, norm=t_nm, rough=d.solid((0.6, 0.6, 0.6)))
    d.instance(d.add_blas([d.mesh(gp, gi, gn, gu * 4.0, tangents=gt, material=gmat)]))
    # emissive quad inside a rotated + scaled instance
    lp, li, ln, lu = quad((-0.5, 0, -0.5), (1, 0, 0), (0, 0, 1))
    emit = d.solid((1.0, 0.9, 0.75), scale=(12, 12, 12))
    lb = d.add_blas([d.mesh(lp, li, ln, lu, material=d.diffuse(d.solid((0, 0, 0))), emissive_tex=emit)])
    d.instance(lb, trs((0.3, 2.6, 0.2), 1.4, (1, 0, 0.3), 0.5))
    if sky:
        # baked_sky: the device reads the sky from the baked table (tex >= 0); the colours stay in the descriptor for the reference,
        # whose FunctionInfiniteLight evaluates the same function in closed form
        tex = d.image_f32(baked_gradient_table((1.0, 1.0, 1.0), (0.5, 0.7, 1.0), 0.6)) if baked_sky else -1
        d.light(capi.LIGHT_GRADIENT_INFINITE, color=(1.0, 1.0, 1.0), color_b=(0.5, 0.7, 1.0), scale=0.6, tex=tex)
    if point_light:
        d.light(capi.LIGHT_POINT, color=(6, 5, 4), v=(-1.5, 2.0, 1.0), power_func=1)
    if distant_light:
        d.light(capi.LIGHT_DISTANT, color=(1.5, 1.4, 1.2), v=(0.4, 1.0, 0.3))
    cam = camera_desc((0.0, 2.2, 5.2), (0.0, 0.5, 0.0), fov=0.9)
    return d, cam


def medium_scene(subdiv=2, g=0.75, density=0.1):
    """C4-style: the materials scene inside a global HomogeneusMedium (Scene::SetMedium), rendered
    with VolPathIntegrator: NEE + MIS through the medium."""
    d, cam = materials_scene(subdiv, sky=True)
    d.scene_medium = d.medium((0.5, 0.5, 0.5), (2.0, 2.2, 2.6), g, density=density)
    return d, cam


def alpha_scene(mode=capi.ALPHA_MASK, four_channel=False):
    """Cornell box with alpha-tested cut-out cards (Material::Alpha inside traversal, Primitive.cpp:6-26):
    one untransformed card and one inside a rotated instance.  mode = Mask (deterministic) or Blend
    (stochastic).  four_channel: the opacity comes from the albedo image's 4th channel
    (Texture::alpha) instead of a separate alpha texture."""
    d, cam = cornell_scene()
    n = 32
    y, x = np.mgrid[0:n, 0:n].astype(np.float32) / n
    opacity = np.clip(0.5 + 0.5 * np.sin(6.28 * 2 * x) * np.cos(6.28 * 2 * y), 0.0, 1.0)
    rgb = np.stack([0.2 + 0.6 * x, 0.7 - 0.4 * y, 0.3 + 0 * x], -1)
    to8 = lambda a: np.ascontiguousarray(np.clip(np.round(a * 255), 0, 255).astype(np.uint8))
    if four_channel:
        tex = d.image_u8(to8(np.concatenate([rgb, opacity[..., None]], -1)))
        mat = d.diffuse(tex, alpha_mode=mode, alpha_cutoff=0.5)
    else:
        tex = d.image_u8(to8(rgb))
        a_tex = d.image_u8(to8(np.repeat(opacity[..., None], 3, -1)))
        mat = d.diffuse(tex, alpha=a_tex, alpha_mode=mode, alpha_cutoff=0.5)
    glass = d.dielectric(1.5, d.solid((1.0, 1.0, 1.0)), alpha=d.checker(d.solid((1, 1, 1)), d.solid((0.25, 0.25, 0.25)), (0.25, 0.25)),
                         alpha_mode=mode)
    q = quad((0.1, 0.15, 0.9), (0.45, 0, 0), (0, 0.6, 0))
    d.instance(d.add_blas([d.mesh(q[0], q[1], q[2], q[3], material=mat)]))
    q2 = quad((-0.5, -0.5, 0.0), (1, 0, 0), (0, 1, 0))
    b2 = d.add_blas([d.mesh(q2[0], q2[1], q2[2], q2[3], material=glass)])
    d.instance(b2, trs((0.72, 0.5, 1.1), 0.4, (0, 1, 0), -0.5))
    return d, cam


# ---- BASELINE configs[2..4] at full scale ----------------------------------------------------

def displaced_sphere(subdiv: int, seed: int, amp: float = 0.08):
    """An icosphere with smooth radial noise and recomputed vertex normals: a 'model' BLAS."""
    pos, idx, nrm, uv = icosphere(subdiv)
    rng = np.random.default_rng(seed)
    k = rng.normal(size=(4, 3)) * 3.0
    ph = rng.random(4) * 6.28
    r = 1.0 + amp * sum(np.sin(pos @ k[j] + ph[j]) for j in range(4)) / 2.0
    pos = (pos * r[:, None]).astype(np.float32)
    fn = np.cross(pos[idx[:, 1]] - pos[idx[:, 0]], pos[idx[:, 2]] - pos[idx[:, 0]])
    vn = np.zeros_like(pos, dtype=np.float64)
    for c in range(3):
        np.add.at(vn, idx[:, c], fn)
    vn /= np.maximum(np.linalg.norm(vn, axis=1, keepdims=True), 1e-20)
    return pos, idx, vn.astype(np.float32), uv


def c3_scene(subdiv: int = 6, n_instances: int = 122, seed: int = 3):
    """C3: instanced scene, ~10 M triangles through the TLAS (122 x 81 920-triangle BLASes at the
    defaults), materials round-robin {rough dielectric, rough metal, Lambertian, specular conductor},
    checkered ground, gradient sky + a quad light."""
    d = SceneDesc()
    mats = [d.dielectric(1.5, d.solid((0.95, 0.97, 1.0)), rough=d.solid((0.15, 0.15, 0.15))),
            d.diffuse(d.solid((0.95, 0.64, 0.54)), rough=d.solid((0.3, 0.3, 0.3)), metal=d.solid((1, 1, 1))),
            d.diffuse(d.solid((0.2, 0.5, 0.8))),
            d.conductor((0.9, 0.85, 0.6))]
    blas = []
    for k, m in enumerate(mats):
        pos, idx, nrm, uv = displaced_sphere(subdiv, seed + k)
        blas.append(d.add_blas([d.mesh(pos, idx, nrm, uv, material=m)]))
    rng = np.random.default_rng(seed)
    side = int(math.ceil(math.sqrt(n_instances)))
    for k in range(n_instances):
        gx, gz = k % side, k // side
        M = trs(((gx - side / 2 + 0.5) * 2.6 + rng.normal() * 0.2, 1.0 + rng.random() * 1.5, (gz - side / 2 + 0.5) * 2.6 + rng.normal() * 0.2),
                0.8 + 0.4 * rng.random(), rng.normal(size=3), rng.random() * 6.28)
        d.instance(blas[k % len(blas)], M)
    ext = side * 2.6
    gp, gi, gn, gu = quad((-ext, 0, -ext), (0, 0, 2 * ext), (2 * ext, 0, 0))
    ground = d.diffuse(d.checker(d.solid((0.75, 0.75, 0.75)), d.solid((0.2, 0.25, 0.3)), (0.02, 0.02)))
    d.instance(d.add_blas([d.mesh(gp, gi, gn, gu, material=ground)]))
    lp, li, ln, lu = quad((-3, 0, -3), (6, 0, 0), (0, 0, 6))
    emit = d.solid((1.0, 0.9, 0.75), scale=(20, 20, 20))
    lb = d.add_blas([d.mesh(lp, li, ln, lu, material=d.diffuse(d.solid((0, 0, 0))), emissive_tex=emit)])
    d.instance(lb, trs((0.0, 9.0, 0.0), 1.0, (1, 0, 0), math.pi))       # facing down
    d.light(capi.LIGHT_GRADIENT_INFINITE, color=(1.0, 0.85, 0.55), color_b=(0.45, 0.65, 1.0), scale=1.5)
    cam = camera_desc((0.0, 7.0, ext * 0.95), (0.0, 1.0, 0.0), fov=1.0)
    return d, cam


def c4_scene(tex_size: int = 1024, subdiv: int = 5):
    """C4: the materials scene with generated tex_size^2 albedo / roughness-metallic / normal-map images
    inside a global HomogeneusMedium (HG g = 0.75, density 0.1; main.cpp:174,266-268): NEE + MIS through media."""
    d = SceneDesc()
    alb, orm, nm = procedural_textures(tex_size)
    t_alb, t_orm, t_nm = d.image_u8(alb), d.image_u8(orm), d.image_u8(nm)
    mats = [d.diffuse(t_alb, norm=t_nm, rough=t_orm, metal=t_orm),
            d.dielectric(1.5, d.solid((0.95, 0.97, 1.0)), rough=d.solid((0.1, 0.1, 0.1))),
            d.diffuse(t_alb, rough=d.solid((0.5, 0.5, 0.5))),
            d.conductor((0.9, 0.85, 0.6))]
    rng = np.random.default_rng(5)
    for k in range(16):
        pos, idx, nrm, uv = displaced_sphere(subdiv, 20 + k)
        b = d.add_blas([d.mesh(pos, idx, nrm, uv, material=mats[k % 4])])
        ang = 2 * math.pi * k / 16
        rad = 2.2 + 1.6 * (k % 2)
        d.instance(b, trs((rad * math.cos(ang), 0.6 + 0.2 * (k % 3), rad * math.sin(ang)), 0.55, rng.normal(size=3), rng.random() * 6.28))
    gp, gi, gn, gu = quad((-8, 0, -8), (0, 0, 16), (16, 0, 0))
    gt = np.tile(np.array([[0, 0, 1]], np.float32), (4, 1))
    d.instance(d.add_blas([d.mesh(gp, gi, gn, gu * 6.0, tangents=gt, material=d.diffuse(t_alb, norm=t_nm, rough=t_orm))]))
    lp, li, ln, lu = quad((-1, 0, -1), (2, 0, 0), (0, 0, 2))
    emit = d.solid((1.0, 0.9, 0.75), scale=(25, 25, 25))
    lb = d.add_blas([d.mesh(lp, li, ln, lu, material=d.diffuse(d.solid((0, 0, 0))), emissive_tex=emit)])
    d.instance(lb, trs((0.0, 5.0, 0.0), 1.0, (1, 0, 0), math.pi))
    d.light(capi.LIGHT_GRADIENT_INFINITE, color=(1.0, 1.0, 1.0), color_b=(0.5, 0.7, 1.0), scale=0.4)
    d.scene_medium = d.medium((0.5, 0.5, 0.5), (2.0, 2.2, 2.6), 0.75, density=0.1)
    cam = camera_desc((0.0, 2.4, 7.0), (0.0, 0.6, 0.0), fov=0.9)
    return d, cam


def hdr_sky(w: int = 1024, h: int = 512, seed: int = 2):
    """A generated HDR environment (float RGB): gradient sky, a bright sun disc, soft clouds."""
    rng = np.random.default_rng(seed)
    v, u = np.mgrid[0:h, 0:w].astype(np.float32)
    u, v = (u + 0.5) / w, (v + 0.5) / h
    theta = v * math.pi
    up = np.cos(theta)
    sky = np.stack([0.35 + 0.3 * (1 - up), 0.5 + 0.25 * (1 - up), 0.9 + 0 * up], -1) * np.clip(0.4 + 0.6 * up, 0.05, 1)[..., None]
    clouds = sum(np.sin(6.28 * (kx * u + ky * v) + p) for kx, ky, p in rng.normal(size=(6, 3)) * np.array([4, 3, 3]))
    sky *= (1.0 + 0.15 * np.tanh(clouds))[..., None]
    su, sv = 0.3, 0.28
    dsun = np.hypot((u - su) * 2 * np.sin(theta), v - sv)
    sky += (np.exp(-(dsun / 0.01) ** 2) * 400.0)[..., None] * np.array([1.0, 0.9, 0.7], np.float32)
    # as decoded from a Radiance .hdr file: the only way the reference can be given an HDR texture
    return rgbe_roundtrip(sky.astype(np.float32))


def c5_scene(grid: int = 2237, n_textures: int = 64, tex_size: int = 256, seed: int = 8, env: bool = True, n_lights: int = 8, n_objects: int = 24):
    """C5: ~10 M UNIQUE triangles (a grid x grid displaced terrain = 2*(grid-1)^2 triangles, split into
    n_textures patches each with its own generated albedo/roughness images), scattered emissive and glass
    objects, an HDR environment through TextureInfiniteLight."""
    d = SceneDesc()
    pos, tri, uv = terrain(grid, grid, seed, height=0.35, octaves=7)
    pos = (pos * np.array([40.0, 6.0, 40.0], np.float32) - np.array([20.0, 3.0, 20.0], np.float32)).astype(np.float32)
    # smooth normals from the height field
    fn = np.cross(pos[tri[:, 1]] - pos[tri[:, 0]], pos[tri[:, 2]] - pos[tri[:, 0]]).astype(np.float32)
    vn = np.zeros_like(pos)
    for c in range(3):
        np.add.at(vn, tri[:, c], fn)
    vn /= np.maximum(np.linalg.norm(vn, axis=1, keepdims=True), 1e-20)
    rng = np.random.default_rng(seed)
    patches = int(math.isqrt(n_textures))
    cell = (grid - 1) // patches
    quad_x = (np.arange(len(tri)) // 2) % (grid - 1)
    quad_z = (np.arange(len(tri)) // 2) // (grid - 1)
    pid = np.minimum(quad_z // cell, patches - 1) * patches + np.minimum(quad_x // cell, patches - 1)
    meshes = []
    y, x = np.mgrid[0:tex_size, 0:tex_size].astype(np.float32) / tex_size
    for p in range(patches * patches):
        f = rng.random(6) * 5 + 1
        alb = np.stack([0.45 + 0.35 * np.sin(6.28 * f[0] * x) * np.cos(6.28 * f[1] * y), 0.45 + 0.3 * np.cos(6.28 * f[2] * (x + y)),
                        0.35 + 0.25 * np.sin(6.28 * f[3] * y)], -1)
        rough = np.clip(0.4 + 0.4 * np.sin(6.28 * f[4] * x) * np.sin(6.28 * f[5] * y), 0.05, 1.0)
        orm = np.stack([np.ones_like(rough), rough, np.zeros_like(rough)], -1)
        to8 = lambda a: np.ascontiguousarray(np.clip(np.round(a * 255), 0, 255).astype(np.uint8))
        m = d.diffuse(d.image_u8(to8(alb)), rough=d.image_u8(to8(orm)))
        sel = tri[pid == p]
        used, inv = np.unique(sel, return_inverse=True)
        meshes.append(d.mesh(pos[used], inv.reshape(-1, 3).astype(np.uint32), vn[used], (uv[used] * patches * 4).astype(np.float32), material=m))
    d.instance(d.add_blas(meshes))
    glass = d.dielectric(1.5, d.solid((0.97, 0.98, 1.0)))
    metal = d.diffuse(d.solid((0.9, 0.7, 0.4)), rough=d.solid((0.25, 0.25, 0.25)), metal=d.solid((1, 1, 1)))
    sp, si, sn, su = displaced_sphere(5, 99)
    bg, bm = d.add_blas([d.mesh(sp, si, sn, su, material=glass)]), d.add_blas([d.mesh(sp, si, sn, su, material=metal)])
    for k in range(n_objects):
        c = (rng.random(2) - 0.5) * 30
        d.instance(bg if k % 2 else bm, trs((c[0], 2.5 + rng.random() * 2, c[1]), 0.6 + 0.5 * rng.random(), rng.normal(size=3), rng.random() * 6.28))
    lp, li, ln, lu = quad((-0.5, 0, -0.5), (1, 0, 0), (0, 0, 1))
    emit = d.solid((1.0, 0.8, 0.5), scale=(30, 30, 30))
    lb = d.add_blas([d.mesh(lp, li, ln, lu, material=d.diffuse(d.solid((0, 0, 0))), emissive_tex=emit)])
    for k in range(n_lights):
        c = (rng.random(2) - 0.5) * 24
        d.instance(lb, trs((c[0], 5.0 + rng.random(), c[1]), 1.5, (1, 0, 0), math.pi))
    if env:
        d.light(capi.LIGHT_TEXTURE_INFINITE, tex=d.image_f32(hdr_sky()), scale=1.0, power_func=2)
    else:
        d.light(capi.LIGHT_GRADIENT_INFINITE, color=(1.0, 1.0, 1.0), color_b=(0.5, 0.7, 1.0), scale=0.8)
    cam = camera_desc((0.0, 6.0, 24.0), (0.0, 0.5, 0.0), fov=1.0)
    return d, cam


def rgbe_roundtrip(img: np.ndarray) -> np.ndarray:
    """What a float RGB image becomes after being written as Radiance .hdr and loaded back (shared 8-bit
    exponent, 8-bit TRUNCATED mantissas): the reference can only build FloatImageTexture from a file, so
    these are the texels it actually sees.  Mirrors the float32 arithmetic of the stb encoder/decoder the
    reference vendors (checked against it in tests/test_synth_cpu.py)."""
    a = np.ascontiguousarray(img, dtype=np.float32)
    assert a.shape[-1] == 3
    mx = a.max(axis=-1)
    mant, expo = np.frexp(mx.astype(np.float64))
    with np.errstate(divide="ignore", invalid="ignore"):
        normalize = (mant.astype(np.float32) * np.float32(256.0) / mx).astype(np.float32)
    small = mx < np.float32(1e-32)
    bytes_ = np.where(small[..., None], 0, (a * normalize[..., None]).astype(np.float32)).astype(np.int64) & 0xFF
    e = np.where(small, 0, expo + 128).astype(np.int64)
    f1 = np.ldexp(np.float32(1.0), (e - 136).astype(np.int32)).astype(np.float32)
    out = np.where((e != 0)[..., None], bytes_.astype(np.float32) * f1[..., None], np.float32(0.0))
    return np.ascontiguousarray(out.astype(np.float32))


def misc_scene(uv_tex=True, normal_tex=True, thin=True, lens=True, uniform_sky=True, metal_half=True):
    """The corners the other scenes leave out: UniformInfiniteLight, UVTexture / NormalTexture albedos, a
    thin-lens camera (focus angle and distance), ThinDielectric with an albedo texture, a one-channel-scaled
    SolidColor (Texture::colorScale)."""
    d = SceneDesc()
    gp, gi, gn, gu = quad((-3, 0, -3), (0, 0, 6), (6, 0, 0))
    ground = d.uv_texture() if uv_tex else d.solid((0.5, 0.5, 0.5))
    d.instance(d.add_blas([d.mesh(gp, gi, gn, gu, material=d.diffuse(ground))]))
    pos, idx, nrm, uv = icosphere(3)
    plain = d.diffuse(d.solid((0.6, 0.6, 0.6)))
    mats = [d.diffuse(d.normal_texture(), rough=d.solid((0.4, 0.4, 0.4))) if normal_tex else plain,
            d.thin_dielectric(1.5, d.solid((0.8, 0.9, 1.0), scale=(1.0, 0.9, 0.8))) if thin else plain,
            d.diffuse(d.solid((0.9, 0.4, 0.3)), rough=d.solid((0.2, 0.2, 0.2)), metal=d.solid((0.5, 0.5, 0.5))) if metal_half else plain]
    for k, m in enumerate(mats):
        b = d.add_blas([d.mesh(pos, idx, nrm, uv, material=m)])
        d.instance(b, trs((-1.4 + 1.4 * k, 0.6, -0.5 * k), 0.55, (0.3, 1, 0.2), 0.7 * k))
    if uniform_sky:
        d.light(capi.LIGHT_UNIFORM_INFINITE, color=(0.7, 0.8, 1.0))
    else:
        d.light(capi.LIGHT_GRADIENT_INFINITE, color=(0.7, 0.8, 1.0), color_b=(0.7, 0.8, 1.0), scale=1.0)
    cam = camera_desc((0.0, 1.3, 3.2), (0.0, 0.5, 0.0), fov=0.9, focus_angle=0.12 if lens else 0.0, focus_distance=3.3 if lens else 0.0)
    return d, cam


def shapes_scene(mesh=True, quad_light=True, sphere_light=True, one_sided=True, boundary=False, sky=True, n_spheres=9, seed=4):
    """main.cpp-style scene made of analytic GeometricPrimitives: a huge checker ground QuadShape, a row of
    SphereShapes with the README materials, a QuadShape AreaLight (one-sided or not), an emissive SphereShape
    with a textured AreaLight, optionally one transformed triangle mesh in the same TLAS and a material-less
    medium-boundary sphere (VolPathIntegrator)."""
    d = SceneDesc()
    rng = np.random.default_rng(seed)
    ground = d.diffuse(d.checker(d.solid((0.8, 0.8, 0.8)), d.solid((0.2, 0.3, 0.1)), (0.002, 0.002)))
    d.quad((-100, -0.3, -100), (1000, 0, 0), (0, 0, 1000), material=ground)
    mats = [
        d.diffuse(d.solid((0.1, 0.2, 0.5))),
        d.dielectric(1.5, d.solid((1, 1, 1))),
        d.conductor((0.8, 0.6, 0.2)),
        d.diffuse(d.solid((0.95, 0.64, 0.54)), rough=d.solid((0.25, 0.25, 0.25)), metal=d.solid((1, 1, 1))),
        d.dielectric(1.5, d.solid((0.9, 1.0, 0.9)), rough=d.solid((0.2, 0.2, 0.2))),
        d.thin_dielectric(1.5, d.solid((1.0, 0.9, 0.9))),
        d.diffuse(d.checker(d.solid((0.9, 0.1, 0.1)), d.solid((0.9, 0.9, 0.9)), (0.1, 0.1))),
    ]
    for k in range(n_spheres):
        x = -2.4 + 4.8 * (k + 0.5) / n_spheres
        r = 0.18 + 0.12 * rng.random()
        d.sphere((x, -0.3 + r + 0.4 * (k % 3 == 1), -0.6 * (k % 2) + 0.2 * rng.standard_normal()), r, material=mats[k % len(mats)])
    if mesh:
        pos, idx, nrm, uv = icosphere(2, 1.0)
        b = d.add_blas([d.mesh(pos, idx, nrm, uv, material=mats[3])])
        d.instance(b, trs((0.2, 0.25, -1.6), 0.55, (0.2, 1, 0.1), 0.8))
        d.instance(b)
    black = d.diffuse(d.solid((0, 0, 0)))
    if quad_light:
        d.quad((0.5, 1.9, 0.3), (-1.0, 0, 0), (0, 0.3, -0.8), material=black, emissive_tex=d.solid((1, 0.95, 0.85), scale=(14, 14, 14)), one_sided=one_sided)
    if sphere_light:
        glow = d.checker(d.solid((6, 2, 1)), d.solid((1, 3, 6)), (0.125, 0.125))
        d.sphere((-1.6, 1.0, -1.0), 0.3, material=d.diffuse(d.solid((0.1, 0.1, 0.1))), emissive_tex=glow)
    if boundary:
        fog = d.medium((0.01, 0.9, 0.9), (1.0, 0.1, 0.1), 0.8, density=5.0)
        d.sphere((1.3, 0.3, 0.4), 0.6, material=-1, medium=fog)
    if sky:
        d.light(capi.LIGHT_GRADIENT_INFINITE, color=(1.0, 1.0, 1.0), color_b=(0.5, 0.7, 1.0), scale=0.4)
    cam = camera_desc((0.0, 1.1, 4.0), (0.0, 0.3, 0.0), fov=0.9)
    return d, cam


def wrapped_scene(boundary=False, lights=True, seed=12):
    """shapes_scene plus every wrapper the reference can put around a top-level primitive: a
    TransformedPrimitive over a sphere and over a quad GeometricPrimitive (Primitive.cpp:32-73), an
    AnimatedPrimitive (Primitive.cpp:77-96; main.cpp:189) over a plain sphere, over a transformed sphere,
    over a BLAS and over a TransformedPrimitive(BLAS), emissive wrapped shapes (TransformedLight /
    AnimatedLight, Light.cpp:296-365).  Rays need a `time` for the animated ones."""
    d, cam = shapes_scene(boundary=boundary, sphere_light=lights, quad_light=lights)
    blue = d.diffuse(d.solid((0.1, 0.2, 0.5)))
    gold = d.conductor((0.8, 0.6, 0.2))
    red = d.diffuse(d.solid((0.7, 0.15, 0.1)))
    # main.cpp:187-190: a diffuse sphere moving up by 0.2 over the shutter interval
    d.animate(d.sphere((0.0, 0.1, -1.2), 0.5, material=blue), (0.0, 0.2, 0.0), (0.0, 1.0))
    # an ellipsoid (non-uniformly scaled, rotated sphere) and a sheared, rotated quad
    m = trs((1.4, 0.35, -0.9), 1.0, (0.3, 1, 0.2), 0.7) @ np.diag([1.0, 0.45, 0.7, 1.0]).astype(np.float32)
    d.sphere((0.1, 0.0, -0.1), 0.6, material=gold, transform=m)
    d.quad((-0.5, 0.0, -0.5), (1.0, 0, 0), (0.2, 0, 1.0), material=red, transform=trs((-1.3, 0.6, 0.4), 0.8, (1, 0.2, 0.1), 1.1))
    # animated ellipsoid, time bounds that do not start at 0 (the reference clamps time - tb.x to [tb.x, tb.y])
    m2 = trs((-0.4, 1.2, -0.4), 0.6, (0, 0, 1), 0.5) @ np.diag([1.0, 0.5, 1.0, 1.0]).astype(np.float32)
    d.animate(d.sphere((0.0, 0.0, 0.0), 0.5, material=blue, transform=m2), (0.3, -0.1, 0.2), (0.25, 0.75))
    # animated BLAS, with and without its own transform
    pos, idx, nrm, uv = icosphere(1, 1.0)
    b = d.add_blas([d.mesh(pos, idx, nrm, uv, material=red)])
    d.animate(d.instance(b, trs((2.0, 0.5, 0.3), 0.35, (0.1, 1, 0.3), 2.0)), (-0.2, 0.3, 0.1), (0.0, 1.0))
    pos2, idx2, nrm2, uv2 = icosphere(1, 0.3)
    b2 = d.add_blas([d.mesh(pos2 + np.float32((-2.2, 0.4, 0.6)), idx2, nrm2, uv2, material=gold)])
    d.animate(d.instance(b2), (0.0, 0.25, 0.0), (0.0, 0.5))
    if lights:
        warm = d.solid((1.0, 0.8, 0.5), scale=(9, 9, 9))
        black = d.diffuse(d.solid((0, 0, 0)))
        d.quad((-0.3, 0.0, -0.3), (0.6, 0, 0), (0, 0, 0.6), material=black, emissive_tex=warm, one_sided=False,
               transform=trs((1.2, 1.7, -0.3), 1.2, (1, 0, 0.2), 2.8))
        d.animate(d.sphere((-0.9, 1.5, 0.5), 0.12, material=black, emissive_tex=d.solid((0.5, 0.8, 1.0), scale=(12, 12, 12))), (0.4, 0.0, 0.0), (0.0, 1.0))
    return d, cam


def with_times(rays, seed=3, lo=-0.25, hi=1.25):
    """random ray times (the camera's Get1D() shutter sample lives in [0, 1); tests go a little outside)"""
    rng = np.random.default_rng(seed)
    rays = rays.copy()
    rays["time"] = (lo + (hi - lo) * rng.random(len(rays))).astype(np.float32)
    return rays


def alpha_shapes_scene(mode=capi.ALPHA_MASK, wrapped=True):
    """alpha_scene plus alpha-tested analytic shapes (GeometricPrimitive::Intersect[Pred] runs Material::Alpha on the
    shape's own uv, Primitive.cpp:6-26): a sphere whose opacity comes from a checker over GetSphereUV, a quad card with
    a smooth opacity image over (alpha, beta), and -- wrapped -- the same kind of card inside a TransformedPrimitive."""
    d, cam = alpha_scene(mode)
    n = 32
    y, x = np.mgrid[0:n, 0:n].astype(np.float32) / n
    opacity = np.clip(0.5 + 0.5 * np.sin(6.28 * 3 * x) * np.cos(6.28 * 2 * y), 0.0, 1.0)
    to8 = lambda a: np.ascontiguousarray(np.clip(np.round(a * 255), 0, 255).astype(np.uint8))
    a_img = d.image_u8(to8(np.repeat(opacity[..., None], 3, -1)))
    holes = d.checker(d.solid((1, 1, 1)), d.solid((0.2, 0.2, 0.2)), (0.125, 0.125))
    ball = d.diffuse(d.solid((0.8, 0.3, 0.2)), alpha=holes, alpha_mode=mode, alpha_cutoff=0.5)
    card = d.diffuse(d.solid((0.2, 0.4, 0.8)), alpha=a_img, alpha_mode=mode, alpha_cutoff=0.5)
    d.sphere((0.3, 0.6, 1.25), 0.18, material=ball)
    d.quad((0.55, 0.1, 1.3), (0.35, 0, 0.05), (0, 0.45, 0), material=card)
    if wrapped:
        d.quad((-0.5, -0.5, 0.0), (1, 0, 0), (0, 1, 0), material=card, transform=trs((0.2, 0.3, 1.45), 0.3, (0, 1, 0), 0.4))
    return d, cam
