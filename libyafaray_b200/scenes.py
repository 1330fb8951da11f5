"""Synthetic procedural scenes and ray batches.

Everything is generated on the host with numpy from fixed seeds so that the
reference-based checker used by the tests and the CUDA path see byte-identical
inputs.  Triangles are returned as float32 [n_tri, 9] (v0 v1 v2, world space);
triangle i is primitive index i (reference: Scene::updateObjects ordering,
/root/reference/src/scene/scene.cc:1032-1041).  Rays are float32 [n, 8] =
(ox oy oz tmin dx dy dz tmax) with tmax < 0 meaning "unbounded"
(reference Ray: include/geometry/ray.h:37-52).
"""
from __future__ import annotations

import numpy as np

MIN_RAYDIST = 5.0e-5  # reference include/common/yafaray_common.h:28
SHADOW_BIAS = 5.0e-4  # reference include/common/yafaray_common.h:29


def _splitmix64(x: np.ndarray) -> np.ndarray:
    x = (x + np.uint64(0x9E3779B97F4A7C15)).astype(np.uint64)
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def hash01(seed: int, *idx: np.ndarray) -> np.ndarray:
    """Deterministic uniform [0,1) float64 from integer indices (splitmix64 chain)."""
    with np.errstate(over="ignore"):
        h = np.full(np.broadcast(*idx).shape, seed, dtype=np.uint64)
        for a in idx:
            h = _splitmix64(h ^ np.asarray(a).astype(np.uint64))
        return (h >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))


def displaced_sphere(nu: int = 1000, nv: int = 500, seed: int = 0xC3) -> np.ndarray:
    """C3: UV sphere, radius 1 + 0.05 sin(20 theta) cos(20 phi) + 0.01 hash; nu*nv*2 triangles."""
    j = np.arange(nv + 1)[:, None]  # latitude rows
    i = np.arange(nu + 1)[None, :]  # longitude columns (last == first)
    theta = np.pi * j / nv
    phi = 2.0 * np.pi * (i % nu) / nu
    r = 1.0 + 0.05 * np.sin(20.0 * theta) * np.cos(20.0 * phi) + 0.01 * hash01(seed, j % (nv + 1) + 0 * i, i % nu + 0 * j)
    p = np.stack([r * np.sin(theta) * np.cos(phi), r * np.cos(theta) + 0 * phi, r * np.sin(theta) * np.sin(phi)], -1)
    p = p.astype(np.float32)
    a, b, c, d = p[:-1, :-1], p[:-1, 1:], p[1:, :-1], p[1:, 1:]
    t0 = np.concatenate([a, c, b], -1).reshape(-1, 9)
    t1 = np.concatenate([b, c, d], -1).reshape(-1, 9)
    tris = np.empty((t0.shape[0] * 2, 9), np.float32)
    tris[0::2], tris[1::2] = t0, t1
    return np.ascontiguousarray(tris)


def triangle_soup(n_tri: int, seed: int = 0xC5, first: int = 0) -> np.ndarray:
    """C5: random triangles, centres uniform in [0,1]^3, edge length log-uniform in [1e-3, 1e-2].
    `first`: index of the first triangle (elements are hashed from their index, so a large soup can be generated in chunks)."""
    k = np.arange(first, first + n_tri)
    c = np.stack([hash01(seed, k, np.full_like(k, a)) for a in range(3)], -1)
    size = 1e-3 * np.power(10.0, hash01(seed, k, np.full_like(k, 3)))
    out = np.empty((n_tri, 3, 3))
    for v in range(3):
        d = np.stack([hash01(seed, k, np.full_like(k, 4 + 3 * v + a)) for a in range(3)], -1) - 0.5
        out[:, v, :] = c + d * size[:, None]
    return np.ascontiguousarray(out.reshape(n_tri, 9).astype(np.float32))


def primary_rays(width: int, height: int, eye=(0.0, 0.0, 3.0), look=(0.0, 0.0, 0.0), fov_deg: float = 43.6) -> np.ndarray:
    """Pinhole grid: one ray through each pixel centre, tmin=0, tmax=-1 (unbounded)."""
    eye = np.asarray(eye, np.float64)
    fwd = np.asarray(look, np.float64) - eye
    fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, [0.0, 1.0, 0.0])
    right /= np.linalg.norm(right)
    up = np.cross(right, fwd)
    half = np.tan(np.radians(fov_deg) * 0.5)
    x = ((np.arange(width) + 0.5) / width * 2 - 1) * half * (width / height)
    y = (1 - (np.arange(height) + 0.5) / height * 2) * half
    d = fwd[None, None, :] + x[None, :, None] * right + y[:, None, None] * up
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    rays = np.empty((height * width, 8), np.float32)
    rays[:, 0:3] = eye
    rays[:, 3] = 0.0
    rays[:, 4:7] = d.reshape(-1, 3)
    rays[:, 7] = -1.0
    return rays


def hit_points_normals(tris: np.ndarray, rays: np.ndarray, t: np.ndarray, prim: np.ndarray):
    """Hit position and unit geometric normal (faced toward the ray origin) for hit rays."""
    ok = prim >= 0
    p = rays[:, 0:3].astype(np.float64) + rays[:, 4:7].astype(np.float64) * t[:, None].astype(np.float64)
    tr = tris[np.where(ok, prim, 0)].astype(np.float64).reshape(-1, 3, 3)
    n = np.cross(tr[:, 1] - tr[:, 0], tr[:, 2] - tr[:, 0])
    ln = np.linalg.norm(n, axis=-1, keepdims=True)
    n = n / np.where(ln > 0, ln, 1.0)
    flip = np.sum(n * rays[:, 4:7], -1) > 0
    n[flip] *= -1
    return p, n, ok


def shadow_rays(p: np.ndarray, ok: np.ndarray, light=(2.0, 3.0, 4.0)) -> np.ndarray:
    """Shadow rays from hit points toward a point light, reference bias convention
    (tmin = shadow_bias*max(1,|p|), tmax = dist; integrator_montecarlo.cc:105-114)."""
    src = p[ok]
    n = p.shape[0]
    reps = -(-n // max(1, src.shape[0]))
    src = np.tile(src, (reps, 1))[:n]  # misses recycled
    d = np.asarray(light, np.float64) - src
    dist = np.linalg.norm(d, axis=-1)
    d /= dist[:, None]
    rays = np.empty((n, 8), np.float32)
    rays[:, 0:3] = src
    rays[:, 3] = SHADOW_BIAS * np.maximum(1.0, np.linalg.norm(src, axis=-1))
    rays[:, 4:7] = d
    rays[:, 7] = dist
    return rays


def _radical_inverse2(i: np.ndarray) -> np.ndarray:
    i = i.astype(np.uint32)
    i = (i << np.uint32(16)) | (i >> np.uint32(16))
    i = ((i & np.uint32(0x00FF00FF)) << np.uint32(8)) | ((i & np.uint32(0xFF00FF00)) >> np.uint32(8))
    i = ((i & np.uint32(0x0F0F0F0F)) << np.uint32(4)) | ((i & np.uint32(0xF0F0F0F0)) >> np.uint32(4))
    i = ((i & np.uint32(0x33333333)) << np.uint32(2)) | ((i & np.uint32(0xCCCCCCCC)) >> np.uint32(2))
    i = ((i & np.uint32(0x55555555)) << np.uint32(1)) | ((i & np.uint32(0xAAAAAAAA)) >> np.uint32(1))
    return i.astype(np.float64) * (1.0 / 4294967296.0)


def diffuse_rays(p: np.ndarray, n: np.ndarray, ok: np.ndarray, seed: int = 0xD1F) -> np.ndarray:
    """Incoherent cosine-weighted bounce rays from hit points, tmin = ray_min_dist (absolute),
    tmax unbounded (reference integrator_path_tracer.cc:189)."""
    cnt = p.shape[0]
    src, nn = p[ok], n[ok]
    reps = -(-cnt // max(1, src.shape[0]))
    src, nn = np.tile(src, (reps, 1))[:cnt], np.tile(nn, (reps, 1))[:cnt]
    k = np.arange(cnt)
    s1 = _radical_inverse2(k)
    s2 = hash01(seed, k)
    a = np.where(np.abs(nn[:, [0]]) > 0.5, [[0.0, 1.0, 0.0]], [[1.0, 0.0, 0.0]])
    tu = np.cross(nn, a)
    tu /= np.linalg.norm(tu, axis=-1, keepdims=True)
    tv = np.cross(nn, tu)
    r = np.sqrt(1.0 - s1)
    d = (tu * np.cos(2 * np.pi * s2)[:, None] + tv * np.sin(2 * np.pi * s2)[:, None]) * r[:, None] + nn * np.sqrt(s1)[:, None]
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    rays = np.empty((cnt, 8), np.float32)
    rays[:, 0:3] = src
    rays[:, 3] = MIN_RAYDIST
    rays[:, 4:7] = d
    rays[:, 7] = -1.0
    return rays


def incoherent_rays(n: int, seed: int = 0x1C0, lo=0.0, hi=1.0, first: int = 0) -> np.ndarray:
    """C5 stress: uniform origins in the cube, uniform directions; tmin=0, unbounded."""
    k = np.arange(first, first + n)
    o = np.stack([hash01(seed, k, np.full_like(k, a)) for a in range(3)], -1) * (hi - lo) + lo
    z = 2 * hash01(seed, k, np.full_like(k, 3)) - 1
    ph = 2 * np.pi * hash01(seed, k, np.full_like(k, 4))
    s = np.sqrt(np.maximum(0.0, 1 - z * z))
    rays = np.empty((n, 8), np.float32)
    rays[:, 0:3] = o
    rays[:, 3] = 0.0
    rays[:, 4:7] = np.stack([s * np.cos(ph), s * np.sin(ph), z], -1)
    rays[:, 7] = -1.0
    return rays


def segment_rays(n: int, seed: int = 0x5E6, lo=0.0, hi=1.0, first: int = 0) -> np.ndarray:
    """C5 any-hit stress: shadow-style rays between random point pairs (tmin = bias, tmax = dist)."""
    k = np.arange(first, first + n)
    a = np.stack([hash01(seed, k, np.full_like(k, c)) for c in range(3)], -1) * (hi - lo) + lo
    b = np.stack([hash01(seed, k, np.full_like(k, 3 + c)) for c in range(3)], -1) * (hi - lo) + lo
    d = b - a
    dist = np.linalg.norm(d, axis=-1)
    d /= dist[:, None]
    rays = np.empty((n, 8), np.float32)
    rays[:, 0:3] = a
    rays[:, 3] = SHADOW_BIAS * np.maximum(1.0, np.linalg.norm(a, axis=-1))
    rays[:, 4:7] = d
    rays[:, 7] = dist
    return rays
