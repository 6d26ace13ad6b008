"""nn_bvh_b200/rays.py -- deterministic ray batches for the BASELINE.json configs.

The reference never materialises a ray batch on its CPU path (rays are created
one at a time inside Li(), /root/reference/src/pbrt/cpu/integrators.cpp:953-1345),
so the harness regenerates the batches the configs describe:
pinhole camera rays, cosine-weighted diffuse bounces off the primary hits, and
unnormalised shadow rays with tMax = 1 - ShadowEpsilon = 0.9999
(cpu/integrators.h:52-54, util/math.h:42).  Tracing the primary rays is done by
whatever `trace` callable the caller injects (the GPU engine in bench.py, the
CPU oracle in the CPU-only tests) -- this module does no intersection itself.
"""
from __future__ import annotations

import math

import numpy as np

from .capi import MISS, RAY_DTYPE

SHADOW_TMAX = np.float32(0.9999)


def _look_at(eye, look, up):
    eye, look, up = (np.asarray(v, dtype=np.float64) for v in (eye, look, up))
    d = look - eye
    d /= np.linalg.norm(d)
    right = np.cross(up / np.linalg.norm(up), d)
    right /= np.linalg.norm(right)
    newup = np.cross(d, right)
    m = np.eye(4)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = right, newup, d, eye
    return m  # world_from_camera


def camera_rays(cam: dict, xres: int = None, yres: int = None) -> np.ndarray:
    """Pixel-centre pinhole rays, normalised d, tMax = inf (config 1's 1024x1024 @ 1spp).
    pbrt convention: camera looks down +z, fov spans the shorter image axis
    (cameras.cpp PerspectiveCamera)."""
    xres = xres or cam["xres"]
    yres = yres or cam["yres"]
    if "camera_from_world" in cam:
        wfc = np.linalg.inv(np.asarray(cam["camera_from_world"], dtype=np.float64))
    else:
        wfc = _look_at(cam["eye"], cam["look"], cam["up"])
    aspect = xres / yres
    sx, sy = (aspect, 1.0) if aspect > 1 else (1.0, 1.0 / aspect)
    th = math.tan(math.radians(cam["fov"]) / 2)
    px = (np.arange(xres) + 0.5) / xres
    py = (np.arange(yres) + 0.5) / yres
    X, Y = np.meshgrid((2 * px - 1) * sx * th, (1 - 2 * py) * sy * th, indexing="xy")
    dc = np.stack([X, Y, np.ones_like(X)], -1).reshape(-1, 3)
    dw = dc @ wfc[:3, :3].T
    dw /= np.linalg.norm(dw, axis=1, keepdims=True)
    r = np.zeros(dw.shape[0], dtype=RAY_DTYPE)
    r["o"] = wfc[:3, 3].astype(np.float32)
    r["d"] = dw.astype(np.float32)
    r["tMax"] = np.inf
    return r


def _hit_frames(rays, hits, ordered_tris9):
    """World-space hit points and geometric normals facing the incoming ray."""
    ok = hits["prim"] != MISS
    r, h = rays[ok], hits[ok]
    p = r["o"].astype(np.float64) + r["d"].astype(np.float64) * h["t"].astype(np.float64)[:, None]
    tri = ordered_tris9[h["prim"]].astype(np.float64).reshape(-1, 3, 3)
    n = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    n = n / np.where(ln > 0, ln, 1)
    n = np.where((np.sum(n * r["d"], axis=1) > 0)[:, None], -n, n)
    return p, n


def bounce_and_shadow_rays(trace, ordered_tris9, cam, n_rays: int, seed: int = 2,
                           primary_res: int = 1024, bounds=None):
    """Config 2: n_rays incoherent cosine-hemisphere bounce rays (closest-hit, tMax=inf) and
    n_rays shadow rays (any-hit, d = pTo - pFrom, tMax = 0.9999) from primary hit points.
    trace(rays) -> hits (HIT_DTYPE)."""
    prim = camera_rays(cam, primary_res, primary_res)
    hits = trace(prim)
    p, n = _hit_frames(prim, hits, ordered_tris9)
    if len(p) == 0:
        raise RuntimeError("no primary ray hit the scene; cannot build bounce rays")
    v = ordered_tris9.reshape(-1, 3)
    lo, hi = (v.min(0), v.max(0)) if bounds is None else bounds
    eps = 1e-4 * float(np.max(hi - lo))
    rng = np.random.default_rng(seed)
    pick = rng.integers(0, len(p), size=n_rays)
    P, Nn = p[pick], n[pick]
    o = (P + Nn * eps).astype(np.float32)
    # SampleCosineHemisphere about Nn (util/sampling.h:409 semantics: concentric disk + lift)
    u1, u2 = rng.random(n_rays), rng.random(n_rays)
    rad, phi = np.sqrt(u1), 2 * np.pi * u2
    lx, ly, lz = rad * np.cos(phi), rad * np.sin(phi), np.sqrt(np.maximum(0.0, 1 - u1))
    a = np.where(np.abs(Nn[:, [0]]) > 0.9, np.array([[0.0, 1.0, 0.0]]), np.array([[1.0, 0.0, 0.0]]))
    T = np.cross(a, Nn)
    T /= np.linalg.norm(T, axis=1, keepdims=True)
    B = np.cross(Nn, T)
    d = T * lx[:, None] + B * ly[:, None] + Nn * lz[:, None]
    bounce = np.zeros(n_rays, dtype=RAY_DTYPE)
    bounce["o"], bounce["d"], bounce["tMax"] = o, d.astype(np.float32), np.inf
    target = lo + (hi - lo) * rng.random((n_rays, 3))
    shadow = np.zeros(n_rays, dtype=RAY_DTYPE)
    shadow["o"] = o
    shadow["d"] = (target - o.astype(np.float64)).astype(np.float32)
    shadow["tMax"] = SHADOW_TMAX
    return bounce, shadow


def path_bounce_batches(trace, ordered_tris9, cam, res: int = 1024, bounces: int = 4, seed: int = 3):
    """Config 3: batch k = all live rays at bounce k of a diffuse random walk (coherent
    camera rays first, increasingly incoherent afterwards)."""
    rng = np.random.default_rng(seed)
    rays = camera_rays(cam, res, res)
    v = ordered_tris9.reshape(-1, 3)
    eps = 1e-4 * float(np.max(v.max(0) - v.min(0)))
    out = [rays]
    for _ in range(bounces):
        hits = trace(rays)
        p, n = _hit_frames(rays, hits, ordered_tris9)
        if len(p) == 0:
            break
        m = len(p)
        u1, u2 = rng.random(m), rng.random(m)
        rad, phi = np.sqrt(u1), 2 * np.pi * u2
        lx, ly, lz = rad * np.cos(phi), rad * np.sin(phi), np.sqrt(np.maximum(0.0, 1 - u1))
        a = np.where(np.abs(n[:, [0]]) > 0.9, np.array([[0.0, 1.0, 0.0]]), np.array([[1.0, 0.0, 0.0]]))
        T = np.cross(a, n)
        T /= np.linalg.norm(T, axis=1, keepdims=True)
        B = np.cross(n, T)
        d = T * lx[:, None] + B * ly[:, None] + n * lz[:, None]
        rays = np.zeros(m, dtype=RAY_DTYPE)
        rays["o"], rays["d"], rays["tMax"] = (p + n * eps).astype(np.float32), d.astype(np.float32), np.inf
        out.append(rays)
    return out


def random_rays(n: int, seed: int = 6, lo=-0.25, hi=1.25) -> np.ndarray:
    """Config 5: o ~ U[lo,hi)^3, d uniform on the sphere, tMax = inf."""
    rng = np.random.default_rng(seed)
    r = np.zeros(n, dtype=RAY_DTYPE)
    r["o"] = (lo + (hi - lo) * rng.random((n, 3))).astype(np.float32)
    z = 1 - 2 * rng.random(n)
    rad = np.sqrt(np.maximum(0.0, 1 - z * z))
    phi = 2 * np.pi * rng.random(n)
    r["d"] = np.stack([rad * np.cos(phi), rad * np.sin(phi), z], 1).astype(np.float32)
    r["tMax"] = np.inf
    return r


# ---- config 5 on the device: ray i is a pure function of its GLOBAL index ------------------------
def _hash01(idx, stream: int):
    """splitmix64-style hash of (global index, stream) -> float32 in [0,1) with 24 random bits.
    idx: int64 torch tensor (any device)."""
    import torch

    def w(v):   # python int -> the int64 with the same low 64 bits
        v &= (1 << 64) - 1
        return v - (1 << 64) if v >= (1 << 63) else v

    def lsr(v, k):   # logical shift right on int64
        return (v >> k) & ((1 << (64 - k)) - 1)

    x = idx * -7046029254386353131 + w((stream + 1) * 0xBF58476D1CE4E5B9)   # wraps mod 2^64
    x = (x ^ lsr(x, 30)) * -4658895280553007687
    x = (x ^ lsr(x, 27)) * -7723592293110705685
    x = x ^ lsr(x, 31)
    return (lsr(x, 40) & 0xFFFFFF).to(torch.float32) * (1.0 / 16777216.0)


def hashed_random_rays(a: int, b: int, device, lo: float = -0.25, hi: float = 1.25):
    """Config 5's rays [a, b) of the global batch as an (b-a, 8) float32 torch tensor on `device`:
    o ~ U[lo,hi)^3, d uniform on the sphere, tMax = inf.  Every world size and every
    sharding sees the same ray for the same global index, so checksums are comparable across N."""
    import torch
    out = torch.empty((b - a, 8), dtype=torch.float32, device=device)
    step = 1 << 24
    for s in range(a, b, step):
        e = min(b, s + step)
        idx = torch.arange(s, e, dtype=torch.int64, device=device)
        o = out[s - a:e - a]
        for k in range(3):
            o[:, k] = _hash01(idx, k) * (hi - lo) + lo
        z = 1 - 2 * _hash01(idx, 3)
        phi = (2 * np.pi) * _hash01(idx, 4)
        rad = torch.sqrt(torch.clamp(1 - z * z, min=0))
        o[:, 4], o[:, 5], o[:, 6] = rad * torch.cos(phi), rad * torch.sin(phi), z
        o[:, 3] = float("inf")
        o[:, 7] = 0
    return out
