"""nn_bvh_b200/scenes.py -- triangle soups and cameras for the BASELINE.json configs.

Geometry comes from scene_cache/<name>.tris.npy, exported once from the
reference's shipped scenes by tools/export_scenes.py (the GPU box has no
/root/reference; the cache travels with the working tree).  When a cache file
is missing load() raises; only with allow_procedural=True is a deterministic
procedural stand-in of similar size generated, and the returned name says so -- a
bench line always names what it really traced.

Nothing here intersects anything: this is input preparation for
GPUBVHAggregate (the reference's equivalent is its parser + CreateAggregate,
/root/reference/src/pbrt/scene.cpp:1352-1591, out of scope).
"""
from __future__ import annotations

import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CACHE = os.path.join(ROOT, "scene_cache")

# config index (BASELINE.json) -> scene; conference is missing from the reference
# tree (.MISSING_LARGE_BLOBS:12), bathroom is its stand-in.
CONFIG_SCENES = {0: "killeroos", 1: "bathroom", 2: "crown", 3: "coffee_maker", 4: "synthetic"}


def _icosphere(level: int):
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t],
                  [0, -1, -t], [0, 1, -t], [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], float)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4],
                  [11, 10, 2], [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8],
                  [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]])
    tri = v[f]  # (F,3,3)
    for _ in range(level):
        a, b, c = tri[:, 0], tri[:, 1], tri[:, 2]
        ab, bc, ca = a + b, b + c, c + a
        for m in (ab, bc, ca):
            m /= np.linalg.norm(m, axis=1, keepdims=True)
        tri = np.concatenate([np.stack([a, ab, ca], 1), np.stack([b, bc, ab], 1),
                              np.stack([c, ca, bc], 1), np.stack([ab, bc, ca], 1)], 0)
    return tri


def procedural_room(n_objects: int = 60, sphere_level: int = 5, seed: int = 1234):
    """A closed room (tessellated walls) filled with bumpy tessellated blobs:
    ~n_objects * 20 * 4^level triangles of very mixed sizes, like an interior scene."""
    rng = np.random.default_rng(seed)
    parts = []
    # walls / floor / ceiling: 6 grids of 64x64 quads over [-3,3] x [0,3] x [-3,3]
    g = np.linspace(0.0, 1.0, 65)
    u, v = np.meshgrid(g, g, indexing="ij")
    quads = np.stack([np.stack([u[:-1, :-1], v[:-1, :-1]], -1), np.stack([u[1:, :-1], v[1:, :-1]], -1),
                      np.stack([u[1:, 1:], v[1:, 1:]], -1), np.stack([u[:-1, 1:], v[:-1, 1:]], -1)], 2)
    quads = quads.reshape(-1, 4, 2)
    tri2 = np.concatenate([quads[:, [0, 1, 2]], quads[:, [2, 3, 0]]], 0)  # (T,3,2)
    lo, hi = np.array([-3.0, 0.0, -3.0]), np.array([3.0, 3.0, 3.0])
    for axis in range(3):
        for side in (0, 1):
            p = np.zeros(tri2.shape[:2] + (3,))
            o = [a for a in range(3) if a != axis]
            p[..., o[0]] = lo[o[0]] + tri2[..., 0] * (hi[o[0]] - lo[o[0]])
            p[..., o[1]] = lo[o[1]] + tri2[..., 1] * (hi[o[1]] - lo[o[1]])
            p[..., axis] = hi[axis] if side else lo[axis]
            parts.append(p)
    base = _icosphere(sphere_level)
    for _ in range(n_objects):
        c = lo + (hi - lo) * (0.1 + 0.8 * rng.random(3))
        r = 0.08 + 0.35 * rng.random()
        s = r * (0.6 + 0.8 * rng.random(3))
        k = rng.integers(2, 7, size=3)
        ph = rng.random(3) * 6.28
        bump = 1.0 + 0.12 * np.sin(base[..., 0] * k[0] + ph[0]) * np.sin(base[..., 1] * k[1] + ph[1]) \
            * np.sin(base[..., 2] * k[2] + ph[2])
        parts.append(c + base * bump[..., None] * s)
    tris = np.concatenate(parts, 0).astype(np.float32).reshape(-1, 9) + np.float32(0.0)
    cam = {"eye": [2.6, 1.6, 2.6], "look": [-1.0, 1.0, -1.0], "up": [0.0, 1.0, 0.0], "fov": 60.0,
           "xres": 1024, "yres": 1024}
    return tris, cam


def synthetic_soup(n_tris: int, seed: int = 5):
    """Config 5's mesh: centroid ~U[0,1)^3, vertices c + s*U(-1,1)^3, s = 0.5*T^(-1/3)."""
    rng = np.random.default_rng(seed)
    s = 0.5 * n_tris ** (-1.0 / 3.0)
    out = np.empty((n_tris, 3, 3), dtype=np.float32)
    step = 1 << 20
    for a in range(0, n_tris, step):
        b = min(n_tris, a + step)
        c = rng.random((b - a, 1, 3), dtype=np.float32)
        out[a:b] = c + np.float32(s) * (rng.random((b - a, 3, 3), dtype=np.float32) * 2 - 1)
    return out.reshape(-1, 9) + np.float32(0.0)   # no -0.0 (pbrt's mesh transform removes it)


def _camera_from_meta(meta: dict, tris: np.ndarray) -> dict:
    cam = meta.get("camera", {})
    out = {"fov": float(cam.get("fov", 45.0)), "xres": int(cam.get("xres", 1024)),
           "yres": int(cam.get("yres", 1024))}
    if "camera_from_world" in cam:
        out["camera_from_world"] = np.array(cam["camera_from_world"], dtype=np.float64)
    else:
        v = tris.reshape(-1, 3)
        c, ext = (v.min(0) + v.max(0)) / 2, (v.max(0) - v.min(0)).max()
        out.update(eye=(c + np.array([0.9, 0.4, 0.9]) * ext).tolist(), look=c.tolist(),
                   up=[0.0, 1.0, 0.0])
    return out


def load_spheres(name: str):
    """The scene's `Shape "sphere"` primitives (killeroos: the spherical light; crown: the
    environment shell) as nnbvh_sphere records, or None.  They are NOT part of load()'s soup."""
    mp = os.path.join(CACHE, name + ".json")
    if not os.path.exists(mp):
        return None
    with open(mp) as f:
        sph = json.load(f).get("spheres") or []
    if not sph:
        return None
    from .aggregate import make_spheres
    return make_spheres(np.array([s["render_from_object"] for s in sph]),
                        np.array([s["radius"] for s in sph], dtype=np.float32),
                        np.array([s["zmin"] for s in sph], dtype=np.float32),
                        np.array([s["zmax"] for s in sph], dtype=np.float32),
                        np.array([s["phimax"] for s in sph], dtype=np.float32))


def load(name: str, allow_procedural: bool = False):
    """-> (tris9 float32 (T,9), camera dict, workload name actually loaded)."""
    path = os.path.join(CACHE, name + ".tris.npy")
    if os.path.exists(path):
        tris = np.load(path)
        meta = {}
        mp = os.path.join(CACHE, name + ".json")
        if os.path.exists(mp):
            with open(mp) as f:
                meta = json.load(f)
        return np.ascontiguousarray(tris, dtype=np.float32), _camera_from_meta(meta, tris), name
    if name.startswith("synthetic"):
        n = int(name.split(":")[1]) if ":" in name else 1_000_000
        tris = synthetic_soup(n)
        return tris, _camera_from_meta({}, tris), f"synthetic:{n}"
    if not allow_procedural:
        raise FileNotFoundError(
            f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.export_scenes()'` where "
            "/root/reference exists (build() does), or pass allow_procedural=True / --allow-procedural "
            "to trace a procedural stand-in that is NOT a BASELINE.json config")
    tris, cam = procedural_room()
    return tris, cam, f"procedural_room(stand-in for {name}: scene_cache missing)"
