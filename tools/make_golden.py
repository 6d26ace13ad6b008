#!/usr/bin/env python3
"""tools/make_golden.py -- generate tests/golden/*.npz from the UNMODIFIED reference.

Runs only in this container (needs oracle/_ref built from /root/reference).
Every array below is produced by the reference's own code through
oracle/ref_driver.cpp; the fixtures are small so they can be committed.

  tri_vectors.npz   IntersectTriangle known answers: the reference's one golden
                    vector (Triangle.BadCases, shapes_test.cpp:435-449), the
                    Reintersect-style random triangles with coordinates
                    10^U(-8,8) (shapes_test.cpp:22-25,133-154), edge/vertex hits
                    that take the double-precision fallback, degenerate triangles.
  box_vectors.npz   Bounds3f::IntersectP answers incl. zero direction components,
                    origins on slab planes (0*inf = NaN), negative zero.
  scene_small.npz   a 3000-triangle soup: reference SAH/Middle/EqualCounts trees
                    (nodes + order), closest-hit records, IntersectP results,
                    counter totals, for 4096 rays.
  interaction.npz   Triangle::InteractionFromIntersection (shapes.h:884-1010) answers on
                    indexed meshes with / without uv, normals, tangents, reversed
                    orientation, degenerate uv and zero shading normals
                    (--only-interaction regenerates just this file).
  sphere_vectors.npz  Sphere::BasicIntersect (shapes.h:147-232) and Sphere::Bounds answers for
                    4000 (ray, sphere) pairs -- rotated / scaled / translated, full and
                    z-clipped, rays from inside and outside, grazing, short tMax -- and a
                    mixed scene (3000 triangles + 40 spheres): reference tree, hit records,
                    IntersectP results (--only-spheres).
  instanced.npz     TransformedPrimitive instances (cpu/primitive.cpp:112-130): 1500 top-level
                    triangles + 30 instances of 3 objects (one a single triangle): the reference's
                    top-level and per-object trees, TransformedPrimitive::Bounds, closest hits
                    (with the primitive inside the object), IntersectP (--only-instances).
  sphere_interaction.npz  Sphere::Intersect = BasicIntersect + InteractionFromIntersection
                    (shapes.h:130-139, :238-280) for 48 spheres x 64 rays: rotated / scaled /
                    mirrored transforms, z-clipped, reverseOrientation; plus what Interaction's
                    constructor stores as wo (--only-sphere-interaction).
  sphere_phi.npz    Sphere::BasicIntersect for 4000 (ray, sphere) pairs with phimax < 360 (the std::atan2
                    clipping test, shapes.h:196-203), 1500 of them aimed within 1e-7 .. 1e-3 (relative)
                    of the phi = phimax boundary (--only-sphere-phi).
  alpha.npz         GeometricPrimitive with a constant alpha texture (cpu/primitive.cpp:50-84): HashFloat(o, d)
                    for 2000 pairs, the origin of intr.SpawnRay(r.d) for the hits of tri_vectors.npz, and a
                    4000-triangle scene (half the triangles with alpha in [0,1), a tenth with alpha 0): the
                    reference's tree, closest hits and IntersectP (--only-alpha).
  instanced_surface.npz  BVHAggregate::Intersect's SurfaceInteraction on an instanced scene (some
                    instances mirrored, ray directions not normalised): TransformedPrimitive::Intersect's
                    (*renderFromPrimitive)(si->intr) (cpu/primitive.cpp:112-125), 42 words per ray
                    (--only-instanced-surface).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.pyoracle import Reference, as_pbrt_stores  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def p_exp(rng, n):
    """pExp of shapes_test.cpp:22-25: sign * 10^U(-8,8)."""
    logu = rng.uniform(-8, 8, size=n)
    sign = np.where(rng.random(n) < 0.5, -1.0, 1.0)
    return (sign * 10.0 ** logu).astype(np.float32)


def tri_vectors(ref, rng):
    rays, tris = [], []
    # (1) Triangle.BadCases, shapes_test.cpp:435-449 (expects a miss)
    rays.append([-0.36098, 3.2618, 2.6918, np.inf, 0.0061, -0.0014, -0.9999, 0.0])  # placeholder 0
    tris.append([5.0, 8.0, 9.0, 5.0, 8.0, 9.0, 5.0, 8.0, 9.0])
    # (2) random triangles over 16 orders of magnitude, rays aimed at interior points
    n = 3000
    p = p_exp(rng, n * 9).reshape(n, 3, 3)
    b = rng.dirichlet([1, 1, 1], size=n)
    target = (p.astype(np.float64) * b[:, :, None]).sum(1)
    o = p_exp(rng, n * 3).reshape(n, 3).astype(np.float64)
    d = target - o
    for i in range(n):
        tm = np.inf if i % 3 else 1.0
        dd = d[i] / (np.linalg.norm(d[i]) if i % 2 else 1.0)
        rays.append([*o[i], tm, *dd, 0.0])
        tris.append(p[i].ravel())
    # (3) rays through vertices / edge midpoints of well-conditioned triangles (e == 0 paths)
    n = 2000
    p = rng.integers(-8, 9, size=(n, 3, 3)).astype(np.float32)
    for i in range(n):
        k = i % 6
        tgt = p[i, k] if k < 3 else 0.5 * (p[i, k - 3] + p[i, (k - 2) % 3])
        o = rng.integers(-20, 21, size=3).astype(np.float32)
        rays.append([*o, np.inf, *(tgt - o), 0.0])
        tris.append(p[i].ravel())
    # (4) axis-aligned rays (zero direction components), small random scenes
    n = 1000
    p = rng.random((n, 3, 3)).astype(np.float32)
    for i in range(n):
        d = np.zeros(3, np.float32)
        d[i % 3] = 1.0 if (i // 3) % 2 else -1.0
        c = p[i].mean(0)
        o = c - d * 2
        if i % 5 == 0:
            o = o + rng.normal(size=3).astype(np.float32) * 0.3
        rays.append([*o, np.inf, *d, 0.0])
        tris.append(p[i].ravel())
    # (5) degenerate triangles
    for i in range(50):
        a = rng.random(3).astype(np.float32)
        bb = rng.random(3).astype(np.float32)
        tri = [a, bb, a + (bb - a) * np.float32(0.5)] if i % 2 else [a, a, bb]
        rays.append([0, 0, -5, np.inf, *(a - np.array([0, 0, -5], np.float32)), 0.0])
        tris.append(np.concatenate(tri))
    rays = np.array(rays, dtype=np.float32)
    tris = np.array(tris, dtype=np.float32)
    # the real BadCases vector (shapes_test.cpp:438-446)
    tris[0] = [-1113.45459, -79.049614, -56.2431908, -1113.45459, -87.0922699, -56.2431908,
               -1113.45459, -79.2090149, -56.2431908]
    rays[0] = [-1081.47925, 99.9999542, 87.7701111, np.inf, -32.1072998, -183.355865,
               -144.607635, 0.014]
    hit = np.zeros(len(rays), np.uint8)
    out = np.zeros((len(rays), 4), np.float32)
    for i in range(len(rays)):
        h, o4 = ref.triangle(rays[i], tris[i])
        hit[i] = h
        out[i] = o4 if h else 0
    assert hit[0] == 0, "reference BadCases vector must miss"
    np.savez_compressed(os.path.join(OUT, "tri_vectors.npz"), rays=rays, tris=tris, hit=hit, out=out)
    print("tri_vectors", len(rays), "hits", int(hit.sum()))


def box_vectors(ref, rng):
    n = 6000
    lo = rng.integers(-4, 4, size=(n, 3)).astype(np.float32)
    hi = lo + rng.integers(0, 4, size=(n, 3)).astype(np.float32)  # some flat boxes
    o = rng.integers(-6, 7, size=(n, 3)).astype(np.float32)       # often ON a slab plane
    d = rng.integers(-2, 3, size=(n, 3)).astype(np.float32)       # many zero components
    d[np.all(d == 0, axis=1)] = [0, 0, 1]
    neg0 = rng.random((n, 3)) < 0.15
    d = np.where((d == 0) & neg0, np.float32(-0.0), d)
    tmax = np.where(rng.random(n) < 0.5, np.inf, rng.random(n) * 8).astype(np.float32)
    # second half: generic floats
    h = n // 2
    lo[h:] = rng.normal(size=(n - h, 3)).astype(np.float32)
    hi[h:] = lo[h:] + np.abs(rng.normal(size=(n - h, 3))).astype(np.float32)
    o[h:] = rng.normal(size=(n - h, 3)).astype(np.float32) * 3
    d[h:] = rng.normal(size=(n - h, 3)).astype(np.float32)
    aim = rng.random(n) < 0.6   # aim most rays at a point of (or on the surface of) the box
    w = rng.integers(0, 3, size=(n, 3)).astype(np.float32) / 2   # 0, .5, 1 -> faces/edges/corners
    tgt = lo + (hi - lo) * w
    d = np.where(aim[:, None] & np.any(tgt != o, axis=1)[:, None], tgt - o, d).astype(np.float32)
    rays = np.zeros((n, 8), np.float32)
    rays[:, :3], rays[:, 3], rays[:, 4:7] = o, tmax, d
    boxes = np.concatenate([lo, hi], 1)
    hit = np.array([ref.box(rays[i], boxes[i]) for i in range(n)], dtype=np.uint8)
    np.savez_compressed(os.path.join(OUT, "box_vectors.npz"), rays=rays, boxes=boxes, hit=hit)
    print("box_vectors", n, "hits", int(hit.sum()))


def scene_small(ref, rng):
    T = 3000
    # clustered soup with shared vertices (grid patches) + duplicates -> ties, leaves of 1-4
    g = 20
    u, v = np.meshgrid(np.arange(g + 1), np.arange(g + 1), indexing="ij")
    P = np.stack([u, v, np.sin(u * 0.7) * np.cos(v * 0.5) * 2], -1).astype(np.float32) / 10
    q = np.stack([P[:-1, :-1], P[1:, :-1], P[1:, 1:], P[:-1, 1:]], 2).reshape(-1, 4, 3)
    grid = np.concatenate([q[:, [0, 1, 2]], q[:, [2, 3, 0]]], 0).reshape(-1, 9)   # 800 tris
    c = rng.random((T - 2 * len(grid), 1, 3), dtype=np.float32) * 2
    soup = (c + 0.15 * (rng.random((len(c), 3, 3), dtype=np.float32) * 2 - 1)).reshape(-1, 9)
    tris = as_pbrt_stores(np.concatenate([grid, grid, soup], 0))   # grid twice: exact ties
    N = 4096
    rays = np.zeros((N, 8), np.float32)
    rays[:, :3] = rng.random((N, 3)) * 3 - 0.5
    d = rng.normal(size=(N, 3))
    rays[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays[:, 3] = np.inf
    # a block of straight-down rays through grid vertices (shared edges/vertices, duplicates)
    k = 600
    rays[:k, 0] = rng.integers(0, g + 1, size=k) / 10
    rays[:k, 1] = rng.integers(0, g + 1, size=k) / 10
    rays[:k, 2] = 3.0
    rays[:k, 4:7] = [0, 0, -1]
    rays[N // 2:, 3] = rng.random(N - N // 2) * 1.5   # finite tMax on half
    out = {"tris": tris, "rays": rays}
    for split in ("sah", "middle", "equal"):
        sc = ref.scene(tris, 4, split)
        nodes, order = ref.export(sc, len(tris))
        ref.stats()
        hits = ref.intersect(sc, rays, parallel=False)
        st_c = ref.stats()
        occ = ref.intersect_p(sc, rays, parallel=False)
        st_p = ref.stats()
        out.update({f"{split}_nodes": nodes, f"{split}_order": order, f"{split}_hits": hits,
                    f"{split}_occ": occ,
                    f"{split}_stats": np.array([st_c[0], st_c[1], st_p[0], st_p[1]], np.int64)})
        print(split, "nodes", len(nodes), "hit frac", (hits[:, 0] != 0xFFFFFFFF).mean(), st_c, st_p)
    np.savez_compressed(os.path.join(OUT, "scene_small.npz"), **out)


def hlbvh(ref):
    """HLBVH trees of the unmodified reference (cpu/aggregates.cpp:389-503, 626-723): the
    scene_small triangles (duplicates -> equal Morton codes) and a soup wide enough for a few
    thousand treelets, at two leaf sizes.  Own seed, own file, so the other goldens stay put."""
    rng = np.random.default_rng(20261105)
    small = np.load(os.path.join(OUT, "scene_small.npz"))
    c = rng.random((12000, 1, 3), dtype=np.float32)
    soup = as_pbrt_stores((c + 0.01 * (rng.random((12000, 3, 3), dtype=np.float32) * 2 - 1)).reshape(-1, 9))
    out = {"soup": soup}
    for name, tris, mp in (("small4", small["tris"], 4), ("soup4", soup, 4), ("soup1", soup, 1),
                           ("soup255", soup, 255)):
        sc = ref.scene(tris, mp, "hlbvh")
        nodes, order = ref.export(sc, len(tris))
        out[f"{name}_nodes"], out[f"{name}_order"] = nodes, order
        if name == "small4":
            out["small4_hits"] = ref.intersect(sc, small["rays"], parallel=False)
            out["small4_occ"] = ref.intersect_p(sc, small["rays"], parallel=False)
        print("hlbvh", name, "nodes", len(nodes))
    np.savez_compressed(os.path.join(OUT, "hlbvh.npz"), **out)


def interaction_cases(rng):
    """Meshes + queries for InteractionFromIntersection; shared with tests/test_wavefront.py."""
    cases = {}
    V, T, Q = 60, 100, 400
    for name in ("plain", "uv", "uv_n", "n_only", "uv_n_s", "reversed", "degenerate_uv", "zero_ns"):
        verts = as_pbrt_stores((rng.random((V, 3)) * 4 - 2).astype(np.float32))
        idx = np.stack([rng.permutation(V)[:3] for _ in range(T)]).astype(np.int32)
        uv = n = s = None
        if "uv" in name:
            uv = rng.random((V, 2)).astype(np.float32)
        if name == "degenerate_uv":
            uv = np.repeat(rng.random((1, 2)).astype(np.float32), V, axis=0)   # all equal: determinant 0
            uv[::7] += np.float32(1e-6)                                         # and almost equal
        if "_n" in name or name in ("n_only", "reversed", "degenerate_uv", "zero_ns"):
            n = rng.normal(size=(V, 3)).astype(np.float32)
            n /= np.linalg.norm(n, axis=1, keepdims=True).astype(np.float32)
            n = as_pbrt_stores(n.astype(np.float32))
        if name == "zero_ns":
            n[::3] = 0                                                          # LengthSquared(ns) == 0 paths
        if "_s" in name:
            s = as_pbrt_stores(rng.normal(size=(V, 3)).astype(np.float32))
            s[::5] = 0                                                          # ss falls back to dpdu
        tri = rng.integers(0, T, Q).astype(np.int32)
        b = rng.random((Q, 3)).astype(np.float32)
        b /= b.sum(axis=1, keepdims=True)
        b[::11] = [1, 0, 0]
        wo_time = rng.normal(size=(Q, 4)).astype(np.float32)
        cases[name] = dict(verts=verts, idx=idx, uv=uv, n=n, s=s, tri=tri, b=b.astype(np.float32),
                           wo_time=wo_time, reverse=(name == "reversed"))
    return cases


def interaction(ref):
    out = {}
    for name, c in interaction_cases(np.random.default_rng(884)).items():
        out[name + "_want"] = ref.triangle_interaction(c["verts"], c["idx"], c["tri"], c["b"], c["wo_time"],
                                                       normals=c["n"], tangents=c["s"], uvs=c["uv"],
                                                       reverse_orientation=c["reverse"])
        for k, v in c.items():   # the inputs travel with the answers
            if v is not None:
                out[name + "_" + k] = np.asarray(v)
        print("interaction", name, out[name + "_want"].shape)
    np.savez_compressed(os.path.join(OUT, "interaction.npz"), **out)


def random_spheres(rng, n, clip_every=3, scale=1.0, center_box=2.0):
    """n nnbvh_sphere records (28 floats): rotation * uniform scale + translation."""
    out = np.zeros((n, 28), dtype=np.float32)
    for i in range(n):
        q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        m = np.eye(4)
        m[:3, :3] = q * np.exp(rng.uniform(-1, 1))
        m[:3, 3] = rng.uniform(-center_box, center_box, 3)
        rad = scale * np.exp(rng.uniform(-2, 1))
        out[i, :12] = m[:3].ravel()
        out[i, 12:24] = np.linalg.inv(m)[:3].ravel()
        out[i, 24] = rad
        if clip_every and i % clip_every == 0:
            out[i, 25], out[i, 26] = -rad * rng.uniform(0.2, 1.2), rad * rng.uniform(0.2, 1.2)
        else:
            out[i, 25], out[i, 26] = -rad, rad
        out[i, 27] = 360
    return out


def sphere_rays(rng, spheres):
    """One ray per sphere: aimed near it from outside / inside, normalised or not, grazing, short tMax."""
    n = len(spheres)
    rays = np.zeros((n, 8), dtype=np.float32)
    for i, s in enumerate(spheres):
        c = s[[3, 7, 11]].astype(np.float64)
        k = abs(np.linalg.det(s[:12].reshape(3, 4)[:, :3].astype(np.float64))) ** (1 / 3)
        r_world = s[24] * k
        o = c + rng.normal(size=3) * r_world * np.exp(rng.uniform(-1, 1.5))
        aim = 1.0 if i % 7 else 1.0 + rng.normal() * 1e-3            # grazing every 7th
        tgt = c + (lambda v: v / np.linalg.norm(v))(rng.normal(size=3)) * r_world * aim * (rng.uniform(0, 1) if i % 7 else 1)
        d = tgt - o
        if i % 2:
            d /= np.linalg.norm(d)
        rays[i, :3], rays[i, 4:7] = o, d
        rays[i, 3] = np.inf if i % 5 else rng.uniform(0, 3)
        rays[i, 7] = rng.random()
    return rays


def sphere_vectors(ref, rng):
    n = 4000
    spheres = random_spheres(rng, n)
    rays = sphere_rays(rng, spheres)
    hit = np.zeros(n, dtype=np.uint8)
    out = np.zeros((n, 4), dtype=np.float32)
    bounds = np.zeros((n, 6), dtype=np.float32)
    for i in range(n):
        h, o4 = ref.sphere(rays[i], spheres[i])
        hit[i] = h
        out[i] = o4 if h else 0
        bounds[i] = ref.sphere_bounds(spheres[i])
    # a mixed primitive list: 3000 small triangles in [0,1]^3 and 40 spheres among them
    T = 3000
    c = rng.random((T, 1, 3), dtype=np.float32)
    tris = as_pbrt_stores((c + np.float32(0.03) * (rng.random((T, 3, 3), dtype=np.float32) * 2 - 1)).reshape(T, 9))
    msph = random_spheres(rng, 40, scale=0.08, center_box=0.0)
    for k in range(40):                      # centres inside the soup
        m = np.eye(4)
        m[:3] = msph[k, :12].reshape(3, 4)
        m[:3, 3] = rng.random(3)
        msph[k, :12] = m[:3].astype(np.float32).ravel()
        msph[k, 12:24] = np.linalg.inv(m)[:3].astype(np.float32).ravel()
    N = 4096
    mr = np.zeros((N, 8), dtype=np.float32)
    mr[:, :3] = rng.random((N, 3)) * 1.5 - 0.25
    tgt = msph[rng.integers(0, 40, N)][:, [3, 7, 11]] + rng.normal(size=(N, 3)) * 0.05
    d = np.where((np.arange(N) % 2 == 0)[:, None], tgt - mr[:, :3], rng.normal(size=(N, 3)))
    mr[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    mr[:, 3] = np.inf
    mr[N // 2:, 3] = rng.random(N - N // 2) * 1.5
    sc = ref.scene_mixed(tris, msph, 4, "sah")
    nodes, order = ref.export(sc, T + 40)
    hits = ref.intersect(sc, mr, parallel=False)
    occ = ref.intersect_p(sc, mr, parallel=False)
    print("sphere_vectors: pair hits", int(hit.sum()), "of", n, "| mixed scene nodes", len(nodes),
          "sphere hits", int((order[hits[hits[:, 0] != 0xFFFFFFFF, 0]] >= T).sum()), "occluded", int(occ.sum()))
    np.savez_compressed(os.path.join(OUT, "sphere_vectors.npz"), rays=rays, spheres=spheres, hit=hit, out=out,
                        bounds=bounds, mixed_tris=tris, mixed_spheres=msph, mixed_rays=mr, mixed_nodes=nodes,
                        mixed_order=order, mixed_hits=hits, mixed_occ=occ)


def instanced_scene(rng, T=1500, K=30, obj_sizes=(400, 1, 900)):
    """(top-level soup, object soups, inst24 (K,24), inst_object (K,)): rotations, uniform scales,
    translations into the unit cube."""
    def soup(n, s, lo, hi):
        c = lo + (hi - lo) * rng.random((n, 1, 3))
        return as_pbrt_stores((c + s * (rng.random((n, 3, 3)) * 2 - 1)).reshape(n, 9).astype(np.float32))
    top = soup(T, 0.03, 0.0, 1.0)
    objs = [soup(n, 0.05 if n < 600 else 0.03, -0.2, 0.2) for n in obj_sizes]
    inst = np.zeros((K, 24), dtype=np.float32)
    for k in range(K):
        q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        if np.linalg.det(q) < 0:
            q[:, 0] = -q[:, 0]
        m = np.eye(4)
        m[:3, :3] = q * np.exp(rng.uniform(-1, 0.3))
        m[:3, 3] = rng.random(3)
        inst[k, :12] = m[:3].ravel()
        inst[k, 12:] = np.linalg.inv(m)[:3].ravel()
    return top, objs, inst, rng.integers(0, len(objs), K).astype(np.int32)


def instanced(ref, rng):
    top, objs, inst, io = instanced_scene(rng)
    sc = ref.scene_instanced(top, objs, inst, io, 4, "sah")
    N = 4096
    rays = np.zeros((N, 8), dtype=np.float32)
    rays[:, :3] = rng.random((N, 3)) * 1.5 - 0.25
    d = rng.normal(size=(N, 3))
    rays[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays[:, 3] = np.inf
    rays[N // 2:, 3] = rng.random(N - N // 2)
    hits, inner = ref.intersect_instanced(sc["handle"], rays)
    occ = ref.intersect_p(sc["handle"], rays, parallel=False)
    print("instanced: nodes", len(sc["data"]["nodes"]), "instance hits", int((inner != 0xFFFFFFFF).sum()),
          "triangle hits", int(((hits[:, 0] != 0xFFFFFFFF) & (inner == 0xFFFFFFFF)).sum()), "occluded", int(occ.sum()))
    out = dict(top=top, inst24=inst, inst_object=io, rays=rays, hits=hits, inner=inner, occ=occ, order=sc["order"],
               instance_bounds=sc["instance_bounds"], n_objects=np.int32(len(objs)))
    for j, o in enumerate(objs):
        out[f"obj{j}"] = o
        out[f"obj{j}_order"] = sc["obj_orders"][j]
    out.update({"sc_" + k: v for k, v in sc["data"].items()})
    np.savez_compressed(os.path.join(OUT, "instanced.npz"), **out)


def sphere_interaction(ref, rng):
    n_s, n_r = 48, 64
    spheres = random_spheres(rng, n_s)
    for i in range(0, n_s, 4):               # mirrored: the transform swaps handedness
        m = np.eye(4)
        m[:3] = spheres[i, :12].reshape(3, 4).astype(np.float64)
        m[:3, 0] *= -1
        spheres[i, :12] = m[:3].ravel()
        spheres[i, 12:24] = np.linalg.inv(m)[:3].ravel()
    reverse = (np.arange(n_s) % 3 == 0)
    rays = np.zeros((n_s, n_r, 8), dtype=np.float32)
    hit = np.zeros((n_s, n_r), dtype=bool)
    t_hit = np.zeros((n_s, n_r), dtype=np.float32)
    p_obj = np.zeros((n_s, n_r, 3), dtype=np.float32)
    surf = np.zeros((n_s, n_r, 42), dtype=np.float32)
    for i in range(n_s):
        rays[i] = sphere_rays(rng, np.repeat(spheres[i][None], n_r, 0))
        hit[i], t_hit[i], surf[i] = ref.sphere_interaction(spheres[i], reverse[i], rays[i])
        for k in range(n_r):
            h, o4 = ref.sphere(rays[i, k], spheres[i])
            assert h == hit[i, k]
            if h:
                p_obj[i, k] = o4[1:]
    wo = rng.normal(size=(512, 3)).astype(np.float32) * np.exp(rng.uniform(-3, 3, (512, 1))).astype(np.float32)
    print("sphere_interaction:", int(hit.sum()), "hits of", hit.size)
    np.savez_compressed(os.path.join(OUT, "sphere_interaction.npz"), spheres=spheres, reverse=reverse, rays=rays,
                        hit=hit, t_hit=t_hit, p_obj=p_obj, surf=surf.view(np.uint32), wo=wo,
                        wo_stored=ref.interaction_wo(wo))


def instanced_surface(ref, rng):
    top, objs, inst, io = instanced_scene(rng, T=800, K=24, obj_sizes=(300, 1, 500))
    for k in range(0, len(inst), 5):         # mirrored instances: the transform swaps handedness
        m = np.eye(4)
        m[:3] = inst[k, :12].reshape(3, 4).astype(np.float64)
        m[:3, 1] *= -1
        inst[k, :12] = m[:3].ravel()
        inst[k, 12:] = np.linalg.inv(m)[:3].ravel()
    sc = ref.scene_instanced(top, objs, inst, io, 4, "sah")
    N = 3000
    rays = np.zeros((N, 8), dtype=np.float32)
    rays[:, :3] = rng.random((N, 3)) * 1.5 - 0.25
    d = rng.normal(size=(N, 3))
    rays[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True) * np.exp(rng.uniform(-1, 1, (N, 1)))
    rays[:, 3] = np.inf
    rays[:, 7] = rng.random(N)
    hits, inner = ref.intersect_instanced(sc["handle"], rays)
    hit, t_hit, surf = ref.intersect_surface(sc["handle"], rays)
    assert np.array_equal(hit, hits[:, 0] != 0xFFFFFFFF)
    print("instanced_surface: instance hits", int((inner != 0xFFFFFFFF).sum()), "of", int(hit.sum()), "hits")
    out = dict(top=top, inst24=inst, inst_object=io, rays=rays, hits=hits, inner=inner, order=sc["order"],
               t_hit=t_hit, surf=surf.view(np.uint32), n_objects=np.int32(len(objs)))
    for j, o in enumerate(objs):
        out[f"obj{j}"] = o
        out[f"obj{j}_order"] = sc["obj_orders"][j]
    np.savez_compressed(os.path.join(OUT, "instanced_surface.npz"), **out)


def phi_spheres(rng, n):
    """random_spheres with phimax in (-20, 400) (clamped by the constructor) and a few exact values."""
    s = random_spheres(rng, n)
    s[:, 27] = rng.uniform(-20, 400, n).astype(np.float32)
    s[::7, 27] = rng.choice([0, 90, 180, 270, 359.99, 360], len(s[::7]))
    return s


def phi_boundary_rays(rng, spheres):
    """One ray per sphere whose first hit lies at phi = phimax * (1 +- 1e-7 .. 1e-3)."""
    n = len(spheres)
    rays = np.zeros((n, 8), dtype=np.float32)
    for i, s in enumerate(spheres):
        m = np.eye(4)
        m[:3] = s[:12].reshape(3, 4).astype(np.float64)
        r = float(s[24])
        phi = np.radians(np.clip(float(s[27]), 0, 360)) * (1 + rng.choice([-1, 1]) * 10 ** rng.uniform(-7, -3))
        ct = rng.uniform(-0.9, 0.9)
        st = np.sqrt(1 - ct * ct)
        p_obj = np.array([r * st * np.cos(phi), r * st * np.sin(phi), r * ct])
        o_obj = p_obj * rng.uniform(1.5, 4)                   # outside, on the hit point's normal: first hit is p
        o = m[:3, :3] @ o_obj + m[:3, 3]
        p = m[:3, :3] @ p_obj + m[:3, 3]
        d = p - o
        if i % 2:
            d /= np.linalg.norm(d)
        rays[i, :3], rays[i, 4:7], rays[i, 3], rays[i, 7] = o, d, np.inf, rng.random()
    return rays


def sphere_phi(ref, rng):
    n, nb = 2500, 1500
    spheres = phi_spheres(rng, n + nb)
    spheres[n:, 25], spheres[n:, 26] = -spheres[n:, 24], spheres[n:, 24]      # boundary cases: no z clipping
    spheres[n:, 27] = rng.uniform(5, 355, nb).astype(np.float32)
    rays = np.concatenate([sphere_rays(rng, spheres[:n]), phi_boundary_rays(rng, spheres[n:])])
    hit = np.zeros(n + nb, dtype=np.uint8)
    out = np.zeros((n + nb, 4), dtype=np.float32)
    for i in range(n + nb):
        h, o4 = ref.sphere(rays[i], spheres[i])
        hit[i] = h
        out[i] = o4 if h else 0
    print("sphere_phi: hits", int(hit[:n].sum()), "of", n, "; boundary hits", int(hit[n:].sum()), "of", nb)
    np.savez_compressed(os.path.join(OUT, "sphere_phi.npz"), spheres=spheres, rays=rays, hit=hit, out=out)


def alpha(ref, rng):
    od = rng.normal(size=(2000, 6)).astype(np.float32) * np.exp(rng.uniform(-3, 3, (2000, 1))).astype(np.float32)
    hashes = np.array([ref.hash_float(v[:3], v[3:]) for v in od], dtype=np.float32)
    tv = np.load(os.path.join(OUT, "tri_vectors.npz"))
    n = len(tv["rays"])
    spawn_hit = np.zeros(n, dtype=np.uint8)
    spawn = np.zeros((n, 4), dtype=np.float32)
    for i in range(n):
        h, o4 = ref.spawn_origin(tv["rays"][i], tv["tris"][i])
        spawn_hit[i] = h
        spawn[i] = o4 if h else 0
    T = 4000
    c = rng.random((T, 1, 3), dtype=np.float32)
    tris = as_pbrt_stores((c + np.float32(0.05) * (rng.random((T, 3, 3), dtype=np.float32) * 2 - 1)).reshape(T, 9))
    a = np.ones(T, dtype=np.float32)
    a[::2] = rng.random(len(a[::2])).astype(np.float32)
    a[::10] = 0.0
    sc = ref.scene_alpha(tris, a, 4, "sah")
    nodes, order = ref.export(sc, T)
    N = 8192
    rays = np.zeros((N, 8), dtype=np.float32)
    rays[:, :3] = rng.random((N, 3)) * 1.5 - 0.25
    d = rng.normal(size=(N, 3))
    rays[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays[:, 3] = np.inf
    rays[N // 2:, 3] = rng.random(N - N // 2)
    hits = ref.intersect(sc, rays, parallel=False)
    occ = ref.intersect_p(sc, rays, parallel=False)
    print("alpha: spawn hits", int(spawn_hit.sum()), "; scene hits", int((hits[:, 0] != 0xFFFFFFFF).sum()), "occluded", int(occ.sum()))
    np.savez_compressed(os.path.join(OUT, "alpha.npz"), od=od, hashes=hashes, spawn_hit=spawn_hit, spawn=spawn,
                        tris=tris, alpha=a, nodes=nodes, order=order, rays=rays, hits=hits, occ=occ)


def general(ref, rng):
    """tests/golden/general.npz: ONE list with triangles, constant-alpha triangles, spheres and instances -- the
    reference's tree over it and BVHAggregate::Intersect / IntersectP answers (tests/test_general_scene.py)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from test_general_scene import _rays, _scene
    tris, alpha, sph, soups, inst24, inst_obj = _scene(rng, 1500, 16, 2, 10)
    g = ref.scene_general(tris, alpha, sph, soups, inst24, inst_obj, 4, "sah")
    rays = _rays(rng, 8192, sph, inst24)
    hits, inner = ref.intersect_general(g["handle"], rays)
    occ = ref.intersect_p(g["handle"], rays, parallel=False)
    T, S = len(tris), len(sph)
    k = g["order"][hits[hits[:, 0] != 0xFFFFFFFF, 0]]
    print("general: hits on triangles", int((k < T).sum()), "spheres", int(((k >= T) & (k < T + S)).sum()),
          "instances", int((k >= T + S).sum()), "occluded", int(occ.sum()))
    np.savez_compressed(os.path.join(OUT, "general.npz"), nodes=g["nodes"], order=g["order"],
                        ordered_tris9=g["ordered_tris9"], ordered_alpha=g["ordered_alpha"], sphere_of=g["sphere_of"],
                        instance_of=g["instance_of"], spheres28=sph, inst24=inst24, inst_object=inst_obj,
                        obj_nodes=np.concatenate([o["nodes"] for o in g["objects"]]),
                        obj_node_count=np.array([len(o["nodes"]) for o in g["objects"]], dtype=np.int32),
                        obj_tris=np.concatenate([o["ordered_tris9"] for o in g["objects"]]),
                        obj_tri_count=np.array([len(o["ordered_tris9"]) for o in g["objects"]], dtype=np.int32),
                        rays=rays, hits=hits, inner=inner, occ=occ)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    ref = Reference(threads=1)
    if "--only-general" in sys.argv:
        general(ref, np.random.default_rng(211))
        sys.exit(0)
    if "--only-instances" in sys.argv:
        instanced(ref, np.random.default_rng(112))
        sys.exit(0)
    if "--only-spheres" in sys.argv:
        sphere_vectors(ref, np.random.default_rng(147))
        sys.exit(0)
    if "--only-interaction" in sys.argv:
        interaction(ref)
        sys.exit(0)
    if "--only-sphere-interaction" in sys.argv:
        sphere_interaction(ref, np.random.default_rng(238))
        sys.exit(0)
    if "--only-instanced-surface" in sys.argv:
        instanced_surface(ref, np.random.default_rng(415))
        sys.exit(0)
    if "--only-sphere-phi" in sys.argv:
        sphere_phi(ref, np.random.default_rng(196))
        sys.exit(0)
    if "--only-alpha" in sys.argv:
        alpha(ref, np.random.default_rng(606))
        sys.exit(0)
    if "--only-hlbvh" in sys.argv:
        hlbvh(ref)
        sys.exit(0)
    rng = np.random.default_rng(20261104)
    tri_vectors(ref, rng)
    box_vectors(ref, rng)
    scene_small(ref, rng)
    hlbvh(ref)
    interaction(ref)
    sphere_vectors(ref, np.random.default_rng(147))
    instanced(ref, np.random.default_rng(112))
    sphere_interaction(ref, np.random.default_rng(238))
    instanced_surface(ref, np.random.default_rng(415))
    sphere_phi(ref, np.random.default_rng(196))
    alpha(ref, np.random.default_rng(606))
