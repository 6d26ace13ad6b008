"""tests/test_spheres.py -- sphere leaves: Sphere::BasicIntersect / Sphere::Bounds and
BVHAggregate over primitive lists that mix triangles and spheres.

Reference: /root/reference/src/pbrt/shapes.h:107-236 (Sphere), shapes.cpp:37-40 (Bounds),
util/math.h:819-1076 (Interval), util/transform.h:133-176, 272-306.

CPU: the oracle's restatement and the host builder against golden answers produced by the reference's
own code (tools/make_golden.py --only-spheres) and against oracle/_ref live.
GPU: the device SphereTest on the golden pairs, and mixed-scene traversals through the C ABI against the
oracle -- bit-exact prim / tHit / pObj, occlusion and both counters.
"""
import ctypes
import os

import numpy as np
import pytest

from conftest import GOLDEN, rand_rays

MISS = 0xFFFFFFFF


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN, "sphere_vectors.npz"))


def _sphere_of(order, T):
    return np.where(order >= T, order - T, -1).astype(np.int32)


def _ordered(tris, order, T):
    out = np.zeros((len(order), 9), dtype=np.float32)
    m = order < T
    out[m] = tris[order[m]]
    return out


# ---------------------------------------------------------------------------- CPU
def test_oracle_sphere_matches_reference_golden(oracle, gold):
    rays, sph = gold["rays"], gold["spheres"]
    assert gold["hit"].sum() > 3000 and (gold["hit"] == 0).sum() > 300
    for i in range(len(rays)):
        h, o4 = oracle.sphere(rays[i], sph[i])
        assert h == bool(gold["hit"][i]), i
        if h:
            assert np.array_equal(o4.view(np.uint32), gold["out"][i].view(np.uint32)), i
        assert np.array_equal(oracle.sphere_bounds(sph[i]).view(np.uint32), gold["bounds"][i].view(np.uint32)), i


def test_oracle_mixed_scene_matches_reference_golden(oracle, gold):
    tris, sph, T = gold["mixed_tris"], gold["mixed_spheres"], len(gold["mixed_tris"])
    import nn_bvh_b200 as nb
    bounds = np.concatenate([nb.triangle_bounds(tris), np.stack([oracle.sphere_bounds(s) for s in sph])])
    nodes, order = oracle.build_bounds(bounds, 4, "sah")
    assert np.array_equal(nodes, gold["mixed_nodes"]) and np.array_equal(order, gold["mixed_order"])
    so, ot = _sphere_of(order, T), _ordered(tris, order, T)
    hits = oracle.intersect_mixed(nodes, ot, so, sph, gold["mixed_rays"])
    assert np.array_equal(hits, gold["mixed_hits"])
    assert np.array_equal(oracle.intersect_p_mixed(nodes, ot, so, sph, gold["mixed_rays"]), gold["mixed_occ"])
    hit = hits[:, 0] != MISS
    assert (so[hits[hit, 0]] >= 0).sum() > 1000 and (so[hits[hit, 0]] < 0).sum() > 100


def test_host_builder_and_sphere_bounds_match_reference_golden(gold):
    """nnbvh_sphere_bounds + nnbvh_build_bvh_bounds (host code of the product library)."""
    import nn_bvh_b200 as nb
    sph = np.ascontiguousarray(gold["spheres"]).view(nb.SPHERE_DTYPE).reshape(-1)
    assert np.array_equal(nb.sphere_bounds(sph).view(np.uint32), gold["bounds"].view(np.uint32))
    tris = gold["mixed_tris"]
    msph = np.ascontiguousarray(gold["mixed_spheres"]).view(nb.SPHERE_DTYPE).reshape(-1)
    bounds = np.concatenate([nb.triangle_bounds(tris), nb.sphere_bounds(msph)])
    for split in ("sah",):
        nodes, order = nb.build_bvh_bounds(bounds, 4, split, threads=1)
        assert np.array_equal(nodes.view(np.uint32).reshape(-1, 8), gold["mixed_nodes"])
        assert np.array_equal(order.astype(np.int32), gold["mixed_order"])


def test_oracle_sphere_matches_reference_live(oracle, reference):
    from tools.make_golden import random_spheres, sphere_rays
    rng = np.random.default_rng(99)
    sph = random_spheres(rng, 3000)
    rays = sphere_rays(rng, sph)
    nh = 0
    for i in range(len(sph)):
        h1, o1 = oracle.sphere(rays[i], sph[i])
        h2, o2 = reference.sphere(rays[i], sph[i])
        assert h1 == h2 and (not h1 or np.array_equal(o1.view(np.uint32), o2.view(np.uint32))), i
        nh += h1
    assert nh > 2000


@pytest.mark.parametrize("split", ["sah", "hlbvh", "middle", "equal"])
def test_mixed_scene_trees_and_traversal_match_reference_live(oracle, reference, split):
    from oracle.pyoracle import as_pbrt_stores
    from tools.make_golden import random_spheres
    import nn_bvh_b200 as nb
    rng = np.random.default_rng(5 + len(split))
    T, S = 2000, 25
    c = rng.random((T, 1, 3), dtype=np.float32)
    tris = as_pbrt_stores((c + np.float32(0.03) * (rng.random((T, 3, 3), dtype=np.float32) * 2 - 1)).reshape(T, 9))
    sph = random_spheres(rng, S, scale=0.1, center_box=0.0)
    for k in range(S):
        m = np.eye(4)
        m[:3] = sph[k, :12].reshape(3, 4)
        m[:3, 3] = rng.random(3)
        sph[k, :12] = m[:3].astype(np.float32).ravel()
        sph[k, 12:24] = np.linalg.inv(m)[:3].astype(np.float32).ravel()
    sc = reference.scene_mixed(tris, sph, 4, split)
    rnodes, rorder = reference.export(sc, T + S)
    bounds = np.concatenate([nb.triangle_bounds(tris), np.stack([oracle.sphere_bounds(s) for s in sph])])
    nodes, order = oracle.build_bounds(bounds, 4, split)
    assert np.array_equal(nodes, rnodes) and np.array_equal(order, rorder)
    pn, po = nb.build_bvh_bounds(bounds, 4, split, threads=1)          # the product's host builder
    assert np.array_equal(pn.view(np.uint32).reshape(-1, 8), rnodes) and np.array_equal(po.astype(np.int32), rorder)
    rays = rand_rays(rng, 3000)
    rays[1500:, 3] = rng.random(1500)
    so, ot = _sphere_of(order, T), _ordered(tris, order, T)
    assert np.array_equal(oracle.intersect_mixed(nodes, ot, so, sph, rays), reference.intersect(sc, rays, parallel=False))
    assert np.array_equal(oracle.intersect_p_mixed(nodes, ot, so, sph, rays),
                          reference.intersect_p(sc, rays, parallel=False))


def test_partial_phi_spheres_are_rejected_loudly():
    """phimax < 360 clips by std::atan2's value: the library refuses instead of approximating."""
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import capi
    sph = nb.make_spheres(np.eye(4)[None], 1.0, phi_max=180.0)
    # the check sits in the upload path, which needs a device; without one creation already fails loudly
    import torch
    if not torch.cuda.is_available():
        with pytest.raises(capi.NnbvhError):
            nb.GPUBVHAggregate(None, 4, "sah", spheres=sph)
        return
    with pytest.raises(capi.NnbvhError, match="phimax"):
        nb.GPUBVHAggregate(None, 4, "sah", spheres=sph)


# ---------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_device_sphere_test_matches_reference_golden(gold):
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import capi
    agg = nb.GPUBVHAggregate(None)
    rays = np.ascontiguousarray(gold["rays"])
    sph = np.ascontiguousarray(gold["spheres"])
    n = len(rays)
    hit = np.zeros(n, dtype=np.uint8)
    out = np.zeros((n, 4), dtype=np.float32)
    capi.check(capi.lib().nnbvh_test_spheres(agg._h, capi.ptr(rays), capi.ptr(sph), n, capi.ptr(hit), capi.ptr(out)))
    assert np.array_equal(hit, gold["hit"])
    assert np.array_equal(out.view(np.uint32), gold["out"].view(np.uint32))


def _mixed_scene(rng, T, S, scale):
    from tools.make_golden import random_spheres
    c = rng.random((T, 1, 3), dtype=np.float32)
    tris = ((c + np.float32(scale) * (rng.random((T, 3, 3), dtype=np.float32) * 2 - 1)).reshape(T, 9)
            + np.float32(0)).astype(np.float32)
    sph = random_spheres(rng, S, scale=3 * scale, center_box=0.0)
    for k in range(S):
        m = np.eye(4)
        m[:3] = sph[k, :12].reshape(3, 4)
        m[:3, 3] = rng.random(3)
        sph[k, :12] = m[:3].astype(np.float32).ravel()
        sph[k, 12:24] = np.linalg.inv(m)[:3].astype(np.float32).ravel()
    return tris, sph


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", [0, 1, 2])
def test_mixed_scene_gpu_matches_oracle(oracle, gold, kernel):
    """Golden mixed scene + a larger seeded one; AUTO (StepKernel), BASIC, and WIDE (which has no sphere
    leaves and must hand over to StepKernel)."""
    import nn_bvh_b200 as nb
    tris, sph, T = gold["mixed_tris"], gold["mixed_spheres"], len(gold["mixed_tris"])
    agg = nb.GPUBVHAggregate(tris, 4, "sah", kernel=kernel, spheres=sph.view(nb.SPHERE_DTYPE).reshape(-1))
    assert np.array_equal(agg.nodes.view(np.uint32).reshape(-1, 8), gold["mixed_nodes"])
    hits = agg.IntersectBatch(gold["mixed_rays"])
    assert np.array_equal(np.ascontiguousarray(hits).view(np.uint32).reshape(-1, 5), gold["mixed_hits"])
    assert np.array_equal(agg.IntersectPBatch(gold["mixed_rays"]), gold["mixed_occ"])

    rng = np.random.default_rng(11 + kernel)
    tris, sph = _mixed_scene(rng, 40_000, 300, 0.012)
    agg = nb.GPUBVHAggregate(tris, 4, "sah", kernel=kernel, spheres=sph.view(nb.SPHERE_DTYPE).reshape(-1))
    N = 300_000
    rays = rand_rays(rng, N)
    rays[N // 2:, 3] = rng.random(N - N // 2)
    rays[::97, 4] = 0.0                                  # axis-parallel rays take the deferred exact pass
    nodes = agg.nodes.view(np.uint32).reshape(-1, 8)
    so = agg.ordered_sphere_index
    want, wnv, wtt = oracle.intersect_mixed(nodes, agg.ordered_tris9, so, sph, rays, counters=True)
    got = np.ascontiguousarray(agg.IntersectBatch(rays)).view(np.uint32).reshape(-1, 5)
    assert np.array_equal(got, want)
    hit = want[:, 0] != MISS
    assert (so[want[hit, 0]] >= 0).sum() > 2000 and (so[want[hit, 0]] < 0).sum() > 2000
    wocc, onv, ott = oracle.intersect_p_mixed(nodes, agg.ordered_tris9, so, sph, rays, counters=True)
    assert np.array_equal(agg.IntersectPBatch(rays), wocc)
    # counters (BasicKernel with sphere leaves): nTriTests counts triangles only
    h2, nv, tt = agg.IntersectCounters(rays)
    assert np.array_equal(np.ascontiguousarray(h2).view(np.uint32).reshape(-1, 5), want)
    assert np.array_equal(nv, wnv) and np.array_equal(tt, wtt)
    o2, nv, tt = agg.IntersectCounters(rays, any_hit=True)
    assert np.array_equal(o2, wocc) and np.array_equal(nv, onv) and np.array_equal(tt, ott)
    # SoA planes (the wavefront queues' layout) through device memory
    import torch
    planes = [torch.from_numpy(np.ascontiguousarray(rays[:, k])).cuda() for k in (0, 1, 2, 4, 5, 6, 3)]
    dh = torch.zeros((N, 5), dtype=torch.int32, device="cuda")
    agg.IntersectBatchSoA(*planes, hits=dh)
    torch.cuda.synchronize()
    assert np.array_equal(dh.cpu().numpy().view(np.uint32), want)


@pytest.mark.gpu
def test_spheres_only_scene_and_a_single_sphere(oracle):
    import nn_bvh_b200 as nb
    rng = np.random.default_rng(3)
    _, sph = _mixed_scene(rng, 1, 500, 0.02)
    for S in (1, 500):
        agg = nb.GPUBVHAggregate(None, 4, "sah", spheres=sph[:S].view(nb.SPHERE_DTYPE).reshape(-1))
        rays = rand_rays(rng, 100_000)
        nodes = agg.nodes.view(np.uint32).reshape(-1, 8)
        want = oracle.intersect_mixed(nodes, agg.ordered_tris9, agg.ordered_sphere_index, sph[:S], rays)
        got = np.ascontiguousarray(agg.IntersectBatch(rays)).view(np.uint32).reshape(-1, 5)
        assert np.array_equal(got, want) and (want[:, 0] != MISS).sum() > (50 if S == 1 else 5000)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["killeroos", "crown"])
def test_reference_scenes_with_their_spheres(oracle, name):
    """killeroos: the r=3 spherical light; crown: the r=100 environment shell every escaping ray hits
    from inside (scene files: killeroo-simple-v4.pbrt:28, crown.pbrt:108)."""
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import rays as raygen, scenes
    tris, cam, loaded = scenes.load(name, allow_procedural=False)
    sph = scenes.load_spheres(name)
    assert sph is not None and len(sph) == 1
    agg = nb.GPUBVHAggregate(tris, 4, "sah", spheres=sph)
    T = len(tris)
    prim = raygen.camera_rays(cam, 512, 512)
    rng = np.random.default_rng(8)
    # rays aimed at the sphere from around the scene, and the camera rays
    c = sph["renderFromObject"][0].reshape(3, 4)[:, 3]
    N = 200_000
    aimed = np.zeros((N, 8), dtype=np.float32)
    lo, hi = agg.Bounds()[:3], agg.Bounds()[3:]
    aimed[:, :3] = lo + (hi - lo) * rng.random((N, 3))
    tgt = c + rng.normal(size=(N, 3)) * sph["radius"][0]
    d = tgt - aimed[:, :3]
    aimed[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    aimed[:, 3] = np.inf
    rays = np.concatenate([np.ascontiguousarray(prim).view(np.float32).reshape(-1, 8), aimed])
    sub = rng.choice(len(rays), 60_000, replace=False)
    nodes = agg.nodes.view(np.uint32).reshape(-1, 8)
    s28 = np.ascontiguousarray(sph).view(np.float32).reshape(-1, 28)
    got = np.ascontiguousarray(agg.IntersectBatch(rays)).view(np.uint32).reshape(-1, 5)
    want = oracle.intersect_mixed(nodes, agg.ordered_tris9, agg.ordered_sphere_index, s28, rays[sub])
    assert np.array_equal(got[sub], want)
    on_sphere = (got[:, 0] != MISS) & (agg.ordered_sphere_index[np.minimum(got[:, 0], T)] >= 0)
    assert on_sphere.sum() > 1000
    occ = agg.IntersectPBatch(rays)
    assert np.array_equal(occ[sub], oracle.intersect_p_mixed(nodes, agg.ordered_tris9, agg.ordered_sphere_index,
                                                            s28, rays[sub]))
    assert np.array_equal(occ != 0, got[:, 0] != MISS)   # tMax = inf: occluded <=> a closest hit exists
