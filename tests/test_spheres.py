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

from conftest import GOLDEN, ROOT, rand_rays

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


@pytest.fixture(scope="module")
def gold_phi():
    return np.load(os.path.join(GOLDEN, "sphere_phi.npz"))


def test_oracle_partial_phi_matches_reference_golden(oracle, gold_phi):
    """phimax < 360: the clipping test compares std::atan2's value (shapes.h:196-203, :218-222)."""
    g = gold_phi
    for i in range(len(g["rays"])):
        h, o4 = oracle.sphere(g["rays"][i], g["spheres"][i])
        assert h == bool(g["hit"][i]), i
        if h:
            assert np.array_equal(o4.view(np.uint32), g["out"][i].view(np.uint32)), i
    assert 1000 < g["hit"][2500:].sum() < 1400           # the boundary cases fall on both sides


def test_oracle_partial_phi_matches_reference_live(oracle, reference):
    from tools.make_golden import phi_boundary_rays, phi_spheres, sphere_rays
    rng = np.random.default_rng(303)
    sph = phi_spheres(rng, 3000)
    sph[2000:, 25], sph[2000:, 26] = -sph[2000:, 24], sph[2000:, 24]
    sph[2000:, 27] = rng.uniform(1, 359, 1000).astype(np.float32)
    rays = np.concatenate([sphere_rays(rng, sph[:2000]), phi_boundary_rays(rng, sph[2000:])])
    for i in range(len(rays)):
        a, b = reference.sphere(rays[i], sph[i]), oracle.sphere(rays[i], sph[i])
        assert a[0] == b[0] and (not a[0] or np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32))), i


def test_atan2_restatement_matches_host_libm(tmp_path):
    """tools/atan2_check.c: the host twin of csrc/atan2_host.cuh against this machine's atan2f on 3e7
    inputs (arbitrary bit patterns and pairs in [-1,1]^2); 1e9 were run when the file was written."""
    import subprocess
    exe = str(tmp_path / "atan2_check")
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fno-builtin", "-o", exe,
                           os.path.join(ROOT, "tools", "atan2_check.c"), "-lm"])
    out = subprocess.run([exe, "30000000"], capture_output=True, text=True)
    assert out.returncode == 0 and "bad=0" in out.stdout, out.stdout[-600:]


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
    if not os.path.exists(os.path.join(ROOT, "scene_cache", name + ".tris.npy")):
        pytest.skip(f"scene_cache/{name} not present on this machine (build() exports it where "
                    "/root/reference exists)")
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



def _atan2_inputs(rng, n):
    """(y, x): arbitrary bit patterns, pairs in [-1,1]^2, huge / tiny ratios, zeros, infinities, x = 1."""
    y = rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32).view(np.float32).copy()
    x = rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32).view(np.float32).copy()
    q = n // 4
    y[:q], x[:q] = rng.uniform(-1, 1, q), rng.uniform(-1, 1, q)
    y[q:2 * q] = (rng.normal(size=q) * np.exp2(rng.uniform(-40, 40, q))).astype(np.float32)
    x[q:2 * q] = (rng.normal(size=q) * np.exp2(rng.uniform(-40, 40, q))).astype(np.float32)
    sp = np.array([0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, np.nan, 1e-45, -1e-45, 3.4e38], dtype=np.float32)
    y[2 * q:2 * q + 100], x[2 * q:2 * q + 100] = np.repeat(sp, 10), np.tile(sp, 10)
    x[2 * q + 100:2 * q + 5000] = 1.0
    return y, x


@pytest.mark.gpu
def test_device_atan2_matches_host_libm(oracle):
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import capi
    agg = nb.GPUBVHAggregate(None)
    y, x = _atan2_inputs(np.random.default_rng(8), 8_000_000)
    got = np.zeros_like(y)
    capi.check(capi.lib().nnbvh_test_atan2(agg._h, capi.ptr(y), capi.ptr(x), y.size, capi.ptr(got)))
    want = oracle.atan2f(y, x)
    nan = np.isnan(want)
    assert np.array_equal(np.isnan(got), nan)
    assert np.array_equal(got.view(np.uint32)[~nan], want.view(np.uint32)[~nan])


@pytest.mark.gpu
def test_device_partial_phi_spheres_match_reference_golden(gold_phi):
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import capi
    agg = nb.GPUBVHAggregate(None)
    rays, sph = np.ascontiguousarray(gold_phi["rays"]), np.ascontiguousarray(gold_phi["spheres"])
    n = len(rays)
    hit = np.zeros(n, dtype=np.uint8)
    out = np.zeros((n, 4), dtype=np.float32)
    capi.check(capi.lib().nnbvh_test_spheres(agg._h, capi.ptr(rays), capi.ptr(sph), n, capi.ptr(hit), capi.ptr(out)))
    assert np.array_equal(hit, gold_phi["hit"])
    assert np.array_equal(out.view(np.uint32), gold_phi["out"].view(np.uint32))


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", [0, 1])
def test_mixed_scene_with_partial_phi_spheres_matches_oracle(oracle, kernel):
    """Triangles + spheres clipped in z and in phi behind the traversal kernels (StepKernel, BasicKernel)."""
    import nn_bvh_b200 as nb
    rng = np.random.default_rng(29 + kernel)
    tris, sph = _mixed_scene(rng, 20_000, 400, 0.012)
    sph[:, 27] = rng.uniform(20, 400, len(sph)).astype(np.float32)
    agg = nb.GPUBVHAggregate(tris, 4, "sah", kernel=kernel, spheres=sph.view(nb.SPHERE_DTYPE).reshape(-1))
    N = 300_000
    rays = rand_rays(rng, N)
    aim = rng.integers(0, len(sph), N // 2)                      # half the rays aimed at a sphere
    rays[:N // 2, 4:7] = sph[aim][:, [3, 7, 11]] + rng.normal(size=(N // 2, 3)).astype(np.float32) * 0.02 - rays[:N // 2, :3]
    rays[N // 2::2, 3] = rng.random(len(rays[N // 2::2]))
    nodes, so = agg.nodes.view(np.uint32).reshape(-1, 8), agg.ordered_sphere_index
    want = oracle.intersect_mixed(nodes, agg.ordered_tris9, so, sph, rays)
    got = np.ascontiguousarray(agg.IntersectBatch(rays)).view(np.uint32).reshape(-1, 5)
    assert np.array_equal(got, want)
    hit = want[:, 0] != MISS
    assert (so[want[hit, 0]] >= 0).sum() > 20_000
    full = sph.copy()
    full[:, 27] = 360
    assert (oracle.intersect_mixed(nodes, agg.ordered_tris9, so, full, rays) != want).any(axis=1).sum() > 5000   # the clip matters
    assert np.array_equal(agg.IntersectPBatch(rays), oracle.intersect_p_mixed(nodes, agg.ordered_tris9, so, sph, rays))
