"""tests/test_general_scene.py -- triangles + constant-alpha triangles + spheres + instances in ONE primitive list.

The reference puts every primitive kind behind one BVHAggregate (scene.cpp:1352-1591; cpu/primitive.cpp:50-130);
nnbvh_upload_scene_general takes the same list.  The GPU result is compared, bit for bit, with the reference's own
BVHAggregate::Intersect / IntersectP over the same list (oracle/_ref live, built from the unmodified reference
sources), on the reference's own tree and on the library's own build.
"""
import numpy as np
import pytest

from conftest import rand_rays

MISS = 0xFFFFFFFF


def _scene(rng, T=5000, S=40, objs=3, K=25):
    from tools.make_golden import random_spheres
    c = rng.random((T, 1, 3), dtype=np.float32)
    tris = ((c + np.float32(0.03) * (rng.random((T, 3, 3), dtype=np.float32) * 2 - 1)).reshape(T, 9)
            + np.float32(0)).astype(np.float32)
    alpha = np.ones(T, dtype=np.float32)
    alpha[::2] = rng.random(len(alpha[::2])).astype(np.float32)
    alpha[::10] = 0.0
    sph = random_spheres(rng, S, scale=0.08, center_box=0.0)
    for k in range(S):
        m = np.eye(4)
        m[:3] = sph[k, :12].reshape(3, 4)
        m[:3, 3] = rng.random(3)
        sph[k, :12] = m[:3].astype(np.float32).ravel()
        sph[k, 12:24] = np.linalg.inv(m)[:3].astype(np.float32).ravel()
    sph[1::3, 27] = rng.uniform(40, 350, len(sph[1::3])).astype(np.float32)
    soups = []
    for _ in range(objs):
        n = int(rng.integers(200, 600))
        cc = rng.random((n, 1, 3), dtype=np.float32) - 0.5
        soups.append(((cc + np.float32(0.08) * (rng.random((n, 3, 3), dtype=np.float32) * 2 - 1)).reshape(n, 9)
                      + np.float32(0)).astype(np.float32))
    inst24 = np.zeros((K, 24), dtype=np.float32)
    inst_obj = rng.integers(0, objs, K).astype(np.int32)
    for k in range(K):
        a, b, g = rng.random(3) * 6.28
        rz = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
        ry = np.array([[np.cos(b), 0, np.sin(b)], [0, 1, 0], [-np.sin(b), 0, np.cos(b)]])
        m = np.eye(4)
        m[:3, :3] = (rz @ ry) * (0.08 + 0.12 * rng.random())
        m[:3, 3] = rng.random(3)
        m32 = m.astype(np.float32)
        inst24[k, :12] = m32[:3].ravel()
        inst24[k, 12:] = np.linalg.inv(m32.astype(np.float64))[:3].astype(np.float32).ravel()
    return tris, alpha, sph, soups, inst24, inst_obj


def _rays(rng, N, sph, inst24):
    rays = rand_rays(rng, N)
    q = N // 4                                   # aim a quarter at spheres, a quarter at instances
    aim = rng.integers(0, len(sph), q)
    rays[:q, 4:7] = sph[aim][:, [3, 7, 11]] + rng.normal(size=(q, 3)).astype(np.float32) * 0.04 - rays[:q, :3]
    aim = rng.integers(0, len(inst24), q)
    rays[q:2 * q, 4:7] = inst24[aim][:, [3, 7, 11]] + rng.normal(size=(q, 3)).astype(np.float32) * 0.06 - rays[q:2 * q, :3]
    rays[N // 2::2, 3] = rng.random(len(rays[N // 2::2])) * 1.5
    return rays


# kernel "1": StepKernel's general instantiation (spheres + instance slots + alpha); "0": BasicKernel's
@pytest.mark.gpu
@pytest.mark.parametrize("kernel", ["1", "0"])
@pytest.mark.parametrize("split", ["sah", "middle"])
def test_one_list_with_every_primitive_kind_matches_the_reference(reference, split, kernel, monkeypatch):
    import nn_bvh_b200 as nb
    monkeypatch.setenv("NNBVH_INST_STEP", kernel)
    from nn_bvh_b200.aggregate import INSTANCE_DTYPE, SPHERE_DTYPE
    rng = np.random.default_rng(101)
    tris, alpha, sph, soups, inst24, inst_obj = _scene(rng)
    ref = reference.scene_general(tris, alpha, sph, soups, inst24, inst_obj, 4, split)
    spheres = np.ascontiguousarray(sph).view(SPHERE_DTYPE).reshape(-1)
    instances = np.zeros(len(inst24), dtype=INSTANCE_DTYPE)
    instances["renderFromPrimitive"], instances["primitiveFromRender"] = inst24[:, :12], inst24[:, 12:]
    instances["object"] = inst_obj
    N = 120_000
    rays = _rays(rng, N, sph, inst24)
    want, want_inner = reference.intersect_general(ref["handle"], rays)
    want_occ = reference.intersect_p(ref["handle"], rays, parallel=False)
    T, S = len(tris), len(sph)
    kind = np.where(want[:, 0] == MISS, -1, np.select(
        [ref["order"][np.minimum(want[:, 0], len(ref["order"]) - 1)] < T,
         ref["order"][np.minimum(want[:, 0], len(ref["order"]) - 1)] < T + S], [0, 1], 2))
    assert (kind == 0).sum() > 5000 and (kind == 1).sum() > 1000 and (kind == 2).sum() > 1000
    assert not (want[:, 0] == 0xFFFFFFFE).any()
    # (a) on the reference's own tree
    agg = nb.GPUBVHAggregate.General(spheres=spheres, instances=instances, tree=ref)
    hits, inner = agg.IntersectBatchInstanced(rays)
    got = np.ascontiguousarray(hits).view(np.uint32).reshape(-1, 5)
    bad = np.flatnonzero((got != want).any(axis=1))
    assert bad.size == 0, f"{len(bad)} records differ, first ray {bad[0]}: got {got[bad[0]]} want {want[bad[0]]}"
    assert np.array_equal(inner, want_inner)
    assert np.array_equal(agg.IntersectPBatch(rays), want_occ)
    # the alpha test really changed something, and removing it changes the answer
    plain = nb.GPUBVHAggregate.General(spheres=spheres, instances=instances, tree=dict(ref, ordered_alpha=np.ones_like(ref["ordered_alpha"])))
    assert (np.ascontiguousarray(plain.IntersectBatch(rays)).view(np.uint32).reshape(-1, 5) != want).any(axis=1).sum() > 1000
    # (b) on the library's own build of the same list: same primitives hit, same t and b
    own = nb.GPUBVHAggregate.General(tris, alpha, spheres, soups, instances, 4, split)
    assert np.array_equal(own.nodes.view(np.uint32).reshape(-1, 8)[:, :6], ref["nodes"][:, :6])   # same boxes, same shape
    h2, in2 = own.IntersectBatchInstanced(rays)
    g2 = np.ascontiguousarray(h2).view(np.uint32).reshape(-1, 5)
    hit = want[:, 0] != MISS
    assert np.array_equal(g2[:, 0] == MISS, ~hit)
    assert np.array_equal(own.order[g2[hit, 0]], ref["order"][want[hit, 0]])            # caller primitive ids
    assert np.array_equal(g2[hit, 1:], want[hit, 1:])
    assert np.array_equal(own.IntersectPBatch(rays), want_occ)


@pytest.mark.gpu
def test_general_description_with_one_kind_takes_that_kinds_path(reference):
    """spheres only -> the StepKernel sphere path (counters available); nothing special -> the plain path."""
    import nn_bvh_b200 as nb
    from nn_bvh_b200.aggregate import SPHERE_DTYPE
    rng = np.random.default_rng(103)
    tris, alpha, sph, soups, inst24, inst_obj = _scene(rng, 3000, 20, 1, 1)
    spheres = np.ascontiguousarray(sph).view(SPHERE_DTYPE).reshape(-1)
    a = nb.GPUBVHAggregate.General(tris, None, spheres)
    b = nb.GPUBVHAggregate(tris, 4, "sah", spheres=spheres)
    rays = rand_rays(rng, 30_000)
    assert np.array_equal(np.ascontiguousarray(a.IntersectBatch(rays)).view(np.uint32), np.ascontiguousarray(b.IntersectBatch(rays)).view(np.uint32))
    _, nv, tt = a.IntersectCounters(rays)           # per-ray counters exist on the specialised path
    assert nv.sum() > 0
    c = nb.GPUBVHAggregate.General(tris)
    d = nb.GPUBVHAggregate(tris, 4, "sah")
    assert np.array_equal(np.ascontiguousarray(c.IntersectBatch(rays)).view(np.uint32), np.ascontiguousarray(d.IntersectBatch(rays)).view(np.uint32))


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", ["1", "0"])
def test_general_scene_matches_the_reference_golden(kernel, monkeypatch):
    """The same claim without oracle/_ref: answers the reference produced here (tools/make_golden.py --only-general)."""
    import os
    monkeypatch.setenv("NNBVH_INST_STEP", kernel)
    import nn_bvh_b200 as nb
    from conftest import GOLDEN
    from nn_bvh_b200.aggregate import INSTANCE_DTYPE, SPHERE_DTYPE
    g = np.load(os.path.join(GOLDEN, "general.npz"))
    spheres = np.ascontiguousarray(g["spheres28"]).view(SPHERE_DTYPE).reshape(-1)
    instances = np.zeros(len(g["inst24"]), dtype=INSTANCE_DTYPE)
    instances["renderFromPrimitive"], instances["primitiveFromRender"] = g["inst24"][:, :12], g["inst24"][:, 12:]
    instances["object"] = g["inst_object"]
    objects, nb_, tb_ = [], 0, 0
    for nn, nt in zip(g["obj_node_count"], g["obj_tri_count"]):
        objects.append(dict(nodes=g["obj_nodes"][nb_:nb_ + nn], ordered_tris9=g["obj_tris"][tb_:tb_ + nt]))
        nb_, tb_ = nb_ + nn, tb_ + nt
    tree = dict(nodes=g["nodes"], order=g["order"], ordered_tris9=g["ordered_tris9"], ordered_alpha=g["ordered_alpha"],
                sphere_of=g["sphere_of"], instance_of=g["instance_of"], objects=objects)
    agg = nb.GPUBVHAggregate.General(spheres=spheres, instances=instances, tree=tree)
    hits, inner = agg.IntersectBatchInstanced(g["rays"])
    assert np.array_equal(np.ascontiguousarray(hits).view(np.uint32).reshape(-1, 5), g["hits"])
    assert np.array_equal(inner, g["inner"])
    assert np.array_equal(agg.IntersectPBatch(g["rays"]), g["occ"])


@pytest.mark.gpu
def test_ray_ordering_does_not_change_one_thread_per_ray_results(reference, monkeypatch):
    """Batches of >= 2 Mi rays go through the ray-ordering pre-pass on the general / alpha / instanced paths too
    (thread j traces ray perm[j], the record goes to the ray's own slot): forced on and off, same records; and a
    strided subset equals the reference."""
    import nn_bvh_b200 as nb
    from nn_bvh_b200.aggregate import INSTANCE_DTYPE, SPHERE_DTYPE
    rng = np.random.default_rng(107)
    tris, alpha, sph, soups, inst24, inst_obj = _scene(rng, 4000, 30, 3, 20)
    ref = reference.scene_general(tris, alpha, sph, soups, inst24, inst_obj, 4, "sah")
    spheres = np.ascontiguousarray(sph).view(SPHERE_DTYPE).reshape(-1)
    instances = np.zeros(len(inst24), dtype=INSTANCE_DTYPE)
    instances["renderFromPrimitive"], instances["primitiveFromRender"] = inst24[:, :12], inst24[:, 12:]
    instances["object"] = inst_obj
    import torch
    N = (2 << 20) + 12345
    rays = _rays(rng, N, sph, inst24)
    d_rays = torch.from_numpy(rays).cuda()      # resident rays: one launch over the whole batch (host arrays are chunked)
    d_hits = torch.zeros((N, 5), dtype=torch.int32, device="cuda")
    d_inner = torch.zeros(N, dtype=torch.int32, device="cuda")
    d_occ = torch.zeros(N, dtype=torch.uint8, device="cuda")
    out = {}
    monkeypatch.setenv("NNBVH_INST_STEP", "0")          # BasicKernel's general instantiation (one thread per ray)
    for mode in ("0", "1"):
        monkeypatch.setenv("NNBVH_BIN", mode)
        agg = nb.GPUBVHAggregate.General(spheres=spheres, instances=instances, tree=ref)
        agg.IntersectBatchInstanced(d_rays, hits=d_hits, inner=d_inner, n=N)
        agg.IntersectPBatch(d_rays, occluded=d_occ, n=N)
        torch.cuda.synchronize()
        out[mode] = (d_hits.cpu().numpy().view(np.uint32).copy(), d_inner.cpu().numpy().view(np.uint32).copy(),
                     d_occ.cpu().numpy().copy(), agg.launch_count())
        agg.close()
        al = nb.GPUBVHAggregate(tris, 4, "sah", alpha=alpha)          # the alpha-only path
        al.IntersectBatch(d_rays, hits=d_hits, n=N)
        torch.cuda.synchronize()
        out["a" + mode] = d_hits.cpu().numpy().view(np.uint32).copy()
        al.close()
    assert out["1"][3] > out["0"][3] + 10                               # the sort really ran
    for k in range(3):
        assert np.array_equal(out["0"][k], out["1"][k])
    assert np.array_equal(out["a0"], out["a1"])
    sub = np.ascontiguousarray(rays[::97])
    want, want_inner = reference.intersect_general(ref["handle"], sub)
    assert np.array_equal(out["1"][0][::97], want) and np.array_equal(out["1"][1][::97], want_inner)
