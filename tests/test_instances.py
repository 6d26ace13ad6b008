"""tests/test_instances.py -- TransformedPrimitive instances: two-level traversal.

Reference: /root/reference/src/pbrt/cpu/primitive.cpp:112-130 (TransformedPrimitive::Intersect / IntersectP),
cpu/primitive.h:94 (Bounds), util/transform.h:416-430 + util/transform.cpp:263-303 (Transform::ApplyInverse of a
ray: interval error bound of the origin, origin pushed along d, tMax shortened), scene.cpp:1541-1571.

CPU: the oracle restatement and the product's host builders against golden answers produced by the
reference (tools/make_golden.py --only-instances) and against oracle/_ref live.
GPU: instanced scenes through the C ABI against the oracle: top-level slot, t, b, the primitive inside the
object, occlusion and both counters (which include the nested traversals), bit-exact.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN, rand_rays

MISS = 0xFFFFFFFF


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN, "instanced.npz"))


def _sc(g):
    return {k[3:]: g[k] for k in g.files if k.startswith("sc_")}


def test_oracle_instanced_matches_reference_golden(oracle, gold):
    sc = _sc(gold)
    hits, inner = oracle.intersect_instanced(sc, gold["rays"])
    assert np.array_equal(hits, gold["hits"]) and np.array_equal(inner, gold["inner"])
    assert np.array_equal(oracle.intersect_instanced(sc, gold["rays"], any_hit=True), gold["occ"])
    assert (inner != MISS).sum() > 800 and ((hits[:, 0] != MISS) & (inner == MISS)).sum() > 100
    for k in range(len(gold["inst24"])):
        root = sc["obj_nodes"][sc["obj_node_base"][gold["inst_object"][k]]].view(np.float32)[:6]
        assert np.array_equal(oracle.instance_bounds(gold["inst24"][k, :12], root).view(np.uint32),
                              gold["instance_bounds"][k].view(np.uint32))


def test_host_builders_reproduce_the_reference_instanced_trees(gold):
    """nnbvh_build_bvh per object, nnbvh_instance_bounds, nnbvh_build_bvh_bounds for the top level."""
    import nn_bvh_b200 as nb
    sc = _sc(gold)
    J = int(gold["n_objects"])
    roots = []
    for j in range(J):
        nodes, order = nb.build_bvh(gold[f"obj{j}"], 4, "sah", threads=1)
        lo = sc["obj_node_base"][j]
        assert np.array_equal(nodes.view(np.uint32).reshape(-1, 8), sc["obj_nodes"][lo:lo + len(nodes)])
        assert np.array_equal(order.astype(np.int32), gold[f"obj{j}_order"])
        roots.append(nodes[0]["bounds"])
    inst = nb.make_instances(gold["inst24"][:, :12].reshape(-1, 3, 4), gold["inst_object"], gold["inst24"][:, 12:])
    assert np.array_equal(inst["renderFromPrimitive"], gold["inst24"][:, :12])
    ib = nb.instance_bounds(inst, np.stack([roots[j] for j in gold["inst_object"]]))
    assert np.array_equal(ib.view(np.uint32), gold["instance_bounds"].view(np.uint32))
    nodes, order = nb.build_bvh_bounds(np.concatenate([nb.triangle_bounds(gold["top"]), ib]), 4, "sah", threads=1)
    assert np.array_equal(nodes.view(np.uint32).reshape(-1, 8), sc["nodes"])
    assert np.array_equal(order.astype(np.int32), gold["order"])


@pytest.mark.parametrize("seed", [1, 2])
def test_oracle_instanced_matches_reference_live(oracle, reference, seed):
    from tools.make_golden import instanced_scene
    rng = np.random.default_rng(seed)
    top, objs, inst, io = instanced_scene(rng, T=800 * seed, K=20, obj_sizes=(300, 2, 700, 50))
    sc = reference.scene_instanced(top, objs, inst, io, 4, "sah" if seed == 1 else "middle")
    rays = rand_rays(rng, 6000)
    rays[3000:, 3] = rng.random(3000)
    rh, ri = reference.intersect_instanced(sc["handle"], rays)
    assert (rh[:, 0] != 0xFFFFFFFE).all()
    oh, oi = oracle.intersect_instanced(sc["data"], rays)
    assert np.array_equal(oh, rh) and np.array_equal(oi, ri)
    assert np.array_equal(oracle.intersect_instanced(sc["data"], rays, any_hit=True),
                          reference.intersect_p(sc["handle"], rays, parallel=False))


def _oracle_scene(agg):
    """The arrays Oracle.intersect_instanced takes, from the product mirror's own trees."""
    nb_, pb_ = [0], [0]
    for o in agg.objects:
        nb_.append(nb_[-1] + len(o["nodes"]))
        pb_.append(pb_[-1] + len(o["ordered_tris9"]))
    return dict(nodes=agg.nodes.view(np.uint32).reshape(-1, 8), ordered_tris9=agg.ordered_tris9,
                instance_of=agg.ordered_instance_index,
                inst24=np.concatenate([agg.instances["renderFromPrimitive"], agg.instances["primitiveFromRender"]], axis=1),
                inst_object=agg.instances["object"].astype(np.int32),
                obj_nodes=np.concatenate([o["nodes"].view(np.uint32).reshape(-1, 8) for o in agg.objects]),
                obj_node_base=np.array(nb_[:-1], dtype=np.int32),
                obj_tris=np.concatenate([o["ordered_tris9"] for o in agg.objects]),
                obj_prim_base=np.array(pb_[:-1], dtype=np.int32))


# inst_step 1: StepKernel<..., INST> (the objects' nodes and triangles appended to the top level's arrays, the second
# level entered as a pushed sub-state); 0: the one-thread-per-ray BasicKernel with nested traversal
@pytest.mark.gpu
@pytest.mark.parametrize("inst_step", ["1", "0"])
def test_instanced_scene_gpu_matches_reference_golden(gold, inst_step, monkeypatch):
    import nn_bvh_b200 as nb
    monkeypatch.setenv("NNBVH_INST_STEP", inst_step)
    J = int(gold["n_objects"])
    inst = nb.make_instances(gold["inst24"][:, :12].reshape(-1, 3, 4), gold["inst_object"], gold["inst24"][:, 12:])
    assert np.array_equal(inst["renderFromPrimitive"], gold["inst24"][:, :12])
    agg = nb.GPUBVHAggregate(gold["top"], 4, "sah", objects=[gold[f"obj{j}"] for j in range(J)], instances=inst)
    assert np.array_equal(agg.nodes.view(np.uint32).reshape(-1, 8), gold["sc_nodes"])
    hits, inner = agg.IntersectBatchInstanced(gold["rays"])
    assert np.array_equal(np.ascontiguousarray(hits).view(np.uint32).reshape(-1, 5), gold["hits"])
    assert np.array_equal(inner, gold["inner"])
    assert np.array_equal(agg.IntersectPBatch(gold["rays"]), gold["occ"])
    # the wavefront queues' planes (tMax plane given / NULL = Infinity)
    r = gold["rays"]
    pl = [np.ascontiguousarray(r[:, k]) for k in (0, 1, 2, 4, 5, 6, 3)]
    h2, i2 = agg.IntersectBatchInstancedSoA(*pl)
    assert np.array_equal(np.ascontiguousarray(h2).view(np.uint32).reshape(-1, 5), gold["hits"]) and np.array_equal(i2, gold["inner"])
    inf = np.isinf(r[:, 3])
    h3, i3 = agg.IntersectBatchInstancedSoA(*pl[:6])
    assert inf.sum() > 100 and np.array_equal(np.ascontiguousarray(h3).view(np.uint32).reshape(-1, 5)[inf], gold["hits"][inf])
    assert np.array_equal(i3[inf], gold["inner"][inf])
    # the plain closest entry point works too (no inner plane)
    assert np.array_equal(np.ascontiguousarray(agg.IntersectBatch(gold["rays"])).view(np.uint32).reshape(-1, 5), gold["hits"])


@pytest.mark.gpu
@pytest.mark.parametrize("inst_step", ["1", "0"])
def test_instanced_scene_gpu_matches_oracle_at_size(oracle, inst_step, monkeypatch):
    import torch
    import nn_bvh_b200 as nb
    monkeypatch.setenv("NNBVH_INST_STEP", inst_step)
    from tools.make_golden import instanced_scene
    rng = np.random.default_rng(21)
    top, objs, inst24, io = instanced_scene(rng, T=20_000, K=400, obj_sizes=(3000, 1, 12_000, 500))
    # smaller instances so that 400 of them do not fill the cube
    m = inst24[:, :12].reshape(-1, 3, 4).astype(np.float64)
    m[:, :, :3] *= 0.35
    inst = nb.make_instances(m, io)
    agg = nb.GPUBVHAggregate(top, 4, "sah", objects=objs, instances=inst)
    sc = _oracle_scene(agg)
    N = 400_000
    rays = rand_rays(rng, N)
    rays[N // 2:, 3] = rng.random(N - N // 2)
    rays[::101, 5] = 0.0
    (wh, wi, wnv, wtt) = oracle.intersect_instanced(sc, rays, counters=True)
    hits, inner = agg.IntersectBatchInstanced(rays)
    assert np.array_equal(np.ascontiguousarray(hits).view(np.uint32).reshape(-1, 5), wh) and np.array_equal(inner, wi)
    assert (wi != MISS).sum() > 20_000 and ((wh[:, 0] != MISS) & (wi == MISS)).sum() > 20_000
    (wo, onv, ott) = oracle.intersect_instanced(sc, rays, any_hit=True, counters=True)
    assert np.array_equal(agg.IntersectPBatch(rays), wo)
    h2, nv, tt = agg.IntersectCounters(rays)
    assert np.array_equal(np.ascontiguousarray(h2).view(np.uint32).reshape(-1, 5), wh)
    assert np.array_equal(nv, wnv) and np.array_equal(tt, wtt)
    o2, nv, tt = agg.IntersectCounters(rays, any_hit=True)
    assert np.array_equal(o2, wo) and np.array_equal(nv, onv) and np.array_equal(tt, ott)
    # device-resident rays and results
    d = torch.from_numpy(rays).cuda()
    dh = torch.zeros((N, 5), dtype=torch.int32, device="cuda")
    di = torch.zeros(N, dtype=torch.int32, device="cuda")
    agg.IntersectBatchInstanced(d, dh, di, n=N)
    torch.cuda.synchronize()
    assert np.array_equal(dh.cpu().numpy().view(np.uint32), wh) and np.array_equal(di.cpu().numpy().view(np.uint32), wi)
