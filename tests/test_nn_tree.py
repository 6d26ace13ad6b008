"""NN-split trees: the oracle restatement against golden vectors made by the
reference's own Python, the product exporter against the oracle, and (GPU) traversal of exported
trees against the traversal oracle."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, rand_rays, soup

CASES = ["fix", "a", "b", "c", "d"]


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLDEN, "nn_tree.npz"))


@pytest.mark.parametrize("case", CASES)
def test_oracle_matches_reference_python(g, case):
    from oracle import nn_tree_oracle as O
    tris, planes = g[case + "_tris"], g[case + "_planes"]
    root, h = O.build_from_planes(tris, planes)
    off = g[case + "_offsets"]
    for k, nd in enumerate(h):
        assert nd.prims == list(g[case + "_members"][off[k]:off[k + 1]]), (case, k)
        assert np.array_equal(np.array(nd.aabb), g[case + "_aabbs"][k]), (case, k)
        assert nd.is_leaf == bool(g[case + "_is_leaf"][k]), (case, k)
    assert O.sah(root) == pytest.approx(float(g[case + "_sah"]), rel=1e-12)
    assert O.epo(tris, root) == pytest.approx(float(g[case + "_epo"]), rel=1e-12)


def test_reference_fixture_split(g):
    """machine_learning/nn_test.py:48-63: split(Axis.y, 0.75) puts l_prim0 left, the rest right."""
    off = g["fix_offsets"]
    assert list(g["fix_members"][off[1]:off[2]]) == [0] and list(g["fix_members"][off[2]:off[3]]) == [1, 2]


@pytest.mark.parametrize("case", CASES)
def test_product_top_tree_matches_golden(g, case):
    from nn_bvh_b200 import nn_tree
    tris, planes = g[case + "_tris"], g[case + "_planes"]
    top = nn_tree.TopTree(tris, planes)
    off = g[case + "_offsets"]
    for k in range(len(top.members)):
        assert np.array_equal(top.members[k], g[case + "_members"][off[k]:off[k + 1]]), (case, k)
        assert np.array_equal(top.aabbs[k], g[case + "_aabbs"][k]), (case, k)
        assert bool(top.is_leaf[k]) == bool(g[case + "_is_leaf"][k])
    assert top.sah() == pytest.approx(float(g[case + "_sah"]), rel=1e-9)
    assert top.epo() == pytest.approx(float(g[case + "_epo"]), rel=1e-9)


def _check_linear_tree(nodes, order, tris32, nb):
    """Structural validity of an exported LinearBVHNode array (cpu/aggregates.cpp:128-137)."""
    T = len(tris32)
    assert sorted(order.tolist()) == list(range(T))
    ordered = tris32[order].reshape(-1, 3, 3)
    seen = np.zeros(T, bool)

    def rec(i):
        nd = nodes[i]
        if nd["nPrimitives"] > 0:
            a, n = int(nd["offset"]), int(nd["nPrimitives"])
            assert not seen[a:a + n].any()
            seen[a:a + n] = True
            v = ordered[a:a + n].reshape(-1, 3)
            lo, hi = v.min(0), v.max(0)
            end = i + 1
        else:
            assert nd["axis"] in (0, 1, 2)
            lo0, hi0, e0 = rec(i + 1)
            assert int(nd["offset"]) == e0          # second child right after the first subtree
            lo1, hi1, end = rec(e0)
            lo, hi = np.minimum(lo0, lo1), np.maximum(hi0, hi1)
        assert np.all(nd["bounds"][:3] <= lo) and np.all(nd["bounds"][3:] >= hi)
        return nd["bounds"][:3], nd["bounds"][3:], end

    assert rec(0)[2] == len(nodes) and seen.all()
    assert nb.tree_stack_depth(nodes) <= 64


@pytest.mark.parametrize("case", ["a", "b", "c", "d"])
def test_export_linear_is_a_valid_tree(g, case):
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import nn_tree
    tris32 = g[case + "_tris"].astype(np.float32)
    nodes, order = nn_tree.export_linear(tris32, g[case + "_planes"], max_prims_in_node=4)
    _check_linear_tree(nodes, order, tris32, nb)
    # the top of the exported tree IS the plane tree: the root splits on plane 0's axis and its
    # first child holds exactly the primitives the reference puts left
    off = g[case + "_offsets"]
    left = set(g[case + "_members"][off[1]:off[2]].tolist())
    if left and len(left) < len(tris32):
        assert nodes[0]["nPrimitives"] == 0 and nodes[0]["axis"] == int(np.argmax(g[case + "_planes"][0][:3]))
        n_first = int(nodes[0]["offset"]) - 1
        first_prims = set()
        for nd in nodes[1:1 + n_first]:
            if nd["nPrimitives"]:
                first_prims |= set(order[nd["offset"]:nd["offset"] + nd["nPrimitives"]].tolist())
        assert first_prims == left


def test_greedy_sah_planes_and_costs():
    from nn_bvh_b200 import nn_tree
    rng = np.random.default_rng(3)
    tris = (1 + soup(rng, 5000, 0.02)).astype(np.float32)
    planes = nn_tree.greedy_sah_planes(tris, levels=3, candidates=32)
    assert planes.shape == (7, 4) and np.all(planes[:, :3].sum(1) == 1)
    top = nn_tree.TopTree(tris, planes)
    med = nn_tree.TopTree(tris, nn_tree.median_planes(tris, levels=3))
    assert top.sah() <= med.sah() * 1.05     # greedy SAH should not lose to plain medians


@pytest.mark.gpu
@pytest.mark.parametrize("levels", [2, 4])
def test_exported_tree_traversal_matches_oracle(oracle, levels):
    """Same rays through the SAH tree and the NN-split tree: each is bit-identical to the CPU
    traversal of the same node array, and both find the same closest distance."""
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import nn_tree
    rng = np.random.default_rng(17)
    tris = (1 + soup(rng, 30000, 0.01)).astype(np.float32)
    rays = rand_rays(rng, 200_000, lo=0.75, hi=2.25)
    planes = nn_tree.greedy_sah_planes(tris, levels=levels, candidates=16)
    nodes, order = nn_tree.export_linear(tris, planes, max_prims_in_node=4)
    agg = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=tris[order], order=order)
    sah = nb.GPUBVHAggregate.Create(tris, {"maxnodeprims": 4})
    n32 = agg.nodes.view(np.uint32).reshape(-1, 8)
    h = agg.IntersectBatch(rays)
    assert np.array_equal(np.ascontiguousarray(h).view(np.uint32).reshape(-1, 5),
                          oracle.intersect(n32, agg.ordered_tris9, rays))
    assert np.array_equal(agg.IntersectPBatch(rays), oracle.intersect_p(n32, agg.ordered_tris9, rays))
    h2 = sah.IntersectBatch(rays)
    assert np.array_equal(h["prim"] == nb.MISS, h2["prim"] == nb.MISS)
    ok = h["prim"] != nb.MISS
    # different trees may resolve exact ties to different primitives; the distance agrees
    assert np.allclose(h["t"][ok], h2["t"][ok], rtol=1e-5, atol=0)
    same = order[h["prim"][ok]] == sah.order[h2["prim"][ok]]
    assert same.mean() > 0.999
    res = nn_tree.traversal_cost(agg, rays[:50_000])
    assert res["nodes_visited"] > 0 and res["tri_tests"] > 0
