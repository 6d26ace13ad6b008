"""GPU: HLBVH with the data-parallel stages on the device (nnbvh_build_hlbvh_device) against the
trees the unmodified reference produced (tests/golden/hlbvh.npz), the host builder and the oracle.
Byte-identical nodes and primitive order."""
import numpy as np
import pytest

from conftest import rand_rays, soup

pytestmark = pytest.mark.gpu


def _words(nodes):
    return np.ascontiguousarray(nodes).view(np.uint32).reshape(-1, 8)


@pytest.mark.parametrize("name,mp", [("small4", 4), ("soup4", 4), ("soup1", 1), ("soup255", 255)])
def test_device_hlbvh_reference_golden(golden, name, mp):
    import nn_bvh_b200 as nb
    g = golden["hlbvh"]
    tris = golden["scene_small"]["tris"] if name == "small4" else g["soup"]
    nodes, order, _ = nb.build_hlbvh_device(tris, mp)
    assert np.array_equal(_words(nodes), g[f"{name}_nodes"])
    assert np.array_equal(order, g[f"{name}_order"].astype(np.uint32))


def _cases():
    rng = np.random.default_rng(77)
    yield "one", soup(rng, 1, 0.3), 4
    yield "two", soup(rng, 2, 0.3), 4
    yield "tile-1", soup(rng, 2047, 0.02), 4          # radix tile boundaries
    yield "tile", soup(rng, 2048, 0.02), 4
    yield "tile+1", soup(rng, 2049, 0.02), 1
    yield "dups", np.tile(soup(rng, 300, 0.05), (40, 1)), 4      # 40 copies: equal codes, big leaves
    big = soup(rng, 400_000, 0.003)
    yield "big", big, 4
    # clustered: 90% of the primitives in 1% of the volume -> a few very large treelets
    cl = big.copy()
    cl[:360_000] = (cl[:360_000].reshape(-1, 3, 3) * 0.1 + 0.45).reshape(-1, 9)
    yield "clustered", cl, 8
    flat = big[:50_000].copy()
    flat[:, 2::3] = 0.25   # planar scene: one centroid axis is degenerate (Offset's pMax > pMin test)
    yield "flat", flat, 4


@pytest.mark.parametrize("case", list(_cases()), ids=lambda c: c[0])
def test_device_hlbvh_equals_host(case):
    import nn_bvh_b200 as nb
    _, tris, mp = case
    hn, ho = nb.build_bvh(tris, mp, "hlbvh")
    dn, do, ms = nb.build_hlbvh_device(tris, mp)
    assert np.array_equal(_words(dn), _words(hn))
    assert np.array_equal(do, ho)
    assert ms > 0


def test_hlbvh_aggregate_traces_like_the_reference(oracle, golden):
    """GPUBVHAggregate(split_method="hlbvh") builds on the device and traverses; hits equal the
    reference's on its own HLBVH tree."""
    import nn_bvh_b200 as nb
    g, s = golden["hlbvh"], golden["scene_small"]
    agg = nb.GPUBVHAggregate(s["tris"], 4, "hlbvh")
    assert np.array_equal(_words(agg.nodes), g["small4_nodes"])
    hits = agg.IntersectBatch(s["rays"])
    assert np.array_equal(np.ascontiguousarray(hits).view(np.uint32).reshape(-1, 5), g["small4_hits"])
    assert np.array_equal(agg.IntersectPBatch(s["rays"]), g["small4_occ"])
    rng = np.random.default_rng(5)
    tris = soup(rng, 100_000, 0.01)
    agg = nb.GPUBVHAggregate.Create(tris, {"splitmethod": "hlbvh", "maxnodeprims": 4})
    rays = rand_rays(rng, 50_000)
    want = oracle.intersect(_words(agg.nodes), tris[agg.order], rays)
    got = np.ascontiguousarray(agg.IntersectBatch(rays)).view(np.uint32).reshape(-1, 5)
    assert np.array_equal(got, want)


def test_device_hlbvh_errors():
    import nn_bvh_b200 as nb
    with pytest.raises(nb.NnbvhError):
        nb.build_hlbvh_device(np.zeros((0, 9), np.float32), 4)
    # 70000 identical primitives: one Morton code, one leaf of >= 65536 -> the reference CHECKs
    one = np.array([[0, 0, 0, 1, 0, 0, 0, 1, 0]], np.float32)
    with pytest.raises(nb.NnbvhError):
        nb.build_hlbvh_device(np.tile(one, (70_000, 1)), 4)
    with pytest.raises(nb.NnbvhError):
        nb.build_bvh(np.tile(one, (70_000, 1)), 4, "hlbvh")


def test_device_hlbvh_many_small_inputs():
    """200 tiny soups (1..300 triangles; duplicates, planar, collinear, identical, 1e-6..1e6 scales) at
    several leaf sizes: device == host, byte for byte; where the host refuses (the reference's CHECKs)
    the device refuses too."""
    import nn_bvh_b200 as nb
    rng = np.random.default_rng(4052)
    for case in range(200):
        T = int(rng.integers(1, 301))
        scale = np.float32(10.0 ** rng.integers(-6, 7))
        tris = (rng.random((T, 9), dtype=np.float32) * scale).astype(np.float32)
        kind = case % 6
        if kind == 1:
            tris[T // 2:] = tris[: T - T // 2]
        elif kind == 2:
            tris[:, 2::3] = tris[0, 2]
        elif kind == 3:
            tris[:, 1::3] = tris[0, 1]
            tris[:, 2::3] = tris[0, 2]
        elif kind == 4:
            tris[:] = tris[0]
        mp = int(rng.choice([1, 2, 4, 8, 255]))
        try:
            hn, ho = nb.build_bvh(tris, mp, "hlbvh")
        except nb.NnbvhError:
            with pytest.raises(nb.NnbvhError):
                nb.build_hlbvh_device(tris, mp)
            continue
        dn, do, _ = nb.build_hlbvh_device(tris, mp)
        assert np.array_equal(_words(dn), _words(hn)), (case, T, mp)
        assert np.array_equal(do, ho), (case, T, mp)
