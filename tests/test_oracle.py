"""CPU: the oracle restatement against the reference's golden vectors and, where
oracle/_ref is built (this container), against the unmodified reference itself."""
import numpy as np
import pytest

from conftest import rand_rays, soup


def test_oracle_triangle_golden(oracle, golden):
    g = golden["tri_vectors"]
    for i in range(len(g["rays"])):
        hit, out = oracle.triangle(g["rays"][i], g["tris"][i])
        assert hit == bool(g["hit"][i]), i
        if hit:
            assert np.array_equal(out.view(np.uint32), g["out"][i].view(np.uint32)), i


def test_oracle_badcases_vector_misses(oracle, golden):
    """Triangle.BadCases, /root/reference/src/pbrt/shapes_test.cpp:435-449."""
    g = golden["tri_vectors"]
    assert not oracle.triangle(g["rays"][0], g["tris"][0])[0]


def test_oracle_box_golden(oracle, golden):
    g = golden["box_vectors"]
    got = np.array([oracle.box(g["rays"][i], g["boxes"][i]) for i in range(len(g["rays"]))])
    assert np.array_equal(got.astype(np.uint8), g["hit"])


@pytest.mark.parametrize("split", ["sah", "middle", "equal"])
def test_oracle_scene_golden(oracle, golden, split):
    g = golden["scene_small"]
    nodes, order = oracle.build_bvh(g["tris"], 4, split)
    assert np.array_equal(nodes, g[f"{split}_nodes"])
    assert np.array_equal(order, g[f"{split}_order"])
    ot = g["tris"][order]
    hits, nv, tt = oracle.intersect(nodes, ot, g["rays"], counters=True)
    assert np.array_equal(hits, g[f"{split}_hits"])
    occ, nvp, ttp = oracle.intersect_p(nodes, ot, g["rays"], counters=True)
    assert np.array_equal(occ, g[f"{split}_occ"])
    st = g[f"{split}_stats"]
    assert (int(nv.sum()), int(tt.sum()), int(nvp.sum()), int(ttp.sum())) == tuple(st)


HLBVH_CASES = [("small4", 4), ("soup4", 4), ("soup1", 1), ("soup255", 255)]


@pytest.mark.parametrize("name,mp", HLBVH_CASES)
def test_oracle_hlbvh_golden(oracle, golden, name, mp):
    """buildHLBVH / emitLBVH / buildUpperSAH, cpu/aggregates.cpp:389-503, 626-723."""
    g = golden["hlbvh"]
    tris = golden["scene_small"]["tris"] if name == "small4" else g["soup"]
    nodes, order = oracle.build_bvh(tris, mp, "hlbvh")
    assert np.array_equal(nodes, g[f"{name}_nodes"])
    assert np.array_equal(order, g[f"{name}_order"])
    if name == "small4":
        rays = golden["scene_small"]["rays"]
        assert np.array_equal(oracle.intersect(nodes, tris[order], rays), g["small4_hits"])
        assert np.array_equal(oracle.intersect_p(nodes, tris[order], rays), g["small4_occ"])


def test_oracle_vs_reference_live(oracle, reference):
    """Fresh seeded inputs through both (only where oracle/_ref exists)."""
    rng = np.random.default_rng(99)
    for T, s, mp, split in [(5000, 0.03, 4, "sah"), (5000, 0.03, 1, "sah"), (4000, 0.05, 8, "middle"),
                            (4000, 0.05, 4, "equal"), (30000, 0.004, 4, "hlbvh"), (37, 0.3, 1, "hlbvh")]:
        tris = soup(rng, T, s)
        sc = reference.scene(tris, mp, split)
        rn, ro = reference.export(sc, T)
        on, oo = oracle.build_bvh(tris, mp, split)
        assert np.array_equal(on, rn) and np.array_equal(oo, ro)
        rays = rand_rays(rng, 20000)
        rays[10000:, 3] = rng.random(10000)
        reference.stats()
        rh = reference.intersect(sc, rays, parallel=False)
        st = reference.stats()
        oh, nv, tt = oracle.intersect(on, tris[oo], rays, counters=True)
        assert np.array_equal(oh, rh)
        assert (int(nv.sum()), int(tt.sum())) == st[:2]
        assert np.array_equal(oracle.intersect_p(on, tris[oo], rays), reference.intersect_p(sc, rays))


def test_oracle_threads_agree(oracle, golden):
    g = golden["scene_small"]
    ot = g["tris"][g["sah_order"]]
    rays = np.tile(g["rays"], (4, 1))
    a = oracle.intersect(g["sah_nodes"], ot, rays, threads=1)
    b = oracle.intersect(g["sah_nodes"], ot, rays, threads=4)
    assert np.array_equal(a, b)
