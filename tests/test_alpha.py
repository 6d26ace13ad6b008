"""tests/test_alpha.py -- alpha-tested primitives, the checker's side.

GeometricPrimitive::Intersect with an alpha texture (/root/reference/src/pbrt/cpu/primitive.cpp:50-84): a hit
is ignored when HashFloat(r.o, r.d) > alpha (always, for alpha <= 0), and the primitive is asked again with
si->intr.SpawnRay(r.d); IntersectP is Intersect(...).has_value().  The oracle restates it for constant alpha
values -- MurmurHash64A / HashFloat (util/hash.h), the triangle interaction's pHit / pError / n and
OffsetRayOrigin (shapes.h:924-941, ray.h:75-92) -- and is pinned here to the reference, golden and live.
The device side (csrc/alpha.cuh) is checked against this oracle in tests/test_alpha_gpu.py.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN, rand_rays

MISS = 0xFFFFFFFF


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN, "alpha.npz"))


def test_hash_float_matches_reference_golden(oracle, gold):
    got = np.array([oracle.hash_float(v[:3], v[3:]) for v in gold["od"]], dtype=np.float32)
    assert np.array_equal(got.view(np.uint32), gold["hashes"].view(np.uint32))
    assert 0.45 < got.mean() < 0.55 and got.min() >= 0 and got.max() < 1


def test_spawn_origin_matches_reference_golden(oracle, gold, golden):
    """The re-traced ray's origin, bit for bit, on every hit of the triangle known-answer set (edge hits,
    slivers, huge and tiny coordinates)."""
    tv = golden["tri_vectors"]
    n_hit = 0
    for i in range(len(tv["rays"])):
        h, o4 = oracle.spawn_origin(tv["rays"][i], tv["tris"][i])
        assert h == bool(gold["spawn_hit"][i]), i
        if h:
            n_hit += 1
            assert np.array_equal(o4.view(np.uint32), gold["spawn"][i].view(np.uint32)), i
    assert n_hit > 3000


def test_oracle_alpha_scene_matches_reference_golden(oracle, gold):
    order = gold["order"]
    ot, oa = gold["tris"][order], gold["alpha"][order]
    got = oracle.intersect_alpha(gold["nodes"], ot, oa, gold["rays"])
    assert np.array_equal(got, gold["hits"])
    assert np.array_equal(oracle.intersect_alpha(gold["nodes"], ot, oa, gold["rays"], any_hit=True), gold["occ"])
    # the alpha test matters: the same tree without it answers differently for many rays
    plain = oracle.intersect(gold["nodes"], ot, gold["rays"])
    assert (plain != gold["hits"]).any(axis=1).sum() > 500
    # alpha = 0 primitives are never hit, alpha >= 1 ones are unaffected
    hit = gold["hits"][:, 0] != MISS
    assert (oa[gold["hits"][hit, 0]] > 0).all()


@pytest.mark.parametrize("seed", [1, 2])
def test_oracle_alpha_scene_matches_reference_live(oracle, reference, seed):
    from tools.make_golden import as_pbrt_stores
    rng = np.random.default_rng(seed)
    T = 3000
    c = rng.random((T, 1, 3), dtype=np.float32)
    tris = as_pbrt_stores((c + np.float32(0.06) * (rng.random((T, 3, 3), dtype=np.float32) * 2 - 1)).reshape(T, 9))
    a = rng.random(T).astype(np.float32)
    a[::3] = 1.0
    a[1::7] = 0.0
    sc = reference.scene_alpha(tris, a, 4, ("sah", "middle")[seed % 2])
    nodes, order = reference.export(sc, T)
    rays = rand_rays(rng, 6000)
    rays[3000:, 3] = rng.random(3000)
    rays[:, 4:7] *= np.exp(rng.uniform(-1, 1, (6000, 1))).astype(np.float32)      # the hash sees d as given
    assert np.array_equal(oracle.intersect_alpha(nodes, tris[order], a[order], rays),
                          reference.intersect(sc, rays, parallel=False))
    assert np.array_equal(oracle.intersect_alpha(nodes, tris[order], a[order], rays, any_hit=True),
                          reference.intersect_p(sc, rays, parallel=False))
