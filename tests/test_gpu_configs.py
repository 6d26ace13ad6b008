"""GPU: the BASELINE.json configs at (or near) full size.

Real scenes come from scene_cache/ (exported from the reference's shipped scenes by
tools/export_scenes.py; skipped when the cache did not travel).  Each case is checked
  * against the oracle on a seeded subset the oracle finishes in seconds, bit for bit, and
  * at full size through size-independent properties: product kernel == BasicKernel record for
    record, chunked == unchunked (the host pipeline), and IntersectP(ray) == (Intersect(ray)
    found something) -- both traversals accept their first primitive at the same point of the
    same order (cpu/aggregates.cpp:547-555 vs :597-601), so the two must agree exactly.
"""
import os

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _words(h):
    return np.ascontiguousarray(h).view(np.uint32).reshape(-1, 5)


def _scene(name):
    from nn_bvh_b200 import scenes
    if not os.path.exists(os.path.join(ROOT, "scene_cache", name + ".tris.npy")):
        pytest.skip(f"scene_cache/{name} not present on this machine")
    return scenes.load(name, allow_procedural=False)


def _check(agg, oracle, rays, n_oracle, tag):
    import nn_bvh_b200 as nb
    nodes = agg.nodes.view(np.uint32).reshape(-1, 8)
    r8 = np.ascontiguousarray(rays).view(np.float32).reshape(-1, 8)
    hits = agg.IntersectBatch(r8)
    occ = agg.IntersectPBatch(r8)
    # oracle on a strided subset
    step = max(1, len(r8) // n_oracle)
    sub = np.ascontiguousarray(r8[::step])
    assert np.array_equal(_words(hits)[::step], oracle.intersect(nodes, agg.ordered_tris9, sub)), tag
    assert np.array_equal(occ[::step], oracle.intersect_p(nodes, agg.ordered_tris9, sub)), tag
    # full-size properties
    assert np.array_equal(occ != 0, hits["prim"] != nb.MISS), tag + ": IntersectP != (Intersect hit)"
    ok = hits["prim"] != nb.MISS
    assert np.all(hits["prim"][ok] < len(agg.ordered_tris9))
    # t may exceed the tMax it was accepted against by rounding (shapes.cpp:239-247): a few ulps
    assert np.all(hits["t"][ok] <= r8[ok, 3] * np.float32(1.000001)) and np.all(hits["t"][ok] > 0)
    b = hits["b"][ok].astype(np.float64).sum(1)
    assert np.all(np.abs(b - 1) < 1e-4)
    basic = nb.GPUBVHAggregate(nodes=agg.nodes, ordered_tris9=agg.ordered_tris9, kernel=1)
    assert np.array_equal(_words(basic.IntersectBatch(r8)), _words(hits)), tag + ": step != basic"
    assert np.array_equal(basic.IntersectPBatch(r8), occ), tag
    basic.close()
    return hits, occ


@pytest.mark.parametrize("name", ["killeroos", "coffee_maker", "bathroom", "crown"])
def test_scene_primary_bounce_shadow(oracle, name):
    """Configs 0-3: camera rays (coherent), diffuse-bounce rays (incoherent) and shadow rays
    (unnormalised d, tMax = 0.9999) over the reference's own scenes and SAH trees."""
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import rays as R
    tris, cam, _ = _scene(name)
    agg = nb.GPUBVHAggregate.Create(tris, {"maxnodeprims": 4, "splitmethod": "sah"})
    prim = R.camera_rays(cam, 1024, 1024)
    _check(agg, oracle, prim, 100_000, name + "/primary")
    n = 2 << 20
    bounce, shadow = R.bounce_and_shadow_rays(lambda r: agg.IntersectBatch(r), agg.ordered_tris9, cam, n)
    _check(agg, oracle, bounce, 200_000, name + "/bounce")
    _check(agg, oracle, shadow, 200_000, name + "/shadow")
    # host pipeline: a chunk size that does not divide the batch gives the same records
    os.environ["NNBVH_CHUNK_RAYS"] = "300007"
    try:
        small = nb.GPUBVHAggregate(nodes=agg.nodes, ordered_tris9=agg.ordered_tris9)
    finally:
        del os.environ["NNBVH_CHUNK_RAYS"]
    assert np.array_equal(_words(small.IntersectBatch(bounce)), _words(agg.IntersectBatch(bounce)))
    small.close()
    agg.close()


def test_path_bounce_batches(oracle):
    """Config 2 (crown in BASELINE.json; any cached scene here): batches of a diffuse random
    walk, bounce 0 coherent ... bounce 4 incoherent."""
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import rays as R
    tris, cam, name = _scene("coffee_maker")
    agg = nb.GPUBVHAggregate.Create(tris, {"maxnodeprims": 4})
    batches = R.path_bounce_batches(lambda r: agg.IntersectBatch(r), agg.ordered_tris9, cam, res=768, bounces=4)
    assert len(batches) >= 3
    for k, b in enumerate(batches):
        _check(agg, oracle, b, 50_000, f"{name}/bounce{k}")
    agg.close()


def test_synthetic_soup_many_rays(oracle):
    """Config 4, scaled to what the test budget allows for the host side (1M triangles; the bench
    runs the 10M-triangle tree): 64M random rays generated on the device, traced resident,
    checksummed against a chunked second pass, oracle on a strided 100k subset."""
    import torch
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import scenes
    tris = scenes.synthetic_soup(1_000_000)
    agg = nb.GPUBVHAggregate.Create(tris, {"maxnodeprims": 4})
    n = 64 << 20
    g = torch.Generator(device="cuda").manual_seed(6)
    rays = torch.empty((n, 8), dtype=torch.float32, device="cuda")
    rays[:, 0:3] = torch.rand((n, 3), generator=g, device="cuda") * 1.5 - 0.25
    z = 1 - 2 * torch.rand(n, generator=g, device="cuda")
    phi = 2 * np.pi * torch.rand(n, generator=g, device="cuda")
    rad = torch.sqrt(torch.clamp(1 - z * z, min=0))
    rays[:, 4], rays[:, 5], rays[:, 6] = rad * torch.cos(phi), rad * torch.sin(phi), z
    rays[:, 3] = float("inf")
    rays[:, 7] = 0
    hits = torch.zeros((n, 5), dtype=torch.int32, device="cuda")
    occ = torch.zeros(n, dtype=torch.uint8, device="cuda")
    agg.IntersectBatch(rays, hits=hits)
    agg.IntersectPBatch(rays, occluded=occ)
    torch.cuda.synchronize()
    assert torch.equal(occ != 0, hits[:, 0] != -1)
    # same records when the batch is traced as 5 ragged sub-batches (what a sharded run does)
    hits2 = torch.zeros_like(hits)
    edges = [0, 1, n // 3 + 5, n // 2, n - 7, n]
    for a, b in zip(edges[:-1], edges[1:]):
        agg.IntersectBatch(rays[a:b], hits=hits2[a:b], n=b - a)
    torch.cuda.synchronize()
    assert torch.equal(hits, hits2)
    step = n // 100_000
    sub = rays[::step].cpu().numpy()
    nodes = agg.nodes.view(np.uint32).reshape(-1, 8)
    want = oracle.intersect(nodes, agg.ordered_tris9, sub)
    assert np.array_equal(hits[::step].cpu().numpy().view(np.uint32), want)
    agg.close()
