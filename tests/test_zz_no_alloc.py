"""tests/test_zz_no_alloc.py -- nnbvh_intersect_closest_noalloc (include/nnbvh.h).

The entry point has no counterpart in the reference (added for
callers that cannot call cudaMalloc); its results must be those of nnbvh_intersect_closest, bit for bit
(the tolerance of every traversal parity test here: integer / bit equality of the 20-byte hit records).
Named to be collected last.
"""
import ctypes

import numpy as np
import pytest

from conftest import rand_rays, soup

pytestmark = pytest.mark.gpu


def _free_bytes(torch):
    torch.cuda.synchronize()
    return torch.cuda.mem_get_info()[0]


def test_noalloc_matches_the_allocating_entry_point_twice_in_a_row(oracle):
    import torch
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import capi
    rng = np.random.default_rng(77)
    T = 50_000
    tris = soup(rng, T, 0.03)
    tris[::997, 3:6] = tris[::997, 0:3]            # some degenerate triangles
    agg = nb.GPUBVHAggregate(tris, 4, "sah", device=0)
    dev = torch.device("cuda", 0)
    L = capi.lib()
    sizes = (300_001, capi.NOALLOC_MAX_RAYS)        # ragged, then the stated maximum
    # everything the test itself allocates exists before the calls that are checked for allocations
    d_rays = [torch.empty((n, 8), dtype=torch.float32, device=dev) for n in sizes]
    d_want = [torch.zeros((n, 5), dtype=torch.int32, device=dev) for n in sizes]
    d_got = [torch.full((n, 5), -7, dtype=torch.int32, device=dev) for n in sizes]
    rays = []
    for k, n in enumerate(sizes):
        r = rand_rays(rng, n, tmax=np.inf)
        r[: n // 3, 3] = rng.random(n // 3).astype(np.float32)    # short rays
        r[5::1013, 4] = 0.0                                        # axis-parallel: the exact (deferred) launch
        r[7::1019, 5] = -0.0
        rays.append(r)
        d_rays[k].copy_(torch.from_numpy(r))
        agg.IntersectBatch(d_rays[k], hits=d_want[k], n=n)        # the existing entry point, device pointers
    torch.cuda.synchronize()
    deferred = [int((r[:, 4:7] == 0).any(axis=1).sum()) for r in rays]
    assert min(deferred) > 100

    free_before = None
    for k, n in enumerate(sizes):                  # two calls in a row on the shared static scratch
        if k == 1:
            free_before = _free_bytes(torch)
        capi.check(L.nnbvh_intersect_closest_noalloc(agg._h, capi.ptr(d_rays[k]), n, capi.ptr(d_got[k])))
    free_after = _free_bytes(torch)
    assert free_after == free_before, "the second no-alloc call changed the device's free memory"
    for k, n in enumerate(sizes):
        assert torch.equal(d_got[k], d_want[k]), f"call {k}: records differ from nnbvh_intersect_closest"
    # ... and a third, repeating the first batch after the larger one left its contents in the scratch
    # (the engine launches on its own stream: order torch's fill before it and the compare after it)
    d_got[0].fill_(-7)
    torch.cuda.synchronize()
    agg.IntersectBatchNoAlloc(d_rays[0], d_got[0], n=sizes[0])
    torch.cuda.synchronize()
    assert torch.equal(d_got[0], d_want[0])
    # anchor: the existing entry point itself against the oracle on a subset
    sub = np.ascontiguousarray(rays[0][:20_000])
    nodes = agg.nodes.view(np.uint32).reshape(-1, 8)
    want = oracle.intersect(nodes, agg.ordered_tris9, sub)
    got = d_got[0][:20_000].cpu().numpy().view(np.uint32)
    assert np.array_equal(got, want)
    assert (want[:, 0] != 0xFFFFFFFF).sum() > 2000


def test_noalloc_rejects_what_it_cannot_do():
    import torch
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import capi
    rng = np.random.default_rng(78)
    agg = nb.GPUBVHAggregate(soup(rng, 2000, 0.05), 4, "sah", device=0)
    dev = torch.device("cuda", 0)
    L = capi.lib()
    n = capi.NOALLOC_MAX_RAYS + 1
    d_rays = torch.zeros((n, 8), dtype=torch.float32, device=dev)
    d_hits = torch.zeros((n, 5), dtype=torch.int32, device=dev)
    rc = L.nnbvh_intersect_closest_noalloc(agg._h, capi.ptr(d_rays), n, capi.ptr(d_hits))
    assert rc == -1 and b"NNBVH_NOALLOC_MAX_RAYS" in L.nnbvh_last_error()
    host = np.zeros((16, 8), dtype=np.float32)
    rc = L.nnbvh_intersect_closest_noalloc(agg._h, capi.ptr(host), 16, capi.ptr(d_hits))
    assert rc == -1 and b"device pointers" in L.nnbvh_last_error()
    assert L.nnbvh_intersect_closest_noalloc(agg._h, None, 0, None) == 0      # empty batch
    fresh = ctypes.c_void_p()
    capi.check(L.nnbvh_create(None, ctypes.byref(fresh)))
    assert L.nnbvh_intersect_closest_noalloc(fresh, capi.ptr(d_rays), 16, capi.ptr(d_hits)) == -3   # no scene
    L.nnbvh_destroy(fresh)
    alpha = nb.GPUBVHAggregate(soup(rng, 500, 0.05), 4, "sah", device=0, alpha=np.full(500, 0.5, np.float32))
    rc = L.nnbvh_intersect_closest_noalloc(alpha._h, capi.ptr(d_rays), 16, capi.ptr(d_hits))
    assert rc == -1 and b"alpha" in L.nnbvh_last_error()
