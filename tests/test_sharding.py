"""CPU: the N>1 host logic (shard ranges, padded gather / scatter of records) under a
world-size-2 and -3 `gloo` group.  The per-shard tracer injected here is the oracle (tests may
use it); on the GPU box bench.py injects GPUBVHAggregate and NCCL does the same gather."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, rand_rays, soup


def test_shard_ranges_tile_exactly():
    from nn_bvh_b200.sharding import max_shard, shard_range
    for n in (0, 1, 2, 7, 8, 9, 1000, 8388608, 256_000_000):
        for world in (1, 2, 3, 4, 8):
            edges = [shard_range(n, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[r][1] == edges[r + 1][0] for r in range(world - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1 and max(sizes) == max_shard(n, world)
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, out_dir):
    import sys
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from nn_bvh_b200 import sharding
        from oracle.pyoracle import Oracle
        rng = np.random.default_rng(42)            # same scene and batch on every rank
        tris = soup(rng, 3000, 0.05)
        orc = Oracle()
        nodes, order = orc.build_bvh(tris, 4, "sah")
        ordered = tris[order]
        rays = rand_rays(rng, n)

        def closest(r):
            h = orc.intersect(nodes, ordered, r.numpy(), threads=1)
            return torch.from_numpy(h.view(np.int32))

        def shadow(r):
            return torch.from_numpy(orc.intersect_p(nodes, ordered, r.numpy(), threads=1))

        full = torch.from_numpy(rays)
        hits = sharding.trace_sharded(closest, full, dst=0)
        occ = sharding.trace_sharded(shadow, full, dst=0)
        # rank 0 also scatters the batch and the others trace what they received
        mine = sharding.scatter_rays(full if rank == 0 else None, n, src=0)
        a, b = sharding.shard_range(n, rank, world)
        assert mine.shape[0] == b - a and torch.equal(mine, full[a:b])
        hits2 = sharding.gather_records(closest(mine), n, dst=0)
        if rank == 0:
            want = orc.intersect(nodes, ordered, rays, threads=1).view(np.int32)
            wocc = orc.intersect_p(nodes, ordered, rays, threads=1)
            assert hits.shape == (n, 5) and np.array_equal(hits.numpy(), want)
            assert np.array_equal(hits2.numpy(), want)
            assert np.array_equal(occ.numpy(), wocc)
            open(os.path.join(out_dir, f"ok_{world}_{n}"), "w").write("ok")
        else:
            assert hits is None and occ is None and hits2 is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 10001), (2, 4096), (2, 1), (3, 2), (2, 0)])
def test_sharded_trace_matches_single_process(tmp_path, world, n):
    """Results gathered from `world` ranks are identical, record for record and in global ray
    order, to tracing the whole batch in one process (incl. ragged, n < world and empty)."""
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n, str(tmp_path)), nprocs=world, join=True)
    assert os.path.exists(tmp_path / f"ok_{world}_{n}")


def test_weighted_ranges_tile_the_batch():
    from nn_bvh_b200 import sharding
    for n, w in ((100, [0.5, 1, 1, 1]), (10, [1, 1, 1]), (7, [0, 1]), (268435456, [0.7] + [1.0] * 7), (3, [1] * 8)):
        r = sharding.weighted_ranges(n, w)
        assert len(r) == len(w) and r[0][0] == 0 and r[-1][1] == n
        assert all(a <= b for a, b in r) and all(r[i][1] == r[i + 1][0] for i in range(len(r) - 1))
    eq = sharding.weighted_ranges(1000, [1] * 7)
    assert max(b - a for a, b in eq) - min(b - a for a, b in eq) <= 1
    small = sharding.weighted_ranges(1 << 20, [0.5, 1, 1, 1])
    assert abs((small[0][1] - small[0][0]) * 2 - (small[1][1] - small[1][0])) <= 2
    with pytest.raises(ValueError):
        sharding.weighted_ranges(10, [0, 0])
