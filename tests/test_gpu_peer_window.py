"""GPU, two devices: the sharded exchange step through a peer-mapped result window
(sharding.PeerWindow: CUDA IPC + copy engines over NVLink).  Skipped on a one-GPU box; the N>1
bench lines assert the same property on every run."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, rand_rays, soup

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, out_dir):
    import sys
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import nn_bvh_b200 as nb
        from nn_bvh_b200 import sharding
        rng = np.random.default_rng(42)            # same scene and batch on every rank
        tris = soup(rng, 20_000, 0.02)
        rays = rand_rays(rng, n)
        agg = nb.GPUBVHAggregate(tris, 4, "sah", device=rank)
        stream = torch.cuda.Stream()
        torch.cuda.set_stream(stream)
        agg.set_stream(stream.cuda_stream)
        a, b = sharding.shard_range(n, rank, world)          # ragged: n is odd
        d_rays = torch.from_numpy(rays[a:b]).cuda()
        d_hits = torch.zeros((b - a, 5), dtype=torch.int32, device="cuda")
        d_occ = torch.zeros(b - a, dtype=torch.uint8, device="cuda")
        wh = sharding.PeerWindow(n, (5,), torch.int32, dst=0)
        wo = sharding.PeerWindow(n, (), torch.uint8, dst=0)
        for _ in range(3):                                    # repeated deliveries into the same window
            wh.wait_on(stream)
            wo.wait_on(stream)
            agg.IntersectBatch(d_rays, hits=d_hits, n=b - a)
            wh.deliver(d_hits, stream)
            agg.IntersectPBatch(d_rays, occluded=d_occ, n=b - a)
            wo.deliver(d_occ, stream)
            wh.wait_on(stream)
            wo.wait_on(stream)
        torch.cuda.synchronize()
        dist.barrier()
        if rank == 0:
            full = nb.GPUBVHAggregate(nodes=agg.nodes, ordered_tris9=agg.ordered_tris9, device=0)
            want_h = np.ascontiguousarray(full.IntersectBatch(rays)).view(np.int32).reshape(-1, 5)
            want_o = full.IntersectPBatch(rays)
            ok = bool(np.array_equal(wh.tensor.cpu().numpy(), want_h) and
                      np.array_equal(wo.tensor.cpu().numpy(), want_o))
            with open(os.path.join(out_dir, "ok"), "w") as f:
                f.write("1" if ok else "0")
        dist.barrier()
        wh.close()
        wo.close()
    finally:
        dist.destroy_process_group()


def test_peer_window_delivers_every_ranks_slice(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    mp.spawn(_worker, args=(2, _free_port(), 300_001, str(tmp_path)), nprocs=2, join=True)
    assert open(tmp_path / "ok").read() == "1"
