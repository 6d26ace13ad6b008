#!/usr/bin/env python3
"""tools/config5.py -- BASELINE.json configs[4]: synthetic 10M-triangle mesh, 256M random rays,
sharded over the ranks it is launched with (torchrun); the hit records are returned to rank 0 inside
the timed region (--gather peer: copy engines into a peer-mapped window over NVLink, chunk by chunk while
the next chunk is traced; --gather direct: the traversal kernel itself stores each hit record into
rank 0's window through the peer mapping (SM stores over NVLink, no copy at all); --gather nccl: one
NCCL gather after the kernel; none: kernel only).

Ray i is a pure function of its GLOBAL index (counter-based hash), so every world size traces the
same 256M rays; rank g takes the contiguous slice shard_range(N, g, G).  The
run prints Mrays/s (device timed, max over ranks) and three checksums of the hit records that
must be identical for every G:  number of hits, sum of primitive indices, xor of the t bit
patterns.  A strided subset is checked against the oracle on rank 0 (tools may use the oracle).
usage: [torchrun ...] config5.py [--tris 10000000] [--rays 268435456] [--oracle 200000]
"""
import argparse
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def gen_rays(a: int, b: int, dev) -> torch.Tensor:
    from nn_bvh_b200 import rays as R
    return R.hashed_random_rays(a, b, dev)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=10_000_000)
    ap.add_argument("--rays", type=int, default=256 << 20)
    ap.add_argument("--oracle", type=int, default=200_000)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--gather", default="peer", choices=["peer", "direct", "nccl", "none"])
    ap.add_argument("--chunks", type=int, default=3, help="chunks per rank (delivery overlaps the next chunk)")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import scenes, sharding
    tris = scenes.synthetic_soup(a.tris)
    t0 = time.time()
    nodes, order = nb.build_bvh(tris, 4, "sah")
    t_build = time.time() - t0
    agg = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=tris[order], order=order, device=local)
    lo, hi = sharding.shard_range(a.rays, rank, world)
    rays = gen_rays(lo, hi, dev)
    hits = torch.zeros((hi - lo, 5), dtype=torch.int32, device=dev)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    agg.set_stream(stream.cuda_stream)
    best = best_step = 1e30
    gather = a.gather if world > 1 else "none"
    win = sharding.PeerWindow(a.rays, (5,), torch.int32, dst=0) if gather in ("peer", "direct") else None
    direct = gather == "direct"
    glob = torch.zeros((a.rays, 5), dtype=torch.int32, device=dev) if (gather == "nccl" and rank == 0) else None
    n_loc = hi - lo
    nch = a.chunks if gather == "peer" else 1
    bounds = [n_loc * c // nch for c in range(nch + 1)]
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(a.reps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ev0.record(stream)
        kms = 0.0
        for c in range(len(bounds) - 1):
            s0, s1 = bounds[c], bounds[c + 1]
            if s1 == s0:
                continue
            if direct:   # the kernel's StoreHit goes straight into rank 0's memory (peer mapping)
                agg.IntersectBatch(rays[s0:s1], hits=win.base + (lo + s0) * win.row_bytes, n=s1 - s0)
                continue
            agg.IntersectBatch(rays[s0:s1], hits=hits[s0:s1], n=s1 - s0)
            if win is not None:
                win.deliver(hits[s0:s1], stream, at=s0)
        if win is not None and not direct:
            win.wait_on(stream)
        elif gather == "nccl":
            sharding.gather_records(hits, a.rays, dst=0, out=glob)
        ev1.record(stream)
        torch.cuda.synchronize()
        t = torch.tensor([agg.last_kernel_ms() if len(bounds) == 2 else 0.0, ev0.elapsed_time(ev1)],
                         dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best, best_step = min(best, t[0].item()), min(best_step, t[1].item())
    if direct:   # checksums below read this rank's records: fetch them back from the window (untimed)
        from nn_bvh_b200 import capi
        import ctypes
        if world > 1:
            dist.barrier()
        capi.check(capi.lib().nnbvh_peer_copy(capi.ptr(hits), ctypes.c_void_p(win.base + lo * win.row_bytes),
                                              n_loc * win.row_bytes, ctypes.c_void_p(stream.cuda_stream)))
        torch.cuda.synchronize()
    if len(bounds) != 2:
        best = best_step   # chunked: the kernels' own time is not separable from the overlapped delivery
    hit = hits[:, 0] != -1
    xor = torch.tensor(0, dtype=torch.int64, device=dev)
    tb = hits[:, 1].to(torch.int64)
    while tb.numel() > 1:   # xor-reduce
        if tb.numel() % 2:
            tb = torch.cat([tb, tb.new_zeros(1)])
        tb = tb[0::2] ^ tb[1::2]
    xor = tb[0]
    cs = torch.stack([hit.sum(), hits[:, 0].to(torch.int64)[hit].sum(), xor])
    if world > 1:
        parts = [torch.empty_like(cs) for _ in range(world)]
        dist.all_gather(parts, cs)
        tot = torch.stack(parts)
        cs = torch.stack([tot[:, 0].sum(), tot[:, 1].sum(), torch.tensor(
            int(np.bitwise_xor.reduce(tot[:, 2].cpu().numpy())), device=dev)])
    gathered_ok = None
    if rank == 0 and gather != "none":
        g = win.tensor if win is not None else glob
        ghit = g[:, 0] != -1
        gathered_ok = bool(int(ghit.sum()) == int(cs[0]) and
                           int(g[:, 0].to(torch.int64)[ghit].sum()) == int(cs[1]) and
                           bool(torch.equal(g[lo:hi], hits)))
    if rank == 0:
        ok = None
        if a.oracle:
            from oracle.pyoracle import Oracle
            stepn = max(1, (hi - lo) // a.oracle)
            sub = rays[::stepn].cpu().numpy()
            want = Oracle().intersect(agg.nodes.view(np.uint32).reshape(-1, 8), agg.ordered_tris9, sub)
            ok = bool(np.array_equal(hits[::stepn].cpu().numpy().view(np.uint32), want))
        print({"config": f"synthetic {a.tris} triangles, {a.rays} random rays", "n_gpus": world,
               "bvh_nodes": int(len(nodes)), "build_s": round(t_build, 1), "kernel_ms": round(best, 2),
               "step_ms_with_gather": round(best_step, 2), "gather": gather,
               "gathered_records_match": gathered_ok, "Mrays_s": round(a.rays / best_step / 1e3, 1), "hits": int(cs[0]), "sum_prim": int(cs[1]),
               "xor_t_bits": int(cs[2]), "oracle_subset_bit_exact": ok}, flush=True)
    if win is not None:
        win.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
