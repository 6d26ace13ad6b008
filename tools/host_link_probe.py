#!/usr/bin/env python3
"""tools/host_link_probe.py -- where does end-to-end scaling go at N ranks on one host?

Run under torchrun (one rank per GPU).  Every rank times pinned-host <-> device copies of 512 MiB
(CUDA events): H2D, D2H and both at once; first each rank ALONE (the others idle at a barrier), then ALL
ranks together.  If the concurrent per-rank rate falls well below the solo rate the limit is a shared
host resource (root complex / PCIe switch uplinks / host DRAM), not this library's pipeline.  Also:
a host-side memcpy rate per rank (all ranks at once), the NUMA layout the process sees, and the same
H2D copy from write-combined pinned memory.
usage: torchrun --nproc-per-node N tools/host_link_probe.py
"""
import ctypes
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
rt = ctypes.CDLL("libcudart.so.12")
N = 512 << 20


def host_alloc(flags):
    p = ctypes.c_void_p()
    rc = rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(N), ctypes.c_uint(flags))
    assert rc == 0, rc
    ctypes.memset(p, 1, N)   # first touch on this rank's cores
    return p


def dev_alloc():
    p = ctypes.c_void_p()
    assert rt.cudaMalloc(ctypes.byref(p), ctypes.c_size_t(N)) == 0
    return p


def stream():
    s = ctypes.c_void_p()
    assert rt.cudaStreamCreateWithFlags(ctypes.byref(s), 1) == 0
    return s


def event():
    e = ctypes.c_void_p()
    assert rt.cudaEventCreate(ctypes.byref(e)) == 0
    return e


h_in, h_out, h_wc = host_alloc(0), host_alloc(0), host_alloc(4)   # 4 = cudaHostAllocWriteCombined
d_in, d_out = dev_alloc(), dev_alloc()
s1, s2 = stream(), stream()
e0, e1, e2, e3 = event(), event(), event(), event()


def timed(kind, reps=4):
    """-> GB/s per direction of this rank."""
    best = 0.0
    for _ in range(reps):
        rt.cudaDeviceSynchronize()
        if kind in ("h2d", "both", "h2d_wc"):
            rt.cudaEventRecord(e0, s1)
            rt.cudaMemcpyAsync(d_in, h_wc if kind == "h2d_wc" else h_in, ctypes.c_size_t(N), 1, s1)
            rt.cudaEventRecord(e1, s1)
        if kind in ("d2h", "both"):
            rt.cudaEventRecord(e2, s2)
            rt.cudaMemcpyAsync(h_out, d_out, ctypes.c_size_t(N), 2, s2)
            rt.cudaEventRecord(e3, s2)
        rt.cudaDeviceSynchronize()
        ms = ctypes.c_float(0)
        rates = []
        if kind in ("h2d", "both", "h2d_wc"):
            rt.cudaEventElapsedTime(ctypes.byref(ms), e0, e1)
            rates.append(N / ms.value / 1e6)
        if kind in ("d2h", "both"):
            rt.cudaEventElapsedTime(ctypes.byref(ms), e2, e3)
            rates.append(N / ms.value / 1e6)
        best = max(best, min(rates))
    return best


def barrier():
    if world > 1:
        dist.barrier()


def gather(v):
    t = torch.tensor([v], dtype=torch.float64, device=dev)
    if world == 1:
        return [v]
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return [round(o.item(), 1) for o in out]


res = {"world": world, "bytes_per_copy": N}
for kind in ("h2d", "d2h", "both", "h2d_wc"):
    solo = 0.0
    for r in range(world):     # one rank at a time
        barrier()
        if r == rank:
            solo = timed(kind)
        barrier()
    barrier()
    conc = timed(kind)         # everybody at once
    barrier()
    res[kind] = {"solo_GBs_per_rank": gather(solo), "concurrent_GBs_per_rank": gather(conc)}
    res[kind]["concurrent_sum_GBs"] = round(sum(res[kind]["concurrent_GBs_per_rank"]), 1)
# host memcpy, all ranks at once
a = np.ones(N // 8, dtype=np.float64)
b = np.empty_like(a)
barrier()
t0 = time.perf_counter()
for _ in range(3):
    np.copyto(b, a)
dt = time.perf_counter() - t0
res["host_memcpy_concurrent_GBs_per_rank"] = gather(3 * 2 * N / dt / 1e9)   # read + write
if rank == 0:
    try:
        nodes = sorted(d for d in os.listdir("/sys/devices/system/node") if d.startswith("node"))
        res["numa_nodes"] = nodes
        res["cpus_allowed"] = len(os.sched_getaffinity(0))
        for nd in nodes:
            with open(f"/sys/devices/system/node/{nd}/cpulist") as f:
                res[nd + "_cpus"] = f.read().strip()
    except OSError as e:
        res["numa"] = str(e)
    print(json.dumps(res), flush=True)
if world > 1:
    dist.destroy_process_group()
