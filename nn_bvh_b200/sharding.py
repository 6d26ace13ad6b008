"""nn_bvh_b200/sharding.py -- ray-batch sharding across ranks (one process per GPU).

The reference has no multi-device path at all (no NCCL/MPI; `--gpu-device`
picks one GPU, /root/reference/src/pbrt/pbrt.cpp:68).  Rays are independent, so the path
shards with no data-path collective: the scene is replicated on every rank, rank g of G
traces the contiguous slice [g*N/G, (g+1)*N/G) of the batch, and the one
exchange step is returning the 20-byte hit records / occlusion bytes to the caller's rank
(an NCCL gather over NVLink on GPUs; gloo in the CPU tests).

Nothing here intersects anything: `trace` callables are injected (GPUBVHAggregate methods in
bench.py, anything with the same contract in tests).
"""
from __future__ import annotations

from typing import Callable, Optional

import torch
import torch.distributed as dist


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous slice of rank `rank`: [rank*n//world, (rank+1)*n//world).  Slices tile
    [0, n) exactly, differ in size by at most one, and may be empty when n < world."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    return (rank * n) // world, ((rank + 1) * n) // world


def weighted_ranges(n: int, weights) -> list:
    """Contiguous slices of [0, n) in rank order with sizes proportional to `weights` (one per rank): the
    slices tile [0, n) exactly (equal weights: sizes differ by at most one, like shard_range).  Used when one rank has other work
    in the step -- rank 0 of a sharded batch takes every other rank's hit records in over its NVLink
    ports while it traces -- and should trace a smaller share."""
    w = [float(x) for x in weights]
    if not w or min(w) < 0 or sum(w) <= 0:
        raise ValueError("weights must be non-negative and not all zero")
    tot, acc, cuts = sum(w), 0.0, [0]
    for x in w:
        acc += x
        cuts.append(int(round(n * acc / tot)))
    cuts[-1] = n
    return [(cuts[i], max(cuts[i], cuts[i + 1])) for i in range(len(w))]


def max_shard(n: int, world: int) -> int:
    return max(shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world))


def gather_records(local: torch.Tensor, n_total: int, dst: int = 0, group=None,
                   out: Optional[torch.Tensor] = None, async_op: bool = False):
    """Gather per-rank result rows (hit records (m,5) int32 or occlusion bytes (m,)) into one
    (n_total, ...) tensor on rank `dst`, in global ray order; returns it on `dst`, None elsewhere.

    Equal shards (n_total % world == 0) are received straight into slices of `out` -- no
    staging copy -- and with async_op=True the call returns (out_or_None, work) without
    waiting, so the caller can overlap the transfer with its next kernel.  Ragged shards are
    padded to the largest shard (NCCL gather needs same-shape contributions); the padding is
    dropped on `dst`."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    m = max_shard(n_total, world)
    a, b = shard_range(n_total, rank, world)
    if local.shape[0] != b - a:
        raise ValueError(f"rank {rank}: local result has {local.shape[0]} rows, shard has {b - a}")
    tail = tuple(local.shape[1:])
    if rank == dst and out is None:
        out = local.new_empty((n_total,) + tail)
    if m == 0:
        return ((out if rank == dst else None), None) if async_op else (out if rank == dst else None)
    if n_total % world == 0:
        parts = list(out.view((world, m) + tail).unbind(0)) if rank == dst else None
        work = dist.gather(local.contiguous(), parts, dst=dst, group=group, async_op=async_op)
        res = out if rank == dst else None
        return (res, work) if async_op else res
    if async_op:
        raise ValueError("async gather needs equal shards (n_total % world == 0)")
    send = local.new_zeros((m,) + tail)
    send[: b - a] = local
    parts = [torch.empty_like(send) for _ in range(world)] if rank == dst else None
    dist.gather(send, parts, dst=dst, group=group)
    if rank != dst:
        return None
    for r, part in enumerate(parts):
        ra, rb = shard_range(n_total, r, world)
        out[ra:rb] = part[: rb - ra]
    return out


def scatter_rays(rays: Optional[torch.Tensor], n_total: int, src: int = 0, group=None,
                 device=None) -> torch.Tensor:
    """Distribute an (n_total, 8) float32 ray batch held by rank `src` so that each rank
    ends up with its shard_range slice (padded scatter, padding dropped)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    m = max_shard(n_total, world)
    a, b = shard_range(n_total, rank, world)
    if device is None:
        device = rays.device if rays is not None else torch.device("cpu")
    recv = torch.zeros((m, 8), dtype=torch.float32, device=device)
    if m == 0:
        return recv
    parts = None
    if rank == src:
        parts = []
        for r in range(world):
            ra, rb = shard_range(n_total, r, world)
            part = torch.zeros((m, 8), dtype=torch.float32, device=device)
            part[: rb - ra] = rays[ra:rb]
            parts.append(part)
    dist.scatter(recv, parts, src=src, group=group)
    return recv[: b - a]


def trace_sharded(trace: Callable[[torch.Tensor], torch.Tensor], rays_full: torch.Tensor,
                  dst: int = 0, group=None) -> Optional[torch.Tensor]:
    """Every rank holds (or can regenerate) the full batch: trace the own slice with `trace`
    (rays (m,8) -> result rows (m,...)) and gather the records to `dst` in global order."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    n = rays_full.shape[0]
    a, b = shard_range(n, rank, world)
    local = trace(rays_full[a:b])
    return gather_records(local, n, dst=dst, group=group)


class PeerWindow:
    """The global result array of a sharded batch, living on rank `dst`'s GPU and written by every
    rank directly: rank r delivers its slice with a copy-engine transfer over NVLink (CUDA IPC
    mapping of `dst`'s allocation, libnnbvh nnbvh_peer_*), ordered after the kernel that produced
    it and overlapping whatever the rank launches next.  No collective kernel runs, so nothing
    competes with the persistent traversal CTAs for SMs -- an NCCL gather issued next to them only
    starts once they drain (N=8: 7.30 vs 6.18 ms per step, profiles/r01j_*).

    rows: (n_total,) + tail of `dtype`; equal or ragged shards (shard_range).  `tensor` is the
    array on `dst` (None elsewhere).  GPU only; the gloo tests cover gather_records instead."""

    def __init__(self, n_total: int, tail: tuple, dtype: torch.dtype, dst: int = 0, group=None,
                 copy_streams: int = 1, ranges=None):
        import ctypes
        from . import capi
        self._capi, self._ct = capi, ctypes
        L = capi.lib()
        self.world, self.rank, self.dst = dist.get_world_size(group), dist.get_rank(group), dst
        self._group = group
        self.n_total, self.tail, self.dtype = n_total, tuple(tail), dtype
        # ranges: [(a, b)] per rank (weighted_ranges); default = equal contiguous slices (shard_range)
        self.ranges = list(ranges) if ranges is not None else [shard_range(n_total, r, self.world) for r in range(self.world)]
        self.row_bytes = torch.empty((), dtype=dtype).element_size()
        for t in self.tail:
            self.row_bytes *= t
        nbytes = n_total * self.row_bytes
        handle = (ctypes.c_ubyte * 64)()
        p = ctypes.c_void_p()
        box = [None]
        self.base = 0
        self._owner = self.rank == dst
        err = None
        if self._owner:   # whatever happens here, the broadcast below must still be reached
            try:
                capi.check(L.nnbvh_peer_alloc(nbytes, ctypes.byref(p), handle))
                box[0] = bytes(handle)
                self.base = p.value
            except Exception as e:   # noqa: BLE001
                err = e
        dist.broadcast_object_list(box, src=dst, group=group)
        if box[0] is None:
            raise RuntimeError(f"PeerWindow: rank {dst} could not allocate / export the window") from err
        if not self._owner:
            handle = (ctypes.c_ubyte * 64).from_buffer_copy(box[0])
            capi.check(L.nnbvh_peer_open(handle, ctypes.byref(p)))
        self.base = p.value
        self.tensor = None
        if self._owner:
            shape = (n_total,) + self.tail
            iface = {"shape": shape, "typestr": {torch.int32: "<i4", torch.uint8: "|u1", torch.float32: "<f4"}[dtype],
                     "data": (self.base, False), "version": 3, "strides": None}
            holder = type("PeerWindowMemory", (), {"__cuda_array_interface__": iface})()
            self.tensor = torch.as_tensor(holder, device=torch.device("cuda", torch.cuda.current_device()))
        # copy_streams > 1: a delivery is split into that many pieces on as many streams, i.e. copy engines
        # (one engine does not fill an NVLink 5 port: 2.7 GB took 8.6 ms = 312 GB/s, r02k)
        self.streams = [torch.cuda.Stream() for _ in range(max(1, copy_streams))]
        self.stream = self.streams[0]
        self._ready = torch.cuda.Event()
        self._dones = [torch.cuda.Event() for _ in self.streams]
        self._done = self._dones[0]

    def deliver(self, local: torch.Tensor, producer: "torch.cuda.Stream", at: int = None):
        """Enqueue the copy of this rank's slice, ordered after everything `producer` holds now.
        Returns at once; wait_on(stream) makes a stream wait for the delivery.
        at: deliver `local` as rows [at, at + len(local)) of this rank's shard (a batch traced in chunks
        delivers each chunk while the next one is traced); None = the whole shard."""
        a, b = self.ranges[self.rank]
        if at is None:
            if local.shape[0] != b - a:
                raise ValueError(f"rank {self.rank}: local result has {local.shape[0]} rows, shard has {b - a}")
            at = 0
        elif at < 0 or at + local.shape[0] > b - a:
            raise ValueError(f"rank {self.rank}: rows [{at}, {at + local.shape[0]}) outside a shard of {b - a}")
        src = local.contiguous()           # a temporary only for a non-contiguous input
        self._ready.record(producer)
        rows = local.shape[0]
        k = len(self.streams) if rows >= (1 << 16) else 1
        for i in range(k):
            r0, r1 = rows * i // k, rows * (i + 1) // k
            st = self.streams[i]
            src.record_stream(st)          # the caching allocator must not reuse the block before the copy has read it
            st.wait_event(self._ready)
            if r1 > r0:
                self._capi.check(self._capi.lib().nnbvh_peer_copy(
                    self._ct.c_void_p(self.base + (a + at + r0) * self.row_bytes),
                    self._ct.c_void_p(src.data_ptr() + r0 * self.row_bytes),
                    (r1 - r0) * self.row_bytes, self._ct.c_void_p(st.cuda_stream)))
            self._dones[i].record(st)

    def wait_on(self, stream: "torch.cuda.Stream"):
        """Make `stream` wait for THIS rank's last delivery.  It orders nothing across ranks: before
        rank `dst` reads `tensor` every rank must have waited for its own delivery and the ranks must
        have met (dist.barrier(), or any collective ordered after the wait) -- bench.py's barrier()."""
        for ev in self._dones:
            stream.wait_event(ev)

    def close(self, collective: bool = True):
        """Collective (default): every rank of the group calls it.  The mappings are closed before the owner
        frees the allocation (freeing exported memory that a peer still maps is undefined).
        collective=False: no barrier -- the caller orders the non-owners' close before the owner's."""
        L = self._capi.lib()
        torch.cuda.synchronize()
        if self.base and not self._owner:
            L.nnbvh_peer_close(self._ct.c_void_p(self.base))
            self.base = 0
        if self.world > 1 and collective:
            dist.barrier(self._group)
        if self.base and self._owner:
            self.tensor = None
            L.nnbvh_peer_free(self._ct.c_void_p(self.base))
            self.base = 0
