#!/usr/bin/env python3
"""tools/e2e_probe.py -- host-buffer path: raw PCIe copy rates and the C-ABI pipeline under chunk / slot knobs."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nn_bvh_b200 as nb
from nn_bvh_b200 import rays as R, scenes

n = 8 << 20
h = torch.empty((n, 8), dtype=torch.float32).pin_memory()
d = torch.empty((n, 8), dtype=torch.float32, device="cuda")
for name, a, b in (("H2D", d, h), ("D2H", h, d)):
    for _ in range(2):
        torch.cuda.synchronize(); t = time.perf_counter(); a.copy_(b, non_blocking=True); torch.cuda.synchronize()
        dt = time.perf_counter() - t
    print(f"{name} {n * 32 / dt / 1e9:.1f} GB/s")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
torch.cuda.synchronize(); t = time.perf_counter()
with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
h2 = torch.empty((n, 8), dtype=torch.float32).pin_memory(); d2 = torch.empty((n, 8), dtype=torch.float32, device="cuda")
torch.cuda.synchronize(); t = time.perf_counter()
with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize(); dt = time.perf_counter() - t
print(f"bidirectional: {n * 32 / dt / 1e9:.1f} GB/s each way")

tris, cam, name = scenes.load("bathroom")
nodes, order = nb.build_bvh(tris, 4, "sah")
base = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=tris[order])
bounce, shadow = R.bounce_and_shadow_rays(lambda r: base.IntersectBatch(r), tris[order], cam, n)
hb = torch.from_numpy(bounce.view(np.float32).reshape(-1, 8)).pin_memory()
hs = torch.from_numpy(shadow.view(np.float32).reshape(-1, 8)).pin_memory()
hh = torch.zeros((n, 5), dtype=torch.int32).pin_memory()
ho = torch.zeros(n, dtype=torch.uint8).pin_memory()
pb = [torch.from_numpy(np.ascontiguousarray(hb.numpy()[:, k])).pin_memory() for k in (0, 1, 2, 4, 5, 6)]
ps = [torch.from_numpy(np.ascontiguousarray(hs.numpy()[:, k])).pin_memory() for k in (0, 1, 2, 4, 5, 6, 3)]
from concurrent.futures import ThreadPoolExecutor
pool = ThreadPoolExecutor(2)
for chunk, slots, ramp in [(1 << 20, 4, 0), (1 << 20, 4, 1), (1 << 21, 4, 1), (1 << 21, 3, 1), (1 << 19, 6, 1), (1 << 20, 6, 1),
                           (1 << 21, 4, 0), (1 << 22, 3, 1)]:
    os.environ["NNBVH_CHUNK_RAYS_SOA"] = str(chunk); os.environ["NNBVH_SLOTS"] = str(slots)
    os.environ["NNBVH_CHUNK_RAMP"] = str(ramp)
    agg = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=tris[order])
    agg2 = agg.clone()
    one = two = 1e9
    for it in range(4):
        torch.cuda.synchronize(); t = time.perf_counter()
        agg.IntersectBatchSoA(*pb, None, hits=hh)
        agg.IntersectBatchSoA(*ps, hits=ho, any_hit=True)
        dt = time.perf_counter() - t
        if it:
            one = min(one, dt)
    for it in range(4):
        torch.cuda.synchronize(); t = time.perf_counter()
        f1 = pool.submit(agg.IntersectBatchSoA, *pb, None, hits=hh)
        f2 = pool.submit(agg2.IntersectBatchSoA, *ps, hits=ho, any_hit=True)
        f1.result(); f2.result()
        dt = time.perf_counter() - t
        if it:
            two = min(two, dt)
    print(f"SoA chunk {chunk:8d} slots {slots} ramp {ramp}: one lane {2 * n / one / 1e6:7.1f} Mrays/s   two lanes {2 * n / two / 1e6:7.1f} Mrays/s", flush=True)
    agg2.close(); agg.close()
