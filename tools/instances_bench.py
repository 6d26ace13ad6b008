#!/usr/bin/env python3
"""tools/instances_bench.py -- throughput of the one-thread-per-ray paths (instanced / general scenes) with and
without the ray-ordering pre-pass.  Scene: K instances of one object (the killeroos soup, 66 532 triangles) placed in
a box over a tessellated floor; 8 Mi incoherent rays (random origins in the box, random directions).
usage: instances_bench.py [K=64] [rays=8388608]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nn_bvh_b200 as nb
from nn_bvh_b200 import scenes
from nn_bvh_b200.aggregate import make_instances

K = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n = int(sys.argv[2]) if len(sys.argv) > 2 else 8 << 20
obj, _, _ = scenes.load("killeroos")
v = obj.reshape(-1, 3)
c, ext = (v.min(0) + v.max(0)) / 2, (v.max(0) - v.min(0)).max()
obj = ((obj.reshape(-1, 3, 3) - c) / ext).reshape(-1, 9).astype(np.float32)     # unit-sized object at the origin
rng = np.random.default_rng(3)
g = np.linspace(0, 1, 65, dtype=np.float32)
u, w = np.meshgrid(g, g, indexing="ij")
q = np.stack([np.stack([u[:-1, :-1], w[:-1, :-1]], -1), np.stack([u[1:, :-1], w[1:, :-1]], -1),
              np.stack([u[1:, 1:], w[1:, 1:]], -1), np.stack([u[:-1, 1:], w[:-1, 1:]], -1)], 2).reshape(-1, 4, 2)
t2 = np.concatenate([q[:, [0, 1, 2]], q[:, [2, 3, 0]]], 0)
floor = np.zeros((len(t2), 3, 3), dtype=np.float32)
floor[..., 0], floor[..., 2] = t2[..., 0], t2[..., 1]
floor = floor.reshape(-1, 9)
mats = np.tile(np.eye(4), (K, 1, 1))
for k in range(K):
    a = rng.random() * 6.28
    s = 0.08 + 0.08 * rng.random()
    mats[k, :3, :3] = np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]]) * s
    mats[k, :3, 3] = [rng.random(), 0.05 + 0.5 * rng.random(), rng.random()]
inst = make_instances(mats, np.zeros(K, dtype=np.uint32))
rays = np.zeros((n, 8), dtype=np.float32)
rays[:, :3] = rng.random((n, 3), dtype=np.float32) * [1, 0.7, 1]
d = rng.normal(size=(n, 3)).astype(np.float32)
rays[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
rays[:, 3] = np.inf
dr = torch.from_numpy(rays).cuda()
hits = torch.zeros((n, 5), dtype=torch.int32, device="cuda")
inner = torch.zeros(n, dtype=torch.int32, device="cuda")
ref = None
for step, mode in (("0", "0"), ("0", "2"), ("1", "0"), ("1", "2")):
    os.environ["NNBVH_BIN"] = mode
    os.environ["NNBVH_INST_STEP"] = step
    agg = nb.GPUBVHAggregate(floor, 4, "sah", objects=[obj], instances=inst)
    best = 1e9
    for it in range(4):
        agg.IntersectBatchInstanced(dr, hits=hits, inner=inner, n=n)
        ms = agg.last_kernel_ms()
        if it:
            best = min(best, ms)
    h = hits.cpu().numpy()
    if ref is None:
        ref = h
    print(f"instances {K} x {len(obj)} triangles, {n} rays, {'StepKernel<INST>' if step == '1' else 'BasicKernel<INST>'}, NNBVH_BIN={mode}: {n / best / 1e3:8.1f} Mrays/s "
          f"({best:.2f} ms incl. pre-pass), hit fraction {(h[:, 0] != -1).mean():.3f}, same records as unsorted: {np.array_equal(h, ref)}")
    agg.close()
