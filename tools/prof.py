#!/usr/bin/env python3
"""tools/prof.py [scene] [nrays] [kind] [iters] -- run a few launches of one ray kind (for ncu / timing).
kind: bounce | shadow | primary.  Knobs via NNBVH_* env vars."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nn_bvh_b200 as nb
from nn_bvh_b200 import rays as R, scenes

scene = sys.argv[1] if len(sys.argv) > 1 else "bathroom"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 8 << 20
kind = sys.argv[3] if len(sys.argv) > 3 else "bounce"
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
tris, cam, name = scenes.load(scene)
nodes, order = nb.build_bvh(tris, 4, "sah")
ordered = tris[order]
agg = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=ordered)
if kind == "random":
    r = R.random_rays(n)
elif kind == "primary":
    res = int(np.ceil(np.sqrt(n)))
    r = R.camera_rays(cam, res, res)[:n]
else:
    b, s = R.bounce_and_shadow_rays(lambda x: agg.IntersectBatch(x), ordered, cam, n)
    r = b if kind == "bounce" else s
n = len(r)
d = torch.from_numpy(r.view(np.float32).reshape(-1, 8)).cuda()
d_hits = torch.zeros((n, 5), dtype=torch.int32, device="cuda")
d_occ = torch.zeros(n, dtype=torch.uint8, device="cuda")
for _ in range(iters):
    if kind == "shadow":
        agg.IntersectPBatch(d, occluded=d_occ, n=n)
    else:
        agg.IntersectBatch(d, hits=d_hits, n=n)
    ms = agg.last_kernel_ms()
    print(f"{name} {kind} {n} rays: {ms:.3f} ms  {n / ms / 1e3:.1f} Mrays/s", flush=True)
if os.environ.get("PROF_COUNTERS"):
    m = min(n, 1 << 20)
    _, nv, tt = agg.IntersectCounters(r[:m], any_hit=(kind == "shadow"))
    io = 33 if kind == "shadow" else 52
    bpr = io + 32.0 * nv.mean() + 36.0 * tt.mean()
    print(f"nodesVisited {nv.mean():.2f} triTests {tt.mean():.2f} algorithmic B/ray {bpr:.0f} -> "
          f"{bpr * n / ms / 1e6:.0f} GB/s algorithmic; scene {len(nodes) * 32 / 1e6:.0f} MB nodes32 + "
          f"{len(ordered) * 48 / 1e6:.0f} MB tris48 + {(len(nodes) // 2) * 64 / 1e6:.0f} MB wide", flush=True)
agg.close()
