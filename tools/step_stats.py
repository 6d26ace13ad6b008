#!/usr/bin/env python3
"""tools/step_stats.py -- lane occupancy of the traversal kernel (GPU box).

Runs the counting instantiation of StepKernel (NNBVH_STEP_STATS=1) on a scene's bounce / shadow /
primary batches and prints, per phase-A iteration, how many of the 32 lanes were at a node, parked
at a leaf, or without a ray, and how full the phase-B (triangle) passes were.
usage: step_stats.py [scene] [nrays] [K=V,K=V ...]   (knobs: the NNBVH_* env vars)"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["NNBVH_STEP_STATS"] = "1"
import nn_bvh_b200 as nb  # noqa: E402
from nn_bvh_b200 import rays as R, scenes  # noqa: E402

scene = sys.argv[1] if len(sys.argv) > 1 else "bathroom"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4 << 20
configs = sys.argv[3:] or [""]
tris, cam, name = scenes.load(scene)
nodes, order = nb.build_bvh(tris, 4, "sah")
ordered = tris[order]
base = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=ordered)
bounce, shadow = R.bounce_and_shadow_rays(lambda r: base.IntersectBatch(r), ordered, cam, n)
prim = R.camera_rays(cam, 2048, 2048)[:n]
sets = {"bounce": bounce, "shadow": shadow, "primary": prim}
dev = {k: torch.from_numpy(v.view(np.float32).reshape(-1, 8)).cuda() for k, v in sets.items()}
d_hits = torch.zeros((n, 5), dtype=torch.int32, device="cuda")
d_occ = torch.zeros(n, dtype=torch.uint8, device="cuda")
print(f"scene {name} tris {len(tris)} rays {n}")
for c in configs:
    for k in list(os.environ):
        if k.startswith("NNBVH_") and k != "NNBVH_STEP_STATS":
            del os.environ[k]
    if c:
        os.environ.update(dict(kv.split("=") for kv in c.split(",")))
    agg = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=ordered)
    for kind in ("bounce", "shadow", "primary"):
        m = len(sets[kind])
        if kind == "shadow":
            agg.IntersectPBatch(dev[kind], occluded=d_occ, n=m)
        else:
            agg.IntersectBatch(dev[kind], hits=d_hits, n=m)
        s = agg.step_stats()
        ia, ib = max(s["iters_a"], 1), max(s["iters_b"], 1)
        print(f"{c or 'default':40s} {kind:8s} A: {ia / m:6.2f} warp-iters/ray  node {s['lanes_node'] / ia:5.2f} "
              f"parked {s['lanes_parked'] / ia:5.2f} empty {s['lanes_empty'] / ia:5.2f} | "
              f"B: {ib / m:5.2f} passes/ray  parked {s['lanes_b_parked'] / ib:5.2f} testing {s['lanes_b_testing'] / ib:5.2f} | "
              f"refills/ray {s['refills'] / m:.3f}  node-steps/ray {s['lanes_node'] / m:.1f}", flush=True)
    agg.close()
