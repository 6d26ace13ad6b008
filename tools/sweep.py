#!/usr/bin/env python3
"""tools/sweep.py -- time kernel variants / launch knobs on one scene (GPU box).
usage: sweep.py [scene] [nrays] ; knobs are the NNBVH_* env vars read by nnbvh_create."""
import itertools
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nn_bvh_b200 as nb
from nn_bvh_b200 import rays as R, scenes

scene = sys.argv[1] if len(sys.argv) > 1 else "bathroom"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4 << 20
configs = sys.argv[3:] or None
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
want = {}


def run(env):
    for k in list(os.environ):
        if k.startswith("NNBVH_"):
            del os.environ[k]
    os.environ.update({k: str(v) for k, v in env.items()})
    agg = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=ordered)
    out = {}
    for kind in ("bounce", "shadow", "primary"):
        m = len(sets[kind])
        best = 1e9
        for it in range(4):
            if kind == "shadow":
                agg.IntersectPBatch(dev[kind], occluded=d_occ, n=m)
            else:
                agg.IntersectBatch(dev[kind], hits=d_hits, n=m)
            ms = agg.last_kernel_ms()
            if it:
                best = min(best, ms)
        res = (d_occ if kind == "shadow" else d_hits)[:m].cpu().numpy()
        if kind not in want:
            want[kind] = res
        ok = np.array_equal(res, want[kind])
        out[kind] = (m / best / 1e3, ok)
    agg.close()
    return out


grid = []
grid.append({"NNBVH_KERNEL": 1})
for minb, refill, quorum in itertools.product((4, 3), (32, 24, 16), (1, 8, 12, 16, 20, 24, 28, 32)):
    grid.append({"NNBVH_KERNEL": 2, "NNBVH_WIDE_MINB": minb, "NNBVH_REFILL_BELOW": refill,
                 "NNBVH_LEAF_QUORUM": quorum})
if configs:
    grid = [dict(kv.split("=") for kv in c.split(",")) for c in configs]
print(f"scene {name} tris {len(tris)} rays {n}")
for env in grid:
    o = run(env)
    print(" ".join(f"{k[6:]}={v}" for k, v in env.items()).ljust(60),
          " ".join(f"{k}:{v[0]:8.1f}M/s{'' if v[1] else '(!)'}" for k, v in o.items()), flush=True)
