#!/usr/bin/env python3
"""tools/scenes_table.py [scene ...] -- per-scene throughput of the device-resident path (GPU box).

For each cached scene (reference SAH tree, maxnodeprims 4): 8 Mi camera / diffuse-bounce / shadow rays,
best of 3 timed batches (ms from the engine's CUDA events, ray-ordering pre-pass included), the
reference's counting rules for nodesVisited / triTests on a 1 Mi-ray subset, and the algorithmic
bytes per ray of SURVEY 8(d).
--spheres: the scene's `Shape "sphere"` primitives join the primitive list (killeroos: the spherical
light, crown: the environment shell); the same rays (generated on the triangle-only scene) are traced."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nn_bvh_b200 as nb  # noqa: E402
from nn_bvh_b200 import rays as R, scenes  # noqa: E402

N = 8 << 20
WITH_SPHERES = "--spheres" in sys.argv
ARGS = [a for a in sys.argv[1:] if not a.startswith("--")]
for scene in ARGS or ["killeroos", "coffee_maker", "bathroom", "crown"]:
    tris, cam, name = scenes.load(scene, allow_procedural=False)
    nodes, order = nb.build_bvh(tris, 4, "sah")
    ordered = tris[order]
    agg = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=ordered)
    bounce, shadow = R.bounce_and_shadow_rays(lambda r: agg.IntersectBatch(r), ordered, cam, N)
    res = int(np.ceil(np.sqrt(N)))
    sets = {"primary": R.camera_rays(cam, res, res)[:N], "bounce": bounce, "shadow": shadow}
    sph = scenes.load_spheres(scene) if WITH_SPHERES else None
    if sph is not None:
        agg.close()
        agg = nb.GPUBVHAggregate(tris, 4, "sah", spheres=sph)
        nodes = agg.nodes
        name += f"+{len(sph)}sphere"
    for kind, r in sets.items():
        n = len(r)
        d = torch.from_numpy(r.view(np.float32).reshape(-1, 8)).cuda()
        hits = torch.zeros((n, 5), dtype=torch.int32, device="cuda")
        occ = torch.zeros(n, dtype=torch.uint8, device="cuda")
        best = 1e9
        for it in range(4):
            if kind == "shadow":
                agg.IntersectPBatch(d, occluded=occ, n=n)
            else:
                agg.IntersectBatch(d, hits=hits, n=n)
            ms = agg.last_kernel_ms()
            if it:
                best = min(best, ms)
        _, nv, tt = agg.IntersectCounters(r[: 1 << 20], any_hit=(kind == "shadow"))
        io = 33 if kind == "shadow" else 52
        bpr = io + 32 * float(nv.mean()) + 36 * float(tt.mean())
        print(f"{name} {kind} {n} rays: {best:.3f} ms  {n / best / 1e3:.1f} Mrays/s | nodesVisited {nv.mean():.2f} "
              f"triTests {tt.mean():.2f} algorithmic B/ray {bpr:.0f} -> {bpr * n / best / 1e6:.0f} GB/s algorithmic; "
              f"scene {len(nodes) * 32 >> 20} MB nodes32 + {len(tris) * 48 >> 20} MB tris48", flush=True)
    agg.close()
