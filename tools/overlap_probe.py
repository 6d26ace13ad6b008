#!/usr/bin/env python3
"""tools/overlap_probe.py [scene] -- do a closest batch and a shadow batch overlap usefully on two streams?
Two engines (two streams), same scene; K steps of (8 Mi bounce closest + 8 Mi shadow) issued back to back
(a) on one stream, (b) closest on stream A and shadow on stream B.  Wall time by CUDA events."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nn_bvh_b200 as nb  # noqa: E402
from nn_bvh_b200 import rays as R, scenes  # noqa: E402

scene = sys.argv[1] if len(sys.argv) > 1 else "bathroom"
N, K = 8 << 20, 20
tris, cam, name = scenes.load(scene, allow_procedural=False)
nodes, order = nb.build_bvh(tris, 4, "sah")
ordered = tris[order]
A = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=ordered)
B = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=ordered)
bounce, shadow = R.bounce_and_shadow_rays(lambda r: A.IntersectBatch(r), ordered, cam, N)
db = torch.from_numpy(bounce.view(np.float32).reshape(-1, 8)).cuda()
ds = torch.from_numpy(shadow.view(np.float32).reshape(-1, 8)).cuda()
hits = torch.zeros((N, 5), dtype=torch.int32, device="cuda")
occ = torch.zeros(N, dtype=torch.uint8, device="cuda")
sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
A.set_stream(sa.cuda_stream)


def run(two):
    B.set_stream((sb if two else sa).cuda_stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(sa)
    if two:
        sb.wait_event(e0)
    for _ in range(K):
        A.IntersectBatch(db, hits=hits, n=N)
        B.IntersectPBatch(ds, occluded=occ, n=N)
    if two:
        j = torch.cuda.Event()
        j.record(sb)
        sa.wait_event(j)
    e1.record(sa)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / K


for _ in range(2):
    run(False), run(True)
one = min(run(False) for _ in range(3))
two = min(run(True) for _ in range(3))
print(f"{name}: one stream {one:.3f} ms/step ({2 * N / one / 1e3:.0f} Mrays/s), two streams {two:.3f} ms/step "
      f"({2 * N / two / 1e3:.0f} Mrays/s), {100 * (one / two - 1):+.1f} %")
