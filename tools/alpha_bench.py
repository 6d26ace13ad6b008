#!/usr/bin/env python3
"""tools/alpha_bench.py -- constant-alpha triangles: AlphaKernel (one thread per ray) vs StepKernel<..., ALPHA>.
bathroom with a random constant alpha on half of its triangles, 8 Mi bounce rays, device-timed incl. the pre-pass."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nn_bvh_b200 as nb
from nn_bvh_b200 import rays as R, scenes

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8 << 20
tris, cam, name = scenes.load("bathroom")
rng = np.random.default_rng(9)
alpha = np.ones(len(tris), dtype=np.float32)
alpha[::2] = rng.random(len(alpha[::2])).astype(np.float32)
plain = nb.GPUBVHAggregate(tris, 4, "sah")
bounce, _ = R.bounce_and_shadow_rays(lambda r: plain.IntersectBatch(r), plain.ordered_tris9, cam, n)
d = torch.from_numpy(bounce.view(np.float32).reshape(-1, 8)).cuda()
hits = torch.zeros((n, 5), dtype=torch.int32, device="cuda")
ref = None
for step in ("0", "1"):
    os.environ["NNBVH_INST_STEP"] = step
    agg = nb.GPUBVHAggregate(tris, 4, "sah", alpha=alpha)
    best = 1e9
    for it in range(4):
        agg.IntersectBatch(d, hits=hits, n=n)
        ms = agg.last_kernel_ms()
        if it:
            best = min(best, ms)
    h = hits.cpu().numpy()
    ref = h if ref is None else ref
    print(f"{name} with alpha on half the triangles, {n} bounce rays, {'StepKernel<ALPHA>' if step == '1' else 'AlphaKernel'}: "
          f"{n / best / 1e3:8.1f} Mrays/s ({best:.2f} ms), same records: {np.array_equal(h, ref)}")
    agg.close()
