#!/usr/bin/env python3
"""tools/config3.py -- BASELINE.json configs[2]: crown, mixed coherent / incoherent path-trace ray batches.

Batch k = all live rays at bounce k of a diffuse random walk started from the scene's camera (SURVEY 8(d):
the fallback to capturing VolPathIntegrator's rays, which needs the full renderer): bounce 0 is the
coherent camera batch, later bounces are increasingly incoherent and shorter (escaped paths drop out).
Every batch is traced resident on the GPU (closest hit), timed by the engine's CUDA events (ray-ordering
pre-pass included), with the reference's per-ray counters on a subset and the reference's own CPU
`BVHAggregate::Intersect` (oracle/_ref, all host cores) on a subset of the same rays.
--spheres: the scene's environment shell joins the primitive list (paths are generated on the
triangle-only scene either way, so both variants trace identical rays).
usage: config3.py [--scene crown] [--res 2048] [--bounces 4] [--spheres]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nn_bvh_b200 as nb  # noqa: E402
from nn_bvh_b200 import rays as R, scenes  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="crown")
    ap.add_argument("--res", type=int, default=2048)
    ap.add_argument("--bounces", type=int, default=4)
    ap.add_argument("--spheres", action="store_true")
    ap.add_argument("--cpu-rays", type=int, default=1 << 20)
    a = ap.parse_args()
    tris, cam, name = scenes.load(a.scene, allow_procedural=False)
    nodes, order = nb.build_bvh(tris, 4, "sah")
    ordered = tris[order]
    agg = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=ordered)
    batches = R.path_bounce_batches(lambda r: agg.IntersectBatch(r), ordered, cam, res=a.res, bounces=a.bounces)
    sph = scenes.load_spheres(a.scene) if a.spheres else None
    if sph is not None:
        agg.close()
        agg = nb.GPUBVHAggregate(tris, 4, "sah", spheres=sph)
        name += f"+{len(sph)}sphere"
    ref = None
    if sph is None:
        try:
            from oracle.pyoracle import Reference, ref_available
            if ref_available():
                ref = Reference(threads=0)
                ref_scene = ref.scene(tris, 4, "sah")
        except Exception as e:   # noqa: BLE001
            print("cpu reference unavailable:", e, file=sys.stderr)
    tot_rays = tot_ms = 0.0
    for k, r in enumerate(batches):
        n = len(r)
        d = torch.from_numpy(r.view(np.float32).reshape(-1, 8)).cuda()
        hits = torch.zeros((n, 5), dtype=torch.int32, device="cuda")
        best = 1e9
        for it in range(4):
            agg.IntersectBatch(d, hits=hits, n=n)
            ms = agg.last_kernel_ms()
            if it:
                best = min(best, ms)
        m = min(n, 1 << 20)
        _, nv, tt = agg.IntersectCounters(r[:m])
        bpr = 52 + 32 * float(nv.mean()) + 36 * float(tt.mean())
        row = {"scene": name, "bounce": k, "rays": n, "ms": round(best, 3), "Mrays_s": round(n / best / 1e3, 1),
               "hit_fraction": round(float((hits[:, 0] != -1).float().mean()), 4),
               "nodes_visited": round(float(nv.mean()), 2), "tri_tests": round(float(tt.mean()), 2),
               "algorithmic_B_per_ray": round(bpr), "algorithmic_GB_s": round(bpr * n / best / 1e6)}
        if ref is not None:
            c = min(n, a.cpu_rays)
            sub = np.ascontiguousarray(r[:c]).view(np.float32).reshape(-1, 8)
            ref.intersect_only(ref_scene, sub[: 1 << 16])
            t0 = time.time()
            ref.intersect_only(ref_scene, sub)
            dt = time.time() - t0
            row["cpu_reference_Mrays_s"] = round(c / dt / 1e6, 2)
            row["cpu_threads"] = ref.threads
            row["cpu_sample_rays"] = c
        print(json.dumps(row), flush=True)
        tot_rays += n
        tot_ms += best
    print(json.dumps({"scene": name, "all_bounces": True, "rays": int(tot_rays), "ms": round(tot_ms, 3),
                      "Mrays_s": round(tot_rays / tot_ms / 1e3, 1)}), flush=True)


if __name__ == "__main__":
    main()
