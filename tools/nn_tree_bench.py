#!/usr/bin/env python3
"""tools/nn_tree_bench.py -- the fork's nn_tree_bench, rewritten to call the GPU engine.

Replaces /root/reference/machine_learning/nn_tree_bench.py:81-139 (which does not parse as
shipped and only prints static SAH / EPO): for each scene it builds
  (i)  pbrt's SAH tree                      (cpu/aggregates.cpp:192-387 via nnbvh_build_bvh)
  (ii) an NN-split tree: `--levels` levels of level-order (axis, offset) planes exported to the
       same LinearBVHNode layout, cells finished by SAH (nn_bvh_b200.nn_tree.export_linear)
and traces the SAME rays through both on the GPU, reporting static cost (nn_loss.SAH / EPO
formulas) next to MEASURED cost: mean nodesVisited / triTests per ray and Mrays/s.

Plane source: --planes file.npy (a (2^L-1, 4) array as nss_model_test.py:13-37 / predict_tree_EPO
would produce), else the reference's greedy SAH search restated (no trained checkpoint and no
TensorFlow here -- stated in the output).

usage: nn_tree_bench.py [--scenes bathroom,coffee_maker] [--rays 8388608] [--levels 4]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed(agg, rays, shadow, reps=3):
    import torch
    n = len(rays)
    d = torch.from_numpy(rays.view(np.float32).reshape(-1, 8)).cuda()
    hits = torch.zeros((n, 5), dtype=torch.int32, device="cuda")
    occ = torch.zeros(n, dtype=torch.uint8, device="cuda")
    best = 1e30
    for _ in range(reps + 1):
        if shadow:
            agg.IntersectPBatch(d, occluded=occ, n=n)
        else:
            agg.IntersectBatch(d, hits=hits, n=n)
        best = min(best, agg.last_kernel_ms())
    return n / best / 1e3, (occ if shadow else hits).cpu().numpy()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scenes", default="bathroom,coffee_maker")
    ap.add_argument("--rays", type=int, default=8 << 20)
    ap.add_argument("--levels", type=int, default=4)   # nss_global_config.py:15 lvls=4
    ap.add_argument("--planes", default=None)
    ap.add_argument("--candidates", type=int, default=64)
    a = ap.parse_args()
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import nn_tree, rays as R, scenes

    for name in a.scenes.split(","):
        tris, cam, loaded = scenes.load(name)
        t0 = time.time()
        sah_nodes, sah_order = nb.build_bvh(tris, 4, "sah")
        t_sah = time.time() - t0
        if a.planes:
            planes, src = np.load(a.planes), f"file {a.planes}"
        else:
            t0 = time.time()
            planes = nn_tree.greedy_sah_planes(tris, a.levels, a.candidates)
            src = (f"greedy SAH search restated from nn_BVH.py:193-260 ({time.time() - t0:.1f} s); "
                   "no trained checkpoint / TensorFlow available")
        nn_nodes, nn_order = nn_tree.export_linear(tris, planes, 4, "sah")
        top = nn_tree.TopTree(tris, planes)
        trees = {"sah": (sah_nodes, sah_order), "nn_split": (nn_nodes, nn_order)}
        aggs = {k: nb.GPUBVHAggregate(nodes=n_, ordered_tris9=tris[o_], order=o_) for k, (n_, o_) in trees.items()}
        bounce, shadow = R.bounce_and_shadow_rays(lambda r: aggs["sah"].IntersectBatch(r),
                                                  tris[sah_order], cam, a.rays)
        out = {"scene": loaded, "triangles": int(len(tris)), "rays": a.rays, "levels": a.levels,
               "plane_source": src, "top_tree": {"SAH": top.sah(), "EPO": top.epo()}, "trees": {}}
        ref_t = None
        for k, agg in aggs.items():
            m = min(a.rays, 1 << 20)
            cb = nn_tree.traversal_cost(agg, bounce[:m])
            cs = nn_tree.traversal_cost(agg, shadow[:m], any_hit=True)
            mb, hb = timed(agg, bounce, False)
            ms, hs = timed(agg, shadow, True)
            t = hb.view(np.float32)[:, 1]
            if ref_t is None:
                ref_t, ref_occ = t, hs
            else:   # same rays, different tree: same visibility, same distances (ties aside)
                assert np.array_equal(hs, ref_occ) and np.allclose(t, ref_t, rtol=1e-5)
            out["trees"][k] = {"nodes": int(len(trees[k][0])), "stack_depth": nb.tree_stack_depth(trees[k][0]),
                               "static_SAH": nn_tree.linear_sah_cost(trees[k][0]),
                               "closest": {"nodes_visited": cb["nodes_visited"], "tri_tests": cb["tri_tests"], "Mrays_s": mb},
                               "shadow": {"nodes_visited": cs["nodes_visited"], "tri_tests": cs["tri_tests"], "Mrays_s": ms}}
        print(json.dumps(out), flush=True)
        for agg in aggs.values():
            agg.close()


if __name__ == "__main__":
    main()
