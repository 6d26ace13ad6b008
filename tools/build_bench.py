#!/usr/bin/env python3
"""tools/build_bench.py -- host tree construction: the library's builders beside the reference's.

For each scene and split method: wall time of nnbvh_build_bvh (include/nnbvh.h; the BVHAggregate
constructor restated, cpu/aggregates.cpp:150-214) and of the reference's own BVHAggregate constructor
(oracle/_ref, timed around the constructor alone), on the same host threads, and whether the two flattened
trees are byte-identical.  CPU only; needs scene_cache/ (tools/export_scenes.py) and oracle/_ref.
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("scenes", nargs="*", default=["bathroom", "crown"])
    ap.add_argument("--threads", type=int, default=os.cpu_count())
    ap.add_argument("--splits", default="sah,hlbvh,middle,equal")
    a = ap.parse_args()
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import scenes
    from oracle import pyoracle
    ref = pyoracle.Reference(threads=a.threads)
    print(f"# maxPrimsInNode 4; wall times in ms; T = {a.threads} host threads (library: best of 5; reference: one build)")
    print("# identical: the library's tree (any thread count) against the reference's single-threaded build, byte for byte --")
    print("# with several threads the reference places subtrees by atomic counters (cpu/aggregates.cpp:263-330) and its")
    print("# primitive order varies from run to run")
    print(f"{'scene':10s} {'triangles':>9s} {'split':>7s} {'lib 1':>8s} {'ref 1':>8s} {'lib T':>8s} {'ref T':>8s} {'ref/lib T':>9s}  identical")
    for name in a.scenes:
        tris, _, _ = scenes.load(name, allow_procedural=False)
        tris = np.ascontiguousarray(tris + np.float32(0))  # -0.0 -> +0.0, as pbrt's own loaders store it
        for split in a.splits.split(","):
            lib = {}
            for th in (1, a.threads):
                best = 1e9
                for _ in range(5):
                    t0 = time.perf_counter()
                    nodes, order = nb.build_bvh(tris, 4, split, threads=th)
                    best = min(best, time.perf_counter() - t0)
                lib[th] = (best, nodes, order)
            rt = {}
            for th in (a.threads, 1):              # scenes stay allocated: the driver has no scene destructor
                ref.set_threads(th)
                sc = ref.scene(tris, 4, split)
                rt[th] = ref.build_seconds(sc)
            rn, ro = ref.export(sc, len(tris))     # the single-threaded build
            same = all(np.array_equal(lib[th][1].view(np.uint32).reshape(-1, 8), rn) and
                       np.array_equal(lib[th][2].astype(np.int32), ro) for th in lib)
            print(f"{name:10s} {len(tris):9d} {split:>7s} {lib[1][0] * 1e3:8.1f} {rt[1] * 1e3:8.1f} "
                  f"{lib[a.threads][0] * 1e3:8.1f} {rt[a.threads] * 1e3:8.1f} {rt[a.threads] / lib[a.threads][0]:9.2f}  {same}")


if __name__ == "__main__":
    main()
