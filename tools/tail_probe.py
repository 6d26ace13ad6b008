#!/usr/bin/env python3
"""tools/tail_probe.py -- what bounds a SMALL batch (BASELINE.json configs[2], later bounces)?

Traces one late-bounce batch of the crown random walk (tools/config3.py) and the same batch (a) without its
longest rays (by the reference's per-ray nodesVisited counter), (b) longest rays first / last, (c) repeated k
times in one launch, (d) as two halves on two engine clones at once.  If the batch is bound by the latency of
its longest rays, (a) and (b, first) shorten it and (c) costs far less than k times.
usage: tail_probe.py [--scene crown] [--bounce 3]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nn_bvh_b200 as nb  # noqa: E402
from nn_bvh_b200 import rays as R, scenes  # noqa: E402


def timed(agg, d, n, reps=5):
    hits = torch.zeros((n, 5), dtype=torch.int32, device="cuda")
    best = 1e9
    for it in range(reps):
        agg.IntersectBatch(d, hits=hits, n=n)
        if it:
            best = min(best, agg.last_kernel_ms())
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="crown")
    ap.add_argument("--bounce", type=int, default=3)
    a = ap.parse_args()
    tris, cam, name = scenes.load(a.scene, allow_procedural=False)
    nodes, order = nb.build_bvh(tris, 4, "sah")
    ordered = tris[order]
    agg = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=ordered)
    batches = R.path_bounce_batches(lambda r: agg.IntersectBatch(r), ordered, cam, res=2048, bounces=a.bounce)
    r = np.ascontiguousarray(batches[a.bounce]).view(np.float32).reshape(-1, 8)
    n = len(r)
    _, nv, tt = agg.IntersectCounters(r)
    out = {"scene": name, "bounce": a.bounce, "rays": n,
           "nodes_visited": {"mean": round(float(nv.mean()), 1), "p99": int(np.percentile(nv, 99)),
                             "p99.9": int(np.percentile(nv, 99.9)), "max": int(nv.max())}}
    d = torch.from_numpy(r).cuda()
    out["ms_as_is"] = round(timed(agg, d, n), 3)
    by = np.argsort(nv, kind="stable")
    for drop in (0.001, 0.01):
        keep = np.sort(by[: int(n * (1 - drop))])
        out[f"ms_without_longest_{drop}"] = round(timed(agg, torch.from_numpy(r[keep]).cuda(), len(keep)), 3)
    out["ms_longest_first"] = round(timed(agg, torch.from_numpy(np.ascontiguousarray(r[by[::-1]])).cuda(), n), 3)
    out["ms_longest_last"] = round(timed(agg, torch.from_numpy(np.ascontiguousarray(r[by])).cuda(), n), 3)
    for k in (2, 4):
        out[f"ms_repeated_x{k}"] = round(timed(agg, torch.from_numpy(np.tile(r, (k, 1))).cuda(), k * n), 3)
    # two independent batches (here: the two halves, and the whole batch twice) on two clones at once
    clone = agg.clone()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    agg.set_stream(s1.cuda_stream)
    clone.set_stream(s2.cuda_stream)
    for label, (da, db) in {"halves": (d[: n // 2].contiguous(), d[n // 2:].contiguous()), "whole_twice": (d, d.clone())}.items():
        ha = torch.zeros((len(da), 5), dtype=torch.int32, device="cuda")
        hb = torch.zeros((len(db), 5), dtype=torch.int32, device="cuda")
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e9
        for it in range(5):
            torch.cuda.synchronize()
            e0.record(s1)
            s2.wait_event(e0)
            agg.IntersectBatch(da, hits=ha, n=len(da))
            clone.IntersectBatch(db, hits=hb, n=len(db))
            done = torch.cuda.Event()
            done.record(s2)
            s1.wait_event(done)
            e1.record(s1)
            torch.cuda.synchronize()
            if it:
                best = min(best, e0.elapsed_time(e1))
        out[f"ms_two_clones_{label}"] = round(best, 3)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
