"""tools/hlbvh_bench.py -- HLBVH build time: device stages (nnbvh_build_hlbvh_device) against the
host builder (nnbvh_build_bvh, all host threads), with the outputs compared byte for byte.

  python tools/hlbvh_bench.py [--tris 1000000 4000000] [--scene crown]
Prints one JSON line per case: device_ms = CUDA-event time of the device stages (bounds, Morton,
4 radix passes, treelet emission, compaction), device_wall_ms = the whole call (H2D of the soup,
D2H of nodes + order, upper SAH and final layout on the host), host_ms = the host builder.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import nn_bvh_b200 as nb  # noqa: E402
from nn_bvh_b200 import scenes  # noqa: E402


def run(name, tris, mp, reps):
    nb.build_hlbvh_device(tris[:1000], mp)   # context + module load
    best = None
    for _ in range(reps):
        t = time.perf_counter()
        dn, do, ms = nb.build_hlbvh_device(tris, mp)
        wall = (time.perf_counter() - t) * 1e3
        if best is None or wall < best[1]:
            best = (ms, wall)
    t = time.perf_counter()
    hn, ho = nb.build_bvh(tris, mp, "hlbvh")
    host = (time.perf_counter() - t) * 1e3
    t = time.perf_counter()
    nb.build_bvh(tris, mp, "sah")
    sah = (time.perf_counter() - t) * 1e3
    same = bool(np.array_equal(dn, hn) and np.array_equal(do, ho))
    print(json.dumps({"case": name, "tris": len(tris), "maxnodeprims": mp, "nodes": len(dn),
                      "device_ms": round(best[0], 3), "device_wall_ms": round(best[1], 2),
                      "host_hlbvh_ms": round(host, 2), "host_sah_ms": round(sah, 2),
                      "host_threads": os.cpu_count(), "identical": same}), flush=True)
    return same


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, nargs="*", default=[1_000_000, 4_000_000])
    ap.add_argument("--scene", nargs="*", default=[])
    ap.add_argument("--maxnodeprims", type=int, default=4)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    ok = True
    for T in a.tris:
        rng = np.random.default_rng(T)
        c = rng.random((T, 1, 3), dtype=np.float32)
        tris = (c + np.float32(0.002) * (rng.random((T, 3, 3), dtype=np.float32) * 2 - 1)).reshape(T, 9)
        ok &= run(f"soup{T}", tris, a.maxnodeprims, a.reps)
    for s in a.scene:
        ok &= run(s, scenes.load(s, allow_procedural=False)[0], a.maxnodeprims, a.reps)
    sys.exit(0 if ok else 1)
