#!/usr/bin/env python3
"""bench.py -- Mrays/s of the closest-hit + shadow hot path on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            (torchrun for N > 1)
  python bench.py --impl reference ...                      (the reference's CPU path)

A "step" is one pass of the hot path over one batch: `--rays` incoherent
diffuse-bounce rays through BVHAggregate::Intersect (closest hit) plus `--rays`
shadow rays through IntersectP (any hit) -- BASELINE.json configs[1] (default
8M + 8M).  The conference scene is missing from the reference tree
(.MISSING_LARGE_BLOBS:12); the workload is its stand-in, the reference's own
bathroom scene (592 188 triangles) from scene_cache/ (exported by
__graft_entry__.build()).  A missing scene file is an error unless
--allow-procedural is given; config.workload names what was really traced.

value  : whole-job Mrays/s, rays resident in HBM, CUDA-event timed (max over ranks).
e2e    : same metric through the C ABI with pinned HOST buffers (H2D/D2H inside).
roofline: closest-hit traversal kernel, algorithmic bytes / event-timed duration; `bound`
         is decided from data (scene size vs the L2, committed ncu DRAM bytes vs algorithmic).
configs.synthetic10M: BASELINE.json configs[4] in the same line -- 10 M-triangle soup (HBM-resident
         scene), ONE 256 Mi-ray batch sharded over the N ranks (strong scaling), hit records
         delivered to rank 0 inside the timed region; scene built once on rank 0 and broadcast.
cpu_baseline / --impl reference: the unmodified reference TUs (oracle/_ref) if built,
         else the oracle port, on the host cores; cpu_baseline.lean = the {prim,t,b}-only loop
         (the oracle port of aggregates.cpp:529-624) on the same cores.
Only the cpu_baseline / reference legs touch oracle/; the product path never does, and the
reference arm never loads libnnbvh.so.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mrays/s closest-hit+shadow"
OUT = sys.stdout   # replaced by the real stdout in __main__ (see _quiet_stdout)
UNIT = "Mrays/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scene", default="bathroom")
    ap.add_argument("--rays", type=int, default=8 << 20, help="rays per kind per step per GPU")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--allow-procedural", action="store_true",
                    help="trace a procedural stand-in when scene_cache/<scene> is missing (NOT a BASELINE.json "
                         "config; without this flag a missing scene is an error)")
    ap.add_argument("--workload", default="all", choices=["all", "config2", "config5"],
                    help="config2 = the metric's workload (BASELINE configs[1]); config5 = the synthetic "
                         "10M-triangle strong-scaling batch only (profiling); all = both in one line")
    ap.add_argument("--c5-tris", type=int, default=10_000_000)
    ap.add_argument("--c5-rays", type=int, default=256 << 20, help="config 5: rays of the ONE batch all ranks share")
    ap.add_argument("--c5-steps", type=int, default=5, help="config 5: at most this many timed steps")
    ap.add_argument("--c5-chunks", default="0.4,0.4,0.2",
                    help="config 5, N>1: fractions of a rank's shard traced and delivered one after another "
                         "(a chunk's delivery overlaps the next chunk's traversal)")
    ap.add_argument("--c5-rank0-share", type=float, default=1.0,
                    help="config 5, N>1: rank 0's share of the batch relative to the other ranks' (1 = equal)")
    ap.add_argument("--c5-copy-streams", type=int, default=4,
                    help="config 5, N>1: copy engines a rank's delivery is spread over")
    ap.add_argument("--c5-lanes", type=int, default=2, choices=[1, 2],
                    help="config 5, N>1: chunks alternate between this many engines / streams (nnbvh_clone), so that "
                         "one chunk's ray-ordering pre-pass and first CTAs fill the SMs the previous chunk's "
                         "persistent CTAs free as it drains")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N>1: how hit records reach rank 0: copy engines into a peer-mapped window, or an NCCL gather")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-i", str(self.index), "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.t = threading.Thread(target=self._read, daemon=True)
        self.t.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def load_scene(args):
    """-> (tris9, camera, name actually loaded); a missing scene file is an error."""
    from nn_bvh_b200 import scenes
    try:
        return scenes.load(args.scene, allow_procedural=args.allow_procedural)
    except FileNotFoundError as e:
        raise SystemExit(f"bench.py: {e}")


def load_workload(args, trace, rank):
    """GPU arm -> dict(tris, cam, nodes, order, ordered, name); the tree is the product's host build."""
    import nn_bvh_b200 as nb
    tris, cam, name = load_scene(args)
    nodes, order = nb.build_bvh(tris, 4, "sah")     # reference defaults (aggregates.cpp:727,742)
    ordered = tris[order]
    return dict(tris=tris, cam=cam, nodes=nodes, order=order, ordered=ordered, name=name)


def make_rays(w, trace, n, rank):
    from nn_bvh_b200 import rays as R
    return R.bounce_and_shadow_rays(trace, w["ordered"], w["cam"], n, seed=2 + 1000 * rank)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------
# CPU legs (the only code in this file that may touch oracle/)
# ---------------------------------------------------------------------------
class CpuPath:
    """The reference's own CPU implementation: oracle/_ref (unmodified reference TUs,
    ParallelFor over rays like CPUAggregate::IntersectClosest, wavefront/aggregate.cpp:42-68)
    when it was built in the dev container, else the oracle port on std::threads."""

    def __init__(self, w):
        from oracle import pyoracle
        self.w = w
        self.cores = os.cpu_count() or 1
        if pyoracle.ref_available():
            self.kind = "reference"
            self.ref = pyoracle.Reference(threads=0)
            self.cores = self.ref.threads
            # the reference builds its own tree from the same soup (parallel build: leaf
            # order may differ from ours, the hit SET does not matter for timing)
            self.scene = self.ref.scene(w["tris"], 4, "sah")
        else:
            self.kind = "port"
            self.orc = pyoracle.Oracle()
            self.nodes = np.ascontiguousarray(w["nodes"]).view(np.uint32).reshape(-1, 8)

    def closest(self, rays):
        r = np.ascontiguousarray(rays).view(np.float32).reshape(-1, 8)
        if self.kind == "reference":
            return self.ref.intersect_only(self.scene, r)
        return self.orc.intersect(self.nodes, self.w["ordered"], r, threads=0)

    def shadow(self, rays):
        r = np.ascontiguousarray(rays).view(np.float32).reshape(-1, 8)
        if self.kind == "reference":
            return self.ref.intersect_p(self.scene, r, parallel=True)
        return self.orc.intersect_p(self.nodes, self.w["ordered"], r, threads=0)

    def timed(self, bounce, shadow, budget_s, chunk=1 << 18, lean=False):
        """Trace chunks of (closest + shadow) until ~budget_s; -> (Mrays/s, rays, seconds).
        lean: the oracle port's {prim,t,b}-only loops instead of the reference's Primitive::Intersect
        (which builds a SurfaceInteraction per hit, shapes.cpp:320-338)."""
        closest, shadow_fn = (self.lean_closest, self.lean_shadow) if lean else (self.closest, self.shadow)
        closest(bounce[:chunk // 8])  # warm-up
        done, t0 = 0, time.perf_counter()
        n = min(len(bounce), len(shadow))
        while done < n and time.perf_counter() - t0 < budget_s:
            e = min(n, done + chunk)
            closest(bounce[done:e])
            shadow_fn(shadow[done:e])
            done = e
        dt = time.perf_counter() - t0
        return 2 * done / dt / 1e6, 2 * done, dt

    def _lean_setup(self):
        if not hasattr(self, "orc"):
            from oracle import pyoracle
            self.orc = pyoracle.Oracle()
        if not hasattr(self, "nodes"):
            self.nodes = np.ascontiguousarray(self.w["nodes"]).view(np.uint32).reshape(-1, 8)

    def lean_closest(self, rays):
        self._lean_setup()
        r = np.ascontiguousarray(rays).view(np.float32).reshape(-1, 8)
        return self.orc.intersect(self.nodes, self.w["ordered"], r, threads=0)

    def lean_shadow(self, rays):
        self._lean_setup()
        r = np.ascontiguousarray(rays).view(np.float32).reshape(-1, 8)
        return self.orc.intersect_p(self.nodes, self.w["ordered"], r, threads=0)


def run_reference(args):
    """--impl reference: the reference CPU path on this box's host cores; rank 0 only.
    Nothing of the product is loaded here: the tree is the reference's own build (oracle/_ref,
    exported), primary rays for the ray generator are traced by the oracle port, and
    nn_bvh_b200.{scenes,rays} are numpy-only input preparation (no libnnbvh.so)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import pyoracle
    orc = pyoracle.Oracle()
    tris, cam, name = load_scene(args)
    if pyoracle.ref_available():
        ref = pyoracle.Reference(threads=0)
        nodes32, order = ref.export(ref.scene(tris, 4, "sah"), len(tris))
    else:
        nodes32, order = orc.build_bvh(tris, 4, "sah")
    nodes32 = np.ascontiguousarray(nodes32).view(np.uint32).reshape(-1, 8)
    ordered = np.ascontiguousarray(tris[np.asarray(order).astype(np.int64)])
    w = dict(tris=tris, cam=cam, nodes=nodes32, ordered=ordered, name=name)
    hit_dtype = np.dtype([("prim", "<u4"), ("t", "<f4"), ("b", "<f4", (3,))])

    def trace(r):   # primary rays for the ray generator: oracle port, all threads
        h = orc.intersect(nodes32, ordered, np.ascontiguousarray(r).view(np.float32).reshape(-1, 8))
        return h.view(hit_dtype).reshape(-1)

    cpu = CpuPath(w)
    # The GPU arm's step is args.rays closest + args.rays shadow rays.  The same count is traced here
    # unless a short probe says a step would take more than ~8 s on this box's cores; then a bounded
    # sample of the same batch is used and config.sample says so.
    probe_b, probe_s = make_rays(w, trace, 1 << 17, 0)
    mr, _, _ = cpu.timed(probe_b, probe_s, 2.0, chunk=1 << 16)
    full_step_s = 2 * args.rays / (mr * 1e6)
    per_step = args.rays if full_step_s <= 8.0 else int(max(1 << 16, mr * 1e6 * 4.0 / 2))
    bounce, shadow = make_rays(w, trace, per_step, 0)
    for _ in range(max(args.warmup, 1)):
        cpu.closest(bounce[: per_step // 8])
        cpu.shadow(shadow[: per_step // 8])
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu.closest(bounce)
        cpu.shadow(shadow)
    dt = time.perf_counter() - t0
    val = 2 * per_step * args.steps / dt / 1e6
    same = per_step == args.rays
    sample = (f"{per_step} closest + {per_step} shadow rays per step" +
              ("" if same else f" (bounded sample of {args.rays}+{args.rays}: a full step would take "
                               f"{full_step_s:.0f} s on these cores)"))
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic rays over " + name,
            "config": {"workload": f"{name}: {per_step} incoherent diffuse-bounce closest-hit rays + "
                                   f"{per_step} shadow any-hit rays per step",
                       "triangles": int(len(tris)), "sample": sample, "same_rays_per_step_as_gpu_arm": same,
                       "tree": "the reference's own SAH build (maxnodeprims 4)"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cpu.cores, "kind": cpu.kind,
                             "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), file=OUT, flush=True)



def l2_bytes(dev):
    import torch
    return int(torch.cuda.get_device_properties(dev).L2_cache_size)


def ncu_capture(key):
    """profiles/ncu_traffic.json: {workload key: counters of ONE launch of the closest-hit StepKernel from a
    committed `ncu --set full` capture of bench.py on that workload}.  None when no capture matches the
    workload that was actually loaded."""
    tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(tp):
        return None
    with open(tp) as f:
        return json.load(f).get(key)


def make_roofline(agg, dev, sample_rays, n, traversal_ms, n_nodes, n_tris, key):
    """The `roofline` object for one launch of the closest-hit traversal over n rays.
    algorithmic bytes/ray = 52 + 32*nodesVisited + 36*triTests; the counts come from the
    engine's counting kernel (the reference's counting rules) on a sample of the same batch.
    `bound` is decided from data: "l2" when the scene (64-byte wide nodes + 48-byte triangles) fits the
    L2, or a committed ncu capture of this workload shows DRAM traffic below half the algorithmic bytes;
    else "hbm".  achieved/peak/frac are stated against that roof; the other roof is reported beside it."""
    _, nv, tt = agg.IntersectCounters(sample_rays)
    bytes_per_ray = 52.0 + 32.0 * float(nv.mean()) + 36.0 * float(tt.mean())
    alg_gb = bytes_per_ray * n / 1e9
    achieved = alg_gb / (traversal_ms / 1e3)
    hbm_peak, hbm_src = peaks()
    scene_bytes = 64 * (n_nodes // 2 + 1) + 48 * n_tris
    l2 = l2_bytes(dev)
    cap = ncu_capture(key)
    traffic = (cap["dram_bytes_read"] + cap["dram_bytes_write"]) / 1e9 if cap else None
    dram_share = traffic / alg_gb if traffic is not None else None
    fits = scene_bytes < l2
    bound = "l2" if (fits or (dram_share is not None and dram_share < 0.5)) else "hbm"
    # the library's own read probe (256-bit read-only loads): an L2-sized buffer, and a 4 GiB one
    l2_gbs = agg.read_bandwidth_probe(min(max(scene_bytes, 16 << 20), (l2 * 3) // 4), 40)
    hbm_probe_gbs = agg.read_bandwidth_probe(4 << 30, 2)
    roofs = {"hbm": {"peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak, "source": hbm_src,
                     "same_probe_as_l2": hbm_probe_gbs},
             "l2": {"peak": l2_gbs, "unit": "GB/s", "frac": achieved / l2_gbs,
                    "source": "nnbvh_read_bandwidth_probe (this run): an L2-resident buffer of "
                              f"{min(max(scene_bytes, 16 << 20), (l2 * 3) // 4) >> 20} MB streamed with the traversal "
                              "kernel's 256-bit read-only loads"}}
    why = (f"scene is {scene_bytes / 1e6:.0f} MB (64-byte wide nodes + 48-byte triangles) against an L2 of "
           f"{l2 / 1e6:.0f} MB: " + ("it fits" if fits else "it does not fit"))
    if dram_share is not None:
        why += (f"; committed ncu capture: DRAM read+write {traffic:.2f} GB per launch = {dram_share:.2f} x the "
                f"algorithmic {alg_gb:.2f} GB")
    else:
        why += "; no committed ncu capture for this workload (traffic null)"
    r = {"bound": bound, "kernel": "StepKernel<closest>", "achieved": achieved, "peak": roofs[bound]["peak"],
         "unit": "GB/s", "frac": roofs[bound]["frac"], "peak_source": roofs[bound]["source"],
         "bound_reason": why, "traffic": traffic, "traffic_unit": "GB per launch",
         "traffic_source": cap["source"] if cap else None,
         "algorithmic_gb_per_launch": alg_gb, "algorithmic_bytes_per_ray": bytes_per_ray, "rays_per_launch": n,
         "kernel_ms": traversal_ms, "nodes_visited_per_ray": float(nv.mean()), "tri_tests_per_ray": float(tt.mean()),
         "hbm": roofs["hbm"], "l2": roofs["l2"], "scene_bytes": scene_bytes, "l2_bytes": l2}
    if cap:   # what the capture says limits the kernel: measured values, not prose
        r["ncu"] = {k: cap[k] for k in ("issue_active_pct", "threads_per_inst", "l1_hit_pct", "l2_hit_pct",
                                        "dram_throughput_pct", "l1tex_throughput_pct", "kernel_ms_under_ncu",
                                        "warp_inst_per_ray") if k in cap}
    return r


# ---------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; nn_bvh_b200 has no CPU path to time")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's version / debug banner goes to stderr
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    line = None
    if args.workload in ("all", "config2"):
        line = run_config2(args, world, rank, local, dev)
    if args.workload in ("all", "config5"):
        torch.cuda.set_stream(torch.cuda.default_stream(dev))
        c5 = run_config5(args, world, rank, local, dev)
        if rank == 0:
            if line is None:   # --workload config5: the line IS config 5 (profiling runs)
                line = {"metric": METRIC + " (config 5: closest-hit only)", "value": c5["value"], "unit": UNIT,
                        "n_gpus": world, "steps": c5["steps"], "warmup": c5["warmup"],
                        "ms_per_step": c5["ms_per_step"], "higher_is_better": True, "scaling": "strong",
                        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": {"workload": c5["workload"]},
                        "gpu_launches": c5["gpu_launches"], "roofline": c5["roofline"]}
            line.setdefault("configs", {})["synthetic10M"] = c5
    if rank == 0:
        print(json.dumps(line), file=OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_config2(args, world, rank, local, dev):
    """BASELINE.json configs[1] (the metric's workload): weak scaling, args.rays closest + args.rays shadow
    rays per GPU per step.  Returns the JSON line (rank 0) or None."""
    import torch
    import torch.distributed as dist
    import nn_bvh_b200 as nb

    w = load_workload(args, None, rank)
    agg = nb.GPUBVHAggregate(nodes=w["nodes"], ordered_tris9=w["ordered"], order=w["order"],
                             device=local)
    bounce, shadow = make_rays(w, lambda r: agg.IntersectBatch(r), args.rays, rank)
    n = args.rays

    # ---- device-resident buffers; engine launches on torch's current stream ----
    # a non-default torch stream: handle 0 would mean "engine's own stream" to the C ABI
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    agg.set_stream(stream.cuda_stream)
    # Two lanes: closest-hit batches on `stream` through `agg`, shadow batches on `stream2` through a clone
    # that shares the uploaded scene (nnbvh_clone).  The two kinds of batch are independent (as bounce k's
    # shadow rays and bounce k+1's closest rays are in an integrator), so the shadow lane's pre-pass and first
    # CTAs take the SMs the closest lane's persistent CTAs free while it drains, and vice versa.
    stream2 = torch.cuda.Stream(device=dev)
    agg2 = agg.clone()
    agg2.set_stream(stream2.cuda_stream)
    d_b = torch.from_numpy(bounce.view(np.float32).reshape(-1, 8)).to(dev)
    d_s = torch.from_numpy(shadow.view(np.float32).reshape(-1, 8)).to(dev)
    d_hits = torch.zeros((n, 5), dtype=torch.int32, device=dev)
    d_occ = torch.zeros(n, dtype=torch.uint8, device=dev)
    from nn_bvh_b200 import sharding
    # N>1: hit records of all ranks on rank 0, global ray order, inside the timed step.
    # "peer": rank 0 owns the array, every rank copies its slice into it with a copy engine over
    # NVLink while its shadow kernel runs (sharding.PeerWindow); "nccl": an NCCL gather -- its
    # kernel needs SMs the persistent traversal CTAs hold, so it only starts when they drain.
    g_hits = g_occ = win_hits = win_occ = None
    if world > 1 and args.exchange == "peer":
        ok = 1
        try:
            win_hits = sharding.PeerWindow(world * n, (5,), torch.int32, dst=0)
            win_occ = sharding.PeerWindow(world * n, (), torch.uint8, dst=0)
        except Exception as e:   # noqa: BLE001 -- e.g. CUDA IPC not permitted in this container
            print(f"bench.py rank {rank}: peer window unavailable ({e}); using the NCCL gather", file=sys.stderr)
            ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)   # every rank takes the same path
        if flag.item() == 0:
            for owner in (False, True):   # mappings first, then (after a barrier) the owner's allocations
                for wdw in (win_hits, win_occ):
                    if wdw is not None and wdw._owner == owner:
                        wdw.close(collective=False)
                if not owner:
                    dist.barrier()
            win_hits = win_occ = None
            args.exchange = "nccl"
        else:
            g_hits, g_occ = win_hits.tensor, win_occ.tensor
    if world > 1 and args.exchange == "nccl" and rank == 0:
        g_hits = torch.empty((world * n, 5), dtype=torch.int32, device=dev)
        g_occ = torch.empty(world * n, dtype=torch.uint8, device=dev)

    lane_sync = torch.cuda.Event()

    def step(ev_mid=None):
        if win_hits is not None:   # the previous step's delivery must not be overtaken by this step's writes
            win_hits.wait_on(stream)
            win_occ.wait_on(stream2)
        agg.IntersectBatch(d_b, hits=d_hits, n=n)
        if ev_mid is not None:
            ev_mid.record(stream)
        w1 = None
        if win_hits is not None:
            win_hits.deliver(d_hits, stream)
        elif world > 1:
            _, w1 = sharding.gather_records(d_hits, world * n, dst=0, out=g_hits, async_op=True)
        agg2.IntersectPBatch(d_s, occluded=d_occ, n=n)
        if win_hits is not None:
            win_occ.deliver(d_occ, stream2)
        elif world > 1:
            lane_sync.record(stream2)
            stream.wait_event(lane_sync)   # NCCL runs on torch's current stream
            _, w2 = sharding.gather_records(d_occ, world * n, dst=0, out=g_occ, async_op=True)
            w1.wait()
            w2.wait()

    def join():
        """`stream` waits for everything the shadow lane and the result deliveries hold."""
        lane_sync.record(stream2)
        stream.wait_event(lane_sync)
        if win_hits is not None:
            win_hits.wait_on(stream)
            win_occ.wait_on(stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    l0 = agg.launch_count() + agg2.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3 * args.steps + 2)]
    barrier()
    ev[0].record(stream)
    stream2.wait_event(ev[0])
    for k in range(args.steps):
        step(ev[3 * k + 1])
        ev[3 * k + 3].record(stream)
    join()   # the K steps end when both lanes and every delivery are done
    ev[3 * args.steps + 1].record(stream)
    barrier()
    launches = agg.launch_count() + agg2.launch_count() - l0
    total_ms = ev[0].elapsed_time(ev[3 * args.steps + 1])
    closest_ms = float(np.mean([ev[3 * k].elapsed_time(ev[3 * k + 1])
                                for k in range(args.steps)]))
    clk = clocks.stop() if rank == 0 else None
    if world > 1:
        # the gathered records must be every rank's own records, in global ray order
        mine = torch.stack([d_hits.to(torch.int64).sum(), d_occ.to(torch.int64).sum()])
        sums = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(sums, mine)
        if rank == 0:
            assert torch.equal(g_hits[:n], d_hits) and torch.equal(g_occ[:n], d_occ)
            for r in range(world):
                got = torch.stack([g_hits[r * n:(r + 1) * n].to(torch.int64).sum(),
                                   g_occ[r * n:(r + 1) * n].to(torch.int64).sum()])
                assert torch.equal(got, sums[r]), f"gathered records of rank {r} differ"
    t = torch.tensor([total_ms, closest_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, closest_ms = t.tolist()
    value = world * 2 * n * args.steps / (total_ms / 1e3) / 1e6
    # The dominant kernel alone, for the roofline: the closest-hit traversal launches of a batch,
    # CUDA events recorded by the engine on the launching stream right around them (the ray
    # ordering pre-pass, which closest_ms above includes, is outside).  L2 is flushed between
    # launches by the shadow batch's 700 MB of traffic, exactly as in the timed loop.
    # (one lane here, batches back to back: in the timed loop above the two lanes overlap, so a batch's own
    # duration is not separable there)
    trav, batch = [], []
    for _ in range(5):
        agg.IntersectBatch(d_b, hits=d_hits, n=n)
        trav.append(agg.last_traversal_ms())
        batch.append(agg.last_kernel_ms())
        agg.IntersectPBatch(d_s, occluded=d_occ, n=n)
    traversal_ms = float(np.mean(trav))
    closest_ms = float(np.mean(batch))

    # ---- e2e: pinned host buffers through the C ABI, copies inside the timed region ----
    # Layout: the SoA planes of the reference's only batch call site, the wavefront queues
    # (SOA<RayWorkItem>: o.x[], o.y[], o.z[], d.x[], d.y[], d.z[]; ShadowRayWorkItem adds tMax[];
    # wavefront/aggregate.cpp:34-68) -> nnbvh_intersect_{closest,any}_soa.  The 32-byte AoS form
    # is timed too and reported as e2e.aos_value.
    agg.set_stream(0)   # engine's own streams for the chunked H2D / kernel / D2H pipeline
    agg2.set_stream(0)
    from concurrent.futures import ThreadPoolExecutor
    pool = ThreadPoolExecutor(2)   # the two lanes again: two host threads, one engine each (ctypes drops the GIL)
    bf = bounce.view(np.float32).reshape(-1, 8)
    sf = shadow.view(np.float32).reshape(-1, 8)
    pb = [torch.from_numpy(np.ascontiguousarray(bf[:, k])).pin_memory() for k in (0, 1, 2, 4, 5, 6)]
    ps = [torch.from_numpy(np.ascontiguousarray(sf[:, k])).pin_memory() for k in (0, 1, 2, 4, 5, 6, 3)]
    h_b = torch.from_numpy(bf).pin_memory()
    h_s = torch.from_numpy(sf).pin_memory()
    h_hits = torch.zeros((n, 5), dtype=torch.int32).pin_memory()
    h_occ = torch.zeros(n, dtype=torch.uint8).pin_memory()
    e2e_steps = max(2, min(args.steps, 10))

    def e2e_soa():
        # closest and shadow batches in flight together: the shadow lane's uploads fill the link while the
        # closest lane's hit records come back, and neither pipeline's start-up / drain leaves it idle
        f1 = pool.submit(agg.IntersectBatchSoA, *pb, None, hits=h_hits)           # tMax plane NULL = Infinity
        f2 = pool.submit(agg2.IntersectBatchSoA, *ps, hits=h_occ, any_hit=True)
        f1.result()
        f2.result()

    def e2e_soa_one_lane():
        agg.IntersectBatchSoA(*pb, None, hits=h_hits)
        agg.IntersectBatchSoA(*ps, hits=h_occ, any_hit=True)

    def e2e_aos():
        agg.IntersectBatch(h_b, hits=h_hits, n=n)
        agg.IntersectPBatch(h_s, occluded=h_occ, n=n)

    def time_e2e(fn):
        for _ in range(2):
            fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            fn()
        torch.cuda.synchronize()
        tt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return world * 2 * n * e2e_steps / tt.item() / 1e6

    e2e_aos_val = time_e2e(e2e_aos)
    assert torch.equal(h_hits, d_hits.cpu()) and torch.equal(h_occ, d_occ.cpu())
    h_hits.zero_(); h_occ.zero_()
    e2e_one_lane_val = time_e2e(e2e_soa_one_lane)
    assert torch.equal(h_hits, d_hits.cpu()) and torch.equal(h_occ, d_occ.cpu())
    h_hits.zero_(); h_occ.zero_()
    e2e_val = time_e2e(e2e_soa)
    # device results and e2e results must agree (same engine, same rays)
    assert torch.equal(h_hits, d_hits.cpu()) and torch.equal(h_occ, d_occ.cpu())

    for wdw in (win_hits, win_occ):
        if wdw is not None:
            wdw.close()
    pool.shutdown()
    if rank != 0:
        agg2.close()
        agg.close()
        return None

    # ---- roofline of the dominant kernel (closest-hit traversal) ----
    cn = min(n, 1 << 21)
    _, nvs, tts = agg.IntersectCounters(shadow[:cn], any_hit=True)
    roofline = make_roofline(agg, dev, bounce[:cn], n, traversal_ms, len(w["nodes"]), len(w["tris"]),
                             f"{w['name']}:{n}")
    roofline.update({"closest_batch_ms": closest_ms,
                     "prepass": "closest_batch_ms - kernel_ms = ray-ordering pre-pass (sample, keys, 3 radix passes)",
                     "kernel_ms_source": "nnbvh_last_traversal_ms (CUDA events recorded by the engine on its launching "
                                         "stream around the traversal launches) over 5 one-lane batches AFTER the timed "
                                         "region, shadow batches in between as in the timed loop; inside the timed "
                                         "region the two lanes overlap and one batch's own duration is not separable",
                     "shadow_nodes_visited_per_ray": float(nvs.mean()),
                     "shadow_tri_tests_per_ray": float(tts.mean())})

    cpu_baseline = None
    if not args.no_cpu_baseline:
        cpu = CpuPath(w)
        mr, rays_done, secs = cpu.timed(bounce, shadow, args.cpu_seconds)
        cpu_baseline = {"value": mr, "unit": UNIT, "cores": cpu.cores, "kind": cpu.kind,
                        "sample": f"first {rays_done // 2} closest + {rays_done // 2} shadow rays "
                                  f"of the same batch, {secs:.1f} s"}
        # the same traversal returning {prim,t,b} only, without the per-hit
        # SurfaceInteraction of shapes.cpp:320-338 -- the oracle port of aggregates.cpp:529-624 on all cores
        lm, lrays, lsecs = cpu.timed(bounce, shadow, max(3.0, args.cpu_seconds / 2), lean=True)
        cpu_baseline["lean"] = {"value": lm, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port",
                                "sample": f"first {lrays // 2} closest + {lrays // 2} shadow rays of the same "
                                          f"batch, {lsecs:.1f} s; {{prim,t,b}}-only loops (no SurfaceInteraction)"}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic rays over " + w["name"],
            "config": {"workload": f"{w['name']}: {n} incoherent diffuse-bounce closest-hit rays + "
                                   f"{n} shadow any-hit rays per GPU per step",
                       "triangles": int(len(w["tris"])), "bvh_nodes": int(len(w["nodes"])),
                       "tree": "reference SAH, maxnodeprims 4", "parallelism": f"rays sharded x{world}, scene replicated; closest and shadow batches on two streams "
                                      "(two engines sharing one uploaded scene, nnbvh_clone)"
                       + ("" if world == 1 else ", hit records to rank 0 inside the step by " +
                          ("copy engines into a peer-mapped window (NVLink)" if args.exchange == "peer"
                           else "an NCCL gather")),
                       "l2": f"ray+hit buffers ({(n * 64 + n * 21) >> 20} MB per step) exceed the "
                             f"{l2_bytes(dev) >> 20} MB L2 between timed iterations (no explicit flush needed)"},
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": n * 24 + n * 28,
                    "d2h_bytes_per_step": n * 20 + n, "steps": e2e_steps,
                    "layout": "SoA planes (wavefront queue layout): 6 float planes closest, 7 shadow; "
                              "20-byte hit records + occlusion bytes back",
                    "lanes": "closest and shadow batches issued together from two host threads, one engine each "
                             "(nnbvh_clone: one uploaded scene)",
                    "one_lane_value": e2e_one_lane_val,
                    "aos_value": e2e_aos_val, "aos_h2d_bytes_per_step": 2 * n * 32},
            "gpu_launches": int(launches), "clocks": clk, "roofline": roofline,
            "cpu_baseline": cpu_baseline}
    agg2.close()
    agg.close()
    return line



# the default config-5 batch (10 M triangles seed 5, 256 Mi hashed rays): number of hits, sum of ordered
# primitive indices, xor of the t bit patterns -- identical at every N (profiles/r01o_config5_scaling.txt)
C5_EXPECTED = {(10_000_000, 256 << 20): [118559821, 591969347588707, 1010546677]}


def run_config5(args, world, rank, local, dev):
    """BASELINE.json configs[4]: synthetic 10 M-triangle soup, ONE batch of 256 Mi hashed random rays
    (closest hit) sharded over the ranks -- strong scaling -- with every 20-byte hit record delivered to
    rank 0's GPU inside the timed region (copy engines into a peer-mapped window over NVLink, chunk by
    chunk while the next chunk is traced).  The scene is built ONCE, on rank 0's host, and replicated by
    an NCCL broadcast of the flattened tree + ordered triangles.  Returns a dict (rank 0)."""
    import torch
    import torch.distributed as dist
    import nn_bvh_b200 as nb
    from nn_bvh_b200 import rays as R
    from nn_bvh_b200 import scenes, sharding

    T, NR = args.c5_tris, args.c5_rays
    t_setup = time.perf_counter()
    build_s = 0.0
    meta = torch.zeros(1, dtype=torch.int64, device=dev)
    if rank == 0:
        tris = scenes.synthetic_soup(T)
        tb = time.perf_counter()
        nodes, order = nb.build_bvh(tris, 4, "sah")
        build_s = time.perf_counter() - tb
        ordered = tris[order]
        del tris
        meta[0] = len(nodes)
    if world > 1:
        dist.broadcast(meta, 0)
    M = int(meta.item())
    d_nodes = torch.empty((M, 8), dtype=torch.int32, device=dev)
    d_tris = torch.empty((T, 9), dtype=torch.float32, device=dev)
    if rank == 0:
        d_nodes.copy_(torch.from_numpy(nodes.view(np.int32).reshape(-1, 8)))
        d_tris.copy_(torch.from_numpy(ordered))
        del nodes, order, ordered
    tb = time.perf_counter()
    if world > 1:
        dist.broadcast(d_nodes, 0)
        dist.broadcast(d_tris, 0)
    torch.cuda.synchronize()
    bcast_s = time.perf_counter() - tb
    agg = nb.GPUBVHAggregate.FromDevice(d_nodes, M, d_tris, T, device=local)
    del d_nodes, d_tris
    torch.cuda.empty_cache()

    # rank 0 also takes in every other rank's records over its NVLink ports: it may trace a smaller share
    ranges = sharding.weighted_ranges(NR, [args.c5_rank0_share if r == 0 and world > 1 else 1.0 for r in range(world)])
    lo, hi = ranges[rank]
    n_loc = hi - lo
    rays = R.hashed_random_rays(lo, hi, dev)
    hits = torch.zeros((n_loc, 5), dtype=torch.int32, device=dev)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    agg.set_stream(stream.cuda_stream)
    win = sharding.PeerWindow(NR, (5,), torch.int32, dst=0, copy_streams=args.c5_copy_streams, ranges=ranges) if world > 1 else None
    if win is not None and rank == 0:
        hits = win.tensor[lo:hi]   # rank 0 traces straight into its slice of the window: nothing to deliver
    fr = [float(x) for x in args.c5_chunks.split(",")] if world > 1 else [1.0]
    cum = np.cumsum([0.0] + fr) / sum(fr)
    bounds = sorted(set(int(round(c * n_loc)) for c in cum))
    # lanes: chunk c runs on engine / stream c % lanes (a clone shares the uploaded scene)
    lanes = [(agg, stream)]
    if world > 1 and args.c5_lanes == 2 and len(bounds) > 2:
        stream2 = torch.cuda.Stream(device=dev)
        agg2 = agg.clone()
        agg2.set_stream(stream2.cuda_stream)
        lanes.append((agg2, stream2))
    lane_done = torch.cuda.Event()
    setup_s = time.perf_counter() - t_setup

    def step():
        for c, (s0, s1) in enumerate(zip(bounds[:-1], bounds[1:])):
            a, st = lanes[c % len(lanes)]
            a.IntersectBatch(rays[s0:s1], hits=hits[s0:s1], n=s1 - s0)
            if win is not None and rank != 0:
                win.deliver(hits[s0:s1], st, at=s0)
        if len(lanes) > 1:
            lane_done.record(lanes[1][1])
            stream.wait_event(lane_done)
        if win is not None:
            win.wait_on(stream)   # the step ends when this rank's records are on rank 0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    warm = 2
    steps = max(1, min(args.steps, args.c5_steps))
    for _ in range(warm):
        step()
    barrier()
    l0 = sum(a.launch_count() for a, _ in lanes)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    if len(lanes) > 1:
        lanes[1][1].wait_event(ev0)
    for _ in range(steps):
        step()
    ev1.record(stream)
    barrier()
    launches = sum(a.launch_count() for a, _ in lanes) - l0
    t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = t.item()

    # ---- checksums: the same for every N; the gathered array must be every rank's own records ----
    hit = hits[:, 0] != -1
    tb_ = hits[:, 1].to(torch.int64)
    while tb_.numel() > 1:   # xor-reduce
        if tb_.numel() % 2:
            tb_ = torch.cat([tb_, tb_.new_zeros(1)])
        tb_ = tb_[0::2] ^ tb_[1::2]
    cs = torch.stack([hit.sum(), hits[:, 0].to(torch.int64)[hit].sum(), tb_[0]])
    if world > 1:
        parts = [torch.empty_like(cs) for _ in range(world)]
        dist.all_gather(parts, cs)
        tot = torch.stack(parts)
        cs = torch.stack([tot[:, 0].sum(), tot[:, 1].sum(),
                          torch.tensor(int(np.bitwise_xor.reduce(tot[:, 2].cpu().numpy())), device=dev)])
    checksum = [int(v) for v in cs.tolist()]
    gathered_ok = None
    if rank == 0 and win is not None:
        g = win.tensor
        gh = g[:, 0] != -1
        gathered_ok = bool(int(gh.sum()) == checksum[0] and int(g[:, 0].to(torch.int64)[gh].sum()) == checksum[1]
                           and bool(torch.equal(g[lo:hi], hits)))

    # ---- the kernel alone (roofline): this rank's whole shard as one batch, nothing delivered ----
    trav = []
    for _ in range(2):
        agg.IntersectBatch(rays, hits=hits, n=n_loc)
        trav.append(agg.last_traversal_ms())
    batch_ms = agg.last_kernel_ms()
    t = torch.tensor([min(trav), batch_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    traversal_ms, batch_ms = t.tolist()
    out = None
    if rank == 0:
        sample = rays[: 1 << 20].cpu().numpy()
        roofline = make_roofline(agg, dev, sample, n_loc, traversal_ms, M, T, f"synthetic:{T}:{n_loc}")
        roofline["batch_ms_with_prepass"] = batch_ms
        exp = C5_EXPECTED.get((T, NR))
        out = {"workload": f"synthetic:{T} triangles, ONE batch of {NR} hashed random closest-hit rays sharded over "
                           f"{world} rank(s)", "metric": "Mrays/s closest-hit", "value": NR * steps / (total_ms / 1e3) / 1e6,
               "unit": UNIT, "scaling": "strong", "n_gpus": world, "steps": steps, "warmup": warm,
               "ms_per_step": total_ms / steps, "triangles": T, "bvh_nodes": M, "rays_total": NR,
               "rays_per_rank": [b - a for a, b in ranges], "gpu_launches": int(launches),
               "exchange": (None if world == 1 else
                            f"each rank traces its shard in {len(bounds) - 1} chunks (fractions {args.c5_chunks}) "
                            f"alternating between {len(lanes)} engine(s) / stream(s), and "
                            "delivers every chunk's hit records into rank 0's peer-mapped window with a copy engine "
                            "over NVLink while the next chunk is traced; the step ends when all records are on rank 0 "
                            f"({NR * 20 / 1e9:.1f} GB)"),
               "scene_replication": ("built on rank 0's host only, replicated by an NCCL broadcast of the flattened "
                                     "tree and ordered triangles, uploaded from device pointers") if world > 1 else "single GPU",
               "host_build_s": build_s, "broadcast_s": bcast_s, "setup_s": setup_s,
               "checksum": {"hits": checksum[0], "sum_prim": checksum[1], "xor_t_bits": checksum[2],
                            "expected": exp, "ok": (checksum == exp) if exp else None,
                            "gathered_records_match": gathered_ok},
               "kernel_only_ms": traversal_ms, "roofline": roofline}
    del rays, hits
    if win is not None:
        win.close()
        win = None
    for a, _ in reversed(lanes):
        a.close()
    torch.cuda.empty_cache()
    return out


def _quiet_stdout():
    """stdout must carry exactly ONE JSON line: whatever a library prints to file descriptor 1 (NCCL's version
    banner does, whatever NCCL_DEBUG_FILE says) goes to stderr instead; the line is written to the real stdout."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


if __name__ == "__main__":
    OUT = _quiet_stdout()
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
