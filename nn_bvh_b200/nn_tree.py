"""nn_bvh_b200/nn_tree.py -- NN-split trees in the engine's `LinearBVHNode` layout.

The fork's `machine_learning/` half predicts, per level-order node, a split axis (one-hot) and
an offset, builds a Python `BVHNode` tree from them and prints its static SAH / EPO cost
(/root/reference/machine_learning/nn_tree_bench.py:44-77, 105-118).  It never produces a tree a
renderer could traverse.  This module is that missing bridge:

  TopTree        the reference's plane tree, vectorised: membership per hierarchy slot with the
                 split rule of nn_BVH.py:32-105, tight AABBs (nn_AABB.py:61-91), the early-leaf
                 flags exactly as nn_tree_bench.py:66-77 sets them, SAH (nn_loss.py:165-192)
                 and EPO (nn_loss.py:52-135).
  export_linear  planes -> flattened `LinearBVHNode[]` (cpu/aggregates.cpp:128-137) + leaf order:
                 the top levels are the NN planes, every bottom cell is finished by pbrt's own
                 builder (nnbvh_build_bvh: SAH, maxnodeprims), so the result can be uploaded
                 next to a pure SAH tree and traced with the same rays.
  greedy_sah_planes / median_planes   plane sources when no trained checkpoint is available
                 (TensorFlow and the Keras checkpoints are absent here): the reference's greedy
                 per-node SAH search (nn_BVH.py:193-260 evaluates SAH_single_node,
                 nn_loss.py:194-222, for every distinct extent-midpoint) restated with sorted
                 prefix/suffix bounds, and plain medians.
  traversal_cost the MEASURED cost the reference never had: mean nodesVisited / triTests per ray
                 from the engine's counters (reference counting rules).

Host-side numpy only; all ray queries go through GPUBVHAggregate.
"""
from __future__ import annotations

import sys

import numpy as np

from .capi import NODE_DTYPE

MAX_PRIMITIVES_PER_LEAF = 2   # nn_BVH.py:9
C_INN, C_TRI = 1.2, 1.0       # nn_loss.py:111-115


def _split_mask(T: np.ndarray, idx: np.ndarray, axis: int, pos: float) -> np.ndarray:
    """True = left.  nn_BVH.py:45-71: max <= pos -> left; min > pos -> right; a straddler goes
    left iff the part left of the plane is at least as long as the part right of it."""
    c = T[idx, :, axis]
    mn, mx = c.min(1), c.max(1)
    return (mx <= pos) | ((mn <= pos) & ((pos - mn) >= (mx - pos)))


def _aabb(T: np.ndarray, idx: np.ndarray) -> np.ndarray:
    """nn_AABB.py:61-91, including its seeds: maxima start at sys.float_info.min (smallest
    positive double) and an empty list gives the all-zero box."""
    if len(idx) == 0:
        return np.zeros(6)
    v = T[idx].reshape(-1, 3)
    return np.concatenate([np.minimum(v.min(0), sys.float_info.max),
                           np.maximum(v.max(0), sys.float_info.min)])


def _surface(b: np.ndarray) -> float:
    x, y, z = b[3] - b[0], b[4] - b[1], b[5] - b[2]
    return float(2 * (x * y + x * z + y * z))


def _plane_axes(planes: np.ndarray) -> list[int]:
    axes = []
    for row in np.asarray(planes, dtype=np.float64):
        oh = row[:3]
        if not (np.sort(oh).tolist() == [0.0, 0.0, 1.0]):
            raise ValueError("Invalid split axis.")   # nn_tree_bench.py:56
        axes.append(int(np.argmax(oh)))
    return axes


class TopTree:
    """Level-order hierarchy of the plane tree: slot k has children 2k+1, 2k+2."""

    def __init__(self, tris9: np.ndarray, planes: np.ndarray):
        self.T = np.asarray(tris9, dtype=np.float64).reshape(-1, 3, 3)
        self.planes = np.asarray(planes, dtype=np.float64).reshape(-1, 4)
        n_planes = len(self.planes)
        self.levels = int(np.log2(n_planes + 1))
        if 2 ** self.levels - 1 != n_planes:
            raise ValueError("planes must describe a complete tree (2^L - 1 rows)")
        self.axes = _plane_axes(self.planes)
        n_slots = 2 ** (self.levels + 1) - 1
        self.members: list[np.ndarray] = [None] * n_slots
        self.members[0] = np.arange(len(self.T))
        self.is_leaf = np.zeros(n_slots, dtype=bool)
        for i in range(n_planes):
            idx = self.members[i]
            left = _split_mask(self.T, idx, self.axes[i], float(self.planes[i, 3])) if len(idx) \
                else np.zeros(0, bool)
            self.members[2 * i + 1], self.members[2 * i + 2] = idx[left], idx[~left]
            # nn_tree_bench.py:72-75: only the LEFT child is ever flagged here
            if len(self.members[2 * i + 1]) <= MAX_PRIMITIVES_PER_LEAF:
                self.is_leaf[2 * i + 1] = True
        self.is_leaf[2 ** self.levels - 1:] = True   # :76-77
        self.aabbs = np.stack([_aabb(self.T, m) for m in self.members])

    def _to_list(self):
        """nn_BVH.py:108-127: DFS that stops at flagged leaves."""
        inner, leaves, todo = [], [], [0]
        while todo:
            cur = todo.pop()
            while not self.is_leaf[cur]:
                inner.append(cur)
                todo.append(2 * cur + 2)
                cur = 2 * cur + 1
            leaves.append(cur)
        return inner, leaves

    def sah(self) -> float:
        inner, leaves = self._to_list()
        s = C_INN * sum(_surface(self.aabbs[k]) for k in inner)
        s += C_TRI * sum(_surface(self.aabbs[k]) * len(self.members[k]) for k in leaves)
        return s / _surface(self.aabbs[0])

    def epo(self) -> float:
        """(S \\ Q(n)) intersected with n, approximated as the reference does (nn_loss.py:31-50):
        a primitive counts with its WHOLE area when at least one vertex lies inside n's box.
        The reference reaches those primitives by walking overlapping sibling subtrees down to
        flagged leaves; every primitive belongs to exactly one such leaf and a box holding a
        vertex of it overlaps every ancestor box of that leaf, so the walk finds exactly the
        primitives outside n's own list with a vertex inside n's box."""
        T = self.T
        area = 0.5 * np.linalg.norm(np.cross(T[:, 1] - T[:, 0], T[:, 2] - T[:, 0]), axis=1)
        inner, leaves = self._to_list()
        tot = 0.0
        for k, w in [(k, C_INN) for k in inner] + [(k, C_TRI) for k in leaves]:
            if k == 0:
                continue
            b = self.aabbs[k]
            inside = np.all((T >= b[:3]) & (T <= b[3:]), axis=2).any(axis=1)
            inside[self.members[k]] = False
            tot += w * float(area[inside].sum())
        return tot / float(area.sum())


def export_linear(tris9: np.ndarray, planes: np.ndarray, max_prims_in_node: int = 4,
                  split_method: str = "sah"):
    """-> (nodes NODE_DTYPE[M], order uint32[T]): pre-order flattened tree whose top levels are
    the planes and whose bottom cells are finished by the reference builder.  A split that
    leaves one side empty is skipped (a `LinearBVHNode` cannot express an empty child)."""
    from .aggregate import build_bvh
    tris32 = np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9)
    T = tris32.astype(np.float64).reshape(-1, 3, 3)
    planes = np.asarray(planes, dtype=np.float64).reshape(-1, 4)
    axes = _plane_axes(planes)
    n_planes = len(planes)

    def rec(slot: int, idx: np.ndarray):
        if slot >= n_planes or len(idx) <= 1:
            nodes, order = build_bvh(tris32[idx], max_prims_in_node, split_method)
            return nodes, idx[order]
        left = _split_mask(T, idx, axes[slot], float(planes[slot, 3]))
        if left.all() or not left.any():
            return rec(2 * slot + 1 if left.all() else 2 * slot + 2, idx)
        ln, lo = rec(2 * slot + 1, idx[left])
        rn, ro = rec(2 * slot + 2, idx[~left])
        nodes = np.zeros(1 + len(ln) + len(rn), dtype=NODE_DTYPE)
        nodes[1:1 + len(ln)] = ln
        nodes[1 + len(ln):] = rn
        # child subtrees carry offsets relative to their own start: rebase
        for sub, nbase, pbase in ((nodes[1:1 + len(ln)], 1, 0), (nodes[1 + len(ln):], 1 + len(ln), len(lo))):
            leaf = sub["nPrimitives"] > 0
            sub["offset"][leaf] += pbase
            sub["offset"][~leaf] += nbase
        root = nodes[0]
        root["bounds"][:3] = np.minimum(ln[0]["bounds"][:3], rn[0]["bounds"][:3])
        root["bounds"][3:] = np.maximum(ln[0]["bounds"][3:], rn[0]["bounds"][3:])
        root["offset"] = 1 + len(ln)
        root["nPrimitives"] = 0
        root["axis"] = axes[slot]
        return nodes, np.concatenate([lo, ro])

    nodes, order = rec(0, np.arange(len(tris32)))
    return nodes, order.astype(np.uint32)


def _best_split(T: np.ndarray, idx: np.ndarray, candidates):
    """Greedy step of nn_BVH.py:193-260: the (axis, offset) minimising SAH_single_node
    (nn_loss.py:194-222) = (SA_l*n_l + SA_r*n_r) / SA_p over offsets just above each distinct
    extent midpoint (get_all_split_offsets, nn_BVH.py:147-161)."""
    best = (np.inf, 0, 0.0)
    if len(idx) < 2:
        return best
    lo, hi = T[idx].min(1), T[idx].max(1)       # per-primitive bounds (n,3)
    sa_p = _surface(np.concatenate([lo.min(0), hi.max(0)]))
    n = len(idx)
    for axis in range(3):
        c = (lo[:, axis] + hi[:, axis]) / 2
        o = np.argsort(c, kind="stable")
        cs, los, his = c[o], lo[o], hi[o]
        pl, ph = np.minimum.accumulate(los, 0), np.maximum.accumulate(his, 0)
        sl, sh = np.minimum.accumulate(los[::-1], 0)[::-1], np.maximum.accumulate(his[::-1], 0)[::-1]
        s = np.nonzero(cs[1:] > cs[:-1])[0] + 1   # left = first s primitives, s at value changes
        if len(s) == 0:
            continue
        if candidates and len(s) > candidates:
            s = s[np.linspace(0, len(s) - 1, candidates).astype(int)]
        dl, dr = ph[s - 1] - pl[s - 1], sh[s] - sl[s]
        sa_l = 2 * (dl[:, 0] * dl[:, 1] + dl[:, 0] * dl[:, 2] + dl[:, 1] * dl[:, 2])
        sa_r = 2 * (dr[:, 0] * dr[:, 1] + dr[:, 0] * dr[:, 2] + dr[:, 1] * dr[:, 2])
        cost = (sa_l * s + sa_r * (n - s)) / sa_p if sa_p > 0 else (s * 0.0 + np.abs(2 * s - n))
        k = int(np.argmin(cost))
        if cost[k] < best[0]:
            best = (float(cost[k]), axis, float(np.nextafter(cs[s[k] - 1], np.inf)))
    return best


def _planes_from(T: np.ndarray, levels: int, choose):
    sets = [np.arange(len(T))]
    rows = []
    for i in range(2 ** levels - 1):
        idx = sets[i]
        axis, off = choose(idx)
        row = [0.0, 0.0, 0.0, off]
        row[axis] = 1.0
        rows.append(row)
        left = _split_mask(T, idx, axis, off) if len(idx) else np.zeros(0, bool)
        sets += [idx[left], idx[~left]]
    return np.array(rows, dtype=np.float64)


def greedy_sah_planes(tris9: np.ndarray, levels: int = 4, candidates: int | None = 64) -> np.ndarray:
    """Level-order planes [one-hot axis, offset] of the greedy SAH search (lvls=4 in the
    reference config, nss_global_config.py:15)."""
    T = np.asarray(tris9, dtype=np.float64).reshape(-1, 3, 3)

    def choose(idx):
        _, axis, off = _best_split(T, idx, candidates)
        return axis, off

    return _planes_from(T, levels, choose)


def median_planes(tris9: np.ndarray, levels: int = 4) -> np.ndarray:
    """Longest axis of the midpoint bounds, offset at the median midpoint."""
    T = np.asarray(tris9, dtype=np.float64).reshape(-1, 3, 3)

    def choose(idx):
        if len(idx) == 0:
            return 0, 0.0
        c = (T[idx].min(1) + T[idx].max(1)) / 2
        axis = int(np.argmax(c.max(0) - c.min(0)))
        return axis, float(np.median(c[:, axis]))

    return _planes_from(T, levels, choose)


def linear_sah_cost(nodes: np.ndarray) -> float:
    """nn_loss.SAH's formula (C_inn=1.2, C_tri=1.0) over a whole flattened tree: comparable
    between a pure SAH tree and an exported NN-split tree."""
    b = nodes["bounds"].astype(np.float64)
    d = b[:, 3:] - b[:, :3]
    sa = 2 * (d[:, 0] * d[:, 1] + d[:, 0] * d[:, 2] + d[:, 1] * d[:, 2])
    leaf = nodes["nPrimitives"] > 0
    return float((C_INN * sa[~leaf].sum() + C_TRI * (sa[leaf] * nodes["nPrimitives"][leaf]).sum()) / sa[0])


def traversal_cost(agg, rays, any_hit: bool = False) -> dict:
    """Measured traversal cost per ray with the reference's counting rules
    (cpu/aggregates.cpp:541,592; shapes.cpp:322,343)."""
    _, nv, tt = agg.IntersectCounters(rays, any_hit=any_hit)
    return {"nodes_visited": float(nv.mean()), "tri_tests": float(tt.mean()), "rays": int(len(nv))}
