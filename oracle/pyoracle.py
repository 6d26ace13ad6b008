"""oracle/pyoracle.py -- TEST INFRASTRUCTURE (the checker), not product code.

ctypes bindings for
  * oracle/_build/libnnbvh_oracle.so : our CPU restatement (oracle/nnbvh_oracle.cpp)
  * oracle/_ref/libnnbvh_ref.so      : the UNMODIFIED reference translation units
                                       behind oracle/ref_driver.cpp (optional)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module.  The product package
(nn_bvh_b200/) must never do so.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "_build", "libnnbvh_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libnnbvh_ref.so")
BINDING_SO = os.path.join(HERE, "_ref", "libnnbvh_ref_binding.so")
MISS = 0xFFFFFFFF

SPLIT = {"sah": 0, "hlbvh": 1, "middle": 2, "equal": 3}


def build(ref: bool = True) -> None:
    """Compile the restatement (always) and oracle/_ref (when /root/reference exists)."""
    targets = ["oracle"] + (["ref", "binding"] if ref else [])
    subprocess.check_call(["make", "-s", "-C", HERE, "-j8"] + targets)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class Oracle:
    """The CPU restatement."""

    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        L = self.lib = ctypes.CDLL(ORACLE_SO)
        L.nnbvh_oracle_build.restype = ctypes.c_int64
        L.nnbvh_oracle_build.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32,
                                         ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p]
        for f in (L.nnbvh_oracle_intersect, L.nnbvh_oracle_intersectP):
            f.restype = None
            f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                          ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32]
        L.nnbvh_oracle_triangle.argtypes = [ctypes.c_void_p] * 3
        L.nnbvh_oracle_box.argtypes = [ctypes.c_void_p] * 2

    def build_bvh(self, tris9: np.ndarray, max_prims: int = 4, split: str = "sah"):
        """tris9: (T,9) float32 in original order -> (nodes (M,8) uint32 view, order (T,) int32)."""
        tris9 = np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9)
        T = tris9.shape[0]
        nodes = np.zeros((2 * T, 8), dtype=np.uint32)
        order = np.zeros(T, dtype=np.int32)
        n = self.lib.nnbvh_oracle_build(_p(tris9), T, max_prims, SPLIT[split], _p(nodes), _p(order))
        if n < 0:
            raise ValueError("oracle build failed")
        return nodes[:n].copy(), order

    def intersect(self, nodes, ordered_tris9, rays8, counters=False, threads=0):
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32).reshape(-1, 8)
        n = rays8.shape[0]
        hits = np.zeros((n, 5), dtype=np.uint32)
        nv = np.zeros(n, dtype=np.uint32) if counters else None
        tt = np.zeros(n, dtype=np.uint32) if counters else None
        self.lib.nnbvh_oracle_intersect(_p(nodes), _p(ordered_tris9), _p(rays8), n, _p(hits),
                                        _p(nv) if counters else None,
                                        _p(tt) if counters else None, threads)
        return (hits, nv, tt) if counters else hits

    def intersect_p(self, nodes, ordered_tris9, rays8, counters=False, threads=0):
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32).reshape(-1, 8)
        n = rays8.shape[0]
        occ = np.zeros(n, dtype=np.uint8)
        nv = np.zeros(n, dtype=np.uint32) if counters else None
        tt = np.zeros(n, dtype=np.uint32) if counters else None
        self.lib.nnbvh_oracle_intersectP(_p(nodes), _p(ordered_tris9), _p(rays8), n, _p(occ),
                                         _p(nv) if counters else None,
                                         _p(tt) if counters else None, threads)
        return (occ, nv, tt) if counters else occ

    # ---- primitive lists that mix triangles and spheres ----
    def build_bounds(self, bounds6, max_prims: int = 4, split: str = "sah"):
        bounds6 = np.ascontiguousarray(bounds6, dtype=np.float32).reshape(-1, 6)
        P = bounds6.shape[0]
        nodes = np.zeros((2 * P, 8), dtype=np.uint32)
        order = np.zeros(P, dtype=np.int32)
        f = self.lib.nnbvh_oracle_build_bounds
        f.restype = ctypes.c_int64
        f.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p,
                      ctypes.c_void_p]
        n = f(_p(bounds6), P, max_prims, SPLIT[split], _p(nodes), _p(order))
        if n < 0:
            raise ValueError("oracle build failed")
        return nodes[:n].copy(), order

    def _mixed(self, fn, out, nodes, ordered_tris9, sphere_of, spheres28, rays8, counters, threads):
        n = rays8.shape[0]
        nv = np.zeros(n, dtype=np.uint32) if counters else None
        tt = np.zeros(n, dtype=np.uint32) if counters else None
        fn.restype = None
        fn.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int64] + [ctypes.c_void_p] * 3 + [ctypes.c_int32]
        fn(_p(nodes), _p(ordered_tris9), _p(sphere_of), _p(spheres28), _p(rays8), n, _p(out),
           _p(nv) if counters else None, _p(tt) if counters else None, threads)
        return (out, nv, tt) if counters else out

    def intersect_mixed(self, nodes, ordered_tris9, sphere_of, spheres28, rays8, counters=False, threads=0):
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32).reshape(-1, 8)
        sphere_of = np.ascontiguousarray(sphere_of, dtype=np.int32)
        spheres28 = np.ascontiguousarray(spheres28, dtype=np.float32).reshape(-1, 28)
        return self._mixed(self.lib.nnbvh_oracle_intersect_mixed, np.zeros((rays8.shape[0], 5), dtype=np.uint32),
                           nodes, np.ascontiguousarray(ordered_tris9, dtype=np.float32), sphere_of, spheres28,
                           rays8, counters, threads)

    def intersect_p_mixed(self, nodes, ordered_tris9, sphere_of, spheres28, rays8, counters=False, threads=0):
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32).reshape(-1, 8)
        sphere_of = np.ascontiguousarray(sphere_of, dtype=np.int32)
        spheres28 = np.ascontiguousarray(spheres28, dtype=np.float32).reshape(-1, 28)
        return self._mixed(self.lib.nnbvh_oracle_intersectP_mixed, np.zeros(rays8.shape[0], dtype=np.uint8),
                           nodes, np.ascontiguousarray(ordered_tris9, dtype=np.float32), sphere_of, spheres28,
                           rays8, counters, threads)

    # ---- instanced scenes (TransformedPrimitive over per-object BVHAggregates) ----
    def intersect_instanced(self, sc: dict, rays8, any_hit=False, counters=False, threads=0):
        """sc: dict with nodes, ordered_tris9, instance_of, inst24, inst_object, obj_nodes, obj_node_base,
        obj_tris, obj_prim_base (numpy arrays).  -> (hits5, inner) or occluded [, nodesVisited, triTests]."""
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32).reshape(-1, 8)
        n = rays8.shape[0]
        hits = np.zeros((n, 5), dtype=np.uint32)
        inner = np.zeros(n, dtype=np.uint32)
        occ = np.zeros(n, dtype=np.uint8)
        nv = np.zeros(n, dtype=np.uint32)
        tt = np.zeros(n, dtype=np.uint32)
        f = self.lib.nnbvh_oracle_intersect_instanced
        f.restype = None
        f.argtypes = [ctypes.c_void_p] * 10 + [ctypes.c_int64] + [ctypes.c_void_p] * 5 + [ctypes.c_int32]
        a = {k: np.ascontiguousarray(v) for k, v in sc.items()}
        f(_p(a["nodes"]), _p(a["ordered_tris9"]), _p(a["instance_of"]), _p(a["inst24"]), _p(a["inst_object"]),
          _p(a["obj_nodes"]), _p(a["obj_node_base"]), _p(a["obj_tris"]), _p(a["obj_prim_base"]), _p(rays8), n,
          _p(hits), _p(inner), _p(occ) if any_hit else None, _p(nv), _p(tt), threads)
        res = (occ,) if any_hit else (hits, inner)
        return res + (nv, tt) if counters else (res[0] if any_hit else res)

    def instance_bounds(self, render_from_primitive12, object_bounds6):
        m = np.ascontiguousarray(render_from_primitive12, dtype=np.float32)
        b = np.ascontiguousarray(object_bounds6, dtype=np.float32)
        out = np.zeros(6, dtype=np.float32)
        self.lib.nnbvh_oracle_instance_bounds.argtypes = [ctypes.c_void_p] * 3
        self.lib.nnbvh_oracle_instance_bounds.restype = None
        self.lib.nnbvh_oracle_instance_bounds(_p(m), _p(b), _p(out))
        return out

    def sphere(self, ray8, sphere28):
        ray8 = np.ascontiguousarray(ray8, dtype=np.float32)
        sphere28 = np.ascontiguousarray(sphere28, dtype=np.float32)
        out = np.zeros(4, dtype=np.float32)
        self.lib.nnbvh_oracle_sphere.argtypes = [ctypes.c_void_p] * 3
        hit = self.lib.nnbvh_oracle_sphere(_p(ray8), _p(sphere28), _p(out))
        return bool(hit), out

    def spawn_origin(self, ray8, p9):
        """(hit, [o.xyz of intr.SpawnRay(ray.d), tHit]) for one (ray, triangle) pair."""
        out = np.zeros(4, dtype=np.float32)
        f = self.lib.nnbvh_oracle_spawn_origin
        f.argtypes = [ctypes.c_void_p] * 3
        hit = f(_p(np.ascontiguousarray(ray8, dtype=np.float32)), _p(np.ascontiguousarray(p9, dtype=np.float32)), _p(out))
        return bool(hit), out

    def intersect_alpha(self, nodes, ordered_tris9, ordered_alpha, rays8, any_hit=False, threads=0):
        """Triangles behind GeometricPrimitives with constant alpha: hits5, or occluded with any_hit."""
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32).reshape(-1, 8)
        n = rays8.shape[0]
        hits = np.zeros((n, 5), dtype=np.uint32)
        occ = np.zeros(n, dtype=np.uint8)
        f = self.lib.nnbvh_oracle_intersect_alpha
        f.restype = None
        f.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32]
        a = np.ascontiguousarray(ordered_alpha, dtype=np.float32)
        t = np.ascontiguousarray(ordered_tris9, dtype=np.float32)
        f(_p(nodes), _p(t), _p(a), _p(rays8), n, _p(hits), _p(occ) if any_hit else None, threads)
        return occ if any_hit else hits

    def hash_float(self, o3, d3) -> float:
        f = self.lib.nnbvh_oracle_hash_float
        f.restype = ctypes.c_float
        f.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        return f(_p(np.ascontiguousarray(o3, dtype=np.float32)), _p(np.ascontiguousarray(d3, dtype=np.float32)))

    def atan2f(self, y, x):
        """The host libm's atan2f, elementwise."""
        y = np.ascontiguousarray(y, dtype=np.float32)
        x = np.ascontiguousarray(x, dtype=np.float32)
        out = np.zeros_like(y)
        f = self.lib.nnbvh_oracle_atan2f
        f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
        f.restype = None
        f(_p(y), _p(x), y.size, _p(out))
        return out

    def sphere_bounds(self, sphere28):
        sphere28 = np.ascontiguousarray(sphere28, dtype=np.float32)
        out = np.zeros(6, dtype=np.float32)
        self.lib.nnbvh_oracle_sphere_bounds.argtypes = [ctypes.c_void_p] * 2
        self.lib.nnbvh_oracle_sphere_bounds.restype = None
        self.lib.nnbvh_oracle_sphere_bounds(_p(sphere28), _p(out))
        return out

    def triangle(self, ray8, p9):
        ray8 = np.ascontiguousarray(ray8, dtype=np.float32)
        p9 = np.ascontiguousarray(p9, dtype=np.float32)
        out = np.zeros(4, dtype=np.float32)
        hit = self.lib.nnbvh_oracle_triangle(_p(ray8), _p(p9), _p(out))
        return bool(hit), out

    def box(self, ray8, box6):
        ray8 = np.ascontiguousarray(ray8, dtype=np.float32)
        box6 = np.ascontiguousarray(box6, dtype=np.float32)
        return bool(self.lib.nnbvh_oracle_box(_p(ray8), _p(box6)))


def as_pbrt_stores(tris9: np.ndarray) -> np.ndarray:
    """Vertex positions as the reference's TriangleMesh holds them.  Its constructor
    pushes every vertex through renderFromObject (util/mesh.cpp:36-39); even for the
    identity that computes 0*x + 0*y + 1*z + 0, which turns -0.0 into +0.0.  Parity is
    on the geometry the reference actually traces, so fixtures store this form."""
    return (np.asarray(tris9, dtype=np.float32) + np.float32(0.0)).astype(np.float32)


def ref_available() -> bool:
    return os.path.exists(REF_SO)


class Reference:
    """The unmodified reference code (oracle/_ref), when it has been built."""

    _inited = False

    def __init__(self, threads: int = 0, so: str = REF_SO):
        if not os.path.exists(so):
            raise FileNotFoundError(so)
        L = self.lib = ctypes.CDLL(so)
        L.nnbvh_ref_scene_create.restype = ctypes.c_void_p
        L.nnbvh_ref_scene_create.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                             ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.nnbvh_ref_scene_num_nodes.argtypes = [ctypes.c_void_p]
        L.nnbvh_ref_scene_export.argtypes = [ctypes.c_void_p] * 3
        L.nnbvh_ref_intersect.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                          ctypes.c_void_p, ctypes.c_int]
        L.nnbvh_ref_intersect_only.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                               ctypes.c_void_p]
        L.nnbvh_ref_intersectP.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                           ctypes.c_void_p, ctypes.c_int]
        L.nnbvh_ref_triangle.argtypes = [ctypes.c_void_p] * 3
        L.nnbvh_ref_box.argtypes = [ctypes.c_void_p] * 2
        L.nnbvh_ref_stats.argtypes = [ctypes.c_void_p] * 3
        L.nnbvh_ref_loop_subdivide.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                               ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                               ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.nnbvh_ref_free.argtypes = [ctypes.c_void_p]
        L.nnbvh_ref_set_threads(threads)

    def set_threads(self, n: int):
        self.lib.nnbvh_ref_set_threads(n)

    @property
    def threads(self) -> int:
        return self.lib.nnbvh_ref_threads()

    def scene(self, tris9: np.ndarray, max_prims: int = 4, split: str = "sah"):
        """Build the reference BVHAggregate over a de-indexed triangle soup."""
        tris9 = np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9)
        if np.any(np.signbit(tris9) & (tris9 == 0)):
            raise ValueError("triangle soup holds -0.0: pass as_pbrt_stores(tris) to both sides")
        T = tris9.shape[0]
        verts = tris9.reshape(-1, 3)
        idx = np.arange(3 * T, dtype=np.int32)
        h = self.lib.nnbvh_ref_scene_create(_p(verts), 3 * T, _p(idx), T, max_prims, SPLIT[split])
        return ctypes.c_void_p(h)

    def spawn_origin(self, ray8, p9):
        """(hit, [o.xyz of intr.SpawnRay(ray.d), tHit]) for one (ray, triangle) pair."""
        out = np.zeros(4, dtype=np.float32)
        f = self.lib.nnbvh_ref_spawn_origin
        f.argtypes = [ctypes.c_void_p] * 3
        hit = f(_p(np.ascontiguousarray(ray8, dtype=np.float32)), _p(np.ascontiguousarray(p9, dtype=np.float32)), _p(out))
        return bool(hit), out

    def scene_alpha(self, tris9, alpha, max_prims: int = 4, split: str = "sah"):
        """BVHAggregate over triangles behind GeometricPrimitives with a constant alpha each (>= 1: none)."""
        tris9 = np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9)
        alpha = np.ascontiguousarray(alpha, dtype=np.float32)
        T = tris9.shape[0]
        verts = tris9.reshape(-1, 3)
        idx = np.arange(3 * T, dtype=np.int32)
        f = self.lib.nnbvh_ref_scene_create_alpha
        f.restype = ctypes.c_void_p
        f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                      ctypes.c_int]
        return ctypes.c_void_p(f(_p(verts), 3 * T, _p(idx), T, _p(alpha), max_prims, SPLIT[split]))

    def hash_float(self, o3, d3) -> float:
        f = self.lib.nnbvh_ref_hash_float
        f.restype = ctypes.c_float
        f.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        return f(_p(np.ascontiguousarray(o3, dtype=np.float32)), _p(np.ascontiguousarray(d3, dtype=np.float32)))

    def build_seconds(self, scene) -> float:
        """Wall time of the BVHAggregate constructor that made this scene."""
        f = self.lib.nnbvh_ref_scene_build_seconds
        f.restype = ctypes.c_double
        f.argtypes = [ctypes.c_void_p]
        return f(scene)

    def scene_mixed(self, tris9, spheres28, max_prims: int = 4, split: str = "sah"):
        """BVHAggregate over [triangles..., spheres...] (SimplePrimitive each)."""
        tris9 = np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9)
        spheres28 = np.ascontiguousarray(spheres28, dtype=np.float32).reshape(-1, 28)
        if np.any(np.signbit(tris9) & (tris9 == 0)):
            raise ValueError("triangle soup holds -0.0: pass as_pbrt_stores(tris) to both sides")
        T = tris9.shape[0]
        verts = np.ascontiguousarray(tris9.reshape(-1, 3))
        idx = np.arange(3 * T, dtype=np.int32)
        f = self.lib.nnbvh_ref_scene_create_mixed
        f.restype = ctypes.c_void_p
        f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                      ctypes.c_int, ctypes.c_int, ctypes.c_int]
        self.lib.nnbvh_ref_scene_num_prims.argtypes = [ctypes.c_void_p]
        return ctypes.c_void_p(f(_p(verts), 3 * T, _p(idx), T, _p(spheres28), spheres28.shape[0],
                                 max_prims, SPLIT[split]))

    def scene_instanced(self, tris9, obj_soups, inst24, inst_object, max_prims: int = 4, split: str = "sah"):
        """BVHAggregate over [top-level triangles..., TransformedPrimitive(object BVHAggregate) per instance].
        Returns a dict with the reference's trees in the layout Oracle.intersect_instanced takes."""
        tris9 = np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9)
        T = tris9.shape[0]
        obj_soups = [np.ascontiguousarray(o, dtype=np.float32).reshape(-1, 9) for o in obj_soups]
        obj_all = np.ascontiguousarray(np.concatenate(obj_soups))
        start = np.concatenate([[0], np.cumsum([len(o) for o in obj_soups])]).astype(np.int32)
        inst24 = np.ascontiguousarray(inst24, dtype=np.float32).reshape(-1, 24)
        inst_object = np.ascontiguousarray(inst_object, dtype=np.int32)
        K = len(inst24)
        L = self.lib
        L.nnbvh_ref_scene_create_instanced.restype = ctypes.c_void_p
        L.nnbvh_ref_scene_create_instanced.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                                       ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                                       ctypes.c_int, ctypes.c_int]
        h = ctypes.c_void_p(L.nnbvh_ref_scene_create_instanced(_p(tris9), T, _p(obj_all), _p(start), len(obj_soups),
                                                               _p(inst24), _p(inst_object), K, max_prims, SPLIT[split]))
        L.nnbvh_ref_scene_num_prims.argtypes = [ctypes.c_void_p]
        nodes, order = self.export(h, T + K)
        L.nnbvh_ref_scene_object_num_nodes.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.nnbvh_ref_scene_export_object.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        L.nnbvh_ref_instance_bounds.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        L.nnbvh_ref_instance_bounds.restype = None
        obj_nodes, obj_tris, obj_orders, nbase, pbase = [], [], [], [0], [0]
        for j, soup in enumerate(obj_soups):
            m = L.nnbvh_ref_scene_object_num_nodes(h, j)
            nd = np.zeros((m, 8), dtype=np.uint32)
            od = np.zeros(len(soup), dtype=np.int32)
            L.nnbvh_ref_scene_export_object(h, j, _p(nd), _p(od))
            obj_nodes.append(nd)
            obj_tris.append(soup[od])
            obj_orders.append(od)
            nbase.append(nbase[-1] + m)
            pbase.append(pbase[-1] + len(soup))
        ib = np.zeros((K, 6), dtype=np.float32)
        for k in range(K):
            L.nnbvh_ref_instance_bounds(h, k, _p(ib[k]))
        ordered = np.zeros((T + K, 9), dtype=np.float32)
        is_tri = order < T
        ordered[is_tri] = tris9[order[is_tri]]
        return dict(handle=h, order=order, obj_orders=obj_orders, instance_bounds=ib,
                    data=dict(nodes=nodes, ordered_tris9=ordered,
                              instance_of=np.where(order >= T, order - T, -1).astype(np.int32),
                              inst24=inst24, inst_object=inst_object,
                              obj_nodes=np.concatenate(obj_nodes), obj_node_base=np.array(nbase[:-1], dtype=np.int32),
                              obj_tris=np.ascontiguousarray(np.concatenate(obj_tris)),
                              obj_prim_base=np.array(pbase[:-1], dtype=np.int32)))

    def tag_materials(self, on: bool):
        """Scenes created while off hold null materials only (needed by the wavefront queue harnesses); the
        intersect_* calls that name the primitive hit need the tags."""
        self.lib.nnbvh_ref_tag_materials(int(on))

    def scene_general(self, tris9, alpha, spheres28, obj_soups, inst24, inst_object, max_prims: int = 4,
                      split: str = "sah"):
        """BVHAggregate over ONE list: triangles (GeometricPrimitive with constant alpha where alpha < 1), then
        spheres, then TransformedPrimitive instances of per-object aggregates.  Returns the handle and the
        reference's trees / orders in the layout nnbvh_upload_scene_general takes."""
        tris9 = np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9)
        T = tris9.shape[0]
        alpha = np.ascontiguousarray(alpha, dtype=np.float32).reshape(-1)
        spheres28 = np.ascontiguousarray(spheres28, dtype=np.float32).reshape(-1, 28)
        S = len(spheres28)
        obj_soups = [np.ascontiguousarray(o, dtype=np.float32).reshape(-1, 9) for o in obj_soups]
        obj_all = np.ascontiguousarray(np.concatenate(obj_soups))
        start = np.concatenate([[0], np.cumsum([len(o) for o in obj_soups])]).astype(np.int32)
        inst24 = np.ascontiguousarray(inst24, dtype=np.float32).reshape(-1, 24)
        inst_object = np.ascontiguousarray(inst_object, dtype=np.int32)
        K = len(inst24)
        L = self.lib
        f = L.nnbvh_ref_scene_create_general
        f.restype = ctypes.c_void_p
        f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                      ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                      ctypes.c_int]
        h = ctypes.c_void_p(f(_p(tris9), T, _p(alpha), _p(spheres28), S, _p(obj_all), _p(start), len(obj_soups),
                              _p(inst24), _p(inst_object), K, max_prims, SPLIT[split]))
        nodes, order = self.export(h, T + S + K)
        L.nnbvh_ref_scene_object_num_nodes.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.nnbvh_ref_scene_export_object.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        objects = []
        for j, soup in enumerate(obj_soups):
            m = L.nnbvh_ref_scene_object_num_nodes(h, j)
            nd = np.zeros((m, 8), dtype=np.uint32)
            od = np.zeros(len(soup), dtype=np.int32)
            L.nnbvh_ref_scene_export_object(h, j, _p(nd), _p(od))
            objects.append(dict(nodes=nd, order=od, ordered_tris9=np.ascontiguousarray(soup[od])))
        ordered = np.zeros((T + S + K, 9), dtype=np.float32)
        is_tri = order < T
        ordered[is_tri] = tris9[order[is_tri]]
        ordered_alpha = np.ones(T + S + K, dtype=np.float32)
        ordered_alpha[is_tri] = alpha[order[is_tri]]
        return dict(handle=h, nodes=nodes, order=order, ordered_tris9=ordered, ordered_alpha=ordered_alpha,
                    sphere_of=np.where((order >= T) & (order < T + S), order - T, -1).astype(np.int32),
                    instance_of=np.where(order >= T + S, order - T - S, -1).astype(np.int32), objects=objects)

    def intersect_general(self, handle, rays8):
        """BVHAggregate::Intersect over a scene_general handle -> (hits (n,5) uint32, inner (n,) uint32)."""
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32).reshape(-1, 8)
        n = rays8.shape[0]
        hits = np.zeros((n, 5), dtype=np.uint32)
        inner = np.zeros(n, dtype=np.uint32)
        f = self.lib.nnbvh_ref_intersect_general
        f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p]
        f.restype = None
        f(handle, _p(rays8), n, _p(hits), _p(inner))
        return hits, inner

    def intersect_instanced(self, handle, rays8):
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32).reshape(-1, 8)
        n = rays8.shape[0]
        hits = np.zeros((n, 5), dtype=np.uint32)
        inner = np.zeros(n, dtype=np.uint32)
        self.lib.nnbvh_ref_intersect_instanced.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                                           ctypes.c_void_p, ctypes.c_void_p]
        self.lib.nnbvh_ref_intersect_instanced.restype = None
        self.lib.nnbvh_ref_intersect_instanced(handle, _p(rays8), n, _p(hits), _p(inner))
        return hits, inner

    def sphere(self, ray8, sphere28):
        ray8 = np.ascontiguousarray(ray8, dtype=np.float32)
        sphere28 = np.ascontiguousarray(sphere28, dtype=np.float32)
        out = np.zeros(4, dtype=np.float32)
        self.lib.nnbvh_ref_sphere.argtypes = [ctypes.c_void_p] * 3
        hit = self.lib.nnbvh_ref_sphere(_p(ray8), _p(sphere28), _p(out))
        return bool(hit), out

    def sphere_bounds(self, sphere28):
        sphere28 = np.ascontiguousarray(sphere28, dtype=np.float32)
        out = np.zeros(6, dtype=np.float32)
        self.lib.nnbvh_ref_sphere_bounds.argtypes = [ctypes.c_void_p] * 2
        self.lib.nnbvh_ref_sphere_bounds.restype = None
        self.lib.nnbvh_ref_sphere_bounds(_p(sphere28), _p(out))
        return out

    def sphere_interaction(self, sphere28, reverse: bool, rays8):
        """Sphere::Intersect per ray: (hit, tHit, 42 words = the 39 of triangle_interaction + wo)."""
        sphere28 = np.ascontiguousarray(sphere28, dtype=np.float32)
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32)
        n = rays8.shape[0]
        hit = np.zeros(n, dtype=np.int32)
        t_hit = np.zeros(n, dtype=np.float32)
        out = np.zeros((n, 42), dtype=np.float32)
        f = self.lib.nnbvh_ref_sphere_interaction
        f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int64] + [ctypes.c_void_p] * 3
        f(_p(sphere28), int(bool(reverse)), _p(rays8), n, _p(hit), _p(t_hit), _p(out))
        return hit.astype(bool), t_hit, out

    def intersect_surface(self, scene, rays8):
        """BVHAggregate::Intersect's SurfaceInteraction per ray: (hit, tHit, 42 words) on any scene handle."""
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32)
        n = rays8.shape[0]
        hit = np.zeros(n, dtype=np.int32)
        t_hit = np.zeros(n, dtype=np.float32)
        out = np.zeros((n, 42), dtype=np.float32)
        f = self.lib.nnbvh_ref_intersect_surface
        f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64] + [ctypes.c_void_p] * 3
        f.restype = None
        f(scene, _p(rays8), n, _p(hit), _p(t_hit), _p(out))
        return hit.astype(bool), t_hit, out

    # ---- the reference's own wavefront queues (oracle/ref_queues.cpp) ----
    def ray_queue(self, rays8, max_queue_size=None):
        """A real pbrt::RayQueue filled by PushCameraRay -> (handle, [7 plane addresses], size)."""
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32).reshape(-1, 8)
        L = self.lib
        L.nnbvh_ref_rayqueue_create.restype = ctypes.c_void_p
        L.nnbvh_ref_rayqueue_create.argtypes = [ctypes.c_int]
        L.nnbvh_ref_rayqueue_push.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        L.nnbvh_ref_rayqueue_planes.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        q = ctypes.c_void_p(L.nnbvh_ref_rayqueue_create(max_queue_size or len(rays8)))
        L.nnbvh_ref_rayqueue_push(q, _p(rays8), len(rays8))
        planes = (ctypes.c_void_p * 7)()
        size = L.nnbvh_ref_rayqueue_planes(q, planes)
        return q, [int(a) for a in planes], size

    def ray_queue_get(self, q, i):
        out = np.zeros(8, dtype=np.float32)
        self.lib.nnbvh_ref_rayqueue_get.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        self.lib.nnbvh_ref_rayqueue_get.restype = None
        self.lib.nnbvh_ref_rayqueue_get(q, i, _p(out))
        return out

    def ray_queue_destroy(self, q):
        self.lib.nnbvh_ref_rayqueue_destroy.argtypes = [ctypes.c_void_p]
        self.lib.nnbvh_ref_rayqueue_destroy.restype = None
        self.lib.nnbvh_ref_rayqueue_destroy(q)

    def shadow_queue(self, rays8, max_queue_size=None):
        """A real pbrt::ShadowRayQueue filled by Push -> (handle, [8 plane addresses], size)."""
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32).reshape(-1, 8)
        L = self.lib
        L.nnbvh_ref_shadowqueue_create.restype = ctypes.c_void_p
        L.nnbvh_ref_shadowqueue_create.argtypes = [ctypes.c_int]
        L.nnbvh_ref_shadowqueue_push.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        L.nnbvh_ref_shadowqueue_planes.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        q = ctypes.c_void_p(L.nnbvh_ref_shadowqueue_create(max_queue_size or len(rays8)))
        L.nnbvh_ref_shadowqueue_push(q, _p(rays8), len(rays8))
        planes = (ctypes.c_void_p * 8)()
        size = L.nnbvh_ref_shadowqueue_planes(q, planes)
        return q, [int(a) for a in planes], size

    def shadow_queue_destroy(self, q):
        self.lib.nnbvh_ref_shadowqueue_destroy.argtypes = [ctypes.c_void_p]
        self.lib.nnbvh_ref_shadowqueue_destroy.restype = None
        self.lib.nnbvh_ref_shadowqueue_destroy(q)

    def interaction_wo(self, wo3):
        """What Interaction's constructor stores for a given wo (interaction.h:30-32)."""
        wo3 = np.ascontiguousarray(wo3, dtype=np.float32)
        out = np.zeros_like(wo3)
        f = self.lib.nnbvh_ref_interaction_wo
        f.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
        f.restype = None
        f(_p(wo3), wo3.shape[0], _p(out))
        return out

    def export(self, scene, n_tris: int):
        n = self.lib.nnbvh_ref_scene_num_nodes(scene)
        nodes = np.zeros((n, 8), dtype=np.uint32)
        order = np.zeros(n_tris, dtype=np.int32)
        self.lib.nnbvh_ref_scene_export(scene, _p(nodes), _p(order))
        return nodes, order

    def intersect(self, scene, rays8, parallel=True):
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32).reshape(-1, 8)
        hits = np.zeros((rays8.shape[0], 5), dtype=np.uint32)
        self.lib.nnbvh_ref_intersect(scene, _p(rays8), rays8.shape[0], _p(hits), int(parallel))
        return hits

    def intersect_only(self, scene, rays8):
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32).reshape(-1, 8)
        hit = np.zeros(rays8.shape[0], dtype=np.uint8)
        self.lib.nnbvh_ref_intersect_only(scene, _p(rays8), rays8.shape[0], _p(hit))
        return hit

    def intersect_p(self, scene, rays8, parallel=True):
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32).reshape(-1, 8)
        occ = np.zeros(rays8.shape[0], dtype=np.uint8)
        self.lib.nnbvh_ref_intersectP(scene, _p(rays8), rays8.shape[0], _p(occ), int(parallel))
        return occ

    def triangle(self, ray8, p9):
        ray8 = np.ascontiguousarray(ray8, dtype=np.float32)
        p9 = np.ascontiguousarray(p9, dtype=np.float32)
        out = np.zeros(4, dtype=np.float32)
        hit = self.lib.nnbvh_ref_triangle(_p(ray8), _p(p9), _p(out))
        return bool(hit), out

    def box(self, ray8, box6):
        ray8 = np.ascontiguousarray(ray8, dtype=np.float32)
        box6 = np.ascontiguousarray(box6, dtype=np.float32)
        return bool(self.lib.nnbvh_ref_box(_p(ray8), _p(box6)))

    def triangle_interaction(self, verts, indices, tri, b, wo_time, normals=None, tangents=None,
                             uvs=None, reverse_orientation=False):
        """Triangle::InteractionFromIntersection (shapes.h:884-1010) for queries (tri[i], b[i],
        wo_time[i]) on an indexed mesh; 39 words per query (see oracle/ref_driver.cpp)."""
        f32 = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float32)
        verts, normals, tangents, uvs = f32(verts), f32(normals), f32(tangents), f32(uvs)
        indices = np.ascontiguousarray(indices, dtype=np.int32)
        tri = np.ascontiguousarray(tri, dtype=np.int32)
        b, wo_time = f32(b), f32(wo_time)
        out = np.zeros((tri.shape[0], 39), dtype=np.uint32)
        fn = self.lib.nnbvh_ref_triangle_interaction
        fn.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int] + \
            [ctypes.c_void_p] * 3 + [ctypes.c_int] + [ctypes.c_void_p] * 3 + [ctypes.c_int64,
                                                                            ctypes.c_void_p]
        q = lambda a: None if a is None else _p(a)
        fn(_p(verts), verts.shape[0], _p(indices), indices.size // 3, q(normals), q(tangents),
           q(uvs), int(reverse_orientation), _p(tri), _p(b), _p(wo_time), tri.shape[0], _p(out))
        return out

    def stats(self):
        """(nodesVisited, triTests, triHits) of single-threaded calls since the last read."""
        a, b, c = ctypes.c_int64(), ctypes.c_int64(), ctypes.c_int64()
        self.lib.nnbvh_ref_stats(ctypes.byref(a), ctypes.byref(b), ctypes.byref(c))
        return a.value, b.value, c.value

    def loop_subdivide(self, levels: int, indices: np.ndarray, verts: np.ndarray):
        indices = np.ascontiguousarray(indices, dtype=np.int32).ravel()
        verts = np.ascontiguousarray(verts, dtype=np.float32).reshape(-1, 3)
        oi, ov = ctypes.c_void_p(), ctypes.c_void_p()
        ni, nv = ctypes.c_int(), ctypes.c_int()
        self.lib.nnbvh_ref_loop_subdivide(levels, _p(indices), indices.size, _p(verts),
                                          verts.shape[0], ctypes.byref(oi), ctypes.byref(ni),
                                          ctypes.byref(ov), ctypes.byref(nv))
        idx = np.ctypeslib.as_array(ctypes.cast(oi, ctypes.POINTER(ctypes.c_int32)),
                                    (ni.value,)).copy()
        v = np.ctypeslib.as_array(ctypes.cast(ov, ctypes.POINTER(ctypes.c_float)),
                                  (nv.value * 3,)).copy().reshape(-1, 3)
        self.lib.nnbvh_ref_free(oi)
        self.lib.nnbvh_ref_free(ov)
        return idx, v


def binding_available() -> bool:
    return os.path.exists(BINDING_SO)


class Binding(Reference):
    """oracle/_ref/libnnbvh_ref_binding.so: the reference driver plus integration/pbrt/gpubvh.h over
    libnnbvh.so -- pbrt::GPUBVHAggregate built on the very Primitive list a reference scene holds."""

    def __init__(self, threads: int = 1):
        super().__init__(threads, so=BINDING_SO)
        L = self.lib
        L.nnbvh_binding_last_error.restype = ctypes.c_char_p
        L.nnbvh_binding_create.restype = ctypes.c_void_p
        L.nnbvh_binding_create.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        L.nnbvh_binding_destroy.argtypes = [ctypes.c_void_p]
        L.nnbvh_binding_destroy.restype = None
        for f in (L.nnbvh_binding_intersect_surface, L.nnbvh_binding_intersect_surface_batch):
            f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64] + [ctypes.c_void_p] * 3
        L.nnbvh_binding_intersect_p.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
        L.nnbvh_binding_bounds.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.nnbvh_binding_bounds.restype = None

    def create(self, scene, max_prims: int = 4, split: str = "sah"):
        h = self.lib.nnbvh_binding_create(scene, max_prims, SPLIT[split])
        if not h:
            raise RuntimeError(self.lib.nnbvh_binding_last_error().decode())
        return ctypes.c_void_p(h)

    def destroy(self, b):
        self.lib.nnbvh_binding_destroy(b)

    def surface(self, b, rays8, batch: bool, parallel: bool = False):
        """parallel: the single-ray calls are made from pbrt's ParallelFor workers (set_threads > 1) at once."""
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32)
        n = rays8.shape[0]
        hit = np.zeros(n, dtype=np.int32)
        t_hit = np.zeros(n, dtype=np.float32)
        out = np.zeros((n, 42), dtype=np.float32)
        f = self.lib.nnbvh_binding_intersect_surface_batch if batch else self.lib.nnbvh_binding_intersect_surface
        if parallel:
            f = self.lib.nnbvh_binding_intersect_surface_parallel
            f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64] + [ctypes.c_void_p] * 3
        rc = f(b, _p(rays8), n, _p(hit), _p(t_hit), _p(out))
        if rc:
            raise RuntimeError(self.lib.nnbvh_binding_last_error().decode())
        return hit.astype(bool), t_hit, out

    def intersect_p_binding(self, b, rays8):
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32)
        occ = np.zeros(rays8.shape[0], dtype=np.uint8)
        rc = self.lib.nnbvh_binding_intersect_p(b, _p(rays8), rays8.shape[0], _p(occ))
        if rc:
            raise RuntimeError(self.lib.nnbvh_binding_last_error().decode())
        return occ

    def bounds(self, b):
        out = np.zeros(6, dtype=np.float32)
        self.lib.nnbvh_binding_bounds(b, _p(out))
        return out

    def wavefront_closest(self, scene, rays8, use_binding: bool, max_prims: int = 4, split: str = "sah"):
        """One IntersectClosest over a real RayQueue -> (kind per ray: 0 escaped / 1 continued, continued rays)."""
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32)
        n = rays8.shape[0]
        kind = np.full(n, -2, dtype=np.int8)
        nxt = np.zeros((n, 8), dtype=np.float32)
        f = self.lib.nnbvh_binding_wavefront_closest
        f.argtypes = [ctypes.c_void_p] + [ctypes.c_int] * 3 + [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        rc = f(scene, max_prims, SPLIT[split], int(use_binding), _p(rays8), n, _p(kind), _p(nxt))
        if rc:
            raise RuntimeError(self.lib.nnbvh_binding_last_error().decode())
        return kind, nxt

    def wavefront_shadow(self, scene, rays8, use_binding: bool, max_prims: int = 4, split: str = "sah"):
        """One IntersectShadow over a real ShadowRayQueue -> PixelSampleState::L[.][0] per ray."""
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32)
        n = rays8.shape[0]
        out = np.zeros(n, dtype=np.float32)
        f = self.lib.nnbvh_binding_wavefront_shadow
        f.argtypes = [ctypes.c_void_p] + [ctypes.c_int] * 3 + [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        rc = f(scene, max_prims, SPLIT[split], int(use_binding), _p(rays8), n, _p(out))
        if rc:
            raise RuntimeError(self.lib.nnbvh_binding_last_error().decode())
        return out


    def wavefront_shadow_tr(self, scene, rays8, use_binding: bool, max_prims: int = 4, split: str = "sah"):
        """One IntersectShadowTr over a real ShadowRayQueue -> (PixelSampleState::L[.][0] per ray, closest-hit
        queries per ray (reference leg only, else zeros), GPU batches issued (binding leg only))."""
        rays8 = np.ascontiguousarray(rays8, dtype=np.float32)
        n = rays8.shape[0]
        out = np.zeros(n, dtype=np.float32)
        seg = np.zeros(n, dtype=np.int32)
        batches = ctypes.c_int(0)
        f = self.lib.nnbvh_binding_wavefront_shadow_tr
        f.argtypes = [ctypes.c_void_p] + [ctypes.c_int] * 3 + [ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 3
        rc = f(scene, max_prims, SPLIT[split], int(use_binding), _p(rays8), n, _p(out), _p(seg), ctypes.byref(batches))
        if rc:
            raise RuntimeError(self.lib.nnbvh_binding_last_error().decode())
        return out, seg, batches.value

    def wavefront_one_random(self, scene, seg6, use_binding: bool, max_prims: int = 4, split: str = "sah"):
        """One IntersectOneRandom over a real SubsurfaceScatterQueue of probes p0 -> p1 ((n,6)) ->
        (reservoirPDF per probe, chosen interaction's p and n (n,6), GPU batches issued (binding leg only))."""
        seg6 = np.ascontiguousarray(seg6, dtype=np.float32)
        n = seg6.shape[0]
        pdf = np.zeros(n, dtype=np.float32)
        pn = np.zeros((n, 6), dtype=np.float32)
        batches = ctypes.c_int(0)
        f = self.lib.nnbvh_binding_wavefront_one_random
        f.argtypes = [ctypes.c_void_p] + [ctypes.c_int] * 3 + [ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 3
        rc = f(scene, max_prims, SPLIT[split], int(use_binding), _p(seg6), n, _p(pdf), _p(pn), ctypes.byref(batches))
        if rc:
            raise RuntimeError(self.lib.nnbvh_binding_last_error().decode())
        return pdf, pn, batches.value
