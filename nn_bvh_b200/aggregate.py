"""nn_bvh_b200/aggregate.py -- host-side mirror of the reference's BVHAggregate.

Same surface as `class BVHAggregate` (/root/reference/src/pbrt/cpu/aggregates.h:28-70):
constructor (primitives, maxPrimsInNode, splitMethod), `Create(prims, parameters)`,
`Bounds()`, `Intersect(ray, tMax)`, `IntersectP(ray, tMax)` -- plus the batch
forms the reference only has in `CPUAggregate::IntersectClosest/IntersectShadow`
(wavefront/aggregate.cpp:34-68).  Every query runs in libnnbvh.so on the GPU;
this file holds no intersection arithmetic.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np

from . import capi
from .capi import HIT_DTYPE, MISS, NODE_DTYPE, RAY_DTYPE


def build_bvh(tris9: np.ndarray, max_prims_in_node: int = 4, split_method: str = "sah",
              threads: int = 0):
    """BVHAggregate ctor + flattenBVH on the host (libnnbvh nnbvh_build_bvh).

    tris9: (T,9) float32, original order.  Returns (nodes: NODE_DTYPE[M], order: uint32[T])
    where order[i] is the original index of ordered primitive i."""
    tris9 = np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9)
    T = tris9.shape[0]
    nodes = np.zeros(max(2 * T - 1, 1), dtype=NODE_DTYPE)
    order = np.zeros(T, dtype=np.uint32)
    n = ctypes.c_uint32(0)
    capi.check(capi.lib().nnbvh_build_bvh(capi.ptr(tris9), T, max_prims_in_node,
                                          capi.SPLIT[split_method], threads, capi.ptr(nodes),
                                          ctypes.byref(n), capi.ptr(order)))
    return nodes[:n.value].copy(), order


SPHERE_DTYPE = np.dtype([("renderFromObject", "<f4", (12,)), ("objectFromRender", "<f4", (12,)),
                         ("radius", "<f4"), ("zMin", "<f4"), ("zMax", "<f4"), ("phiMax", "<f4")])
assert SPHERE_DTYPE.itemsize == 112   # nnbvh_sphere


def make_spheres(render_from_object, radius, z_min=None, z_max=None, phi_max=360.0) -> np.ndarray:
    """nnbvh_sphere records from (S,4,4) or (S,3,4) renderFromObject matrices and radii: what
    `Shape "sphere"` under a CTM hands to Sphere::Sphere (shapes.h:117-128).  objectFromRender is
    the float32 rounding of the float64 inverse (pbrt computes its own Inverse(m) in float; a caller
    that needs pbrt's exact matrices fills the records itself)."""
    m = np.asarray(render_from_object, dtype=np.float64).reshape(-1, *np.shape(render_from_object)[-2:])
    full = np.tile(np.eye(4), (m.shape[0], 1, 1))
    full[:, :m.shape[1], :] = m
    radius = np.broadcast_to(np.asarray(radius, dtype=np.float32), (m.shape[0],))
    s = np.zeros(m.shape[0], dtype=SPHERE_DTYPE)
    s["renderFromObject"] = full[:, :3, :].reshape(-1, 12)
    s["objectFromRender"] = np.linalg.inv(full)[:, :3, :].reshape(-1, 12)
    s["radius"] = radius
    s["zMin"] = -radius if z_min is None else z_min
    s["zMax"] = radius if z_max is None else z_max
    s["phiMax"] = phi_max
    return s


INSTANCE_DTYPE = np.dtype([("renderFromPrimitive", "<f4", (12,)), ("primitiveFromRender", "<f4", (12,)),
                           ("object", "<u4"), ("reserved", "<u4", (3,))])
assert INSTANCE_DTYPE.itemsize == 112   # nnbvh_instance


class _Object(ctypes.Structure):          # nnbvh_object
    _fields_ = [("nodes", ctypes.c_void_p), ("nNodes", ctypes.c_uint32),
                ("orderedTriVerts9", ctypes.c_void_p), ("nPrims", ctypes.c_uint32)]


def make_instances(render_from_primitive, objects, primitive_from_render=None) -> np.ndarray:
    """nnbvh_instance records from (K,4,4) / (K,3,4) matrices and the object index of each
    (TransformedPrimitive(object aggregate, renderFromInstance), scene.cpp:1541-1571).
    primitive_from_render: the inverse matrices as the caller's Transform caches them; by default
    the float32 rounding of the float64 inverse."""
    m = np.asarray(render_from_primitive, dtype=np.float64)
    m = m.reshape(-1, *m.shape[-2:])
    full = np.tile(np.eye(4), (m.shape[0], 1, 1))
    full[:, :m.shape[1], :] = m
    out = np.zeros(m.shape[0], dtype=INSTANCE_DTYPE)
    out["renderFromPrimitive"] = full[:, :3, :].reshape(-1, 12)
    if primitive_from_render is not None:
        inv = np.asarray(primitive_from_render, dtype=np.float32).reshape(m.shape[0], -1)[:, :12]
    else:
        inv = np.linalg.inv(full)[:, :3, :].reshape(-1, 12)
    out["primitiveFromRender"] = inv
    out["object"] = np.asarray(objects, dtype=np.uint32)
    return out


def instance_bounds(instances: np.ndarray, object_bounds6: np.ndarray) -> np.ndarray:
    """TransformedPrimitive::Bounds() (cpu/primitive.h:94): (K,6); object_bounds6[k] = root box of
    instance k's object."""
    instances = np.ascontiguousarray(instances).view(INSTANCE_DTYPE).reshape(-1)
    ob = np.ascontiguousarray(object_bounds6, dtype=np.float32).reshape(-1, 6)
    out = np.zeros((len(instances), 6), dtype=np.float32)
    for k in range(len(instances)):
        capi.check(capi.lib().nnbvh_instance_bounds(capi.ptr(instances[k:k + 1]), capi.ptr(ob[k]), capi.ptr(out[k])))
    return out


def sphere_bounds(spheres: np.ndarray) -> np.ndarray:
    """Sphere::Bounds() (shapes.cpp:37-40) for each record: (S,6) float32."""
    spheres = np.ascontiguousarray(spheres).view(SPHERE_DTYPE).reshape(-1)
    out = np.zeros((len(spheres), 6), dtype=np.float32)
    for i in range(len(spheres)):
        capi.check(capi.lib().nnbvh_sphere_bounds(capi.ptr(spheres[i:i + 1]), capi.ptr(out[i])))
    return out


def triangle_bounds(tris9: np.ndarray) -> np.ndarray:
    """Triangle::Bounds() (shapes.cpp:294-301): (T,6) float32."""
    v = np.asarray(tris9, dtype=np.float32).reshape(-1, 3, 3)
    return np.concatenate([v.min(axis=1), v.max(axis=1)], axis=1)


def build_bvh_bounds(bounds6: np.ndarray, max_prims_in_node: int = 4, split_method: str = "sah",
                     threads: int = 0):
    """build_bvh over arbitrary primitives given their Bounds() (libnnbvh nnbvh_build_bvh_bounds)."""
    bounds6 = np.ascontiguousarray(bounds6, dtype=np.float32).reshape(-1, 6)
    P = bounds6.shape[0]
    nodes = np.zeros(max(2 * P - 1, 1), dtype=NODE_DTYPE)
    order = np.zeros(P, dtype=np.uint32)
    n = ctypes.c_uint32(0)
    capi.check(capi.lib().nnbvh_build_bvh_bounds(capi.ptr(bounds6), P, max_prims_in_node,
                                                 capi.SPLIT[split_method], threads, capi.ptr(nodes),
                                                 ctypes.byref(n), capi.ptr(order)))
    return nodes[:n.value].copy(), order


def build_hlbvh_device(tris9: np.ndarray, max_prims_in_node: int = 4, *, engine=None,
                       device: int = -1):
    """build_bvh(..., "hlbvh") with Morton codes, the radix sort and treelet emission on the GPU
    (libnnbvh nnbvh_build_hlbvh_device); byte-identical to the host builder.  `engine`: an engine
    handle to run on (a temporary one is created otherwise).  Returns (nodes, order, device_ms)."""
    L = capi.lib()
    tris9 = np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9)
    T = tris9.shape[0]
    own = engine is None
    if own:
        engine = ctypes.c_void_p()
        opts = capi.Options(device=device, kernel=capi.KERNEL_AUTO, smem_top_nodes=-1)
        capi.check(L.nnbvh_create(ctypes.byref(opts), ctypes.byref(engine)))
    try:
        nodes = np.zeros(max(2 * T - 1, 1), dtype=NODE_DTYPE)
        order = np.zeros(T, dtype=np.uint32)
        n = ctypes.c_uint32(0)
        capi.check(L.nnbvh_build_hlbvh_device(engine, capi.ptr(tris9), T, max_prims_in_node,
                                              capi.ptr(nodes), ctypes.byref(n), capi.ptr(order)))
        ms = ctypes.c_float(0)
        capi.check(L.nnbvh_last_kernel_ms(engine, ctypes.byref(ms)))
    finally:
        if own:
            L.nnbvh_destroy(engine)
    return nodes[:n.value].copy(), order, ms.value


def tree_stack_depth(nodes: np.ndarray) -> int:
    d = ctypes.c_uint32(0)
    capi.check(capi.lib().nnbvh_tree_stack_depth(capi.ptr(nodes), len(nodes), ctypes.byref(d)))
    return d.value


def make_rays(o, d, tmax=np.inf, time=0.0) -> np.ndarray:
    """Pack (N,3) origins/directions into the 32-byte nnbvh_ray layout."""
    o = np.asarray(o, dtype=np.float32).reshape(-1, 3)
    d = np.asarray(d, dtype=np.float32).reshape(-1, 3)
    r = np.zeros(o.shape[0], dtype=RAY_DTYPE)
    r["o"], r["d"], r["tMax"], r["time"] = o, d, tmax, time
    return r


class GPUBVHAggregate:
    """Drop-in batch backend for BVHAggregate::Intersect / IntersectP."""

    SplitMethod = ("sah", "hlbvh", "middle", "equal")

    def __init__(self, tris9: Optional[np.ndarray] = None, max_prims_in_node: int = 1,
                 split_method: str = "sah", *, nodes: Optional[np.ndarray] = None,
                 ordered_tris9: Optional[np.ndarray] = None, order: Optional[np.ndarray] = None,
                 device: int = -1, kernel: int = capi.KERNEL_AUTO, smem_top_nodes: int = -1,
                 spheres: Optional[np.ndarray] = None,
                 ordered_sphere_index: Optional[np.ndarray] = None,
                 objects: Optional[list] = None, instances: Optional[np.ndarray] = None,
                 alpha: Optional[np.ndarray] = None):
        """Either tris9 (a soup to build over, like the reference ctor) or a ready
        flattened tree: nodes + ordered_tris9 (e.g. an exported NN-split tree).

        spheres (SPHERE_DTYPE records, see make_spheres): sphere primitives appended to the
        primitive list after the triangles, i.e. caller primitive T + k is sphere k.  A hit on
        one is reported as {prim, t, b = pObj} (include/nnbvh.h nnbvh_upload_scene_mixed).  With
        a ready tree pass ordered_sphere_index (-1 = triangle) as well.

        objects (list of (T_j,9) soups in object space) + instances (INSTANCE_DTYPE records, see
        make_instances): each object gets its own tree; caller primitive T + k is instance k, a
        TransformedPrimitive over its object's aggregate (nnbvh_upload_scene_instanced).  Use
        IntersectBatchInstanced to get the primitive inside the object as well.

        alpha (one float per caller triangle; >= 1 = none): the constant alpha texture of a
        GeometricPrimitive (cpu/primitive.cpp:50-84): hits are ignored when HashFloat(o, d) > alpha
        (nnbvh_upload_scene_alpha).  Triangle-only scenes."""
        L = capi.lib()
        opts = capi.Options(device=device, kernel=kernel, smem_top_nodes=smem_top_nodes)
        h = ctypes.c_void_p()
        capi.check(L.nnbvh_create(ctypes.byref(opts), ctypes.byref(h)))
        self._h = h
        self.spheres = None
        self.ordered_sphere_index = None
        self.instances = None
        has_sph = spheres is not None and len(spheres) > 0
        has_inst = instances is not None and len(instances) > 0
        if alpha is not None and (has_sph or has_inst):
            raise ValueError("alpha-tested triangles cannot be combined with spheres or instances in one scene "
                             "(nnbvh_upload_scene_alpha takes triangle-only lists)")
        if has_sph and nodes is not None and ordered_sphere_index is None:
            raise ValueError("a ready tree with spheres needs ordered_sphere_index (-1 = triangle)")
        if instances is not None and len(instances):
            if spheres is not None and len(spheres):
                raise ValueError("a scene holds either spheres or instances")
            self.instances = np.ascontiguousarray(instances).view(INSTANCE_DTYPE).reshape(-1)
            self.objects = []
            for soup in objects:
                soup = np.ascontiguousarray(soup, dtype=np.float32).reshape(-1, 9)
                on, oo = build_bvh(soup, max_prims_in_node, split_method)
                self.objects.append(dict(nodes=on, order=oo, ordered_tris9=np.ascontiguousarray(soup[oo])))
            tris9 = (np.zeros((0, 9), dtype=np.float32) if tris9 is None
                     else np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9))
            T = tris9.shape[0]
            roots = np.stack([self.objects[j]["nodes"][0]["bounds"] for j in self.instances["object"]])
            bounds = np.concatenate([triangle_bounds(tris9), instance_bounds(self.instances, roots)])
            nodes, order = build_bvh_bounds(bounds, max_prims_in_node, split_method)
            self.ordered_instance_index = np.where(order >= T, order.astype(np.int64) - T, -1).astype(np.int32)
            ordered_tris9 = np.zeros((len(order), 9), dtype=np.float32)
            is_tri = order < T
            ordered_tris9[is_tri] = tris9[order[is_tri]]
        if spheres is not None and len(spheres):
            self.spheres = np.ascontiguousarray(spheres).view(SPHERE_DTYPE).reshape(-1)
            if nodes is None:
                tris9 = (np.zeros((0, 9), dtype=np.float32) if tris9 is None
                         else np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9))
                T = tris9.shape[0]
                bounds = np.concatenate([triangle_bounds(tris9), sphere_bounds(self.spheres)])
                nodes, order = build_bvh_bounds(bounds, max_prims_in_node, split_method)
                ordered_sphere_index = np.where(order >= T, order.astype(np.int64) - T, -1).astype(np.int32)
                ordered_tris9 = np.zeros((len(order), 9), dtype=np.float32)
                is_tri = order < T
                ordered_tris9[is_tri] = tris9[order[is_tri]]
            self.ordered_sphere_index = np.ascontiguousarray(ordered_sphere_index, dtype=np.int32)
        if nodes is None:
            if tris9 is None or len(tris9) == 0:
                # The reference CHECK(!primitives.empty())s; an empty aggregate here misses.
                nodes = np.zeros(0, dtype=NODE_DTYPE)
                ordered_tris9 = np.zeros((0, 9), dtype=np.float32)
                order = np.zeros(0, dtype=np.uint32)
            else:
                tris9 = np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9)
                if split_method == "hlbvh":   # the one builder with device stages
                    nodes, order, _ = build_hlbvh_device(tris9, max_prims_in_node, engine=h)
                else:
                    nodes, order = build_bvh(tris9, max_prims_in_node, split_method)
                ordered_tris9 = tris9[order]
        self.nodes = np.ascontiguousarray(nodes).view(NODE_DTYPE).reshape(-1)
        self.ordered_tris9 = np.ascontiguousarray(ordered_tris9, dtype=np.float32).reshape(-1, 9)
        self.order = order  # ordered prim -> caller's primitive id (None = identity)
        if self.instances is not None:
            objs = (_Object * len(self.objects))()
            for j, o in enumerate(self.objects):
                objs[j] = _Object(o["nodes"].ctypes.data, len(o["nodes"]), o["ordered_tris9"].ctypes.data,
                                  len(o["ordered_tris9"]))
            capi.check(L.nnbvh_upload_scene_instanced(h, capi.ptr(self.nodes), len(self.nodes),
                                                      capi.ptr(self.ordered_tris9), capi.ptr(self.ordered_instance_index),
                                                      len(self.ordered_tris9), capi.ptr(self.instances),
                                                      len(self.instances), ctypes.cast(objs, ctypes.c_void_p),
                                                      len(self.objects)))
        elif self.spheres is not None:
            capi.check(L.nnbvh_upload_scene_mixed(h, capi.ptr(self.nodes), len(self.nodes),
                                                  capi.ptr(self.ordered_tris9), capi.ptr(self.ordered_sphere_index),
                                                  len(self.ordered_tris9), capi.ptr(self.spheres), len(self.spheres)))
        elif alpha is not None:
            alpha = np.ascontiguousarray(alpha, dtype=np.float32).reshape(-1)
            if len(alpha) != len(self.ordered_tris9):
                raise ValueError("one alpha value per triangle")
            self.ordered_alpha = np.ascontiguousarray(alpha if self.order is None else alpha[self.order])
            capi.check(L.nnbvh_upload_scene_alpha(h, capi.ptr(self.nodes), len(self.nodes),
                                                  capi.ptr(self.ordered_tris9), capi.ptr(self.ordered_alpha),
                                                  len(self.ordered_tris9)))
        else:
            capi.check(L.nnbvh_upload_scene(h, capi.ptr(self.nodes), len(self.nodes),
                                            capi.ptr(self.ordered_tris9), len(self.ordered_tris9)))

    @classmethod
    def General(cls, tris9=None, alpha=None, spheres=None, objects=None, instances=None,
                max_prims_in_node: int = 4, split_method: str = "sah", *, tree: Optional[dict] = None,
                device: int = -1):
        """ONE primitive list holding any mix of kinds (nnbvh_upload_scene_general): caller primitive ids are
        triangles 0..T-1 (alpha[i] < 1: constant-alpha GeometricPrimitive), then spheres, then instances --
        the order pbrt's scene code appends them in.  objects: (T_j,9) soups in object space.
        tree: instead of building here, a ready flattened tree as a dict(nodes, ordered_tris9, sphere_of,
        instance_of, ordered_alpha, objects=[dict(nodes, ordered_tris9)], order) -- e.g. the reference's own."""
        self = object.__new__(cls)
        L = capi.lib()
        opts = capi.Options(device=device, kernel=capi.KERNEL_AUTO, smem_top_nodes=-1)
        h = ctypes.c_void_p()
        capi.check(L.nnbvh_create(ctypes.byref(opts), ctypes.byref(h)))
        self._h = h
        self.spheres = None if spheres is None else np.ascontiguousarray(spheres).view(SPHERE_DTYPE).reshape(-1)
        self.instances = None if instances is None else np.ascontiguousarray(instances).view(INSTANCE_DTYPE).reshape(-1)
        S = 0 if self.spheres is None else len(self.spheres)
        K = 0 if self.instances is None else len(self.instances)
        if tree is None:
            tris9 = (np.zeros((0, 9), dtype=np.float32) if tris9 is None
                     else np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9))
            T = len(tris9)
            objs = []
            for soup in (objects or []):
                soup = np.ascontiguousarray(soup, dtype=np.float32).reshape(-1, 9)
                on, oo = build_bvh(soup, max_prims_in_node, split_method)
                objs.append(dict(nodes=on, order=oo, ordered_tris9=np.ascontiguousarray(soup[oo])))
            parts = [triangle_bounds(tris9)]
            if S:
                parts.append(sphere_bounds(self.spheres))
            if K:
                roots = np.stack([objs[j]["nodes"][0]["bounds"] for j in self.instances["object"]])
                parts.append(instance_bounds(self.instances, roots))
            nodes, order = build_bvh_bounds(np.concatenate(parts), max_prims_in_node, split_method)
            o64 = order.astype(np.int64)
            ordered = np.zeros((len(order), 9), dtype=np.float32)
            is_tri = o64 < T
            ordered[is_tri] = tris9[o64[is_tri]]
            oa = np.ones(len(order), dtype=np.float32)
            if alpha is not None:
                oa[is_tri] = np.asarray(alpha, dtype=np.float32).reshape(-1)[o64[is_tri]]
            tree = dict(nodes=nodes, order=order, ordered_tris9=ordered, ordered_alpha=oa, objects=objs,
                        sphere_of=np.where((o64 >= T) & (o64 < T + S), o64 - T, -1).astype(np.int32),
                        instance_of=np.where(o64 >= T + S, o64 - T - S, -1).astype(np.int32))
        self.nodes = np.ascontiguousarray(tree["nodes"]).view(NODE_DTYPE).reshape(-1)
        self.ordered_tris9 = np.ascontiguousarray(tree["ordered_tris9"], dtype=np.float32).reshape(-1, 9)
        self.order = tree.get("order")
        self.ordered_sphere_index = np.ascontiguousarray(tree["sphere_of"], dtype=np.int32)
        self.ordered_instance_index = np.ascontiguousarray(tree["instance_of"], dtype=np.int32)
        self.ordered_alpha = np.ascontiguousarray(tree["ordered_alpha"], dtype=np.float32)
        self.objects = [dict(nodes=np.ascontiguousarray(o["nodes"]).view(NODE_DTYPE).reshape(-1),
                             ordered_tris9=np.ascontiguousarray(o["ordered_tris9"], dtype=np.float32).reshape(-1, 9))
                        for o in tree.get("objects", [])]
        objs_c = (_Object * max(1, len(self.objects)))()
        for j, o in enumerate(self.objects):
            objs_c[j] = _Object(o["nodes"].ctypes.data, len(o["nodes"]), o["ordered_tris9"].ctypes.data,
                                len(o["ordered_tris9"]))
        d = capi.SceneDesc(self.nodes.ctypes.data, len(self.nodes), self.ordered_tris9.ctypes.data,
                           len(self.ordered_tris9),
                           self.ordered_sphere_index.ctypes.data if S else None,
                           self.spheres.ctypes.data if S else None, S,
                           self.ordered_instance_index.ctypes.data if K else None,
                           self.instances.ctypes.data if K else None, K,
                           ctypes.cast(objs_c, ctypes.c_void_p) if K else None, len(self.objects) if K else 0,
                           self.ordered_alpha.ctypes.data)
        capi.check(L.nnbvh_upload_scene_general(h, ctypes.byref(d)))
        return self

    @classmethod
    def FromDevice(cls, nodes_dev, n_nodes: int, ordered_tris9_dev, n_prims: int, device: int = -1,
                   kernel: int = capi.KERNEL_AUTO):
        """An aggregate over a flattened tree and ordered triangles that already live in DEVICE memory
        (torch tensors or raw addresses): e.g. a scene one rank built and broadcast over NCCL, so that the
        other ranks neither rebuild nor re-read it.  nnbvh_upload_scene validates the tree and
        derives the wide layout as for host arrays.  `nodes` / `ordered_tris9` stay None on this object."""
        self = object.__new__(cls)
        L = capi.lib()
        opts = capi.Options(device=device, kernel=kernel, smem_top_nodes=-1)
        h = ctypes.c_void_p()
        capi.check(L.nnbvh_create(ctypes.byref(opts), ctypes.byref(h)))
        self._h = h
        self.spheres = self.ordered_sphere_index = self.instances = None
        self.nodes = self.ordered_tris9 = self.order = None
        capi.check(L.nnbvh_upload_scene(h, capi.ptr(nodes_dev), int(n_nodes), capi.ptr(ordered_tris9_dev),
                                        int(n_prims)))
        return self

    @classmethod
    def Create(cls, tris9, parameters: Optional[dict] = None, **kw):
        """BVHAggregate::Create (cpu/aggregates.cpp:725-744): "splitmethod" (default
        "sah"), "maxnodeprims" (default 4)."""
        p = parameters or {}
        name = p.get("splitmethod", "sah")
        if name not in cls.SplitMethod:
            name = "sah"  # the reference warns and falls back to sah (:737-740)
        return cls(tris9, int(p.get("maxnodeprims", 4)), name, **kw)

    def clone(self) -> "GPUBVHAggregate":
        """A second aggregate on the same device sharing this one's uploaded scene (nnbvh_clone), with its
        own stream and scratch: independent batches (closest-hit and shadow) can be in flight together."""
        other = object.__new__(GPUBVHAggregate)
        other.__dict__.update({k: v for k, v in self.__dict__.items() if k != "_h"})
        h = ctypes.c_void_p()
        capi.check(capi.lib().nnbvh_clone(self._h, ctypes.byref(h)))
        other._h = h
        return other

    def close(self):
        if getattr(self, "_h", None) and capi is not None and getattr(capi, "lib", None) is not None:
            capi.lib().nnbvh_destroy(self._h)
            self._h = None

    __del__ = close

    # ---- BVHAggregate surface -------------------------------------------------
    def Bounds(self):
        if self.nodes is None:
            raise RuntimeError("Bounds(): this aggregate was made from device arrays (FromDevice) and keeps no host tree")
        if len(self.nodes) == 0:
            raise RuntimeError("Bounds() of an empty aggregate (the reference CHECK(nodes)s)")
        return self.nodes[0]["bounds"].copy()

    def Intersect(self, ray, tMax=np.inf):
        """Single ray: (o, d) -> None or (prim, t, b0, b1, b2).  One-ray batch on the GPU."""
        h = self.IntersectBatch(make_rays(ray[0], ray[1], tMax))[0]
        if h["prim"] == MISS:
            return None
        return int(h["prim"]), float(h["t"]), *map(float, h["b"])

    def IntersectP(self, ray, tMax=np.inf) -> bool:
        return bool(self.IntersectPBatch(make_rays(ray[0], ray[1], tMax))[0])

    # ---- batch surface --------------------------------------------------------
    def IntersectBatch(self, rays, hits=None, n=None):
        """rays: RAY_DTYPE ndarray / (N,8) float32 / torch tensor (host or CUDA).
        Returns hits (HIT_DTYPE ndarray for numpy input; the tensor passed as `hits`
        otherwise -- a (N,5) int32/float32-viewable 20-byte-record buffer)."""
        n = self._count(rays, n)
        if hits is None:
            if not isinstance(rays, np.ndarray):
                raise ValueError("pass `hits` when rays is a tensor")
            hits = np.zeros(n, dtype=HIT_DTYPE)
        capi.check(capi.lib().nnbvh_intersect_closest(self._h, capi.ptr(rays), n, capi.ptr(hits)))
        return hits

    def IntersectBatchNoAlloc(self, rays, hits, n=None):
        """IntersectBatch for device tensors through nnbvh_intersect_closest_noalloc: at most
        capi.NOALLOC_MAX_RAYS rays, scratch in the library's static device array, no allocation."""
        n = self._count(rays, n)
        capi.check(capi.lib().nnbvh_intersect_closest_noalloc(self._h, capi.ptr(rays), n, capi.ptr(hits)))
        return hits

    def IntersectBatchInstanced(self, rays, hits=None, inner=None, n=None):
        """Instanced scenes: (hits, inner) with inner[i] = ordered primitive inside the hit object
        (MISS for a top-level triangle or a miss)."""
        n = self._count(rays, n)
        if hits is None:
            hits, inner = np.zeros(n, dtype=HIT_DTYPE), np.zeros(n, dtype=np.uint32)
        capi.check(capi.lib().nnbvh_intersect_closest_instanced(self._h, capi.ptr(rays), n, capi.ptr(hits),
                                                                capi.ptr(inner)))
        return hits, inner

    def IntersectPBatch(self, rays, occluded=None, n=None):
        n = self._count(rays, n)
        if occluded is None:
            if not isinstance(rays, np.ndarray):
                raise ValueError("pass `occluded` when rays is a tensor")
            occluded = np.zeros(n, dtype=np.uint8)
        capi.check(capi.lib().nnbvh_intersect_any(self._h, capi.ptr(rays), n, capi.ptr(occluded)))
        return occluded

    def IntersectBatchSoA(self, ox, oy, oz, dx, dy, dz, tMax=None, hits=None, any_hit=False):
        n = int(ox.shape[0])
        if hits is None:
            hits = np.zeros(n, dtype=np.uint8 if any_hit else HIT_DTYPE)
        f = capi.lib().nnbvh_intersect_any_soa if any_hit else capi.lib().nnbvh_intersect_closest_soa
        capi.check(f(self._h, *(capi.ptr(a) for a in (ox, oy, oz, dx, dy, dz, tMax)), n,
                     capi.ptr(hits)))
        return hits

    def IntersectBatchInstancedSoA(self, ox, oy, oz, dx, dy, dz, tMax=None, hits=None, inner=None):
        """IntersectBatchInstanced over the wavefront queues' planes."""
        n = int(ox.shape[0])
        if hits is None:
            hits, inner = np.zeros(n, dtype=HIT_DTYPE), np.zeros(n, dtype=np.uint32)
        capi.check(capi.lib().nnbvh_intersect_closest_instanced_soa(
            self._h, *(capi.ptr(a) for a in (ox, oy, oz, dx, dy, dz, tMax)), n, capi.ptr(hits), capi.ptr(inner)))
        return hits, inner

    def IntersectCounters(self, rays, any_hit=False):
        """Results plus the reference's per-ray statistics (bvhNodesVisited,
        nTriTests).  numpy in, numpy out."""
        n = self._count(rays, None)
        hits = np.zeros(n, dtype=HIT_DTYPE)
        occ = np.zeros(n, dtype=np.uint8)
        nv = np.zeros(n, dtype=np.uint32)
        tt = np.zeros(n, dtype=np.uint32)
        capi.check(capi.lib().nnbvh_intersect_counters(self._h, capi.ptr(rays), n, int(any_hit),
                                                       capi.ptr(hits), capi.ptr(occ),
                                                       capi.ptr(nv), capi.ptr(tt)))
        return (occ if any_hit else hits), nv, tt

    def set_stream(self, cuda_stream: int):
        capi.check(capi.lib().nnbvh_set_stream(self._h, ctypes.c_void_p(cuda_stream)))

    def last_kernel_ms(self) -> float:
        ms = ctypes.c_float(0)
        capi.check(capi.lib().nnbvh_last_kernel_ms(self._h, ctypes.byref(ms)))
        return ms.value

    def last_traversal_ms(self) -> float:
        """Device time of the traversal launches of the last device-pointer batch (no pre-pass)."""
        ms = ctypes.c_float(0)
        capi.check(capi.lib().nnbvh_last_traversal_ms(self._h, ctypes.byref(ms)))
        return ms.value

    def read_bandwidth_probe(self, nbytes: int, iters: int = 20) -> float:
        """GB/s of streaming an nbytes device buffer `iters` times with 256-bit loads (L2 if it fits)."""
        g = ctypes.c_float(0)
        capi.check(capi.lib().nnbvh_read_bandwidth_probe(self._h, nbytes, iters, ctypes.byref(g)))
        return g.value

    def step_stats(self) -> dict:
        """Lane-occupancy counters of the traversal kernel (engine created under NNBVH_STEP_STATS=1)."""
        out = (ctypes.c_uint64 * 8)()
        capi.check(capi.lib().nnbvh_debug_step_stats(self._h, out))
        names = ("iters_a", "lanes_node", "lanes_parked", "lanes_empty", "iters_b", "lanes_b_parked",
                 "lanes_b_testing", "refills")
        return dict(zip(names, (int(v) for v in out)))

    def launch_count(self) -> int:
        c = ctypes.c_uint64(0)
        capi.check(capi.lib().nnbvh_launch_count(self._h, ctypes.byref(c)))
        return c.value

    def test_triangles(self, rays, tris9):
        n = len(rays)
        tris9 = np.ascontiguousarray(tris9, dtype=np.float32).reshape(-1, 9)
        hit = np.zeros(n, dtype=np.uint8)
        out = np.zeros((n, 4), dtype=np.float32)
        capi.check(capi.lib().nnbvh_test_triangles(self._h, capi.ptr(rays), capi.ptr(tris9), n,
                                                   capi.ptr(hit), capi.ptr(out)))
        return hit, out

    def test_boxes(self, rays, boxes6):
        n = len(rays)
        boxes6 = np.ascontiguousarray(boxes6, dtype=np.float32).reshape(-1, 6)
        hit = np.zeros(n, dtype=np.uint8)
        capi.check(capi.lib().nnbvh_test_boxes(self._h, capi.ptr(rays), capi.ptr(boxes6), n,
                                               capi.ptr(hit)))
        return hit

    @staticmethod
    def _count(rays, n):
        if n is not None:
            return int(n)
        if isinstance(rays, np.ndarray) and rays.dtype == RAY_DTYPE:
            return rays.size
        return int(rays.shape[0])
