"""nn_bvh_b200/capi.py -- ctypes binding of libnnbvh.so (include/nnbvh.h).

The library is the product; this module only declares its signatures.  It fails
loudly when the shared object is missing (run `python -c "import
__graft_entry__ as g; g.build()"` or nn_bvh_b200/csrc/build.sh) -- there is no
Python or CPU fallback for any compute entry point.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NNBVH_LIB") or os.path.join(HERE, "libnnbvh.so")   # NNBVH_LIB: an A/B build (tools/)

MISS = 0xFFFFFFFF
NOALLOC_MAX_RAYS = 1 << 20   # NNBVH_NOALLOC_MAX_RAYS
KERNEL_AUTO, KERNEL_BASIC, KERNEL_WIDE, KERNEL_STEP = 0, 1, 2, 3
SPLIT = {"sah": 0, "hlbvh": 1, "middle": 2, "equal": 3}

NODE_DTYPE = np.dtype([("bounds", "<f4", (6,)), ("offset", "<i4"), ("nPrimitives", "<u2"),
                       ("axis", "u1"), ("pad", "u1")])
RAY_DTYPE = np.dtype([("o", "<f4", (3,)), ("tMax", "<f4"), ("d", "<f4", (3,)), ("time", "<f4")])
HIT_DTYPE = np.dtype([("prim", "<u4"), ("t", "<f4"), ("b", "<f4", (3,))])
assert NODE_DTYPE.itemsize == 32 and RAY_DTYPE.itemsize == 32 and HIT_DTYPE.itemsize == 20

# Every symbol include/nnbvh.h declares (tests check the .so exports all of them).
SYMBOLS = [
    "nnbvh_build_bvh", "nnbvh_tree_stack_depth", "nnbvh_create", "nnbvh_destroy",
    "nnbvh_set_stream", "nnbvh_upload_scene", "nnbvh_intersect_closest", "nnbvh_intersect_any",
    "nnbvh_intersect_closest_soa", "nnbvh_intersect_any_soa", "nnbvh_intersect_counters",
    "nnbvh_last_kernel_ms", "nnbvh_launch_count", "nnbvh_test_triangles", "nnbvh_test_boxes",
    "nnbvh_last_error", "nnbvh_version", "nnbvh_build_hlbvh_device", "nnbvh_debug_step_stats", "nnbvh_last_traversal_ms", "nnbvh_peer_alloc", "nnbvh_peer_open",
    "nnbvh_peer_close", "nnbvh_peer_free", "nnbvh_peer_copy", "nnbvh_read_bandwidth_probe",
    "nnbvh_pointer_kind", "nnbvh_copy_to_host", "nnbvh_build_bvh_bounds", "nnbvh_sphere_bounds",
    "nnbvh_upload_scene_mixed", "nnbvh_upload_scene_alpha", "nnbvh_test_spheres", "nnbvh_test_atan2", "nnbvh_instance_bounds",
    "nnbvh_upload_scene_instanced", "nnbvh_intersect_closest_instanced", "nnbvh_intersect_closest_instanced_soa", "nnbvh_clone",
    "nnbvh_intersect_closest_noalloc", "nnbvh_upload_scene_general",
]


class Options(ctypes.Structure):
    _fields_ = [("device", ctypes.c_int32), ("kernel", ctypes.c_int32),
                ("smem_top_nodes", ctypes.c_int32), ("reserved", ctypes.c_int32 * 5)]


class SceneDesc(ctypes.Structure):           # nnbvh_scene_desc
    _fields_ = [("nodes", ctypes.c_void_p), ("nNodes", ctypes.c_uint32),
                ("orderedTriVerts9", ctypes.c_void_p), ("nPrims", ctypes.c_uint32),
                ("orderedSphereIndex", ctypes.c_void_p), ("spheres", ctypes.c_void_p), ("nSpheres", ctypes.c_uint32),
                ("orderedInstanceIndex", ctypes.c_void_p), ("instances", ctypes.c_void_p),
                ("nInstances", ctypes.c_uint32), ("objects", ctypes.c_void_p), ("nObjects", ctypes.c_uint32),
                ("orderedAlpha", ctypes.c_void_p)]


class NnbvhError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libnnbvh error {code}: {msg}")
        self.code = code


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build the CUDA library first "
                          "(__graft_entry__.build()); nn_bvh_b200 has no fallback path")
    L = ctypes.CDLL(LIB_PATH)
    vp, u32, u64, i32 = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_int32
    L.nnbvh_build_bvh.argtypes = [vp, u32, i32, i32, i32, vp, vp, vp]
    L.nnbvh_tree_stack_depth.argtypes = [vp, u32, vp]
    L.nnbvh_build_hlbvh_device.argtypes = [vp, vp, u32, u32, vp, vp, vp]
    L.nnbvh_create.argtypes = [vp, vp]
    L.nnbvh_destroy.argtypes = [vp]
    L.nnbvh_destroy.restype = None
    L.nnbvh_set_stream.argtypes = [vp, vp]
    L.nnbvh_upload_scene.argtypes = [vp, vp, u32, vp, u32]
    L.nnbvh_intersect_closest.argtypes = [vp, vp, u64, vp]
    L.nnbvh_intersect_any.argtypes = [vp, vp, u64, vp]
    L.nnbvh_intersect_closest_noalloc.argtypes = [vp, vp, u64, vp]
    L.nnbvh_upload_scene_general.argtypes = [vp, vp]
    L.nnbvh_intersect_closest_soa.argtypes = [vp] * 8 + [u64, vp]
    L.nnbvh_intersect_any_soa.argtypes = [vp] * 8 + [u64, vp]
    L.nnbvh_intersect_counters.argtypes = [vp, vp, u64, i32, vp, vp, vp, vp]
    L.nnbvh_last_kernel_ms.argtypes = [vp, vp]
    L.nnbvh_launch_count.argtypes = [vp, vp]
    L.nnbvh_debug_step_stats.argtypes = [vp, vp]
    L.nnbvh_last_traversal_ms.argtypes = [vp, vp]
    L.nnbvh_peer_alloc.argtypes = [u64, vp, vp]
    L.nnbvh_peer_open.argtypes = [vp, vp]
    L.nnbvh_peer_close.argtypes = [vp]
    L.nnbvh_peer_free.argtypes = [vp]
    L.nnbvh_peer_copy.argtypes = [vp, vp, u64, vp]
    L.nnbvh_read_bandwidth_probe.argtypes = [vp, u64, u32, vp]
    L.nnbvh_pointer_kind.argtypes = [vp, vp]
    L.nnbvh_build_bvh_bounds.argtypes = [vp, u32, i32, i32, i32, vp, vp, vp]
    L.nnbvh_sphere_bounds.argtypes = [vp, vp]
    L.nnbvh_upload_scene_mixed.argtypes = [vp, vp, u32, vp, vp, u32, vp, u32]
    L.nnbvh_test_spheres.argtypes = [vp, vp, vp, u64, vp, vp]
    L.nnbvh_instance_bounds.argtypes = [vp, vp, vp]
    L.nnbvh_clone.argtypes = [vp, vp]
    L.nnbvh_upload_scene_instanced.argtypes = [vp, vp, u32, vp, vp, u32, vp, u32, vp, u32]
    L.nnbvh_intersect_closest_instanced.argtypes = [vp, vp, u64, vp, vp]
    L.nnbvh_test_atan2.argtypes = [vp, vp, vp, u64, vp]
    L.nnbvh_upload_scene_alpha.argtypes = [vp, vp, ctypes.c_uint32, vp, vp, ctypes.c_uint32]
    L.nnbvh_intersect_closest_instanced_soa.argtypes = [vp] * 8 + [u64, vp, vp]
    L.nnbvh_copy_to_host.argtypes = [vp, vp, u64]
    L.nnbvh_test_triangles.argtypes = [vp, vp, vp, u64, vp, vp]
    L.nnbvh_test_boxes.argtypes = [vp, vp, vp, u64, vp]
    L.nnbvh_last_error.restype = ctypes.c_char_p
    L.nnbvh_version.restype = ctypes.c_char_p
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != 0:
        raise NnbvhError(rc, lib().nnbvh_last_error().decode("utf-8", "replace"))


def ptr(a):
    """Raw address of a numpy array, a torch tensor (host or device), an int or None."""
    if a is None:
        return None
    if isinstance(a, int):
        return ctypes.c_void_p(a)
    if isinstance(a, np.ndarray):
        if not a.flags["C_CONTIGUOUS"]:
            raise ValueError("array must be C-contiguous")
        return ctypes.c_void_p(a.ctypes.data)
    if hasattr(a, "data_ptr"):  # torch.Tensor
        if not a.is_contiguous():
            raise ValueError("tensor must be contiguous")
        return ctypes.c_void_p(a.data_ptr())
    raise TypeError(type(a))
