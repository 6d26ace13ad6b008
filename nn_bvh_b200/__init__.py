"""nn_bvh_b200 -- B200-native batch ray-query engine behind pbrt-v4's BVHAggregate.

Package layout (only what the hot path needs):
  csrc/         CUDA kernels + the C ABI (built into libnnbvh.so; include/nnbvh.h)
  capi.py       ctypes declarations of the C ABI
  aggregate.py  GPUBVHAggregate: host-side mirror of BVHAggregate (+ batch calls)
  scenes.py     triangle soups for the BASELINE.json configs
  rays.py       deterministic ray batches for those configs
"""
from .capi import HIT_DTYPE, MISS, NODE_DTYPE, RAY_DTYPE, NnbvhError  # noqa: F401
from .aggregate import (GPUBVHAggregate, build_bvh, build_hlbvh_device, make_rays,  # noqa: F401
                        tree_stack_depth, make_spheres, sphere_bounds, triangle_bounds,  # noqa: F401
                        build_bvh_bounds, SPHERE_DTYPE, INSTANCE_DTYPE, make_instances,  # noqa: F401
                        instance_bounds)  # noqa: F401
