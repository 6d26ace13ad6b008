// nn_bvh_b200/csrc/alpha.cuh -- alpha-tested triangle primitives, one thread per ray.
//
// GeometricPrimitive::Intersect with an alpha texture (/root/reference/src/pbrt/cpu/primitive.cpp:50-84):
//     si = shape.Intersect(r, tMax);  a = alpha.Evaluate(*si);
//     if (a < 1) { u = (a <= 0) ? 1 : HashFloat(r.o, r.d);
//                  if (u > a) { rNext = si->intr.SpawnRay(r.d);  siNext = Intersect(rNext, tMax - si->tHit);
//                               if (siNext) siNext->tHit += si->tHit;  return siNext; } }
// and IntersectP = Intersect(...).has_value().  Constant alpha values only (FloatConstantTexture); image and
// procedural textures are outside this path.  Restated here operation for operation:
//   MurmurHash64A / Hash / HashFloat          util/hash.h:19-112   (24 bytes: Point3f o, Vector3f d; seed 0)
//   pHit, pError, n of the triangle's hit     shapes.h:924-941
//   Point3fi(pHit, pError), Midpoint, Error   util/math.h:829-853, util/vecmath.h:757
//   OffsetRayOrigin                           ray.h:75-92
// tests/test_alpha.py pins a CPU restatement of the same arithmetic to the reference bit for bit.
//
// Kept apart from kernels.cuh so that the triangle / sphere / instance kernels compile exactly as before.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "alpha_math.cuh"
#include "internal.h"
#include "kernels.cuh"

namespace nnbvh {

// BVHAggregate::Intersect / IntersectP over 32-byte reference nodes (as TraverseRef, kernels.cuh), with the
// alpha test at the triangles that carry one.  alpha: one float per ordered primitive, >= 1 = none.
template <bool ANY>
__global__ void __launch_bounds__(128)
AlphaKernel(SceneDev sc, const float *__restrict__ alpha, RayInput in, uint64_t n, nnbvh_hit *__restrict__ hits,
            uint8_t *__restrict__ occluded) {
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint64_t i = OrderedRay(in, j);
    const RayRegs rr = LoadRay(in, i);
    const RayConst rc = MakeRayConst(rr.ox, rr.oy, rr.oz, rr.dx, rr.dy, rr.dz);
    float tMax = rr.tMax;
    HitRegs hit{NNBVH_MISS, 0.f, 0.f, 0.f, 0.f};
    bool any = false;
    if (sc.nNodes) {
        int stack[kMaxStack];
        int sp = 0, cur = 0;
        const int neg[3] = {rc.negx, rc.negy, rc.negz};
        while (true) {
            const float4 *np = reinterpret_cast<const float4 *>(sc.nodes + cur);
            const float4 na = __ldg(np), nb = __ldg(np + 1);
            const float box[6] = {na.x, na.y, na.z, na.w, nb.x, nb.y};
            const int offset = __float_as_int(nb.z);
            const uint32_t meta = __float_as_uint(nb.w);
            const uint32_t nPrims = meta & 0xffffu, axis = (meta >> 16) & 0xffu;
            if (SlabTest(rc, box, tMax)) {
                if (nPrims > 0) {
                    for (uint32_t k = 0; k < nPrims; ++k) {
                        const uint32_t prim = (uint32_t)offset + k;
                        const float4 *tp = sc.tris + 3 * (size_t)prim;
                        const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
                        if (__float_as_uint(v0.w) & kTriDegenerate) continue;
                        const float a = __ldg(alpha + prim);
                        const float p0[3] = {v0.x, v0.y, v0.z}, p1[3] = {v1.x, v1.y, v1.z}, p2[3] = {v2.x, v2.y, v2.z};
                        float t, b0, b1, b2;
                        const bool got = a < 1 ? AlphaTriangleTest(rc, tMax, p0, p1, p2, a, &t, &b0, &b1, &b2)
                                               : TriangleTest(rc, tMax, v0.x, v0.y, v0.z, v1.x, v1.y, v1.z, v2.x, v2.y,
                                                              v2.z, &t, &b0, &b1, &b2);
                        if (got) {
                            if (ANY) {
                                any = true;
                                goto done;
                            }
                            hit.prim = prim;
                            hit.inner = NNBVH_MISS;
                            hit.t = t;
                            hit.b0 = b0;
                            hit.b1 = b1;
                            hit.b2 = b2;
                            tMax = t;
                        }
                    }
                    if (sp == 0) break;
                    cur = stack[--sp];
                } else if (neg[axis]) {
                    stack[sp++] = cur + 1;
                    cur = offset;
                } else {
                    stack[sp++] = offset;
                    cur = cur + 1;
                }
            } else {
                if (sp == 0) break;
                cur = stack[--sp];
            }
        }
    }
done:
    if (ANY) occluded[i] = any ? 1 : 0;
    else StoreHit(hits, i, hit);
}

}  // namespace nnbvh
