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

#include "exact_math.cuh"
#include "internal.h"
#include "kernels.cuh"
#include "sphere.cuh"

namespace nnbvh {

__device__ __forceinline__ float HashFloatOD(float ox, float oy, float oz, float dx, float dy, float dz) {
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    const int r = 47;
    // the six floats copied back to back, read as three little-endian 64-bit words
    const uint64_t k3[3] = {(uint64_t)__float_as_uint(ox) | ((uint64_t)__float_as_uint(oy) << 32),
                            (uint64_t)__float_as_uint(oz) | ((uint64_t)__float_as_uint(dx) << 32),
                            (uint64_t)__float_as_uint(dy) | ((uint64_t)__float_as_uint(dz) << 32)};
    uint64_t h = 0ull ^ (24ull * m);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        uint64_t k = k3[i];
        k *= m;
        k ^= k >> r;
        k *= m;
        h ^= k;
        h *= m;
    }
    h ^= h >> r;
    h *= m;
    h ^= h >> r;
    return Mul(__uint2float_rn((uint32_t)h), 0x1p-32f);
}

// Origin of si->intr.SpawnRay(w) for the hit (b0, b1, b2) on triangle (p0, p1, p2).  The orientation of n
// does not matter: Abs(n) scales the offset and Dot(w, n) < 0 flips it back.
__device__ __noinline__ void SpawnOrigin(const float *p0, const float *p1, const float *p2, float b0, float b1,
                                         float b2, float wx, float wy, float wz, float *po) {
    float dp02[3], dp12[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        dp02[k] = Sub(p0[k], p2[k]);
        dp12[k] = Sub(p1[k], p2[k]);
    }
    float n[3] = {DifferenceOfProducts(dp02[1], dp12[2], dp02[2], dp12[1]),
                  DifferenceOfProducts(dp02[2], dp12[0], dp02[0], dp12[2]),
                  DifferenceOfProducts(dp02[0], dp12[1], dp02[1], dp12[0])};
    const float len = __fsqrt_rn(Add(Add(Mul(n[0], n[0]), Mul(n[1], n[1])), Mul(n[2], n[2])));
    float mid[3], err[3];
    const float g7 = GammaF(7);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        n[k] = Div(n[k], len);
        const float a0 = Mul(b0, p0[k]), a1 = Mul(b1, p1[k]), a2 = Mul(b2, p2[k]);
        const float pHit = Add(Add(a0, a1), a2);
        const float pErr = Mul(g7, Add(Add(fabsf(a0), fabsf(a1)), fabsf(a2)));
        const Ival pi = FromValueAndError(pHit, pErr);
        mid[k] = Mid(pi);                        // Point3f(pi)
        err[k] = Div(Sub(pi.hi, pi.lo), 2.f);    // pi.Error()
    }
    // SumOfProducts, util/math.h:577-583
    auto sop = [](float a, float b, float c, float d) {
        const float cd = Mul(c, d);
        const float sum = __fmaf_rn(a, b, cd);
        const float error = __fmaf_rn(c, d, -cd);
        return Add(sum, error);
    };
    // Dot(Normal3f, Vector3f) = FMA(n.x, v.x, SumOfProducts(n.y, v.y, n.z, v.z)), vecmath.h:1057-1069
    const float d = __fmaf_rn(fabsf(n[0]), err[0], sop(fabsf(n[1]), err[1], fabsf(n[2]), err[2]));
    float off[3] = {Mul(d, n[0]), Mul(d, n[1]), Mul(d, n[2])};
    if (__fmaf_rn(wx, n[0], sop(wy, n[1], wz, n[2])) < 0) {
        off[0] = -off[0], off[1] = -off[1], off[2] = -off[2];
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float v = Add(mid[k], off[k]);
        if (off[k] > 0) v = NextFloatUp(v);
        else if (off[k] < 0) v = NextFloatDown(v);
        po[k] = v;
    }
}

// GeometricPrimitive::Intersect over a triangle with constant alpha a < 1.  rc: the ray as the caller has it.
// The recursion of the reference is a loop here; tHit sums from the innermost re-trace outwards.
__device__ __noinline__ bool AlphaTriangleTest(const RayConst &rc, float tMax, const float *p0, const float *p1,
                                               const float *p2, float a, float *tOut, float *b0, float *b1,
                                               float *b2) {
    constexpr int kMaxRetrace = 64;  // the checker's bound; a flat triangle is not hit twice in practice
    float ts[kMaxRetrace + 2];
    RayConst cur = rc;
    float tm = tMax;
    int depth = 0;
    while (true) {
        float t, c0, c1, c2;
        if (!TriangleTest(cur, tm, p0[0], p0[1], p0[2], p1[0], p1[1], p1[2], p2[0], p2[1], p2[2], &t, &c0, &c1, &c2))
            return false;
        const float u = (a <= 0) ? 1.f : HashFloatOD(cur.ox, cur.oy, cur.oz, cur.dx, cur.dy, cur.dz);
        if (!(u > a)) {
            float total = t;
            for (int k = depth - 1; k >= 0; --k) total = Add(total, ts[k]);  // siNext->tHit += si->tHit
            *tOut = total, *b0 = c0, *b1 = c1, *b2 = c2;
            return true;
        }
        if (depth > kMaxRetrace) return false;
        ts[depth++] = t;
        float o[3];
        SpawnOrigin(p0, p1, p2, c0, c1, c2, cur.dx, cur.dy, cur.dz, o);
        cur = MakeRayConst(o[0], o[1], o[2], cur.dx, cur.dy, cur.dz);
        tm = Sub(tm, t);
    }
}

// BVHAggregate::Intersect / IntersectP over 32-byte reference nodes (as TraverseRef, kernels.cuh), with the
// alpha test at the triangles that carry one.  alpha: one float per ordered primitive, >= 1 = none.
template <bool ANY>
__global__ void __launch_bounds__(128)
AlphaKernel(SceneDev sc, const float *__restrict__ alpha, RayInput in, uint64_t n, nnbvh_hit *__restrict__ hits,
            uint8_t *__restrict__ occluded) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
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
