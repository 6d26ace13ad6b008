// nn_bvh_b200/csrc/alpha_math.cuh -- the arithmetic of the alpha test (see alpha.cuh): HashFloat(o, d), the
// origin of si->intr.SpawnRay(d), and GeometricPrimitive::Intersect's re-trace over one triangle.
// Device functions only, no kernel: tests/cpp/device_twin.cpp compiles this file (with exact_math.cuh,
// sphere.cuh, atan2_host.cuh) for the HOST, with IEEE-exact stand-ins for the intrinsics, and the CPU suite
// compares it with the checker.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "exact_math.cuh"
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

}  // namespace nnbvh
