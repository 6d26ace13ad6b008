// nn_bvh_b200/csrc/exact_math.cuh
//
// Device restatement of the reference's float arithmetic on the hot path, with
// every rounding pinned by intrinsics (__fmul_rn / __fadd_rn / __fsub_rn /
// __fdiv_rn are never contracted into FMAs, whatever -fmad says), so results are
// bit-identical to the reference built with -ffp-contract=off.  FMAs appear
// only where the reference writes FMA(): DifferenceOfProducts
// (util/math.h:569-575).  Paths relative to /root/reference/src/pbrt.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace nnbvh {

// util/float.h:43,195-197 -- gamma(n) = (n*eps)/(1-n*eps), eps = 2^-24, in float.
__host__ __device__ constexpr float GammaF(int n) {
    return (n * 5.9604644775390625e-08f) / (1 - n * 5.9604644775390625e-08f);
}
// `1 + 2 * gamma(3)` of util/vecmath.h:1583.
__host__ __device__ constexpr float SlabPad() { return 1 + 2 * GammaF(3); }

__device__ __forceinline__ float Mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float Add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float Sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float Div(float a, float b) { return __fdiv_rn(a, b); }

// util/math.h:569-575
__device__ __forceinline__ float DifferenceOfProducts(float a, float b, float c, float d) {
    float cd = Mul(c, d);
    float dop = __fmaf_rn(a, b, -cd);
    float err = __fmaf_rn(-c, d, cd);
    return Add(dop, err);
}

// std::max({a,b,c}) as used by MaxComponentValue (util/vecmath.h:447-451):
// keeps the FIRST largest, `if (largest < x) largest = x`.
__device__ __forceinline__ float Max3First(float a, float b, float c) {
    float m = a;
    if (m < b) m = b;
    if (m < c) m = c;
    return m;
}

// Per-ray constants: everything in Intersect/IntersectP/IntersectTriangle that
// depends on the ray alone, computed once instead of per node / per triangle.
struct RayConst {
    float ox, oy, oz;
    float dx, dy, dz;
    float ix, iy, iz;      // invDir, cpu/aggregates.cpp:534
    int negx, negy, negz;  // dirIsNeg, :535
    // shapes.cpp:186-201: permutation and shear
    int kx, ky, kz;
    float Sx, Sy, Sz;
};

__device__ __forceinline__ float Sel3(int k, float x, float y, float z) {
    return k == 0 ? x : (k == 1 ? y : z);
}

// c ? a : b as ONE select instruction (SASS FSEL).  Written in PTX for the device because, left to
// itself, the compiler turned the nine permutation selects of the triangle test into nine divergent
// branch regions (lanes of a warp hold rays with different kz): 20 % of the leaf phase's issue slots
// (profiles/r02c: 0x3ef0-0x4270 of StepKernel<0,0,5>).
__device__ __forceinline__ float SelF(bool c, float a, float b) {
#ifdef __CUDA_ARCH__
    float r;
    asm("{ .reg .pred p; setp.ne.u32 p, %3, 0; selp.f32 %0, %1, %2, p; }" : "=f"(r) : "f"(a), "f"(b), "r"((unsigned)c));
    return r;
#else
    return c ? a : b;
#endif
}

// Permute(v, kx, ky, kz) of shapes.cpp:186-196 with kx = (kz + 1) % 3, ky = (kx + 1) % 3:
// kz = 0 -> (a1, a2, a0); kz = 1 -> (a2, a0, a1); kz = 2 -> (a0, a1, a2).
__device__ __forceinline__ void Permute3(int kz, float a0, float a1, float a2, float *x, float *y, float *z) {
    const bool k0 = kz == 0, k1 = kz == 1;
    *x = SelF(k0, a1, SelF(k1, a2, a0));
    *y = SelF(k0, a2, SelF(k1, a0, a1));
    *z = SelF(k0, a0, SelF(k1, a1, a2));
}

__device__ __forceinline__ RayConst MakeRayConst(float ox, float oy, float oz, float dx,
                                                 float dy, float dz) {
    RayConst r;
    r.ox = ox; r.oy = oy; r.oz = oz;
    r.dx = dx; r.dy = dy; r.dz = dz;
    r.ix = Div(1.f, dx);
    r.iy = Div(1.f, dy);
    r.iz = Div(1.f, dz);
    r.negx = r.ix < 0;
    r.negy = r.iy < 0;
    r.negz = r.iz < 0;
    // MaxComponentIndex(Abs(d)), util/vecmath.h:453-456
    float ax = fabsf(dx), ay = fabsf(dy), az = fabsf(dz);
    int kz = (ax > ay) ? ((ax > az) ? 0 : 2) : ((ay > az) ? 1 : 2);
    int kx = kz + 1;
    if (kx == 3) kx = 0;
    int ky = kx + 1;
    if (ky == 3) ky = 0;
    r.kx = kx; r.ky = ky; r.kz = kz;
    float pdx = Sel3(kx, dx, dy, dz), pdy = Sel3(ky, dx, dy, dz), pdz = Sel3(kz, dx, dy, dz);
    r.Sx = Div(-pdx, pdz);
    r.Sy = Div(-pdy, pdz);
    r.Sz = Div(1.f, pdz);
    return r;
}

// Bounds3f::IntersectP(o, d, raytMax, invDir, dirIsNeg), util/vecmath.h:1573-1608,
// split in two: everything that does not involve raytMax (returns false on a
// definite miss, else the entry distance tMin), and the final `tMin < raytMax`.
// lo*/hi* are the corner components already selected by dirIsNeg:
// lo = bounds[dirIsNeg[k]], hi = bounds[1 - dirIsNeg[k]].
__device__ __forceinline__ bool SlabStatic(const RayConst &r, float lox, float hix, float loy,
                                           float hiy, float loz, float hiz, float *tEntry) {
    const float pad = SlabPad();
    float tMin = Mul(Sub(lox, r.ox), r.ix);
    float tMax = Mul(Sub(hix, r.ox), r.ix);
    float tyMin = Mul(Sub(loy, r.oy), r.iy);
    float tyMax = Mul(Sub(hiy, r.oy), r.iy);
    tMax = Mul(tMax, pad);
    tyMax = Mul(tyMax, pad);
    if (tMin > tyMax || tyMin > tMax) return false;
    if (tyMin > tMin) tMin = tyMin;
    if (tyMax < tMax) tMax = tyMax;
    float tzMin = Mul(Sub(loz, r.oz), r.iz);
    float tzMax = Mul(Sub(hiz, r.oz), r.iz);
    tzMax = Mul(tzMax, pad);
    if (tMin > tzMax || tzMin > tMax) return false;
    if (tzMin > tMin) tMin = tzMin;
    if (tzMax < tMax) tMax = tzMax;
    *tEntry = tMin;
    return tMax > 0;
}

// Box given as pMin/pMax (the 32-byte node form).
__device__ __forceinline__ bool SlabTest(const RayConst &r, const float b[6], float raytMax) {
    float t;
    bool ok = SlabStatic(r, r.negx ? b[3] : b[0], r.negx ? b[0] : b[3], r.negy ? b[4] : b[1],
                         r.negy ? b[1] : b[4], r.negz ? b[5] : b[2], r.negz ? b[2] : b[5], &t);
    return ok && (t < raytMax);
}

// Degenerate-triangle reject, shapes.cpp:176: LengthSquared(Cross(p2-p0, p1-p0)) == 0.
// Ray independent, so the engine evaluates it once per triangle at upload.
__device__ __forceinline__ bool TriangleDegenerate(float p0x, float p0y, float p0z, float p1x,
                                                   float p1y, float p1z, float p2x, float p2y,
                                                   float p2z) {
    float vx = Sub(p2x, p0x), vy = Sub(p2y, p0y), vz = Sub(p2z, p0z);
    float wx = Sub(p1x, p0x), wy = Sub(p1y, p0y), wz = Sub(p1z, p0z);
    float cx = DifferenceOfProducts(vy, wz, vz, wy);
    float cy = DifferenceOfProducts(vz, wx, vx, wz);
    float cz = DifferenceOfProducts(vx, wy, vy, wx);
    return Add(Add(Mul(cx, cx), Mul(cy, cy)), Mul(cz, cz)) == 0;
}

// IntersectTriangle, shapes.cpp:181-273 (after the degenerate reject).
__device__ __forceinline__ bool TriangleTest(const RayConst &r, float tMax, float p0x, float p0y,
                                             float p0z, float p1x, float p1y, float p1z,
                                             float p2x, float p2y, float p2z, float *tOut,
                                             float *b0Out, float *b1Out, float *b2Out) {
    // :181-183 translate
    float a0 = Sub(p0x, r.ox), a1 = Sub(p0y, r.oy), a2 = Sub(p0z, r.oz);
    float c0 = Sub(p1x, r.ox), c1 = Sub(p1y, r.oy), c2 = Sub(p1z, r.oz);
    float e0_ = Sub(p2x, r.ox), e1_ = Sub(p2y, r.oy), e2_ = Sub(p2z, r.oz);
    // :186-196 permute (kx, ky follow from kz; branch-free)
    float p0tx, p0ty, p0tz, p1tx, p1ty, p1tz, p2tx, p2ty, p2tz;
    Permute3(r.kz, a0, a1, a2, &p0tx, &p0ty, &p0tz);
    Permute3(r.kz, c0, c1, c2, &p1tx, &p1ty, &p1tz);
    Permute3(r.kz, e0_, e1_, e2_, &p2tx, &p2ty, &p2tz);
    // :202-207 shear (separate multiply and add)
    p0tx = Add(p0tx, Mul(r.Sx, p0tz));
    p0ty = Add(p0ty, Mul(r.Sy, p0tz));
    p1tx = Add(p1tx, Mul(r.Sx, p1tz));
    p1ty = Add(p1ty, Mul(r.Sy, p1tz));
    p2tx = Add(p2tx, Mul(r.Sx, p2tz));
    p2ty = Add(p2ty, Mul(r.Sy, p2tz));
    // :210-212 edge functions
    float e0 = DifferenceOfProducts(p1tx, p2ty, p1ty, p2tx);
    float e1 = DifferenceOfProducts(p2tx, p0ty, p2ty, p0tx);
    float e2 = DifferenceOfProducts(p0tx, p1ty, p0ty, p1tx);
    // :215-225 double-precision fallback on an exact zero
    if (e0 == 0.0f || e1 == 0.0f || e2 == 0.0f) {
        double p2txp1ty = __dmul_rn((double)p2tx, (double)p1ty);
        double p2typ1tx = __dmul_rn((double)p2ty, (double)p1tx);
        e0 = __double2float_rn(__dsub_rn(p2typ1tx, p2txp1ty));
        double p0txp2ty = __dmul_rn((double)p0tx, (double)p2ty);
        double p0typ2tx = __dmul_rn((double)p0ty, (double)p2tx);
        e1 = __double2float_rn(__dsub_rn(p0typ2tx, p0txp2ty));
        double p1txp0ty = __dmul_rn((double)p1tx, (double)p0ty);
        double p1typ0tx = __dmul_rn((double)p1ty, (double)p0tx);
        e2 = __double2float_rn(__dsub_rn(p1typ0tx, p1txp0ty));
    }
    // :228-232
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0)) return false;
    float det = Add(Add(e0, e1), e2);
    if (det == 0) return false;
    // :235-242
    p0tz = Mul(p0tz, r.Sz);
    p1tz = Mul(p1tz, r.Sz);
    p2tz = Mul(p2tz, r.Sz);
    float tScaled = Add(Add(Mul(e0, p0tz), Mul(e1, p1tz)), Mul(e2, p2tz));
    float tMaxDet = Mul(tMax, det);
    if (det < 0 && (tScaled >= 0 || tScaled < tMaxDet)) return false;
    else if (det > 0 && (tScaled <= 0 || tScaled > tMaxDet)) return false;
    // :245-247
    float invDet = Div(1.f, det);
    float b0 = Mul(e0, invDet), b1 = Mul(e1, invDet), b2 = Mul(e2, invDet);
    float t = Mul(tScaled, invDet);
    // :250-269 conservative t > deltaT.  Only the OUTCOME of `t <= deltaT` is observable, and for almost every
    // accepted hit t exceeds deltaT by orders of magnitude, so a cheap upper bound U >= deltaT decides first:
    //   deltaX, deltaY <= 2 g5 M,  deltaE <= 2 M^2 (g2 + 4 g5),  deltaZ <= g3 M      (M = largest |coordinate|)
    //   deltaT <= 3 |invDet| M (2 g3 maxE + 2 (g2 + 4 g5) M^2)
    // evaluated in float with a 1 % margin that covers the ~30 roundings of both expressions; t > U implies
    // t > deltaT.  Anything else -- t close to the bound, NaN, magnitudes near overflow (W >= 1e37) or so small
    // that products inside either expression are denormal or underflow (W <= 1e-30: a denormal's rounding error
    // is not relative, and |invDet| can blow it up) -- takes the reference's expression verbatim.  The final
    // `* |invDet|` is monotone in both expressions, so it needs no guard of its own.
    {
        const float M = fmaxf(fmaxf(fmaxf(fmaxf(fabsf(p0tx), fabsf(p1tx)), fabsf(p2tx)),
                                    fmaxf(fmaxf(fabsf(p0ty), fabsf(p1ty)), fabsf(p2ty))),
                              fmaxf(fmaxf(fabsf(p0tz), fabsf(p1tz)), fabsf(p2tz)));
        const float mE = fmaxf(fmaxf(fabsf(e0), fabsf(e1)), fabsf(e2));
        constexpr float c1 = 1.01f * 6.f * GammaF(3), c2 = 1.01f * 6.f * (GammaF(2) + 4.f * GammaF(5));
        // W bounds every intermediate of the reference's expression before its final `* |invDet|`: while W is far
        // from overflow none of them is inf (deltaT = inf rejects whatever t is, and must keep doing so)
        const float W = Mul(M, Add(Mul(c1, mE), Mul(c2, Mul(M, M))));
        const float U = Mul(fabsf(invDet), W);
        if (t > U && W > 1e-30f && W < 1e37f) {
            *tOut = t;
            *b0Out = b0;
            *b1Out = b1;
            *b2Out = b2;
            return true;
        }
    }
    float maxZt = Max3First(fabsf(p0tz), fabsf(p1tz), fabsf(p2tz));
    float deltaZ = Mul(GammaF(3), maxZt);
    float maxXt = Max3First(fabsf(p0tx), fabsf(p1tx), fabsf(p2tx));
    float maxYt = Max3First(fabsf(p0ty), fabsf(p1ty), fabsf(p2ty));
    float deltaX = Mul(GammaF(5), Add(maxXt, maxZt));
    float deltaY = Mul(GammaF(5), Add(maxYt, maxZt));
    float deltaE = Mul(2.f, Add(Add(Mul(Mul(GammaF(2), maxXt), maxYt), Mul(deltaY, maxXt)),
                                Mul(deltaX, maxYt)));
    float maxE = Max3First(fabsf(e0), fabsf(e1), fabsf(e2));
    float deltaT = Mul(Mul(3.f, Add(Add(Mul(Mul(GammaF(3), maxE), maxZt), Mul(deltaE, maxZt)),
                                    Mul(deltaZ, maxE))),
                       fabsf(invDet));
    if (t <= deltaT) return false;
    *tOut = t;
    *b0Out = b0;
    *b1Out = b1;
    *b2Out = b2;
    return true;
}

}  // namespace nnbvh
