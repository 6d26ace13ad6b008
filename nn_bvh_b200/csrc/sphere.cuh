// nn_bvh_b200/csrc/sphere.cuh -- sphere leaves: Sphere::BasicIntersect on the device.
//
// Restates, operation for operation, what the reference's HOST build computes
// (/root/reference/src/pbrt):
//   Sphere::BasicIntersect                 shapes.h:147-232
//   Transform::operator()(Point3fi)        util/transform.h:133-176   (exact input: the IsExact branch)
//   Transform::operator()(Vector3fi)       util/transform.h:272-306
//   Interval and its operators             util/math.h:819-1076
//   NextFloatUp/Down, *RoundUp/*RoundDown  util/float.h:164-320
// On the host the reference rounds outward with NextFloatUp/NextFloatDown of the
// round-to-nearest result (util/float.h:199-320, the #else branches); its own GPU build uses
// __fadd_ru & co. instead, which gives DIFFERENT intervals.  Parity is defined against the CPU
// path, so this file follows the host branches.  std::min / std::max are spelled out as the
// comparisons libstdc++ performs, so NaNs take the same side.
//
// Supported: full, z-clipped and phi-clipped spheres.  phiMax < 360 compares std::atan2's value
// (shapes.h:196-203, :218-222): atan2_host.cuh follows the host libm's algorithm for it.  For
// phiMax = 360 the test `phi > phiMax` can never hold (phiMax = Radians(360) rounds to the float
// 2*Pi, and phi <= 2*Pi by construction), so full spheres skip the atan2.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "atan2_host.cuh"
#include "exact_math.cuh"
#include "internal.h"

namespace nnbvh {

__device__ __forceinline__ float NextFloatUp(float v) {  // util/float.h:164-177
    if (v == __int_as_float(0x7f800000)) return v;
    if (v == -0.f) v = 0.f;
    uint32_t ui = __float_as_uint(v);
    if (v >= 0) ++ui; else --ui;
    return __uint_as_float(ui);
}
__device__ __forceinline__ float NextFloatDown(float v) {  // util/float.h:179-193
    if (v == __int_as_float(0xff800000)) return v;
    if (v == 0.f) v = -0.f;
    uint32_t ui = __float_as_uint(v);
    if (v > 0) --ui; else ++ui;
    return __uint_as_float(ui);
}
__device__ __forceinline__ float StdMin(float a, float b) { return (b < a) ? b : a; }  // std::min(a, b)
__device__ __forceinline__ float StdMax(float a, float b) { return (a < b) ? b : a; }  // std::max(a, b)
__device__ __forceinline__ float StdMin4(float a, float b, float c, float d) {  // std::min({a,b,c,d})
    float m = a;
    if (b < m) m = b;
    if (c < m) m = c;
    if (d < m) m = d;
    return m;
}
__device__ __forceinline__ float StdMax4(float a, float b, float c, float d) {  // std::max({a,b,c,d})
    float m = a;
    if (m < b) m = b;
    if (m < c) m = c;
    if (m < d) m = d;
    return m;
}

struct Ival {
    float lo, hi;
};
// Interval(Float low, Float high), util/math.h:825-826
__device__ __forceinline__ Ival MakeIval(float a, float b) { return Ival{StdMin(a, b), StdMax(a, b)}; }
__device__ __forceinline__ Ival Exact(float v) { return Ival{v, v}; }
__device__ __forceinline__ float Mid(Ival i) { return Div(Add(i.lo, i.hi), 2.f); }  // Midpoint(), :851
// Interval::FromValueAndError, :829-838
__device__ __forceinline__ Ival FromValueAndError(float v, float err) {
    if (err == 0) return Ival{v, v};
    return Ival{NextFloatDown(Sub(v, err)), NextFloatUp(Add(v, err))};
}
__device__ __forceinline__ bool InRange0(Ival i) { return 0.f >= i.lo && 0.f <= i.hi; }  // InRange(0, i)
__device__ __forceinline__ Ival operator+(Ival a, Ival b) {  // :872-875
    return MakeIval(NextFloatDown(Add(a.lo, b.lo)), NextFloatUp(Add(a.hi, b.hi)));
}
__device__ __forceinline__ Ival operator-(Ival a, Ival b) {  // :877-880
    return MakeIval(NextFloatDown(Sub(a.lo, b.hi)), NextFloatUp(Sub(a.hi, b.lo)));
}
__device__ __noinline__ Ival operator*(Ival a, Ival b) {  // :882-890
    const float p0 = Mul(a.lo, b.lo), p1 = Mul(a.hi, b.lo), p2 = Mul(a.lo, b.hi), p3 = Mul(a.hi, b.hi);
    return MakeIval(StdMin4(NextFloatDown(p0), NextFloatDown(p1), NextFloatDown(p2), NextFloatDown(p3)),
                    StdMax4(NextFloatUp(p0), NextFloatUp(p1), NextFloatUp(p2), NextFloatUp(p3)));
}
__device__ __noinline__ Ival operator/(Ival a, Ival b) {  // :961-973
    if (InRange0(b)) return MakeIval(__int_as_float(0xff800000), __int_as_float(0x7f800000));
    const float q0 = Div(a.lo, b.lo), q1 = Div(a.hi, b.lo), q2 = Div(a.lo, b.hi), q3 = Div(a.hi, b.hi);
    return MakeIval(StdMin4(NextFloatDown(q0), NextFloatDown(q1), NextFloatDown(q2), NextFloatDown(q3)),
                    StdMax4(NextFloatUp(q0), NextFloatUp(q1), NextFloatUp(q2), NextFloatUp(q3)));
}
__device__ __forceinline__ Ival Sqr(Ival i) {  // :975-982
    float alow = fabsf(i.lo), ahigh = fabsf(i.hi);
    if (alow > ahigh) {
        const float t = alow;
        alow = ahigh;
        ahigh = t;
    }
    if (InRange0(i)) return MakeIval(0.f, NextFloatUp(Mul(ahigh, ahigh)));
    return MakeIval(NextFloatDown(Mul(alow, alow)), NextFloatUp(Mul(ahigh, ahigh)));
}
__device__ __forceinline__ Ival operator*(float f, Ival i) {  // :995-1000
    if (f > 0) return MakeIval(NextFloatDown(Mul(f, i.lo)), NextFloatUp(Mul(f, i.hi)));
    return MakeIval(NextFloatDown(Mul(f, i.hi)), NextFloatUp(Mul(f, i.lo)));
}
__device__ __forceinline__ Ival Sqrt(Ival i) {  // :1070-1072 with util/float.h:277-299
    const float lo = NextFloatDown(__fsqrt_rn(i.lo));
    return MakeIval((0.f < lo) ? lo : 0.f, NextFloatUp(__fsqrt_rn(i.hi)));
}

// SphereDev: 64 bytes, one per sphere: rows 0..2 of objectFromRender, then radius and the
// constructor's clamped zMin / zMax and phiMax in radians (shapes.h:117-128).
struct alignas(64) SphereDev {
    float m[12];
    float radius, zMin, zMax, phiMax;
};
constexpr float kTwoPiF = 2 * 3.14159265358979323846f;  // 2 * Pi as the reference's float expression

// The clipping test of shapes.h:202-203 / :221-222 on a refined hit point.
__device__ __forceinline__ bool SphereClipped(float radius, float zMin, float zMax, float phiMax, float px, float py,
                                              float pz) {
    if ((zMin > -radius && pz < zMin) || (zMax < radius && pz > zMax)) return true;
    if (!(phiMax < kTwoPiF)) return false;
    float phi = Atan2fHost(py, px);
    if (phi < 0) phi = Add(phi, kTwoPiF);
    return phi > phiMax;
}
static_assert(sizeof(SphereDev) == 64, "SphereDev must be 64 bytes");

struct SphereHit {
    float t, px, py, pz;  // QuadricIntersection{tHit, pObj} (shapes.h:100-104); phi = atan2(pObj.y, pObj.x)
};

// pHit = Point3f(oi) + (Float)tShapeHit * Vector3f(di), refined onto the sphere (shapes.h:189-195).
__device__ __forceinline__ void SpherePoint(float r, float ox, float oy, float oz, float dx, float dy, float dz,
                                            float t, float *px, float *py, float *pz) {
    float x = Add(ox, Mul(t, dx)), y = Add(oy, Mul(t, dy)), z = Add(oz, Mul(t, dz));
    // Distance(pHit, Point3f(0,0,0)) = sqrt(LengthSquared(pHit - 0)), util/vecmath.h:948-955
    const float x0 = Sub(x, 0.f), y0 = Sub(y, 0.f), z0 = Sub(z, 0.f);
    const float s = Div(r, __fsqrt_rn(Add(Add(Mul(x0, x0), Mul(y0, y0)), Mul(z0, z0))));
    x = Mul(x, s);
    y = Mul(y, s);
    z = Mul(z, s);
    if (x == 0 && y == 0) x = Mul(1e-5f, r);
    *px = x, *py = y, *pz = z;
}

// Sphere::BasicIntersect(r, tMax), shapes.h:147-232.  (o, d): the ray in render space.
__device__ __noinline__ bool SphereTest(const SphereDev *__restrict__ sp, float rox, float roy, float roz,
                                        float rdx, float rdy, float rdz, float tMax, SphereHit *out) {
    const float4 r0 = __ldg(reinterpret_cast<const float4 *>(sp)), r1 = __ldg(reinterpret_cast<const float4 *>(sp) + 1),
                 r2 = __ldg(reinterpret_cast<const float4 *>(sp) + 2), r3 = __ldg(reinterpret_cast<const float4 *>(sp) + 3);
    const float radius = r3.x, zMin = r3.y, zMax = r3.z, phiMax = r3.w;
    const float g3 = GammaF(3);
    Ival oix, oiy, oiz, dix, diy, diz;
    {   // Transform::operator()(Point3fi), exact input; Float(p.x) is the midpoint of [x, x]
        const float x = Mid(Exact(rox)), y = Mid(Exact(roy)), z = Mid(Exact(roz));
        const float xp = Add(Add(Mul(r0.x, x), Mul(r0.y, y)), Add(Mul(r0.z, z), r0.w));
        const float yp = Add(Add(Mul(r1.x, x), Mul(r1.y, y)), Add(Mul(r1.z, z), r1.w));
        const float zp = Add(Add(Mul(r2.x, x), Mul(r2.y, y)), Add(Mul(r2.z, z), r2.w));
        const float ex = Mul(g3, Add(Add(Add(fabsf(Mul(r0.x, x)), fabsf(Mul(r0.y, y))), fabsf(Mul(r0.z, z))), fabsf(r0.w)));
        const float ey = Mul(g3, Add(Add(Add(fabsf(Mul(r1.x, x)), fabsf(Mul(r1.y, y))), fabsf(Mul(r1.z, z))), fabsf(r1.w)));
        const float ez = Mul(g3, Add(Add(Add(fabsf(Mul(r2.x, x)), fabsf(Mul(r2.y, y))), fabsf(Mul(r2.z, z))), fabsf(r2.w)));
        oix = FromValueAndError(xp, ex);
        oiy = FromValueAndError(yp, ey);
        oiz = FromValueAndError(zp, ez);
    }
    {   // Transform::operator()(Vector3fi), exact input
        const float x = Mid(Exact(rdx)), y = Mid(Exact(rdy)), z = Mid(Exact(rdz));
        const float ex = Mul(g3, Add(Add(fabsf(Mul(r0.x, x)), fabsf(Mul(r0.y, y))), fabsf(Mul(r0.z, z))));
        const float ey = Mul(g3, Add(Add(fabsf(Mul(r1.x, x)), fabsf(Mul(r1.y, y))), fabsf(Mul(r1.z, z))));
        const float ez = Mul(g3, Add(Add(fabsf(Mul(r2.x, x)), fabsf(Mul(r2.y, y))), fabsf(Mul(r2.z, z))));
        const float xp = Add(Add(Mul(r0.x, x), Mul(r0.y, y)), Mul(r0.z, z));
        const float yp = Add(Add(Mul(r1.x, x), Mul(r1.y, y)), Mul(r1.z, z));
        const float zp = Add(Add(Mul(r2.x, x), Mul(r2.y, y)), Mul(r2.z, z));
        dix = FromValueAndError(xp, ex);
        diy = FromValueAndError(yp, ey);
        diz = FromValueAndError(zp, ez);
    }
    const Ival R = Exact(radius);
    const Ival a = Sqr(dix) + Sqr(diy) + Sqr(diz);                       // shapes.h:157
    const Ival b = 2.f * (dix * oix + diy * oiy + diz * oiz);           // :158
    const Ival c = Sqr(oix) + Sqr(oiy) + Sqr(oiz) - Sqr(R);             // :159
    const Ival f = b / (2.f * a);                                        // :162
    const Ival vx = oix - f * dix, vy = oiy - f * diy, vz = oiz - f * diz;
    const Ival length = Sqrt(Sqr(vx) + Sqr(vy) + Sqr(vz));
    const Ival discrim = 4.f * a * (R + length) * (R - length);         // :164-165
    if (discrim.lo < 0) return false;
    const Ival rootDiscrim = Sqrt(discrim);
    Ival q;
    if (Mid(b) < 0) q = -.5f * (b - rootDiscrim);
    else q = -.5f * (b + rootDiscrim);
    Ival t0 = q / a, t1 = c / q;
    if (t0.lo > t1.lo) {
        const Ival t = t0;
        t0 = t1;
        t1 = t;
    }
    if (t0.hi > tMax || t1.lo <= 0) return false;                       // :183-184
    Ival tShape = t0;
    if (tShape.lo <= 0) {
        tShape = t1;
        if (tShape.hi > tMax) return false;
    }
    const float ox = Mid(oix), oy = Mid(oiy), oz = Mid(oiz), dx = Mid(dix), dy = Mid(diy), dz = Mid(diz);
    float px, py, pz;
    SpherePoint(radius, ox, oy, oz, dx, dy, dz, Mid(tShape), &px, &py, &pz);
    // clipping (:202-222)
    if (SphereClipped(radius, zMin, zMax, phiMax, px, py, pz)) {
        if (tShape.lo == t1.lo && tShape.hi == t1.hi) return false;
        if (t1.hi > tMax) return false;
        tShape = t1;
        SpherePoint(radius, ox, oy, oz, dx, dy, dz, Mid(tShape), &px, &py, &pz);
        if (SphereClipped(radius, zMin, zMax, phiMax, px, py, pz)) return false;
    }
    out->t = Mid(tShape);
    out->px = px, out->py = py, out->pz = pz;
    return true;
}

// ---- instances: TransformedPrimitive (cpu/primitive.cpp:112-130) -------------------------------
// InstanceDev: 64 bytes: rows 0..2 of primitiveFromRender (the mInv of renderFromPrimitive), then where
// the object's tree and ordered triangles start in the concatenated object arrays.
struct alignas(64) InstanceDev {
    float mInv[12];
    uint32_t nodeBase, primBase, nNodes, pad;
};
static_assert(sizeof(InstanceDev) == 64, "InstanceDev must be 64 bytes");

// What StepKernel<..., INST> reads when a lane enters an instance (two float4 loads).
struct alignas(32) InstanceStep {
    float rootBox[6];   // bounds of the object's root node (the reference tests it first, aggregates.cpp:543)
    int32_t rootRef;    // WideNode index (>= 0) or leaf ref (< 0) of the object's root, absolute
    uint32_t primBase;  // first triangle of the object in the combined triangle array
};
static_assert(sizeof(InstanceStep) == 32, "InstanceStep must be 32 bytes");

// Ray Transform::ApplyInverse(const Ray &r, Float *tMax) (util/transform.h:416-430) with
// Point3fi Transform::ApplyInverse(const Point3fi &) on an exact point (util/transform.cpp:263-303).
__device__ __noinline__ void InstanceRay(const InstanceDev *__restrict__ in, float rox, float roy, float roz,
                                         float rdx, float rdy, float rdz, float tMax, float *o3, float *d3,
                                         float *tMaxOut) {
    const float4 r0 = __ldg(reinterpret_cast<const float4 *>(in)), r1 = __ldg(reinterpret_cast<const float4 *>(in) + 1),
                 r2 = __ldg(reinterpret_cast<const float4 *>(in) + 2);
    const float g3 = GammaF(3);
    const float x = Mid(Exact(rox)), y = Mid(Exact(roy)), z = Mid(Exact(roz));
    Ival ox = FromValueAndError(Add(Add(Mul(r0.x, x), Mul(r0.y, y)), Add(Mul(r0.z, z), r0.w)),
                                Mul(g3, Add(Add(fabsf(Mul(r0.x, x)), fabsf(Mul(r0.y, y))), fabsf(Mul(r0.z, z)))));
    Ival oy = FromValueAndError(Add(Add(Mul(r1.x, x), Mul(r1.y, y)), Add(Mul(r1.z, z), r1.w)),
                                Mul(g3, Add(Add(fabsf(Mul(r1.x, x)), fabsf(Mul(r1.y, y))), fabsf(Mul(r1.z, z)))));
    Ival oz = FromValueAndError(Add(Add(Mul(r2.x, x), Mul(r2.y, y)), Add(Mul(r2.z, z), r2.w)),
                                Mul(g3, Add(Add(fabsf(Mul(r2.x, x)), fabsf(Mul(r2.y, y))), fabsf(Mul(r2.z, z)))));
    const float dx = Add(Add(Mul(r0.x, rdx), Mul(r0.y, rdy)), Mul(r0.z, rdz));
    const float dy = Add(Add(Mul(r1.x, rdx), Mul(r1.y, rdy)), Mul(r1.z, rdz));
    const float dz = Add(Add(Mul(r2.x, rdx), Mul(r2.y, rdy)), Mul(r2.z, rdz));
    const float lengthSquared = Add(Add(Mul(dx, dx), Mul(dy, dy)), Mul(dz, dz));
    if (lengthSquared > 0) {
        const float ex = Div(Sub(ox.hi, ox.lo), 2.f), ey = Div(Sub(oy.hi, oy.lo), 2.f), ez = Div(Sub(oz.hi, oz.lo), 2.f);
        const float dt = Div(Add(Add(Mul(fabsf(dx), ex), Mul(fabsf(dy), ey)), Mul(fabsf(dz), ez)), lengthSquared);
        ox = ox + Exact(Mul(dt, dx));  // Point3fi += Vector3f: Interval += Float (math.h:916-917)
        oy = oy + Exact(Mul(dt, dy));
        oz = oz + Exact(Mul(dt, dz));
        tMax = Sub(tMax, dt);
    }
    o3[0] = Mid(ox), o3[1] = Mid(oy), o3[2] = Mid(oz);
    d3[0] = dx, d3[1] = dy, d3[2] = dz;
    *tMaxOut = tMax;
}

}  // namespace nnbvh
