// include/nnbvh_wavefront.hpp -- the wavefront batch backend over libnnbvh.so.
//
// Host-side C++ mirror of the reference's batch caller of the hot path:
//   class WavefrontAggregate            /root/reference/src/pbrt/wavefront/integrator.h:30-53
//   CPUAggregate::IntersectClosest      /root/reference/src/pbrt/wavefront/aggregate.cpp:34-58
//   CPUAggregate::IntersectShadow[Tr]   /root/reference/src/pbrt/wavefront/aggregate.cpp:60-92
// and of what they do with each result:
//   EnqueueWorkAfterMiss                wavefront/intersect.h:16-29
//   RecordShadowRayResult               wavefront/intersect.h:31-46
//   EnqueueWorkAfterIntersection        wavefront/intersect.h:48-163
// The reference's RayQueue / ShadowRayQueue are SOA<RayWorkItem> / SOA<ShadowRayWorkItem>
// generated at build time by `soac` from wavefront/workitems.soa: seven (eight) separate float
// arrays per queue.  RayQueueView / ShadowRayQueueView below are exactly
// those arrays, so a queue is handed to the engine without any gather or repack.
//
// The engine returns {prim, t, b0, b1, b2}; the reference's consumers want a SurfaceInteraction.
// InteractionFromIntersection() below restates Triangle::InteractionFromIntersection
// (shapes.h:884-1010) on the host, operation for operation (compile with -ffp-contract=off):
// tests/test_wavefront.py pins it bit for bit against the reference's own function.  A hit on a
// sphere leaf comes back as {prim, t, pObj}: SphereInteractionFromIntersection() restates
// Sphere::InteractionFromIntersection (shapes.h:238-280) and the Transform::operator()
// (SurfaceInteraction) it ends with (util/transform.cpp:229-261), pinned the same way.
//
// No traversal or intersection arithmetic lives here; every query runs on the GPU.
#ifndef NNBVH_WAVEFRONT_HPP
#define NNBVH_WAVEFRONT_HPP

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "nnbvh.hpp"

namespace nnbvh {
namespace wavefront {

// ---- queue views ---------------------------------------------------------------------------
// soa RayWorkItem { Ray ray; ... } with soa Ray { Point3f o; Vector3f d; Float time; Medium medium; }
// (pbrt.soa, wavefront/workitems.soa).  Pointers may be host, pinned, managed or device memory.
struct RayQueueView {
    const float *ox, *oy, *oz;  // ray.o.x[], ray.o.y[], ray.o.z[]
    const float *dx, *dy, *dz;  // ray.d.x[] ...
    const float *time;          // ray.time[] (may be null: 0)
    int size;                   // WorkQueue::Size(), workqueue.h:36-110
};
// soa ShadowRayWorkItem { Ray ray; Float tMax; ... }
struct ShadowRayQueueView {
    const float *ox, *oy, *oz, *dx, *dy, *dz;
    const float *time;
    const float *tMax;
    int size;
};

// ---- TriangleMesh (util/mesh.h:24-47), the arrays InteractionFromIntersection reads --------
struct TriangleMeshView {
    const int *vertexIndices = nullptr;  // 3 per triangle
    const float *p = nullptr;            // xyz per vertex, render space
    const float *n = nullptr;            // optional shading normals AS TriangleMesh STORES THEM
                                         // (negated when reverseOrientation, util/mesh.cpp:51-55)
    const float *s = nullptr;            // optional tangents
    const float *uv = nullptr;           // optional
    const int *faceIndices = nullptr;    // optional
    bool reverseOrientation = false, transformSwapsHandedness = false;
};

// The geometric content of SurfaceInteraction (interaction.h:131-263) a triangle hit defines.
struct SurfaceHit {
    float p[3], pError[3];        // pHit and its error bound (shapes.h:924-931)
    float piLow[3], piHigh[3];    // Point3fi pi = Point3fi(pHit, pError): what SurfaceInteraction stores
    float n[3];
    float uv[2];
    float dpdu[3], dpdv[3];
    float dndu[3], dndv[3];  // zero for triangles (shapes.h:928)
    struct {
        float n[3], dpdu[3], dpdv[3], dndu[3], dndv[3];
    } shading;
    float wo[3];
    float time;
    int faceIndex;
};

namespace detail {
struct V3 {
    float x, y, z;
};
inline V3 Load3(const float *a, int i) { return V3{a[3 * i], a[3 * i + 1], a[3 * i + 2]}; }
inline void Store3(float *o, V3 v) { o[0] = v.x, o[1] = v.y, o[2] = v.z; }
inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
inline V3 operator*(float s, V3 a) { return V3{s * a.x, s * a.y, s * a.z}; }  // vecmath.h:351-354
inline V3 operator/(V3 a, float d) { return V3{a.x / d, a.y / d, a.z / d}; }  // vecmath.h:364-368
inline V3 Abs(V3 a) { return V3{std::fabs(a.x), std::fabs(a.y), std::fabs(a.z)}; }
// util/math.h:569-575
inline float DoP(float a, float b, float c, float d) {
    float cd = c * d;
    float r = std::fma(a, b, -cd);
    float e = std::fma(-c, d, cd);
    return r + e;
}
// util/math.h:577-583
inline float SoP(float a, float b, float c, float d) {
    float cd = c * d;
    float r = std::fma(a, b, cd);
    float e = std::fma(c, d, -cd);
    return r + e;
}
inline V3 DoP(float a, V3 b, float c, V3 d) {
    return V3{DoP(a, b.x, c, d.x), DoP(a, b.y, c, d.y), DoP(a, b.z, c, d.z)};
}
// vecmath.h:998-1004
inline V3 Cross(V3 v, V3 w) {
    return V3{DoP(v.y, w.z, v.z, w.y), DoP(v.z, w.x, v.x, w.z), DoP(v.x, w.y, v.y, w.x)};
}
inline float LengthSquared(V3 v) { return v.x * v.x + v.y * v.y + v.z * v.z; }  // :948-950
inline V3 Normalize(V3 v) { return v / std::sqrt(LengthSquared(v)); }           // :953-961
// Dot(Normal3, *) (vecmath.h:1057-1075); Dot(Vector3, Vector3) is the plain form (:964-967)
inline float DotN(V3 n, V3 v) { return std::fma(n.x, v.x, SoP(n.y, v.y, n.z, v.z)); }
// vecmath.h:1007-1013
inline void CoordinateSystem(V3 v1, V3 *v2, V3 *v3) {
    float sign = std::copysign(1.f, v1.z);
    float a = -1 / (sign + v1.z);
    float b = v1.x * v1.y * a;
    *v2 = V3{1 + sign * (v1.x * v1.x) * a, sign * b, -sign * v1.x};
    *v3 = V3{b, sign + (v1.y * v1.y) * a, -v1.y};
}
inline constexpr float Gamma(int n) {  // util/float.h:195-197
    return (n * 5.96046448e-08f) / (1 - n * 5.96046448e-08f);
}
// NextFloatUp / NextFloatDown, util/float.h:164-193 (the host build's AddRoundUp / SubRoundDown
// are one float step past the round-to-nearest result, :199-227)
inline float NextFloatUp(float v) {
    if (std::isinf(v) && v > 0.f) return v;
    if (v == -0.f) v = 0.f;
    uint32_t ui;
    std::memcpy(&ui, &v, 4);
    if (v >= 0) ++ui; else --ui;
    std::memcpy(&v, &ui, 4);
    return v;
}
inline float NextFloatDown(float v) {
    if (std::isinf(v) && v < 0.f) return v;
    if (v == 0.f) v = -0.f;
    uint32_t ui;
    std::memcpy(&ui, &v, 4);
    if (v > 0) --ui; else ++ui;
    std::memcpy(&v, &ui, 4);
    return v;
}
// Interval::FromValueAndError, util/math.h:829-838
inline void FromValueAndError(float v, float err, float *low, float *high) {
    if (err == 0)
        *low = *high = v;
    else {
        *low = NextFloatDown(v - err);
        *high = NextFloatUp(v + err);
    }
}
}  // namespace detail

// Triangle::InteractionFromIntersection(mesh, triIndex, ti, time, wo), shapes.h:884-1010.
// Returns false where the reference CHECK-aborts (a triangle whose normal is zero even in double).
inline bool InteractionFromIntersection(const TriangleMeshView &mesh, int triIndex, float b0,
                                        float b1, float b2, float time, const float wo[3],
                                        SurfaceHit *out) {
    using namespace detail;
    const int *v = &mesh.vertexIndices[3 * triIndex];
    V3 p0 = Load3(mesh.p, v[0]), p1 = Load3(mesh.p, v[1]), p2 = Load3(mesh.p, v[2]);
    float uv[3][2] = {{0, 0}, {1, 0}, {1, 1}};  // :896-900
    if (mesh.uv)
        for (int k = 0; k < 3; ++k) uv[k][0] = mesh.uv[2 * v[k]], uv[k][1] = mesh.uv[2 * v[k] + 1];
    float duv02[2] = {uv[0][0] - uv[2][0], uv[0][1] - uv[2][1]};
    float duv12[2] = {uv[1][0] - uv[2][0], uv[1][1] - uv[2][1]};
    V3 dp02 = p0 - p2, dp12 = p1 - p2;
    float determinant = DoP(duv02[0], duv12[1], duv02[1], duv12[0]);

    V3 dpdu{0, 0, 0}, dpdv{0, 0, 0};
    bool degenerateUV = std::fabs(determinant) < 1e-9f;
    if (!degenerateUV) {
        float invdet = 1 / determinant;
        dpdu = invdet * DoP(duv12[1], dp02, duv02[1], dp12);
        dpdv = invdet * DoP(duv02[0], dp12, duv12[0], dp02);
    }
    if (degenerateUV || LengthSquared(Cross(dpdu, dpdv)) == 0) {  // :914-921
        V3 ng = Cross(p2 - p0, p1 - p0);
        if (LengthSquared(ng) == 0) {
            V3 a = p2 - p0, b = p1 - p0;
            double ax = a.x, ay = a.y, az = a.z, bx = b.x, by = b.y, bz = b.z;
            // Cross<double> is DifferenceOfProducts in double as well (vecmath.h:998-1004)
            auto dop = [](double a, double b, double c, double d) {
                double cd = c * d;
                return std::fma(a, b, -cd) + std::fma(-c, d, cd);
            };
            ng = V3{(float)dop(ay, bz, az, by), (float)dop(az, bx, ax, bz), (float)dop(ax, by, ay, bx)};
            if (LengthSquared(ng) == 0) return false;  // CHECK_NE(LengthSquared(ng), 0)
        }
        CoordinateSystem(Normalize(ng), &dpdu, &dpdv);
    }

    V3 pHit = b0 * p0 + b1 * p1 + b2 * p2;  // :924
    float uvHit[2] = {b0 * uv[0][0] + b1 * uv[1][0] + b2 * uv[2][0],
                      b0 * uv[0][1] + b1 * uv[1][1] + b2 * uv[2][1]};
    bool flipNormal = mesh.reverseOrientation ^ mesh.transformSwapsHandedness;
    V3 pAbsSum = Abs(b0 * p0) + Abs(b1 * p1) + Abs(b2 * p2);
    V3 pError = Gamma(7) * pAbsSum;

    Store3(out->p, pHit);
    Store3(out->pError, pError);
    FromValueAndError(pHit.x, pError.x, &out->piLow[0], &out->piHigh[0]);
    FromValueAndError(pHit.y, pError.y, &out->piLow[1], &out->piHigh[1]);
    FromValueAndError(pHit.z, pError.z, &out->piLow[2], &out->piHigh[2]);
    out->uv[0] = uvHit[0], out->uv[1] = uvHit[1];
    Store3(out->wo, Normalize(V3{wo[0], wo[1], wo[2]}));  // Interaction(pi, n, uv, wo, time), interaction.h:30-32
    out->time = time;
    Store3(out->dpdu, dpdu);
    Store3(out->dpdv, dpdv);
    V3 zero{0, 0, 0};
    Store3(out->dndu, zero);
    Store3(out->dndv, zero);
    out->faceIndex = mesh.faceIndices ? mesh.faceIndices[triIndex] : 0;
    // the constructor's normal (interaction.h:166) is overridden at shapes.h:939-941
    V3 n = Normalize(Cross(dp02, dp12));
    if (flipNormal) n = -n;
    V3 sn = n, sdpdu = dpdu, sdpdv = dpdv, sdndu = zero, sdndv = zero;

    if (mesh.n || mesh.s) {  // :943-1005
        V3 ns;
        if (mesh.n) {
            ns = b0 * Load3(mesh.n, v[0]) + b1 * Load3(mesh.n, v[1]) + b2 * Load3(mesh.n, v[2]);
            ns = LengthSquared(ns) > 0 ? Normalize(ns) : n;
        } else
            ns = n;
        V3 ss;
        if (mesh.s) {
            ss = b0 * Load3(mesh.s, v[0]) + b1 * Load3(mesh.s, v[1]) + b2 * Load3(mesh.s, v[2]);
            if (LengthSquared(ss) == 0) ss = dpdu;
        } else
            ss = dpdu;
        V3 ts = Cross(ns, ss);
        if (LengthSquared(ts) > 0)
            ss = Cross(ts, ns);
        else
            CoordinateSystem(ns, &ss, &ts);
        V3 dndu, dndv;
        if (mesh.n) {
            V3 dn1 = Load3(mesh.n, v[0]) - Load3(mesh.n, v[2]);
            V3 dn2 = Load3(mesh.n, v[1]) - Load3(mesh.n, v[2]);
            // the reference compares against the DOUBLE literal 1e-9 here (:977)
            bool degUV = (double)std::fabs(determinant) < 1e-9;
            if (degUV) {
                V3 dn = Cross(Load3(mesh.n, v[2]) - Load3(mesh.n, v[0]),
                              Load3(mesh.n, v[1]) - Load3(mesh.n, v[0]));
                if (LengthSquared(dn) == 0)
                    dndu = dndv = zero;
                else
                    CoordinateSystem(dn, &dndu, &dndv);
            } else {
                float invDet = 1 / determinant;
                dndu = invDet * DoP(duv12[1], dn1, duv02[1], dn2);
                dndv = invDet * DoP(duv02[0], dn2, duv12[0], dn1);
            }
        } else
            dndu = dndv = zero;
        // SetShadingGeometry(ns, ss, ts, dndu, dndv, true), interaction.h:194-214
        sn = ns;
        if (DotN(n, sn) < 0.f) n = -n;
        sdpdu = ss, sdpdv = ts, sdndu = dndu, sdndv = dndv;
        while (LengthSquared(sdpdu) > 1e16f || LengthSquared(sdpdv) > 1e16f) {
            sdpdu = sdpdu / 1e8f;
            sdpdv = sdpdv / 1e8f;
        }
    }
    Store3(out->n, n);
    Store3(out->shading.n, sn);
    Store3(out->shading.dpdu, sdpdu);
    Store3(out->shading.dpdv, sdpdv);
    Store3(out->shading.dndu, sdndu);
    Store3(out->shading.dndv, sdndv);
    return true;
}

// ---- sphere hits ---------------------------------------------------------------------------
// What Sphere keeps beside nnbvh_sphere (shapes.h:117-128): the clamped z range's polar angles, phiMax
// in radians, the two orientation flags.  The same libm (std::acos, std::atan2) as the reference
// build uses is the one linked here: the host side of a drop-in is compiled with the reference.
struct SphereShape {
    const float *renderFromObject;  // 12 floats each, rows 0..2 of the 4x4 (row 3 = 0 0 0 1)
    const float *objectFromRender;
    float radius, thetaZMin, thetaZMax, phiMax;
    bool reverseOrientation, transformSwapsHandedness;
};

namespace detail {
inline float Clamp(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }  // math.h:236-244
// Transform::SwapsHandedness (transform.cpp:145-152) over Determinant(SquareMatrix<3>), math.h:1420-1426
inline bool SwapsHandedness(const float *m) {
    float minor12 = DoP(m[5], m[10], m[6], m[9]);
    float minor02 = DoP(m[4], m[10], m[6], m[8]);
    float minor01 = DoP(m[4], m[9], m[5], m[8]);
    float det = std::fma(m[2], minor01, DoP(m[0], minor12, m[1], minor02));
    return det < 0;
}
// Transform::operator()(Vector3<T>), transform.h:322-326
inline V3 XfVector(const float *m, V3 v) {
    return V3{m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z,
              m[8] * v.x + m[9] * v.y + m[10] * v.z};
}
// Transform::operator()(Normal3<T>): the inverse's transpose, transform.h:329-334
inline V3 XfNormal(const float *mInv, V3 n) {
    return V3{mInv[0] * n.x + mInv[4] * n.y + mInv[8] * n.z, mInv[1] * n.x + mInv[5] * n.y + mInv[9] * n.z,
              mInv[2] * n.x + mInv[6] * n.y + mInv[10] * n.z};
}
inline float DotV(V3 v, V3 w) { return v.x * w.x + v.y * w.y + v.z * w.z; }  // vecmath.h:964-967
}  // namespace detail

inline SphereShape MakeSphereShape(const nnbvh_sphere &s, bool reverseOrientation) {
    using namespace detail;
    SphereShape o;
    o.renderFromObject = s.renderFromObject, o.objectFromRender = s.objectFromRender;
    o.radius = s.radius;
    o.thetaZMin = std::acos(Clamp(std::min(s.zMin, s.zMax) / s.radius, -1, 1));
    o.thetaZMax = std::acos(Clamp(std::max(s.zMin, s.zMax) / s.radius, -1, 1));
    o.phiMax = (3.14159265358979323846f / 180) * Clamp(s.phiMax, 0, 360);  // Radians(), math.h:261-263
    o.reverseOrientation = reverseOrientation;
    o.transformSwapsHandedness = SwapsHandedness(s.renderFromObject);
    return o;
}

// Transform::operator()(const SurfaceInteraction &), util/transform.cpp:229-261, for the fields
// SurfaceHit carries.  m / mInv: rows 0..2 of the transform and of its inverse.
inline void TransformSurfaceHit(const float *m, const float *mInv, SurfaceHit *h) {
    using namespace detail;
    // ret.pi = t(si.pi): Transform::operator()(const Point3fi &), transform.h:133-176
    float x = (h->piLow[0] + h->piHigh[0]) / 2, y = (h->piLow[1] + h->piHigh[1]) / 2,
          z = (h->piLow[2] + h->piHigh[2]) / 2;  // Interval::Midpoint, math.h:851
    float wx = h->piHigh[0] - h->piLow[0], wy = h->piHigh[1] - h->piLow[1], wz = h->piHigh[2] - h->piLow[2];
    bool exact = wx == 0 && wy == 0 && wz == 0;
    float ex = wx / 2, ey = wy / 2, ez = wz / 2;  // Point3fi::Error(), vecmath.h:757
    float pp[3], pe[3];
    for (int r = 0; r < 3; ++r) {
        const float *a = m + 4 * r;
        pp[r] = (a[0] * x + a[1] * y) + (a[2] * z + a[3]);
        float mag = Gamma(3) * (std::fabs(a[0] * x) + std::fabs(a[1] * y) + std::fabs(a[2] * z) + std::fabs(a[3]));
        if (exact)
            pe[r] = mag;
        else
            pe[r] = (Gamma(3) + 1) * (std::fabs(a[0]) * ex + std::fabs(a[1]) * ey + std::fabs(a[2]) * ez) + mag;
    }
    // row 3 is 0 0 0 1: wp == 1 whenever x, y, z are finite
    for (int r = 0; r < 3; ++r) {
        h->p[r] = pp[r], h->pError[r] = pe[r];
        FromValueAndError(pp[r], pe[r], &h->piLow[r], &h->piHigh[r]);
    }
    V3 n = Normalize(XfNormal(mInv, V3{h->n[0], h->n[1], h->n[2]}));
    Store3(h->n, n);
    Store3(h->wo, Normalize(XfVector(m, V3{h->wo[0], h->wo[1], h->wo[2]})));
    Store3(h->dpdu, XfVector(m, V3{h->dpdu[0], h->dpdu[1], h->dpdu[2]}));
    Store3(h->dpdv, XfVector(m, V3{h->dpdv[0], h->dpdv[1], h->dpdv[2]}));
    Store3(h->dndu, XfNormal(mInv, V3{h->dndu[0], h->dndu[1], h->dndu[2]}));
    Store3(h->dndv, XfNormal(mInv, V3{h->dndv[0], h->dndv[1], h->dndv[2]}));
    auto &s = h->shading;
    V3 sn = Normalize(XfNormal(mInv, V3{s.n[0], s.n[1], s.n[2]}));
    Store3(s.dpdu, XfVector(m, V3{s.dpdu[0], s.dpdu[1], s.dpdu[2]}));
    Store3(s.dpdv, XfVector(m, V3{s.dpdv[0], s.dpdv[1], s.dpdv[2]}));
    Store3(s.dndu, XfNormal(mInv, V3{s.dndu[0], s.dndu[1], s.dndu[2]}));
    Store3(s.dndv, XfNormal(mInv, V3{s.dndv[0], s.dndv[1], s.dndv[2]}));
    if (DotN(sn, n) < 0.f) sn = -sn;  // FaceForward(ret.shading.n, ret.n), vecmath.h:1106-1108
    Store3(s.n, sn);
}

// Sphere::InteractionFromIntersection(isect, wo, time), shapes.h:238-280.  pObj is what the engine
// returns in Hit::{b0,b1,b2} for a sphere leaf; wo is in render space and need not be unit length.
inline void SphereInteractionFromIntersection(const SphereShape &sp, const float pObj[3], const float wo[3],
                                              float time, SurfaceHit *out) {
    using namespace detail;
    const float Pi = 3.14159265358979323846f;
    V3 pHit{pObj[0], pObj[1], pObj[2]};
    float phi = std::atan2(pHit.y, pHit.x);  // QuadricIntersection::phi, shapes.h:194-197
    if (phi < 0) phi += 2 * Pi;
    float u = phi / sp.phiMax;
    float cosTheta = pHit.z / sp.radius;
    float theta = std::acos(Clamp(cosTheta, -1, 1));  // SafeACos, math.h:350-353
    float dTheta = sp.thetaZMax - sp.thetaZMin;
    float v = (theta - sp.thetaZMin) / dTheta;
    float zRadius = std::sqrt(pHit.x * pHit.x + pHit.y * pHit.y);
    float cosPhi = pHit.x / zRadius, sinPhi = pHit.y / zRadius;
    V3 dpdu{-sp.phiMax * pHit.y, sp.phiMax * pHit.x, 0};
    float sinTheta = std::sqrt(std::max(0.f, 1 - cosTheta * cosTheta));  // SafeSqrt, math.h:276-279
    V3 dpdv = dTheta * V3{pHit.z * cosPhi, pHit.z * sinPhi, -sp.radius * sinTheta};
    V3 d2Pduu = (-sp.phiMax * sp.phiMax) * V3{pHit.x, pHit.y, 0};
    V3 d2Pduv = (dTheta * pHit.z * sp.phiMax) * V3{-sinPhi, cosPhi, 0};
    V3 d2Pdvv = -(dTheta * dTheta) * pHit;
    float E = DotV(dpdu, dpdu), F = DotV(dpdu, dpdv), G = DotV(dpdv, dpdv);
    V3 nn = Normalize(Cross(dpdu, dpdv));
    float e = DotV(nn, d2Pduu), f = DotV(nn, d2Pduv), g = DotV(nn, d2Pdvv);
    float EGF2 = DoP(E, G, F, F);
    float invEGF2 = (EGF2 == 0) ? 0.f : 1 / EGF2;
    V3 dndu = ((f * F - e * G) * invEGF2) * dpdu + ((e * F - f * E) * invEGF2) * dpdv;
    V3 dndv = ((g * F - f * G) * invEGF2) * dpdu + ((f * F - g * E) * invEGF2) * dpdv;
    V3 pError = Gamma(5) * Abs(pHit);
    bool flipNormal = sp.reverseOrientation ^ sp.transformSwapsHandedness;
    V3 woObject = XfVector(sp.objectFromRender, V3{wo[0], wo[1], wo[2]});

    // SurfaceInteraction(pi, uv, wo, dpdu, dpdv, dndu, dndv, time, flipNormal), interaction.h:151-180
    Store3(out->p, pHit);
    Store3(out->pError, pError);
    FromValueAndError(pHit.x, pError.x, &out->piLow[0], &out->piHigh[0]);
    FromValueAndError(pHit.y, pError.y, &out->piLow[1], &out->piHigh[1]);
    FromValueAndError(pHit.z, pError.z, &out->piLow[2], &out->piHigh[2]);
    out->uv[0] = u, out->uv[1] = v;
    Store3(out->wo, Normalize(woObject));  // Interaction(pi, n, uv, wo, time), interaction.h:30-32
    out->time = time;
    out->faceIndex = 0;
    V3 n = nn;
    if (flipNormal) n = -1.f * n;  // n *= -1
    Store3(out->n, n);
    Store3(out->dpdu, dpdu);
    Store3(out->dpdv, dpdv);
    Store3(out->dndu, dndu);
    Store3(out->dndv, dndv);
    Store3(out->shading.n, n);
    Store3(out->shading.dpdu, dpdu);
    Store3(out->shading.dpdv, dpdv);
    Store3(out->shading.dndu, dndu);
    Store3(out->shading.dndv, dndv);
    TransformSurfaceHit(sp.renderFromObject, sp.objectFromRender, out);
}

// TransformedPrimitive::Intersect (cpu/primitive.cpp:112-125) for a hit on triangle `triIndex` of an
// object's mesh (object space): the object's own Intersect saw the ray renderFromPrimitive->ApplyInverse(r)
// (its direction through mInv, transform.h:401-406), so the interaction's wo is minus that direction;
// the result is carried back by renderFromPrimitive.  dRender: the render-space ray's direction.
inline bool InstanceInteractionFromIntersection(const nnbvh_instance &in, const TriangleMeshView &mesh,
                                                int triIndex, float b0, float b1, float b2, float time,
                                                const float dRender[3], SurfaceHit *out) {
    using namespace detail;
    V3 dl = XfVector(in.primitiveFromRender, V3{dRender[0], dRender[1], dRender[2]});
    const float wl[3] = {-dl.x, -dl.y, -dl.z};
    if (!InteractionFromIntersection(mesh, triIndex, b0, b1, b2, time, wl, out)) return false;
    TransformSurfaceHit(in.renderFromPrimitive, in.primitiveFromRender, out);
    return true;
}

// ---- the aggregate -------------------------------------------------------------------------
// The scene is one TriangleMesh-like soup per primitive list entry: primitive i of the list given
// to the constructor is triangle `triIndex[i]` of mesh `meshIndex[i]` (Triangle::{meshIndex,
// triIndex}, shapes.h:1186-1189).  The Sink plays the part of the reference's queues:
//   sink.EnqueueWorkAfterMiss(int rayIndex)
//   sink.EnqueueWorkAfterIntersection(int rayIndex, uint32_t primitive, float tHit, const SurfaceHit &)
//   sink.RecordShadowRayResult(int rayIndex, bool foundIntersection)
// and is called from `hostThreads` threads at once, each ray index exactly once (the reference
// calls them from ParallelFor, aggregate.cpp:42, 64).
class GPUBVHWavefrontAggregate {
  public:
    struct PrimitiveRef {
        int meshIndex, triIndex;
    };

    GPUBVHWavefrontAggregate(std::vector<TriangleMeshView> meshes, std::vector<PrimitiveRef> prims,
                             int maxPrimsInNode = 4,
                             GPUBVHAggregate::SplitMethod sm = GPUBVHAggregate::SplitMethod::SAH,
                             int device = -1, int hostThreads = 0)
        : meshes_(std::move(meshes)), prims_(std::move(prims)),
          hostThreads_(hostThreads > 0 ? hostThreads
                                       : std::max(1u, std::thread::hardware_concurrency())) {
        agg_.reset(new GPUBVHAggregate(Soup(), maxPrimsInNode, sm, device));
    }

    // The same with `Shape "sphere"` primitives after the triangles: primitive prims.size() + k of the
    // list is spheres[k] (full, z-clipped or phimax-clipped).
    // sphereReverseOrientation[k]: the shape's ReverseOrientation flag (shapes.h:117-123), empty = none.
    GPUBVHWavefrontAggregate(std::vector<TriangleMeshView> meshes, std::vector<PrimitiveRef> prims,
                             std::vector<nnbvh_sphere> spheres, std::vector<uint8_t> sphereReverseOrientation,
                             int maxPrimsInNode = 4,
                             GPUBVHAggregate::SplitMethod sm = GPUBVHAggregate::SplitMethod::SAH,
                             int device = -1, int hostThreads = 0)
        : meshes_(std::move(meshes)), prims_(std::move(prims)), spheres_(std::move(spheres)),
          hostThreads_(hostThreads > 0 ? hostThreads
                                       : std::max(1u, std::thread::hardware_concurrency())) {
        if (!sphereReverseOrientation.empty() && sphereReverseOrientation.size() != spheres_.size())
            throw Error(NNBVH_ERR_INVALID, "one reverseOrientation flag per sphere");
        agg_.reset(new GPUBVHAggregate(Soup(), spheres_, maxPrimsInNode, sm, device));
        for (size_t k = 0; k < spheres_.size(); ++k)  // spheres_ is not resized after this: the pointers hold
            shapes_.push_back(MakeSphereShape(spheres_[k], !sphereReverseOrientation.empty() &&
                                                               sphereReverseOrientation[k] != 0));
    }

    // Object instances (TransformedPrimitive over a per-object BVHAggregate, cpu/primitive.cpp:112-130):
    // objects[j] is object j's own primitive list -- triangles of `meshes`, given in the OBJECT's space --
    // and primitive prims.size() + k of the top-level list is instances[k].
    GPUBVHWavefrontAggregate(std::vector<TriangleMeshView> meshes, std::vector<PrimitiveRef> prims,
                             std::vector<std::vector<PrimitiveRef>> objects, std::vector<nnbvh_instance> instances,
                             int maxPrimsInNode = 4,
                             GPUBVHAggregate::SplitMethod sm = GPUBVHAggregate::SplitMethod::SAH,
                             int device = -1, int hostThreads = 0)
        : meshes_(std::move(meshes)), prims_(std::move(prims)), objects_(std::move(objects)),
          instances_(std::move(instances)),
          hostThreads_(hostThreads > 0 ? hostThreads
                                       : std::max(1u, std::thread::hardware_concurrency())) {
        std::vector<std::vector<float>> objSoups;
        for (const std::vector<PrimitiveRef> &o : objects_) objSoups.push_back(Soup(o));
        agg_.reset(new GPUBVHAggregate(Soup(), objSoups, instances_, maxPrimsInNode, sm, device));
    }

    Bounds3f Bounds() const { return agg_->Bounds(); }  // WavefrontAggregate::Bounds, integrator.h:35
    const GPUBVHAggregate &Aggregate() const { return *agg_; }

    // IntersectClosest(maxRays, rayQ, escapedRayQ, hitAreaLightQ, basicMtlQ, universalMtlQ,
    // mediumSampleQ, nextRayQ), aggregate.cpp:34-58: every output queue is behind the sink.
    template <typename Sink>
    void IntersectClosest(int maxRays, const RayQueueView &q, Sink &sink) const {
        const int n = Count(maxRays, q.size);
        if (n == 0) return;
        hits_.resize((size_t)n);
        if (!instances_.empty()) {
            inner_.resize((size_t)n);
            agg_->IntersectBatchInstancedSoA(q.ox, q.oy, q.oz, q.dx, q.dy, q.dz, /*tMax=*/nullptr, (uint64_t)n,
                                             hits_.data(), inner_.data());
        } else
            agg_->IntersectBatchSoA(q.ox, q.oy, q.oz, q.dx, q.dy, q.dz, /*tMax=*/nullptr, (uint64_t)n,
                                    hits_.data());
        // wo = -ray.d needs the directions on the host; a queue in device memory is staged once.
        HostPlanes d(q.dx, q.dy, q.dz, q.time, n);
        ParallelFor(n, [&](int i) {
            const nnbvh_hit &h = hits_[(size_t)i];
            if (h.prim == NNBVH_MISS) {
                sink.EnqueueWorkAfterMiss(i);  // intersect.h:16-29
                return;
            }
            uint32_t prim = agg_->PrimitiveIndex(h.prim);
            float wo[3] = {-d.x[i], -d.y[i], -d.z[i]};
            SurfaceHit sh;
            if (prim >= prims_.size() && !instances_.empty()) {
                const nnbvh_instance &in = instances_[prim - prims_.size()];
                const PrimitiveRef &pr = objects_[in.object][agg_->Objects()[in.object].order[inner_[(size_t)i]]];
                const float dir[3] = {d.x[i], d.y[i], d.z[i]};
                if (!InstanceInteractionFromIntersection(in, meshes_[pr.meshIndex], pr.triIndex, h.b0, h.b1, h.b2,
                                                         d.t ? d.t[i] : 0.f, dir, &sh))
                    throw Error(NNBVH_ERR_INVALID, "hit on a triangle with a zero normal");
                sink.EnqueueWorkAfterIntersection(i, prim, h.t, sh);
                return;
            }
            if (prim >= prims_.size()) {  // a sphere leaf: {b0,b1,b2} = pObj (Sphere::Intersect, shapes.h:130-139)
                const float pObj[3] = {h.b0, h.b1, h.b2};
                SphereInteractionFromIntersection(shapes_[prim - prims_.size()], pObj, wo, d.t ? d.t[i] : 0.f, &sh);
                sink.EnqueueWorkAfterIntersection(i, prim, h.t, sh);
                return;
            }
            const PrimitiveRef &pr = prims_[prim];
            if (!InteractionFromIntersection(meshes_[pr.meshIndex], pr.triIndex, h.b0, h.b1, h.b2,
                                             d.t ? d.t[i] : 0.f, wo, &sh))
                throw Error(NNBVH_ERR_INVALID, "hit on a triangle with a zero normal");
            sink.EnqueueWorkAfterIntersection(i, prim, h.t, sh);  // intersect.h:48-163
        });
    }

    // IntersectShadow(maxRays, shadowRayQ, pixelSampleState), aggregate.cpp:60-68.
    template <typename Sink>
    void IntersectShadow(int maxRays, const ShadowRayQueueView &q, Sink &sink) const {
        const int n = Count(maxRays, q.size);
        if (n == 0) return;
        occ_.resize((size_t)n);
        agg_->IntersectPBatchSoA(q.ox, q.oy, q.oz, q.dx, q.dy, q.dz, q.tMax, (uint64_t)n, occ_.data());
        ParallelFor(n, [&](int i) { sink.RecordShadowRayResult(i, occ_[(size_t)i] != 0); });
    }

    // IntersectShadowTr (aggregate.cpp:70-92) walks medium interfaces with TraceTransmittance; for
    // a scene without participating media every interface is null and it reduces to IntersectShadow
    // with transmittance 1.  Scenes with media are outside this backend.
    template <typename Sink>
    void IntersectShadowTr(int maxRays, const ShadowRayQueueView &q, Sink &sink) const {
        IntersectShadow(maxRays, q, sink);
    }

  private:
    // the triangles of the primitive list as 9 floats each
    std::vector<float> Soup() const { return Soup(prims_); }
    std::vector<float> Soup(const std::vector<PrimitiveRef> &prims) const {
        std::vector<float> tris9(9 * prims.size());
        for (size_t i = 0; i < prims.size(); ++i) {
            const TriangleMeshView &m = meshes_[prims[i].meshIndex];
            const int *v = &m.vertexIndices[3 * prims[i].triIndex];
            for (int k = 0; k < 3; ++k)
                for (int c = 0; c < 3; ++c) tris9[9 * i + 3 * k + c] = m.p[3 * v[k] + c];
        }
        return tris9;
    }
    static int Count(int maxRays, int size) {
        if (size < 0 || size > maxRays)
            throw Error(NNBVH_ERR_INVALID, "queue size exceeds maxRays");  // the queues are allocated for maxQueueSize
        return size;
    }
    // the four planes the host loop reads, copied to the host when the queue is not host memory
    struct HostPlanes {
        const float *x, *y, *z, *t;
        std::vector<float> store;
        HostPlanes(const float *dx, const float *dy, const float *dz, const float *time, int n) {
            int kind = 0;
            Check(nnbvh_pointer_kind(dx, &kind));
            if (kind != 1) {  // host or managed memory: read in place
                x = dx, y = dy, z = dz, t = time;
                return;
            }
            store.resize((size_t)n * (time ? 4 : 3));
            const float *src[4] = {dx, dy, dz, time};
            for (int k = 0; k < (time ? 4 : 3); ++k)
                Check(nnbvh_copy_to_host(store.data() + (size_t)k * n, src[k], (uint64_t)n * 4));
            x = store.data(), y = x + n, z = y + n, t = time ? z + n : nullptr;
        }
    };
    template <typename F>
    void ParallelFor(int n, F body) const {
        int nt = std::min(hostThreads_, std::max(1, n / 4096));
        if (nt <= 1) {
            for (int i = 0; i < n; ++i) body(i);
            return;
        }
        std::atomic<int> next{0};
        std::atomic<bool> failed{false};
        std::string what;
        int code = 0;
        auto work = [&]() {
            try {
                for (;;) {
                    int b = next.fetch_add(4096);
                    if (b >= n || failed.load()) break;
                    for (int i = b, e = std::min(n, b + 4096); i < e; ++i) body(i);
                }
            } catch (const Error &e) {
                if (!failed.exchange(true)) what = e.what(), code = e.code;
            }
        };
        std::vector<std::thread> th;
        for (int t = 1; t < nt; ++t) th.emplace_back(work);
        work();
        for (auto &t : th) t.join();
        if (failed) throw Error(code, what);
    }

    std::vector<TriangleMeshView> meshes_;
    std::vector<PrimitiveRef> prims_;
    std::vector<nnbvh_sphere> spheres_;
    std::vector<SphereShape> shapes_;
    std::vector<std::vector<PrimitiveRef>> objects_;
    std::vector<nnbvh_instance> instances_;
    int hostThreads_;
    std::unique_ptr<GPUBVHAggregate> agg_;
    mutable std::vector<nnbvh_hit> hits_;
    mutable std::vector<uint32_t> inner_;
    mutable std::vector<uint8_t> occ_;
};

}  // namespace wavefront
}  // namespace nnbvh

#endif  // NNBVH_WAVEFRONT_HPP
