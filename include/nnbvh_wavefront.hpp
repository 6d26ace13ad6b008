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
// tests/test_wavefront.py pins it bit for bit against the reference's own function.
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
    Store3(out->wo, V3{wo[0], wo[1], wo[2]});
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
        std::vector<float> tris9(9 * prims_.size());
        for (size_t i = 0; i < prims_.size(); ++i) {
            const TriangleMeshView &m = meshes_[prims_[i].meshIndex];
            const int *v = &m.vertexIndices[3 * prims_[i].triIndex];
            for (int k = 0; k < 3; ++k)
                for (int c = 0; c < 3; ++c) tris9[9 * i + 3 * k + c] = m.p[3 * v[k] + c];
        }
        agg_.reset(new GPUBVHAggregate(tris9, maxPrimsInNode, sm, device));
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
            const PrimitiveRef &pr = prims_[prim];
            float wo[3] = {-d.x[i], -d.y[i], -d.z[i]};
            SurfaceHit sh;
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
    int hostThreads_;
    std::unique_ptr<GPUBVHAggregate> agg_;
    mutable std::vector<nnbvh_hit> hits_;
    mutable std::vector<uint8_t> occ_;
};

}  // namespace wavefront
}  // namespace nnbvh

#endif  // NNBVH_WAVEFRONT_HPP
