// integration/pbrt/gpubvh_wavefront.h -- the reference-side binding of the wavefront batch backend.
//
// What a maintainer of the reference adds as src/pbrt/wavefront/gpubvh_aggregate.h: a
// WavefrontAggregate (/root/reference/src/pbrt/wavefront/integrator.h:30-53) whose batch calls go to
// libnnbvh.so through include/nnbvh.hpp, and whose per-ray follow-up is the reference's own code:
//   Triangle::InteractionFromIntersection   shapes.h:884-1010   (public static)
//   SurfaceInteraction::SetIntersectionProperties  interaction.h:218-229
//   EnqueueWorkAfterMiss / RecordShadowRayResult / EnqueueWorkAfterIntersection /
//   TraceTransmittance                      wavefront/intersect.h:16-163, :165-240
// It replaces CPUAggregate (wavefront/aggregate.cpp:34-92) where WavefrontPathIntegrator picks its
// aggregate (wavefront/integrator.cpp:157-170).
//
// This file is compiled against the reference's real headers -- including the SOA queue types its
// `soac` generates -- by tests/test_integration_compile.py (type-checked: every `override` below must
// match WavefrontAggregate's signature).  It cannot be linked here: that needs all of pbrt.
//
// Friend access it needs in the reference (three one-line patches):
//   cpu/primitive.h:58  class GeometricPrimitive { friend class GPUBVHWavefrontAggregate; ...
//   cpu/primitive.h:76  class SimplePrimitive    { friend class GPUBVHWavefrontAggregate; ...
//   shapes.h:1179       class Triangle           { friend class GPUBVHWavefrontAggregate; ...
// (the compile check opens them with `#define private public` instead of patching).
//
// Contract: SimplePrimitive / alpha-free GeometricPrimitive over Triangle.  Sphere and
// TransformedPrimitive entries map to nnbvh_sphere / nnbvh_instance as INTEGRATION.md 2.1 says.
#ifndef NNBVH_INTEGRATION_PBRT_GPUBVH_WAVEFRONT_H
#define NNBVH_INTEGRATION_PBRT_GPUBVH_WAVEFRONT_H

#include <pbrt/cpu/aggregates.h>
#include <pbrt/cpu/primitive.h>
#include <pbrt/shapes.h>
#include <pbrt/util/parallel.h>
#include <pbrt/wavefront/integrator.h>
#include <pbrt/wavefront/intersect.h>

#include <memory>
#include <vector>

#include <nnbvh.hpp>

namespace pbrt {

class GPUBVHWavefrontAggregate : public WavefrontAggregate {
  public:
    // prims: the list CPUAggregate would hand to CreateAccelerator (cpu/aggregates.cpp:1163-1178).
    GPUBVHWavefrontAggregate(std::vector<Primitive> prims, int maxPrimsInNode,
                             BVHAggregate::SplitMethod splitMethod, int maxQueueSize, int device = -1)
        : primitives(std::move(prims)), hits(maxQueueSize), occluded(maxQueueSize) {
        std::vector<float> tris9;
        tris9.reserve(9 * primitives.size());
        for (const Primitive &p : primitives) {
            Parts parts = PartsOf(p);
            if (!parts.tri)
                ErrorExit("gpubvh: only (alpha-free) triangle primitives are supported");
            const TriangleMesh *mesh = parts.tri->GetMesh();
            const int *v = &mesh->vertexIndices[3 * parts.tri->triIndex];
            for (int k = 0; k < 3; ++k) {
                Point3f q = mesh->p[v[k]];
                tris9.push_back(q.x), tris9.push_back(q.y), tris9.push_back(q.z);
            }
        }
        // BVHAggregate::SplitMethod { SAH, HLBVH, Middle, EqualCounts } (cpu/aggregates.h:33) has the
        // library's numbering
        agg = std::make_unique<nnbvh::GPUBVHAggregate>(
            tris9, maxPrimsInNode, nnbvh::GPUBVHAggregate::SplitMethod((int)splitMethod), device);
    }

    Bounds3f Bounds() const override {
        nnbvh::Bounds3f b = agg->Bounds();
        return Bounds3f(Point3f(b.pMin[0], b.pMin[1], b.pMin[2]), Point3f(b.pMax[0], b.pMax[1], b.pMax[2]));
    }

    // CPUAggregate::IntersectClosest, wavefront/aggregate.cpp:34-58
    void IntersectClosest(int maxRays, const RayQueue *rayQ, EscapedRayQueue *escapedRayQ,
                          HitAreaLightQueue *hitAreaLightQ, MaterialEvalQueue *basicMtlQ,
                          MaterialEvalQueue *universalMtlQ, MediumSampleQueue *mediumSampleQ,
                          RayQueue *nextRayQ) const override {
        const int n = rayQ->Size();
        if (n == 0)
            return;
        // SOA<RayWorkItem> keeps ray.o.x[], ray.o.y[] ... as separate float arrays: no repack
        agg->IntersectBatchSoA(rayQ->ray.o.x, rayQ->ray.o.y, rayQ->ray.o.z, rayQ->ray.d.x, rayQ->ray.d.y,
                               rayQ->ray.d.z, /*tMax=*/nullptr, (uint64_t)n, hits.data());
        ParallelFor(0, n, [=](int index) {
            const RayWorkItem r = (*rayQ)[index];
            const nnbvh_hit &h = hits[index];
            if (h.prim == NNBVH_MISS)
                EnqueueWorkAfterMiss(r, mediumSampleQ, escapedRayQ);
            else {
                ShapeIntersection si = Resolve(h, r.ray);
                EnqueueWorkAfterIntersection(r, r.ray.medium, si.tHit, si.intr, mediumSampleQ, nextRayQ,
                                             hitAreaLightQ, basicMtlQ, universalMtlQ);
            }
        });
    }

    // CPUAggregate::IntersectShadow, wavefront/aggregate.cpp:60-68
    void IntersectShadow(int maxRays, ShadowRayQueue *shadowRayQueue,
                         SOA<PixelSampleState> *pixelSampleState) const override {
        const int n = shadowRayQueue->Size();
        if (n == 0)
            return;
        const ShadowRayQueue *q = shadowRayQueue;
        agg->IntersectPBatchSoA(q->ray.o.x, q->ray.o.y, q->ray.o.z, q->ray.d.x, q->ray.d.y, q->ray.d.z, q->tMax,
                                (uint64_t)n, occluded.data());
        ParallelFor(0, n, [=](int index) {
            RecordShadowRayResult((*shadowRayQueue)[index], pixelSampleState, occluded[index] != 0);
        });
    }

    // CPUAggregate::IntersectShadowTr, wavefront/aggregate.cpp:70-92.  Scenes without media reduce to
    // IntersectShadow; with media the walk over interfaces is data dependent and goes ray by ray.
    void IntersectShadowTr(int maxRays, ShadowRayQueue *shadowRayQueue,
                           SOA<PixelSampleState> *pixelSampleState) const override {
        ParallelFor(0, shadowRayQueue->Size(), [=](int index) {
            const ShadowRayWorkItem w = (*shadowRayQueue)[index];
            pstd::optional<ShapeIntersection> si;
            TraceTransmittance(
                w, pixelSampleState,
                [&](Ray ray, Float tMax) -> TransmittanceTraceResult {
                    si = IntersectOne(ray, tMax);
                    if (!si)
                        return TransmittanceTraceResult{false, Point3f(), Material()};
                    return TransmittanceTraceResult{true, si->intr.p(), si->intr.material};
                },
                [&](Point3f p) -> Ray { return si->intr.SpawnRayTo(p); });
        });
    }

    // CPUAggregate::IntersectOneRandom, wavefront/aggregate.cpp:94-124
    void IntersectOneRandom(int maxRays, SubsurfaceScatterQueue *subsurfaceScatterQueue) const override {
        ParallelFor(0, subsurfaceScatterQueue->Size(), [=](int index) {
            const SubsurfaceScatterWorkItem &w = (*subsurfaceScatterQueue)[index];
            uint64_t seed = Hash(w.p0, w.p1);
            WeightedReservoirSampler<SubsurfaceInteraction> wrs(seed);
            Interaction base(w.p0, 0.f, Medium());
            while (true) {
                Ray r = base.SpawnRayTo(w.p1);
                if (r.d == Vector3f(0, 0, 0))
                    break;
                pstd::optional<ShapeIntersection> si = IntersectOne(r, 1);
                if (!si)
                    break;
                base = si->intr;
                if (si->intr.material == w.material)
                    wrs.Add(SubsurfaceInteraction(si->intr), 1.f);
            }
            if (wrs.HasSample()) {
                subsurfaceScatterQueue->reservoirPDF[index] = wrs.SampleProbability();
                subsurfaceScatterQueue->ssi[index] = wrs.GetSample();
            } else
                subsurfaceScatterQueue->reservoirPDF[index] = 0;
        });
    }

    // {prim, t, b0, b1, b2} -> what Primitive::Intersect would have returned
    // (SimplePrimitive / GeometricPrimitive::Intersect, cpu/primitive.cpp:50-84, :100-110)
    ShapeIntersection Resolve(const nnbvh_hit &h, const Ray &ray) const {
        Parts parts = PartsOf(primitives[agg->PrimitiveIndex(h.prim)]);
        SurfaceInteraction intr = Triangle::InteractionFromIntersection(
            parts.tri->GetMesh(), parts.tri->triIndex, TriangleIntersection{h.b0, h.b1, h.b2, h.t}, ray.time,
            -ray.d);
        intr.SetIntersectionProperties(parts.material, parts.areaLight, parts.mediumInterface, ray.medium);
        return ShapeIntersection{intr, h.t};
    }

  private:
    struct Parts {
        const Triangle *tri = nullptr;
        Material material;
        Light areaLight;
        const MediumInterface *mediumInterface = nullptr;
    };
    static Parts PartsOf(const Primitive &p) {
        Parts out;
        if (const SimplePrimitive *sp = p.CastOrNullptr<SimplePrimitive>()) {
            out.tri = sp->shape.CastOrNullptr<Triangle>();
            out.material = sp->material;
        } else if (const GeometricPrimitive *gp = p.CastOrNullptr<GeometricPrimitive>()) {
            if (!gp->alpha)
                out.tri = gp->shape.CastOrNullptr<Triangle>();
            out.material = gp->material;
            out.areaLight = gp->areaLight;
            out.mediumInterface = &gp->mediumInterface;
        }
        return out;
    }
    // one ray through the batch entry point (the data-dependent loops above)
    pstd::optional<ShapeIntersection> IntersectOne(const Ray &ray, Float tMax) const {
        nnbvh_ray r{ray.o.x, ray.o.y, ray.o.z, tMax, ray.d.x, ray.d.y, ray.d.z, ray.time};
        nnbvh_hit h;
        agg->IntersectBatch(&r, 1, &h);
        if (h.prim == NNBVH_MISS)
            return {};
        return Resolve(h, ray);
    }

    std::vector<Primitive> primitives;  // caller's order: hit.prim -> primitives[agg->PrimitiveIndex(prim)]
    std::unique_ptr<nnbvh::GPUBVHAggregate> agg;
    mutable std::vector<nnbvh_hit> hits;  // one IntersectClosest / IntersectShadow is in flight at a time
    mutable std::vector<uint8_t> occluded;
};

}  // namespace pbrt

#endif  // NNBVH_INTEGRATION_PBRT_GPUBVH_WAVEFRONT_H
