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
// `soac` generates -- by tests/test_integration_compile.py, and RUN by tests/test_binding.py: inside
// oracle/_ref/libnnbvh_ref_binding.so it is called through the WavefrontAggregate interface with the
// reference's own RayQueue / EscapedRayQueue / ShadowRayQueue / SOA<PixelSampleState>, beside
// CPUAggregate's loops over the reference BVHAggregate, and leaves the same queues and radiance behind.
// (Materials are null there: the full integrator around it needs all of pbrt.)
//
// The primitive list, the hit -> ShapeIntersection step and the friend access that step needs live in
// gpubvh.h (pbrt::GPUBVHAggregate); this class adds the queue plumbing.
//
// Contract: SimplePrimitive / alpha-free GeometricPrimitive over Triangle or Sphere.  TransformedPrimitive
// entries map to nnbvh_instance as INTEGRATION.md 2.1 says.
#ifndef NNBVH_INTEGRATION_PBRT_GPUBVH_WAVEFRONT_H
#define NNBVH_INTEGRATION_PBRT_GPUBVH_WAVEFRONT_H

#include <pbrt/util/parallel.h>
#include <pbrt/wavefront/integrator.h>
#include <pbrt/wavefront/intersect.h>

#include <memory>
#include <vector>

#include "gpubvh.h"

namespace pbrt {

class GPUBVHWavefrontAggregate : public WavefrontAggregate {
  public:
    // prims: the list CPUAggregate would hand to CreateAccelerator (cpu/aggregates.cpp:1163-1178).
    GPUBVHWavefrontAggregate(std::vector<Primitive> prims, int maxPrimsInNode,
                             BVHAggregate::SplitMethod splitMethod, int maxQueueSize, int device = -1)
        : host(std::move(prims), maxPrimsInNode, splitMethod, device), hits(maxQueueSize), occluded(maxQueueSize) {}

    Bounds3f Bounds() const override { return host.Bounds(); }

    // CPUAggregate::IntersectClosest, wavefront/aggregate.cpp:34-58
    void IntersectClosest(int maxRays, const RayQueue *rayQ, EscapedRayQueue *escapedRayQ,
                          HitAreaLightQueue *hitAreaLightQ, MaterialEvalQueue *basicMtlQ,
                          MaterialEvalQueue *universalMtlQ, MediumSampleQueue *mediumSampleQ,
                          RayQueue *nextRayQ) const override {
        const int n = rayQ->Size();
        if (n == 0)
            return;
        CHECK_LE((size_t)n, hits.size());  // the queues are allocated for maxQueueSize
        // SOA<RayWorkItem> keeps ray.o.x[], ray.o.y[] ... as separate float arrays: no repack
        host.Engine().IntersectBatchSoA(rayQ->ray.o.x, rayQ->ray.o.y, rayQ->ray.o.z, rayQ->ray.d.x, rayQ->ray.d.y,
                               rayQ->ray.d.z, /*tMax=*/nullptr, (uint64_t)n, hits.data());
        ParallelFor(0, n, [=](int index) {
            const RayWorkItem r = (*rayQ)[index];
            const nnbvh_hit &h = hits[index];
            if (h.prim == NNBVH_MISS)
                EnqueueWorkAfterMiss(r, mediumSampleQ, escapedRayQ);
            else {
                ShapeIntersection si = host.Resolve(h, r.ray);
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
        CHECK_LE((size_t)n, occluded.size());
        const ShadowRayQueue *q = shadowRayQueue;
        host.Engine().IntersectPBatchSoA(q->ray.o.x, q->ray.o.y, q->ray.o.z, q->ray.d.x, q->ray.d.y, q->ray.d.z, q->tMax,
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
                    si = host.Intersect(ray, tMax);
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
                pstd::optional<ShapeIntersection> si = host.Intersect(r, 1);
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

  private:
    GPUBVHAggregate host;
    mutable std::vector<nnbvh_hit> hits;  // one IntersectClosest / IntersectShadow is in flight at a time
    mutable std::vector<uint8_t> occluded;
};

}  // namespace pbrt

#endif  // NNBVH_INTEGRATION_PBRT_GPUBVH_WAVEFRONT_H
