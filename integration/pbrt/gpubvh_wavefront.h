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

    // What CPUAggregate::IntersectShadowTr does (wavefront/aggregate.cpp:70-92): TraceTransmittance
    // (wavefront/intersect.h:165-240) walks a shadow ray across medium interfaces.  How many surfaces a ray
    // crosses depends on what it meets, so the walk goes ray by ray through the single-ray entry point.
    void IntersectShadowTr(int maxRays, ShadowRayQueue *shadowRayQueue,
                           SOA<PixelSampleState> *pixelSampleState) const override {
        auto walk = [=](int index) {
            pstd::optional<ShapeIntersection> last;  // the surface the walk currently stands on
            auto trace = [&](Ray ray, Float tMax) {
                last = host.Intersect(ray, tMax);
                TransmittanceTraceResult res{last.has_value(), Point3f(), Material()};
                if (last) {
                    res.pHit = last->intr.p();
                    res.material = last->intr.material;
                }
                return res;
            };
            auto spawnTo = [&](Point3f target) { return last->intr.SpawnRayTo(target); };
            TraceTransmittance((*shadowRayQueue)[index], pixelSampleState, trace, spawnTo);
        };
        ParallelFor(0, shadowRayQueue->Size(), walk);
    }

    // What CPUAggregate::IntersectOneRandom does (wavefront/aggregate.cpp:94-124): between p0 and p1 every
    // surface of the probe's material is a candidate of weight 1 for the subsurface exit point; the reservoir
    // is seeded by Hash(p0, p1).
    void IntersectOneRandom(int maxRays, SubsurfaceScatterQueue *subsurfaceScatterQueue) const override {
        SubsurfaceScatterQueue *q = subsurfaceScatterQueue;
        auto probe = [=](int index) {
            const SubsurfaceScatterWorkItem &item = (*q)[index];
            WeightedReservoirSampler<SubsurfaceInteraction> reservoir(Hash(item.p0, item.p1));
            for (Interaction from(item.p0, 0.f, Medium());;) {
                Ray segment = from.SpawnRayTo(item.p1);
                if (segment.d == Vector3f(0, 0, 0))
                    break;
                pstd::optional<ShapeIntersection> next = host.Intersect(segment, 1);
                if (!next)
                    break;
                if (next->intr.material == item.material)
                    reservoir.Add(SubsurfaceInteraction(next->intr), 1.f);
                from = next->intr;
            }
            q->reservoirPDF[index] = reservoir.HasSample() ? reservoir.SampleProbability() : Float(0);
            if (reservoir.HasSample())
                q->ssi[index] = reservoir.GetSample();
        };
        ParallelFor(0, q->Size(), probe);
    }

  private:
    GPUBVHAggregate host;
    mutable std::vector<nnbvh_hit> hits;  // one IntersectClosest / IntersectShadow is in flight at a time
    mutable std::vector<uint8_t> occluded;
};

}  // namespace pbrt

#endif  // NNBVH_INTEGRATION_PBRT_GPUBVH_WAVEFRONT_H
