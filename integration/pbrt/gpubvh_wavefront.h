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

#include <pbrt/materials.h>
#include <pbrt/util/parallel.h>
#include <pbrt/util/sampling.h>
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
        : host(std::move(prims), maxPrimsInNode, splitMethod, device), hits(maxQueueSize), inner(maxQueueSize, NNBVH_MISS), occluded(maxQueueSize) {}

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
        host.IntersectBatchSoA(rayQ->ray.o.x, rayQ->ray.o.y, rayQ->ray.o.z, rayQ->ray.d.x, rayQ->ray.d.y,
                               rayQ->ray.d.z, /*tMax=*/nullptr, (uint64_t)n, hits.data(), inner.data());
        ParallelFor(0, n, [=](int index) {
            const RayWorkItem r = (*rayQ)[index];
            const nnbvh_hit &h = hits[index];
            if (h.prim == NNBVH_MISS)
                EnqueueWorkAfterMiss(r, mediumSampleQ, escapedRayQ);
            else {
                ShapeIntersection si = host.Resolve(h, inner[index], r.ray);
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

    // CPUAggregate::IntersectShadowTr (wavefront/aggregate.cpp:70-92), batched.  The reference walks each shadow
    // ray across the medium interfaces it meets with TraceTransmittance (wavefront/intersect.h:165-275), asking for
    // one closest hit per segment.  Here every still-walking ray's NEXT segment goes to the GPU in one batch per
    // round (as the reference's own GPU backend does per ray inside one launch, gpu/optix.cu:268-296):
    //   round k: each live walk is REPLAYED from its start by the reference's unmodified TraceTransmittance, whose
    //   `trace` callback answers the first k segments from the hits of the earlier rounds; when it asks for a
    //   segment nobody has traced yet, the segment is recorded and the callback answers "opaque surface", which
    //   ends that replay with T_ray = 0 and therefore WITHOUT touching pixelSampleState (intersect.h:263-274).
    //   The replay that gets through with known segments only is the real one and records its radiance once.
    // TraceTransmittance is deterministic in its inputs (its RNG is seeded from the ray, :181), so every replay
    // repeats the same prefix; a walk over s surfaces costs s + 1 rounds and O(s^2) host steps, s being 1-3 in
    // practice.  GPU batches per call = 1 + the longest walk's surface count, not one per segment per ray.
    void IntersectShadowTr(int maxRays, ShadowRayQueue *shadowRayQueue,
                           SOA<PixelSampleState> *pixelSampleState) const override {
        const int n = shadowRayQueue->Size();
        if (n == 0)
            return;
        struct Walk {
            std::vector<nnbvh_hit> known;  // hits of the segments traced so far, in walk order
            std::vector<uint32_t> knownInner;  // ... and the primitive inside the object, for instance hits
            nnbvh_ray query;               // the next segment, when the last replay stopped for it
            bool done = false;
        };
        std::vector<Walk> walks((size_t)n);
        std::vector<int> live((size_t)n);
        for (int i = 0; i < n; ++i)
            live[i] = i;
        std::vector<nnbvh_ray> rays;
        std::vector<nnbvh_hit> segHits;
        std::vector<uint32_t> segInner;
        lastTrBatches = 0;
        while (!live.empty()) {
            const int m = (int)live.size();
            ParallelFor(0, m, [&](int j) {
                Walk &w = walks[live[j]];
                size_t asked = 0;
                bool pending = false;
                pstd::optional<ShapeIntersection> last;  // the surface the walk currently stands on
                auto trace = [&](Ray ray, Float tMax) {
                    if (asked < w.known.size()) {
                        const nnbvh_hit &h = w.known[asked];
                        const uint32_t in = w.knownInner[asked++];
                        last.reset();
                        if (h.prim != NNBVH_MISS)
                            last = host.Resolve(h, in, ray);
                        return TransmittanceTraceResult{last.has_value(), last ? last->intr.p() : Point3f(),
                                                        last ? last->intr.material : Material()};
                    }
                    w.query = nnbvh_ray{ray.o.x, ray.o.y, ray.o.z, tMax, ray.d.x, ray.d.y, ray.d.z, ray.time};
                    pending = true;
                    return TransmittanceTraceResult{true, Point3f(), OpaqueSentinel()};  // ends this replay, no side effect
                };
                auto spawnTo = [&](Point3f target) { return last->intr.SpawnRayTo(target); };
                TraceTransmittance((*shadowRayQueue)[live[j]], pixelSampleState, trace, spawnTo);
                w.done = !pending;
            });
            std::vector<int> next;
            next.reserve(live.size());
            for (int i : live)
                if (!walks[i].done)
                    next.push_back(i);
            live.swap(next);
            if (live.empty())
                break;
            rays.resize(live.size());
            segHits.resize(live.size());
            for (size_t j = 0; j < live.size(); ++j)
                rays[j] = walks[live[j]].query;
            segInner.assign(live.size(), NNBVH_MISS);
            host.IntersectBatch(rays.data(), (uint64_t)rays.size(), segHits.data(), segInner.data());
            ++lastTrBatches;
            for (size_t j = 0; j < live.size(); ++j) {
                walks[live[j]].known.push_back(segHits[j]);
                walks[live[j]].knownInner.push_back(segInner[j]);
            }
        }
    }

    // CPUAggregate::IntersectOneRandom (wavefront/aggregate.cpp:94-124), batched the same way (the reference's GPU
    // backend: gpu/optix.cu:480-573).  Between p0 and p1 every surface of the probe's material is a candidate of
    // weight 1 for the subsurface exit point; the reservoir is seeded by Hash(p0, p1).  The probe loop is short
    // enough to be kept as a per-probe state (the interaction the probe stands on + its reservoir) that advances
    // by one segment per round; one closest-hit batch over the still-walking probes per round.
    void IntersectOneRandom(int maxRays, SubsurfaceScatterQueue *subsurfaceScatterQueue) const override {
        SubsurfaceScatterQueue *q = subsurfaceScatterQueue;
        const int n = q->Size();
        if (n == 0)
            return;
        struct Probe {
            Interaction from;
            WeightedReservoirSampler<SubsurfaceInteraction> reservoir;
            Ray segment;
            bool live = true;
        };
        std::vector<Probe> probes((size_t)n);
        ParallelFor(0, n, [&](int i) {
            const SubsurfaceScatterWorkItem &item = (*q)[i];
            probes[i].from = Interaction(item.p0, 0.f, Medium());
            probes[i].reservoir.Seed(Hash(item.p0, item.p1));
        });
        std::vector<int> live((size_t)n);
        for (int i = 0; i < n; ++i)
            live[i] = i;
        std::vector<nnbvh_ray> rays;
        std::vector<nnbvh_hit> segHits;
        std::vector<uint32_t> segInner;
        lastOneRandomBatches = 0;
        while (true) {
            // the next segment of every live probe (`if (r.d == 0) break`, aggregate.cpp:104-106)
            std::vector<int> next;
            next.reserve(live.size());
            for (int i : live) {
                Probe &p = probes[i];
                p.segment = p.from.SpawnRayTo(Point3f(q->p1[i]));
                if (p.segment.d == Vector3f(0, 0, 0))
                    p.live = false;
                else
                    next.push_back(i);
            }
            live.swap(next);
            if (live.empty())
                break;
            rays.resize(live.size());
            segHits.resize(live.size());
            for (size_t j = 0; j < live.size(); ++j) {
                const Ray &r = probes[live[j]].segment;
                rays[j] = nnbvh_ray{r.o.x, r.o.y, r.o.z, 1.f, r.d.x, r.d.y, r.d.z, r.time};
            }
            segInner.assign(live.size(), NNBVH_MISS);
            host.IntersectBatch(rays.data(), (uint64_t)rays.size(), segHits.data(), segInner.data());
            ++lastOneRandomBatches;
            ParallelFor(0, (int64_t)live.size(), [&](int64_t j) {
                Probe &p = probes[live[j]];
                if (segHits[j].prim == NNBVH_MISS) {
                    p.live = false;
                    return;
                }
                ShapeIntersection si = host.Resolve(segHits[j], segInner[j], p.segment);
                p.from = si.intr;
                if (si.intr.material == Material(q->material[live[j]]))
                    p.reservoir.Add(SubsurfaceInteraction(si.intr), 1.f);
            });
            next.clear();
            for (int i : live)
                if (probes[i].live)
                    next.push_back(i);
            live.swap(next);
        }
        ParallelFor(0, n, [&](int i) {
            Probe &p = probes[i];
            q->reservoirPDF[i] = p.reservoir.HasSample() ? p.reservoir.SampleProbability() : Float(0);
            if (p.reservoir.HasSample())
                q->ssi[i] = p.reservoir.GetSample();
        });
    }

    // GPU batches the last IntersectShadowTr / IntersectOneRandom call issued (tests: <= longest walk + 1)
    int LastShadowTrBatches() const { return lastTrBatches; }
    int LastOneRandomBatches() const { return lastOneRandomBatches; }

  private:
    // A Material that is only ever compared with null: TraceTransmittance reads `result.hit && result.material`
    // (wavefront/intersect.h:192) and never dereferences it.
    static Material OpaqueSentinel() {
        static DiffuseMaterial *const tagOnly = reinterpret_cast<DiffuseMaterial *>(uintptr_t(64));
        return Material(tagOnly);
    }

    GPUBVHAggregate host;
    // One call of this aggregate is in flight at a time (the wavefront render loop is sequential,
    // wavefront/integrator.cpp); the batch buffers and counters below rely on it.
    mutable std::vector<nnbvh_hit> hits;
    mutable std::vector<uint32_t> inner;  // primitive inside the object, for hits on instances
    mutable std::vector<uint8_t> occluded;
    mutable int lastTrBatches = 0, lastOneRandomBatches = 0;
};

}  // namespace pbrt

#endif  // NNBVH_INTEGRATION_PBRT_GPUBVH_WAVEFRONT_H
