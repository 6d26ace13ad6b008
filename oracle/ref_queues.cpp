// oracle/ref_queues.cpp -- TEST INFRASTRUCTURE (checker), not product code.
//
// The reference's own wavefront queue types, instantiated from its own headers:
//   RayQueue        = WorkQueue<RayWorkItem>        /root/reference/src/pbrt/wavefront/workitems.h:330-343
//   ShadowRayQueue  = WorkQueue<ShadowRayWorkItem>  workitems.h:165-172, :386
// SOA<RayWorkItem> / SOA<ShadowRayWorkItem> come from wavefront_workitems_soa.h and pbrt_soa.h, which the
// reference's build generates with its `soac` tool from wavefront/workitems.soa and pbrt.soa
// (CMakeLists.txt:795-813); oracle/Makefile compiles soac from the reference's source and runs it the
// same way (outputs under oracle/_ref/gen).
//
// Rays are pushed with the reference's own PushCameraRay / Push, and the addresses of the planes a
// WavefrontAggregate reads (ray.o.{x,y,z}, ray.d.{x,y,z}, ray.time, tMax) are reported, so the tests
// can hand THE REFERENCE'S queue memory to include/nnbvh_wavefront.hpp's RayQueueView /
// ShadowRayQueueView: "a queue is passed without gather or repack" is checked on the real type.
#include <pbrt/wavefront/workitems.h>

#include <cstddef>
#include <cstdint>

namespace {
pbrt::Ray QueueRay(const float *r) {
    return pbrt::Ray(pbrt::Point3f(r[0], r[1], r[2]), pbrt::Vector3f(r[4], r[5], r[6]), r[7]);
}
}  // namespace

extern "C" {

void *nnbvh_ref_rayqueue_create(int maxQueueSize) {
    return new pbrt::RayQueue(maxQueueSize, pbrt::Allocator());
}
// the planes stay with the default memory resource: these live in a test process
void nnbvh_ref_rayqueue_destroy(void *q) { delete (pbrt::RayQueue *)q; }
void nnbvh_ref_rayqueue_reset(void *q) { ((pbrt::RayQueue *)q)->Reset(); }
// rays8: {o, tMax (unused: closest-hit queues carry none), d, time}.  Returns the queue's size afterwards.
int nnbvh_ref_rayqueue_push(void *queue, const float *rays8, int n) {
    pbrt::RayQueue *q = (pbrt::RayQueue *)queue;
    pbrt::SampledWavelengths lambda = pbrt::SampledWavelengths::SampleUniform(0.5f);
    for (int i = 0; i < n; ++i)
        q->PushCameraRay(QueueRay(rays8 + 8 * (size_t)i), lambda, i);
    return q->Size();
}
// planes7: ray.o.x, ray.o.y, ray.o.z, ray.d.x, ray.d.y, ray.d.z, ray.time
int nnbvh_ref_rayqueue_planes(void *queue, const float **planes7) {
    pbrt::RayQueue *q = (pbrt::RayQueue *)queue;
    planes7[0] = q->ray.o.x, planes7[1] = q->ray.o.y, planes7[2] = q->ray.o.z;
    planes7[3] = q->ray.d.x, planes7[4] = q->ray.d.y, planes7[5] = q->ray.d.z;
    planes7[6] = q->ray.time;
    return q->Size();
}
// What the reference's consumer reads back: (*rayQ)[i].ray (CPUAggregate::IntersectClosest, aggregate.cpp:44-46)
void nnbvh_ref_rayqueue_get(void *queue, int i, float *ray8) {
    pbrt::RayWorkItem r = (*(pbrt::RayQueue *)queue)[i];
    ray8[0] = r.ray.o.x, ray8[1] = r.ray.o.y, ray8[2] = r.ray.o.z, ray8[3] = 0;
    ray8[4] = r.ray.d.x, ray8[5] = r.ray.d.y, ray8[6] = r.ray.d.z, ray8[7] = r.ray.time;
}

void *nnbvh_ref_shadowqueue_create(int maxQueueSize) {
    return new pbrt::ShadowRayQueue(maxQueueSize, pbrt::Allocator());
}
void nnbvh_ref_shadowqueue_destroy(void *q) { delete (pbrt::ShadowRayQueue *)q; }
int nnbvh_ref_shadowqueue_push(void *queue, const float *rays8, int n) {
    pbrt::ShadowRayQueue *q = (pbrt::ShadowRayQueue *)queue;
    pbrt::SampledWavelengths lambda = pbrt::SampledWavelengths::SampleUniform(0.5f);
    for (int i = 0; i < n; ++i) {
        const float *r = rays8 + 8 * (size_t)i;
        pbrt::ShadowRayWorkItem w;
        w.ray = QueueRay(r);
        w.tMax = r[3];
        w.lambda = lambda;
        w.Ld = w.r_u = w.r_l = pbrt::SampledSpectrum(1.f);
        w.pixelIndex = i;
        q->Push(w);
    }
    return q->Size();
}
// planes8: ray.o.{x,y,z}, ray.d.{x,y,z}, ray.time, tMax
int nnbvh_ref_shadowqueue_planes(void *queue, const float **planes8) {
    pbrt::ShadowRayQueue *q = (pbrt::ShadowRayQueue *)queue;
    planes8[0] = q->ray.o.x, planes8[1] = q->ray.o.y, planes8[2] = q->ray.o.z;
    planes8[3] = q->ray.d.x, planes8[4] = q->ray.d.y, planes8[5] = q->ray.d.z;
    planes8[6] = q->ray.time;
    planes8[7] = q->tMax;
    return q->Size();
}

}  // extern "C"
