// integration/pbrt/compile_check.cpp -- type-checks gpubvh.h and gpubvh_wavefront.h against the reference's real headers
// (tests/test_integration_compile.py; -fsyntax-only).  The friend declarations the adapter needs are
// replaced here by opening the classes' private sections.
#include <algorithm>
#include <atomic>
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
#define private public
#include <pbrt/pbrt.h>
#include <pbrt/cpu/aggregates.h>
#include <pbrt/cpu/primitive.h>
#include <pbrt/shapes.h>
#undef private
#include "gpubvh_wavefront.h"

// what CreateAccelerator (cpu/aggregates.cpp:1163-1178) would do with it
pbrt::GPUBVHAggregate *MakeAccelerator(std::vector<pbrt::Primitive> prims, const pbrt::ParameterDictionary &dict) {
    pbrt::GPUBVHAggregate *accel = pbrt::GPUBVHAggregate::Create(std::move(prims), dict);
    pbrt::Ray ray(pbrt::Point3f(0, 0, 0), pbrt::Vector3f(0, 0, 1));
    if (accel->IntersectP(ray, pbrt::Infinity))
        (void)accel->Intersect(ray, pbrt::Infinity)->intr.p();
    return accel;
}

// what wavefront/integrator.cpp:157-170 would do with it
pbrt::WavefrontAggregate *MakeAggregate(std::vector<pbrt::Primitive> prims, int maxQueueSize) {
    return new pbrt::GPUBVHWavefrontAggregate(std::move(prims), 4, pbrt::BVHAggregate::SplitMethod::SAH,
                                              maxQueueSize);
}
