// integration/pbrt/compile_check.cpp -- type-checks gpubvh_wavefront.h against the reference's real headers
// (tests/test_integration_compile.py; -fsyntax-only).  The friend declarations the adapter needs are
// replaced here by opening the three classes' private sections.
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
#include <pbrt/cpu/primitive.h>
#include <pbrt/shapes.h>
#undef private
#include "gpubvh_wavefront.h"

// what wavefront/integrator.cpp:157-170 would do with it
pbrt::WavefrontAggregate *MakeAggregate(std::vector<pbrt::Primitive> prims, int maxQueueSize) {
    return new pbrt::GPUBVHWavefrontAggregate(std::move(prims), 4, pbrt::BVHAggregate::SplitMethod::SAH,
                                              maxQueueSize);
}
