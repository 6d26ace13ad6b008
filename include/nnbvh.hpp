// include/nnbvh.hpp -- C++ host mirror of the reference's BVHAggregate over the C ABI (nnbvh.h).
//
// Header-only; link against libnnbvh.so.  Same surface as `class BVHAggregate`
// (/root/reference/src/pbrt/cpu/aggregates.h:28-70): a constructor taking the
// primitives, maxPrimsInNode and SplitMethod; Bounds(); Intersect(ray, tMax);
// IntersectP(ray, tMax) -- plus the batch forms the reference only has in
// CPUAggregate::IntersectClosest / IntersectShadow (wavefront/aggregate.cpp:34-68).
// No intersection arithmetic lives here: every query runs in libnnbvh.so on the GPU,
// and every failure becomes a nnbvh::Error (the reference LOG_FATALs instead).
//
// What pbrt would pass in: `tris` = the 9 floats of each Triangle's vertices in
// render space (Triangle::GetMesh()->p[v[k]], shapes.cpp:325-327) in the order
// of the std::vector<Primitive> given to BVHAggregate; `PrimitiveIndex(i)` maps a
// hit record back to that order, where the caller re-attaches material / area
// light (SimplePrimitive / GeometricPrimitive, cpu/primitive.cpp:50-109) and
// rebuilds the SurfaceInteraction with Triangle::InteractionFromIntersection
// (shapes.h:884) from {t, b0, b1, b2}.
#ifndef NNBVH_HPP
#define NNBVH_HPP

#include <algorithm>
#include <cstddef>
#include <cmath>
#include <cstdint>
#include <limits>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "nnbvh.h"

namespace nnbvh {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string &what) : std::runtime_error(what), code(c) {}
};

inline void Check(int rc) {
    if (rc != NNBVH_OK) throw Error(rc, std::string("libnnbvh: ") + nnbvh_last_error());
}

// Ray{o, d, time} (ray.h:35-38).
struct Ray {
    float o[3];
    float d[3];
    float time = 0.f;
};

// What survives of ShapeIntersection (shapes.h:93-97) + TriangleIntersection (shapes.h:820-824).
struct Hit {
    uint32_t prim;  // ORDERED primitive index; PrimitiveIndex() maps it back
    float tHit, b0, b1, b2;
};

struct Bounds3f {
    float pMin[3], pMax[3];
};

// nnbvh_sphere takes phimax in degrees, as `Shape "sphere"` gives it; Sphere keeps
// phiMax = Radians(Clamp(phimax, 0, 360)) (shapes.h:117-128, util/math.h:261-263).  A binding that only has
// the Sphere object recovers a degree value that maps to the SAME float: float(Degrees(phiMax)) or one of
// its neighbours (the two grids have nearly the same spacing; a preimage exists because phiMax came from one).
inline float SpherePhiMaxDegrees(float phiMaxRadians) {
    const float kPi = 3.14159265358979323846f;
    auto radians = [kPi](float deg) { return (kPi / 180) * (deg < 0 ? 0.f : (deg > 360 ? 360.f : deg)); };
    float deg = (180 / kPi) * phiMaxRadians;
    if (radians(deg) == phiMaxRadians) return deg;
    float up = deg, down = deg;
    for (int k = 0; k < 16; ++k) {
        up = std::nextafter(up, 1e9f), down = std::nextafter(down, -1e9f);
        if (radians(up) == phiMaxRadians) return up;
        if (radians(down) == phiMaxRadians) return down;
    }
    return deg;  // not produced by Radians(float): the closest value
}

inline nnbvh_ray MakeRay(const Ray &r, float tMax) {
    return nnbvh_ray{r.o[0], r.o[1], r.o[2], tMax, r.d[0], r.d[1], r.d[2], r.time};
}

class GPUBVHAggregate {
  public:
    enum class SplitMethod { SAH = NNBVH_SPLIT_SAH, HLBVH = NNBVH_SPLIT_HLBVH,
                             Middle = NNBVH_SPLIT_MIDDLE, EqualCounts = NNBVH_SPLIT_EQUAL_COUNTS };

    // BVHAggregate(std::vector<Primitive> p, int maxPrimsInNode = 1, SplitMethod = SAH),
    // cpu/aggregates.cpp:140-190.  tris9: 9 floats per triangle, caller's primitive order.
    GPUBVHAggregate(const std::vector<float> &tris9, int maxPrimsInNode = 1,
                    SplitMethod splitMethod = SplitMethod::SAH, int device = -1,
                    int buildThreads = 0) {
        const uint32_t nTris = (uint32_t)(tris9.size() / 9);
        CreateEngine(device);
        if (nTris) {
            nodes_.resize(2 * (size_t)nTris - 1);
            order_.resize(nTris);
            uint32_t nNodes = 0;
            // HLBVH is the builder with device stages (buildHLBVH, cpu/aggregates.cpp:389-449)
            int rc = splitMethod == SplitMethod::HLBVH
                         ? nnbvh_build_hlbvh_device(e_, tris9.data(), nTris, (uint32_t)maxPrimsInNode,
                                                    nodes_.data(), &nNodes, order_.data())
                         : nnbvh_build_bvh(tris9.data(), nTris, maxPrimsInNode, (int)splitMethod,
                                           buildThreads, nodes_.data(), &nNodes, order_.data());
            if (rc != NNBVH_OK) Fail(rc);
            nodes_.resize(nNodes);
            ordered_.resize(9 * (size_t)nTris);
            for (uint32_t i = 0; i < nTris; ++i)
                for (int k = 0; k < 9; ++k)
                    ordered_[9 * (size_t)i + k] = tris9[9 * (size_t)order_[i] + k];
        }
        Upload();
    }

    // A primitive list that mixes triangles and spheres (SimplePrimitive over Triangle / Sphere,
    // shapes.h:107-236): caller primitive i < nTris is triangle i, primitive nTris + k is sphere k.
    // A hit on a sphere comes back as Hit{prim, tHit, b0,b1,b2 = pObj.xyz} (QuadricIntersection,
    // shapes.h:100-104; phi = atan2(pObj.y, pObj.x)); IsSphere() tells the two apart.
    GPUBVHAggregate(const std::vector<float> &tris9, const std::vector<nnbvh_sphere> &spheres,
                    int maxPrimsInNode = 1, SplitMethod splitMethod = SplitMethod::SAH,
                    int device = -1, int buildThreads = 0) {
        const uint32_t nTris = (uint32_t)(tris9.size() / 9), nSph = (uint32_t)spheres.size();
        const uint32_t nPrims = nTris + nSph;
        CreateEngine(device);
        if (nPrims) {
            std::vector<float> bounds(6 * (size_t)nPrims);
            for (uint32_t i = 0; i < nTris; ++i)  // Triangle::Bounds, shapes.cpp:294-301
                for (int k = 0; k < 3; ++k) {
                    const float *p = &tris9[9 * (size_t)i];
                    bounds[6 * (size_t)i + k] = std::min(std::min(p[k], p[3 + k]), p[6 + k]);
                    bounds[6 * (size_t)i + 3 + k] = std::max(std::max(p[k], p[3 + k]), p[6 + k]);
                }
            for (uint32_t k = 0; k < nSph; ++k)  // Sphere::Bounds, shapes.cpp:37-40
                if (int rc = nnbvh_sphere_bounds(&spheres[k], &bounds[6 * (size_t)(nTris + k)])) Fail(rc);
            nodes_.resize(2 * (size_t)nPrims - 1);
            order_.resize(nPrims);
            uint32_t nNodes = 0;
            int rc = nnbvh_build_bvh_bounds(bounds.data(), nPrims, maxPrimsInNode, (int)splitMethod,
                                            buildThreads, nodes_.data(), &nNodes, order_.data());
            if (rc != NNBVH_OK) Fail(rc);
            nodes_.resize(nNodes);
            ordered_.assign(9 * (size_t)nPrims, 0.f);
            sphereOf_.assign(nPrims, -1);
            for (uint32_t i = 0; i < nPrims; ++i) {
                if (order_[i] >= nTris) {
                    sphereOf_[i] = (int32_t)(order_[i] - nTris);
                    continue;
                }
                for (int k = 0; k < 9; ++k) ordered_[9 * (size_t)i + k] = tris9[9 * (size_t)order_[i] + k];
            }
        }
        spheres_ = spheres;
        Upload();
    }

    // TransformedPrimitive instances over per-object BVHAggregates (cpu/primitive.cpp:112-130; what pbrt
    // builds for ObjectInstance, scene.cpp:1541-1571).  objectTris9[j]: object j's triangles in its own
    // space; caller primitive i < nTris is top-level triangle i, primitive nTris + k is instance k.
    // Every tree (each object's and the top level) comes from the library's host builder.
    struct ObjectTree {
        std::vector<nnbvh_node> nodes;
        std::vector<float> orderedTris9;
        std::vector<uint32_t> order;
    };
    GPUBVHAggregate(const std::vector<float> &tris9, const std::vector<std::vector<float>> &objectTris9,
                    const std::vector<nnbvh_instance> &instances, int maxPrimsInNode = 1,
                    SplitMethod splitMethod = SplitMethod::SAH, int device = -1, int buildThreads = 0) {
        const uint32_t nTris = (uint32_t)(tris9.size() / 9), nInst = (uint32_t)instances.size();
        const uint32_t nPrims = nTris + nInst;
        CreateEngine(device);
        if (nInst == 0) FailWith(NNBVH_ERR_INVALID, "GPUBVHAggregate: no instances given");
        objects_.resize(objectTris9.size());
        for (size_t j = 0; j < objectTris9.size(); ++j) {
            const std::vector<float> &src = objectTris9[j];
            const uint32_t n = (uint32_t)(src.size() / 9);
            ObjectTree &o = objects_[j];
            o.nodes.resize(2 * (size_t)std::max(n, 1u));
            o.order.resize(n);
            uint32_t nNodes = 0;
            if (int rc = nnbvh_build_bvh(src.data(), n, maxPrimsInNode, (int)splitMethod, buildThreads,
                                         o.nodes.data(), &nNodes, o.order.data())) Fail(rc);
            o.nodes.resize(nNodes);
            o.orderedTris9.resize(9 * (size_t)n);
            for (uint32_t i = 0; i < n; ++i)
                for (int k = 0; k < 9; ++k) o.orderedTris9[9 * (size_t)i + k] = src[9 * (size_t)o.order[i] + k];
        }
        std::vector<float> bounds(6 * (size_t)nPrims);
        for (uint32_t i = 0; i < nTris; ++i)
            for (int k = 0; k < 3; ++k) {
                const float *p = &tris9[9 * (size_t)i];
                bounds[6 * (size_t)i + k] = std::min(std::min(p[k], p[3 + k]), p[6 + k]);
                bounds[6 * (size_t)i + 3 + k] = std::max(std::max(p[k], p[3 + k]), p[6 + k]);
            }
        for (uint32_t k = 0; k < nInst; ++k) {  // TransformedPrimitive::Bounds, cpu/primitive.h:94
            if (instances[k].object >= objects_.size()) FailWith(NNBVH_ERR_INVALID, "GPUBVHAggregate: instance refers to a missing object");
            if (int rc = nnbvh_instance_bounds(&instances[k], objects_[instances[k].object].nodes[0].bounds,
                                               &bounds[6 * (size_t)(nTris + k)])) Fail(rc);
        }
        nodes_.resize(2 * (size_t)nPrims - 1);
        order_.resize(nPrims);
        uint32_t nNodes = 0;
        if (int rc = nnbvh_build_bvh_bounds(bounds.data(), nPrims, maxPrimsInNode, (int)splitMethod, buildThreads,
                                            nodes_.data(), &nNodes, order_.data())) Fail(rc);
        nodes_.resize(nNodes);
        ordered_.assign(9 * (size_t)nPrims, 0.f);
        instanceOf_.assign(nPrims, -1);
        for (uint32_t i = 0; i < nPrims; ++i) {
            if (order_[i] >= nTris) {
                instanceOf_[i] = (int32_t)(order_[i] - nTris);
                continue;
            }
            for (int k = 0; k < 9; ++k) ordered_[9 * (size_t)i + k] = tris9[9 * (size_t)order_[i] + k];
        }
        instances_ = instances;
        std::vector<nnbvh_object> objs(objects_.size());
        for (size_t j = 0; j < objects_.size(); ++j)
            objs[j] = nnbvh_object{objects_[j].nodes.data(), (uint32_t)objects_[j].nodes.size(),
                                   objects_[j].orderedTris9.data(), (uint32_t)(objects_[j].orderedTris9.size() / 9)};
        if (int rc = nnbvh_upload_scene_instanced(e_, nodes_.data(), (uint32_t)nodes_.size(), ordered_.data(),
                                                  instanceOf_.data(), nPrims, instances_.data(), nInst, objs.data(),
                                                  (uint32_t)objs.size())) Fail(rc);
    }
    // Closest hits of an instanced scene with the primitive inside the object (ordered within the object;
    // Objects()[j].order maps it to the caller's triangle) as a second plane.
    void IntersectBatchInstanced(const nnbvh_ray *rays, uint64_t n, nnbvh_hit *hits, uint32_t *innerPrim) const {
        Check(nnbvh_intersect_closest_instanced(e_, rays, n, hits, innerPrim));
    }
    void IntersectBatchInstancedSoA(const float *ox, const float *oy, const float *oz, const float *dx,
                                    const float *dy, const float *dz, const float *tMax, uint64_t n,
                                    nnbvh_hit *hits, uint32_t *innerPrim) const {
        Check(nnbvh_intersect_closest_instanced_soa(e_, ox, oy, oz, dx, dy, dz, tMax, n, hits, innerPrim));
    }
    // ordered top-level primitive (Hit::prim) -> instance index, or -1
    int32_t InstanceOf(uint32_t ordered) const { return instanceOf_.empty() ? -1 : instanceOf_[ordered]; }
    const std::vector<ObjectTree> &Objects() const { return objects_; }
    const std::vector<int32_t> &OrderedInstanceIndex() const { return instanceOf_; }

    // ONE primitive list with every kind the library traverses, in the CALLER's order (so the tree is the
    // one BVHAggregate builds from the same std::vector<Primitive>): triangles, optionally with a constant
    // alpha (GeometricPrimitive over a FloatConstantTexture, cpu/primitive.cpp:50-84), spheres, and
    // TransformedPrimitive instances (cpu/primitive.cpp:112-130) of objects whose trees are already built --
    // by BuildObject() below, or by the reference itself (a BVHAggregate's flattened nodes have
    // nnbvh_node's layout; integration/pbrt/gpubvh.h hands them over as they are).
    struct GeneralList {
        enum Kind : uint8_t { kTriangle = 0, kSphere = 1, kInstance = 2 };
        std::vector<uint8_t> kind;       // per caller primitive
        std::vector<uint32_t> index;     // ... and its index in tris9 / spheres / instances
        std::vector<float> tris9;
        std::vector<float> alpha;        // empty, or one constant per triangle of tris9 (>= 1: no alpha test)
        std::vector<nnbvh_sphere> spheres;
        std::vector<nnbvh_instance> instances;
        std::vector<ObjectTree> objects;
    };
    static ObjectTree BuildObject(const std::vector<float> &tris9, int maxPrimsInNode = 1,
                                  SplitMethod splitMethod = SplitMethod::SAH, int buildThreads = 0) {
        const uint32_t n = (uint32_t)(tris9.size() / 9);
        ObjectTree o;
        o.nodes.resize(2 * (size_t)std::max(n, 1u));
        o.order.resize(n);
        uint32_t nNodes = 0;
        Check(nnbvh_build_bvh(tris9.data(), n, maxPrimsInNode, (int)splitMethod, buildThreads, o.nodes.data(),
                              &nNodes, o.order.data()));
        o.nodes.resize(nNodes);
        o.orderedTris9.resize(9 * (size_t)n);
        for (uint32_t i = 0; i < n; ++i)
            for (int k = 0; k < 9; ++k) o.orderedTris9[9 * (size_t)i + k] = tris9[9 * (size_t)o.order[i] + k];
        return o;
    }
    GPUBVHAggregate(GeneralList list, int maxPrimsInNode = 1, SplitMethod splitMethod = SplitMethod::SAH,
                    int device = -1, int buildThreads = 0) {
        const uint32_t nPrims = (uint32_t)list.kind.size();
        const uint32_t nTris = (uint32_t)(list.tris9.size() / 9);
        CreateEngine(device);
        if (list.index.size() != nPrims || (!list.alpha.empty() && list.alpha.size() != nTris))
            FailWith(NNBVH_ERR_INVALID, "GPUBVHAggregate: GeneralList arrays disagree in length");
        if (nPrims == 0) {
            Upload();
            return;
        }
        std::vector<float> bounds(6 * (size_t)nPrims);
        for (uint32_t i = 0; i < nPrims; ++i) {
            float *b = &bounds[6 * (size_t)i];
            const uint32_t k = list.index[i];
            if (list.kind[i] == GeneralList::kTriangle) {  // Triangle::Bounds, shapes.cpp:294-301
                if (k >= nTris) FailWith(NNBVH_ERR_INVALID, "GPUBVHAggregate: triangle index out of range");
                const float *p = &list.tris9[9 * (size_t)k];
                for (int a = 0; a < 3; ++a) {
                    b[a] = std::min(std::min(p[a], p[3 + a]), p[6 + a]);
                    b[3 + a] = std::max(std::max(p[a], p[3 + a]), p[6 + a]);
                }
            } else if (list.kind[i] == GeneralList::kSphere) {  // Sphere::Bounds, shapes.cpp:37-40
                if (k >= list.spheres.size()) FailWith(NNBVH_ERR_INVALID, "GPUBVHAggregate: sphere index out of range");
                if (int rc = nnbvh_sphere_bounds(&list.spheres[k], b)) Fail(rc);
            } else if (list.kind[i] == GeneralList::kInstance) {  // TransformedPrimitive::Bounds, cpu/primitive.h:94
                if (k >= list.instances.size() || list.instances[k].object >= list.objects.size() ||
                    list.objects[list.instances[k].object].nodes.empty())
                    FailWith(NNBVH_ERR_INVALID, "GPUBVHAggregate: instance or its object is missing");
                if (int rc = nnbvh_instance_bounds(&list.instances[k],
                                                   list.objects[list.instances[k].object].nodes[0].bounds, b))
                    Fail(rc);
            } else
                FailWith(NNBVH_ERR_INVALID, "GPUBVHAggregate: unknown primitive kind");
        }
        nodes_.resize(2 * (size_t)nPrims - 1);
        order_.resize(nPrims);
        uint32_t nNodes = 0;
        if (int rc = nnbvh_build_bvh_bounds(bounds.data(), nPrims, maxPrimsInNode, (int)splitMethod, buildThreads,
                                            nodes_.data(), &nNodes, order_.data())) Fail(rc);
        nodes_.resize(nNodes);
        ordered_.assign(9 * (size_t)nPrims, 0.f);
        if (!list.spheres.empty()) sphereOf_.assign(nPrims, -1);
        if (!list.instances.empty()) instanceOf_.assign(nPrims, -1);
        if (!list.alpha.empty()) alpha_.assign(nPrims, 1.f);
        for (uint32_t i = 0; i < nPrims; ++i) {
            const uint32_t src = order_[i], k = list.index[src];
            if (list.kind[src] == GeneralList::kSphere)
                sphereOf_[i] = (int32_t)k;
            else if (list.kind[src] == GeneralList::kInstance)
                instanceOf_[i] = (int32_t)k;
            else {
                for (int a = 0; a < 9; ++a) ordered_[9 * (size_t)i + a] = list.tris9[9 * (size_t)k + a];
                if (!alpha_.empty()) alpha_[i] = list.alpha[k];
            }
        }
        spheres_ = std::move(list.spheres);
        instances_ = std::move(list.instances);
        objects_ = std::move(list.objects);
        std::vector<nnbvh_object> objs(objects_.size());
        for (size_t j = 0; j < objects_.size(); ++j)
            objs[j] = nnbvh_object{objects_[j].nodes.data(), (uint32_t)objects_[j].nodes.size(),
                                   objects_[j].orderedTris9.data(), (uint32_t)(objects_[j].orderedTris9.size() / 9)};
        nnbvh_scene_desc d{};
        d.nodes = nodes_.data(), d.nNodes = (uint32_t)nodes_.size();
        d.orderedTriVerts9 = ordered_.data(), d.nPrims = nPrims;
        d.orderedSphereIndex = sphereOf_.empty() ? nullptr : sphereOf_.data();
        d.spheres = spheres_.data(), d.nSpheres = (uint32_t)spheres_.size();
        d.orderedInstanceIndex = instanceOf_.empty() ? nullptr : instanceOf_.data();
        d.instances = instances_.data(), d.nInstances = (uint32_t)instances_.size();
        d.objects = objs.data(), d.nObjects = instances_.empty() ? 0u : (uint32_t)objs.size();
        d.orderedAlpha = alpha_.empty() ? nullptr : alpha_.data();
        if (int rc = nnbvh_upload_scene_general(e_, &d)) Fail(rc);
    }
    bool HasInstances() const { return !instances_.empty(); }

    // A ready flattened tree (e.g. an NN-split tree exported in LinearBVHNode layout).
    GPUBVHAggregate(std::vector<nnbvh_node> nodes, std::vector<float> orderedTris9,
                    std::vector<uint32_t> order = {}, int device = -1)
        : nodes_(std::move(nodes)), ordered_(std::move(orderedTris9)), order_(std::move(order)) {
        Open(device);
    }

    // A second aggregate on the same uploaded scene (nnbvh_clone: nothing is copied on the device) with its
    // own stream and scratch, so that independent batches can be in flight together.
    std::unique_ptr<GPUBVHAggregate> Clone() const {
        return std::unique_ptr<GPUBVHAggregate>(new GPUBVHAggregate(CloneTag{}, *this));
    }

    GPUBVHAggregate(const GPUBVHAggregate &) = delete;
    GPUBVHAggregate &operator=(const GPUBVHAggregate &) = delete;
    ~GPUBVHAggregate() { nnbvh_destroy(e_); }

    // Bounds3f BVHAggregate::Bounds() const, cpu/aggregates.cpp:524-527 (CHECK(nodes) there).
    Bounds3f Bounds() const {
        if (nodes_.empty()) throw Error(NNBVH_ERR_NOSCENE, "Bounds() of an empty aggregate");
        const float *b = nodes_[0].bounds;
        return Bounds3f{{b[0], b[1], b[2]}, {b[3], b[4], b[5]}};
    }

    // pstd::optional<ShapeIntersection> Intersect(const Ray&, Float tMax) const (:529-579).
    std::optional<Hit> Intersect(const Ray &ray,
                                 float tMax = std::numeric_limits<float>::infinity()) const {
        nnbvh_ray r = MakeRay(ray, tMax);
        nnbvh_hit h;
        Check(nnbvh_intersect_closest(e_, &r, 1, &h));
        if (h.prim == NNBVH_MISS) return std::nullopt;
        return Hit{h.prim, h.t, h.b0, h.b1, h.b2};
    }

    // bool IntersectP(const Ray&, Float tMax) const (:581-624).
    bool IntersectP(const Ray &ray,
                    float tMax = std::numeric_limits<float>::infinity()) const {
        nnbvh_ray r = MakeRay(ray, tMax);
        uint8_t occ = 0;
        Check(nnbvh_intersect_any(e_, &r, 1, &occ));
        return occ != 0;
    }

    // Batch forms: n rays in, n results out; pointers may be host or device memory.
    void IntersectBatch(const nnbvh_ray *rays, uint64_t n, nnbvh_hit *hits) const {
        Check(nnbvh_intersect_closest(e_, rays, n, hits));
    }
    void IntersectPBatch(const nnbvh_ray *rays, uint64_t n, uint8_t *occluded) const {
        Check(nnbvh_intersect_any(e_, rays, n, occluded));
    }
    // IntersectBatch for callers that cannot allocate (CUDA graph capture, serving loops): DEVICE pointers,
    // n <= NNBVH_NOALLOC_MAX_RAYS, scratch in the library's static device array; calls that overlap on the
    // GPU race (nnbvh.h nnbvh_intersect_closest_noalloc; no counterpart in the reference).
    void IntersectBatchNoAlloc(const nnbvh_ray *raysDev, uint64_t n, nnbvh_hit *hitsDev) const {
        Check(nnbvh_intersect_closest_noalloc(e_, raysDev, n, hitsDev));
    }
    // The wavefront queues' SoA layout (SOA<RayWorkItem>, wavefront/workitems.soa); tMax may be null.
    void IntersectBatchSoA(const float *ox, const float *oy, const float *oz, const float *dx,
                           const float *dy, const float *dz, const float *tMax, uint64_t n,
                           nnbvh_hit *hits) const {
        Check(nnbvh_intersect_closest_soa(e_, ox, oy, oz, dx, dy, dz, tMax, n, hits));
    }
    void IntersectPBatchSoA(const float *ox, const float *oy, const float *oz, const float *dx,
                            const float *dy, const float *dz, const float *tMax, uint64_t n,
                            uint8_t *occluded) const {
        Check(nnbvh_intersect_any_soa(e_, ox, oy, oz, dx, dy, dz, tMax, n, occluded));
    }
    // With the reference's per-ray statistics (bvhNodesVisited, nTriTests).
    void IntersectCounters(const nnbvh_ray *rays, uint64_t n, bool anyHit, nnbvh_hit *hits,
                           uint8_t *occluded, uint32_t *nodesVisited, uint32_t *triTests) const {
        Check(nnbvh_intersect_counters(e_, rays, n, anyHit, hits, occluded, nodesVisited, triTests));
    }

    // ordered primitive index -> index in the constructor's primitive list.
    uint32_t PrimitiveIndex(uint32_t ordered) const {
        return order_.empty() ? ordered : order_[ordered];
    }
    // ordered primitive index (Hit::prim) -> is it a sphere?
    bool IsSphere(uint32_t ordered) const { return !sphereOf_.empty() && sphereOf_[ordered] >= 0; }
    const std::vector<nnbvh_node> &Nodes() const { return nodes_; }
    const std::vector<float> &OrderedTriangles() const { return ordered_; }
    void SetStream(void *cudaStream) { Check(nnbvh_set_stream(e_, cudaStream)); }
    float LastKernelMs() const {
        float ms = 0;
        Check(nnbvh_last_kernel_ms(e_, &ms));
        return ms;
    }
    nnbvh_engine *Handle() const { return e_; }

  private:
    struct CloneTag {};
    GPUBVHAggregate(CloneTag, const GPUBVHAggregate &o)
        : nodes_(o.nodes_), ordered_(o.ordered_), order_(o.order_), sphereOf_(o.sphereOf_), spheres_(o.spheres_),
          instanceOf_(o.instanceOf_), instances_(o.instances_), objects_(o.objects_), alpha_(o.alpha_) {
        Check(nnbvh_clone(o.e_, &e_));
    }
    void CreateEngine(int device) {
        nnbvh_options o{};
        o.device = device;
        o.kernel = NNBVH_KERNEL_AUTO;
        o.smem_top_nodes = -1;
        Check(nnbvh_create(&o, &e_));
    }
    [[noreturn]] void FailWith(int rc, const char *what) {
        nnbvh_destroy(e_);
        e_ = nullptr;
        throw Error(rc, what);
    }
    [[noreturn]] void Fail(int rc) {  // constructor failure: the destructor will not run
        std::string msg = nnbvh_last_error();
        nnbvh_destroy(e_);
        e_ = nullptr;
        throw Error(rc, "libnnbvh: " + msg);
    }
    void Upload() {
        int rc = spheres_.empty()
                     ? nnbvh_upload_scene(e_, nodes_.data(), (uint32_t)nodes_.size(), ordered_.data(),
                                          (uint32_t)(ordered_.size() / 9))
                     : nnbvh_upload_scene_mixed(e_, nodes_.data(), (uint32_t)nodes_.size(), ordered_.data(),
                                                sphereOf_.data(), (uint32_t)(ordered_.size() / 9),
                                                spheres_.data(), (uint32_t)spheres_.size());
        if (rc != NNBVH_OK) Fail(rc);
    }
    void Open(int device) {
        CreateEngine(device);
        Upload();
    }

    std::vector<nnbvh_node> nodes_;
    std::vector<float> ordered_;
    std::vector<uint32_t> order_;
    std::vector<int32_t> sphereOf_;       // per ordered primitive: -1 triangle, else sphere index
    std::vector<nnbvh_sphere> spheres_;
    std::vector<int32_t> instanceOf_;     // per ordered top-level primitive: -1 triangle, else instance index
    std::vector<nnbvh_instance> instances_;
    std::vector<ObjectTree> objects_;
    std::vector<float> alpha_;            // per ordered primitive: the constant alpha (general lists), or empty
    nnbvh_engine *e_ = nullptr;
};

}  // namespace nnbvh

#endif  // NNBVH_HPP
