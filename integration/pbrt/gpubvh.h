// integration/pbrt/gpubvh.h -- the reference-side binding of the aggregate ("gpubvh" accelerator).
//
// What a maintainer of the reference adds as src/pbrt/cpu/gpubvh.h: a class with BVHAggregate's surface
// (/root/reference/src/pbrt/cpu/aggregates.h:26-60: Create, Bounds, Intersect, IntersectP) plus the batch
// calls, whose traversal runs in libnnbvh.so (include/nnbvh.hpp) and whose per-hit follow-up is the
// reference's own code:
//   Triangle::InteractionFromIntersection          shapes.h:884-1010 (public static)
//   Sphere::InteractionFromIntersection            shapes.h:238-280
//   SurfaceInteraction::SetIntersectionProperties  interaction.h:218-229
//   Transform::ApplyInverse(Ray) / operator()(SurfaceInteraction)   util/transform.h (TransformedPrimitive)
// so that Resolve() returns exactly what SimplePrimitive / GeometricPrimitive / TransformedPrimitive::Intersect
// would have (cpu/primitive.cpp:50-84, :100-110, :112-130).
//
// To make it selectable from a scene file it also has to be listed in Primitive's TaggedPointer
// (cpu/primitive.h:33-35) and named in CreateAccelerator (cpu/aggregates.cpp:1163-1178):
//     else if (name == "gpubvh") accel = GPUBVHAggregate::Create(std::move(prims), parameters);
// Those two edits and the friend declarations (GeometricPrimitive, SimplePrimitive, TransformedPrimitive:
// cpu/primitive.h:58, :76, :87; BVHAggregate: cpu/aggregates.h; Triangle and Sphere: shapes.h) are the whole
// patch.  tests/test_integration_compile.py compiles this file against the reference's real headers (the
// private sections opened by a #define instead).
//
// Contract -- what one list may hold, in any order (the list's order is kept, so the tree is the one
// BVHAggregate builds from it):
//   SimplePrimitive / GeometricPrimitive over Triangle or Sphere;
//   GeometricPrimitive over Triangle with a FloatConstantTexture alpha (the test and the re-trace of
//     cpu/primitive.cpp:62-75 run on the device: include/nnbvh.h, nnbvh_upload_scene_alpha);
//   TransformedPrimitive over a BVHAggregate of alpha-free triangles (what ObjectInstance becomes,
//     scene.cpp:1541-1571): the object's flattened tree goes to the device as the reference built it.
// Anything else (textured alpha, AnimatedPrimitive, other shapes, other inner aggregates) is refused by name.
#ifndef NNBVH_INTEGRATION_PBRT_GPUBVH_H
#define NNBVH_INTEGRATION_PBRT_GPUBVH_H

#include <pbrt/cpu/aggregates.h>
#include <pbrt/cpu/primitive.h>
#include <pbrt/paramdict.h>
#include <pbrt/shapes.h>
#include <pbrt/util/error.h>

#include <cmath>
#include <map>
#include <memory>
#include <mutex>
#include <vector>

#include <nnbvh.hpp>

namespace pbrt {

class GPUBVHAggregate {
  public:
    GPUBVHAggregate(std::vector<Primitive> prims, int maxPrimsInNode, BVHAggregate::SplitMethod splitMethod,
                    int device = -1)
        : primitives(std::move(prims)) {
        // the caller's order is kept (nnbvh::GPUBVHAggregate::GeneralList): primitive i of the library is primitive i here
        nnbvh::GPUBVHAggregate::GeneralList list;
        std::map<const BVHAggregate *, uint32_t> objectIndex;
        for (size_t i = 0; i < primitives.size(); ++i) {
            Parts parts = PartsOf(primitives[i]);
            if (parts.tri) {
                list.kind.push_back(list.kTriangle), list.index.push_back((uint32_t)(list.tris9.size() / 9));
                PushTriangle(*parts.tri, &list.tris9);
                list.alpha.push_back(parts.alpha);
            } else if (parts.sphere) {
                list.kind.push_back(list.kSphere), list.index.push_back((uint32_t)list.spheres.size());
                list.spheres.push_back(SphereRecord(*parts.sphere));
            } else if (parts.instance) {
                const BVHAggregate *object = parts.instance->primitive.CastOrNullptr<BVHAggregate>();
                if (!object)
                    ErrorExit("gpubvh: an object instance must refer to a \"bvh\" aggregate");
                auto at = objectIndex.find(object);
                if (at == objectIndex.end()) {
                    at = objectIndex.emplace(object, (uint32_t)list.objects.size()).first;
                    list.objects.push_back(TreeOf(*object));
                }
                nnbvh_instance in{};
                const SquareMatrix<4> &m = parts.instance->renderFromPrimitive->GetMatrix(),
                                      &mi = parts.instance->renderFromPrimitive->GetInverseMatrix();
                for (int r = 0; r < 3; ++r)
                    for (int c = 0; c < 4; ++c)
                        in.renderFromPrimitive[4 * r + c] = m[r][c], in.primitiveFromRender[4 * r + c] = mi[r][c];
                in.object = at->second;
                list.kind.push_back(list.kInstance), list.index.push_back((uint32_t)list.instances.size());
                list.instances.push_back(in);
            } else
                ErrorExit("gpubvh: primitive %d is not a triangle (alpha-free or with a constant alpha), a sphere or "
                          "an instance of a triangle aggregate", (int)i);
        }
        bool anyAlpha = false;
        for (float a : list.alpha)
            anyAlpha |= a < 1;
        if (!anyAlpha)
            list.alpha.clear();
        // BVHAggregate::SplitMethod { SAH, HLBVH, Middle, EqualCounts } (cpu/aggregates.h:33) has the
        // library's numbering
        auto sm = nnbvh::GPUBVHAggregate::SplitMethod((int)splitMethod);
        agg = std::make_unique<nnbvh::GPUBVHAggregate>(std::move(list), maxPrimsInNode, sm, device);
    }

    // BVHAggregate::Create, cpu/aggregates.cpp:725-744: the same parameters
    static GPUBVHAggregate *Create(std::vector<Primitive> prims, const ParameterDictionary &parameters) {
        std::string name = parameters.GetOneString("splitmethod", "sah");
        BVHAggregate::SplitMethod sm = BVHAggregate::SplitMethod::SAH;
        if (name == "hlbvh")
            sm = BVHAggregate::SplitMethod::HLBVH;
        else if (name == "middle")
            sm = BVHAggregate::SplitMethod::Middle;
        else if (name == "equal")
            sm = BVHAggregate::SplitMethod::EqualCounts;
        else if (name != "sah")
            Warning(R"(BVH split method "%s" unknown.  Using "sah".)", name);
        return new GPUBVHAggregate(std::move(prims), parameters.GetOneInt("maxnodeprims", 4), sm);
    }

    Bounds3f Bounds() const {
        nnbvh::Bounds3f b = agg->Bounds();
        return Bounds3f(Point3f(b.pMin[0], b.pMin[1], b.pMin[2]), Point3f(b.pMax[0], b.pMax[1], b.pMax[2]));
    }

    // The single-ray surface a Primitive must have: a batch of one.  An nnbvh_engine serialises its
    // batches and is NOT re-entrant (include/nnbvh.h, "Threading"), while pbrt calls Primitive::Intersect
    // from every ParallelFor worker: the single-ray calls take a lock.  Correctness only -- anything that
    // has more than a handful of rays uses Engine()'s batch calls (the wavefront subclass does).
    pstd::optional<ShapeIntersection> Intersect(const Ray &ray, Float tMax = Infinity) const {
        nnbvh_ray r{ray.o.x, ray.o.y, ray.o.z, tMax, ray.d.x, ray.d.y, ray.d.z, ray.time};
        nnbvh_hit h;
        uint32_t inner = NNBVH_MISS;
        {
            std::lock_guard<std::mutex> lock(singleRay);
            IntersectBatch(&r, 1, &h, &inner);
        }
        if (h.prim == NNBVH_MISS)
            return {};
        return Resolve(h, inner, ray);
    }
    bool IntersectP(const Ray &ray, Float tMax = Infinity) const {
        nnbvh_ray r{ray.o.x, ray.o.y, ray.o.z, tMax, ray.d.x, ray.d.y, ray.d.z, ray.time};
        uint8_t occluded = 0;
        {
            std::lock_guard<std::mutex> lock(singleRay);
            agg->IntersectPBatch(&r, 1, &occluded);
        }
        return occluded != 0;
    }

    // The batch surface: what an integrator that owns a ray batch calls.  `inner` (n entries) receives, for a
    // hit on an instance, the primitive inside the object -- Resolve() needs it; lists without instances leave
    // it untouched.
    const nnbvh::GPUBVHAggregate &Engine() const { return *agg; }
    void IntersectBatch(const nnbvh_ray *rays, uint64_t n, nnbvh_hit *hits, uint32_t *inner) const {
        if (agg->HasInstances())
            agg->IntersectBatchInstanced(rays, n, hits, inner);
        else
            agg->IntersectBatch(rays, n, hits);
    }
    void IntersectBatchSoA(const float *ox, const float *oy, const float *oz, const float *dx, const float *dy,
                           const float *dz, const float *tMax, uint64_t n, nnbvh_hit *hits, uint32_t *inner) const {
        if (agg->HasInstances())
            agg->IntersectBatchInstancedSoA(ox, oy, oz, dx, dy, dz, tMax, n, hits, inner);
        else
            agg->IntersectBatchSoA(ox, oy, oz, dx, dy, dz, tMax, n, hits);
    }

    // {prim, t, b0, b1, b2} (+ the primitive inside the object) -> what Primitive::Intersect would have returned
    // for this ray
    ShapeIntersection Resolve(const nnbvh_hit &h, const Ray &ray) const { return Resolve(h, NNBVH_MISS, ray); }
    ShapeIntersection Resolve(const nnbvh_hit &h, uint32_t inner, const Ray &ray) const {
        Parts parts = PartsOf(primitives[agg->PrimitiveIndex(h.prim)]);
        SurfaceInteraction intr;
        if (parts.instance) {  // TransformedPrimitive::Intersect, cpu/primitive.cpp:112-130
            const TransformedPrimitive *tp = parts.instance;
            Float tMax = Infinity;  // only shifted by the call; the transformed o and d do not depend on it
            Ray local = tp->renderFromPrimitive->ApplyInverse(ray, &tMax);
            Parts in = PartsOf(tp->primitive.Cast<BVHAggregate>()->primitives[inner]);
            intr = Triangle::InteractionFromIntersection(in.tri->GetMesh(), in.tri->triIndex,
                                                         TriangleIntersection{h.b0, h.b1, h.b2, h.t}, local.time,
                                                         -local.d);
            intr.SetIntersectionProperties(in.material, in.areaLight, in.mediumInterface, local.medium);
            intr = (*tp->renderFromPrimitive)(intr);
            return ShapeIntersection{intr, h.t};
        }
        if (parts.tri)
            intr = Triangle::InteractionFromIntersection(parts.tri->GetMesh(), parts.tri->triIndex,
                                                         TriangleIntersection{h.b0, h.b1, h.b2, h.t}, ray.time,
                                                         -ray.d);
        else {  // a sphere leaf returns QuadricIntersection{tHit, pObj}; phi as shapes.h:196-198
            Float phi = std::atan2(h.b1, h.b0);
            if (phi < 0)
                phi += 2 * Pi;
            intr = parts.sphere->InteractionFromIntersection(
                QuadricIntersection{h.t, Point3f(h.b0, h.b1, h.b2), phi}, -ray.d, ray.time);
        }
        intr.SetIntersectionProperties(parts.material, parts.areaLight, parts.mediumInterface, ray.medium);
        return ShapeIntersection{intr, h.t};
    }

  private:
    struct Parts {
        const Triangle *tri = nullptr;
        const Sphere *sphere = nullptr;
        const TransformedPrimitive *instance = nullptr;
        float alpha = 1;  // the constant of a FloatConstantTexture alpha; 1 = no alpha test
        Material material;
        Light areaLight;
        const MediumInterface *mediumInterface = nullptr;
    };
    static Parts PartsOf(const Primitive &p) {
        Parts out;
        Shape shape;
        if (const SimplePrimitive *sp = p.CastOrNullptr<SimplePrimitive>()) {
            shape = sp->shape;
            out.material = sp->material;
        } else if (const GeometricPrimitive *gp = p.CastOrNullptr<GeometricPrimitive>()) {
            bool known = true;
            if (gp->alpha) {  // only a constant can go to the device; and only on a triangle
                const FloatConstantTexture *c = gp->alpha.CastOrNullptr<FloatConstantTexture>();
                known = c && gp->shape.Is<Triangle>();
                if (known)
                    out.alpha = c->Evaluate(TextureEvalContext());
            }
            if (known)
                shape = gp->shape;
            out.material = gp->material;
            out.areaLight = gp->areaLight;
            out.mediumInterface = &gp->mediumInterface;
        } else if (const TransformedPrimitive *tp = p.CastOrNullptr<TransformedPrimitive>()) {
            out.instance = tp;
            return out;
        }
        out.tri = shape.CastOrNullptr<Triangle>();
        out.sphere = shape.CastOrNullptr<Sphere>();
        return out;
    }
    static void PushTriangle(const Triangle &tri, std::vector<float> *tris9) {
        const TriangleMesh *mesh = tri.GetMesh();
        const int *v = &mesh->vertexIndices[3 * tri.triIndex];
        for (int k = 0; k < 3; ++k) {
            Point3f q = mesh->p[v[k]];
            tris9->push_back(q.x), tris9->push_back(q.y), tris9->push_back(q.z);
        }
    }
    // The object aggregate the reference already built: its flattened nodes (LinearBVHNode,
    // cpu/aggregates.cpp:113-124, is nnbvh_node field for field: bounds, offset, nPrimitives, axis in 32 bytes)
    // and its primitives in the order the nodes index them.
    static nnbvh::GPUBVHAggregate::ObjectTree TreeOf(const BVHAggregate &object) {
        static_assert(sizeof(nnbvh_node) == 32, "nnbvh_node mirrors LinearBVHNode");
        const nnbvh_node *nodes = reinterpret_cast<const nnbvh_node *>(object.nodes);
        // pre-order layout: the node count is one past the largest index reachable from the root
        uint32_t last = 0;
        std::vector<uint32_t> todo{0};
        while (!todo.empty()) {
            uint32_t i = todo.back();
            todo.pop_back();
            last = std::max(last, i);
            if (nodes[i].nPrimitives == 0)
                todo.push_back(i + 1), todo.push_back((uint32_t)nodes[i].offset);
        }
        nnbvh::GPUBVHAggregate::ObjectTree tree;
        tree.nodes.assign(nodes, nodes + last + 1);
        for (size_t i = 0; i < object.primitives.size(); ++i) {
            Parts parts = PartsOf(object.primitives[i]);
            if (!parts.tri || parts.alpha < 1)
                ErrorExit("gpubvh: instanced objects must be made of alpha-free triangles");
            PushTriangle(*parts.tri, &tree.orderedTris9);
            tree.order.push_back((uint32_t)i);
        }
        return tree;
    }
    // Sphere's members (shapes.h:396-400) -> what Sphere::Sphere was given.  zMin / zMax are already
    // clamped and ordered (clamping again changes nothing); phiMax is kept in radians.
    static nnbvh_sphere SphereRecord(const Sphere &s) {
        nnbvh_sphere out;
        const SquareMatrix<4> &m = s.renderFromObject->GetMatrix(), &mi = s.objectFromRender->GetMatrix();
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 4; ++c)
                out.renderFromObject[4 * r + c] = m[r][c], out.objectFromRender[4 * r + c] = mi[r][c];
        out.radius = s.radius, out.zMin = s.zMin, out.zMax = s.zMax;
        out.phiMax = nnbvh::SpherePhiMaxDegrees(s.phiMax);
        return out;
    }

    std::vector<Primitive> primitives;  // the caller's list
    std::unique_ptr<nnbvh::GPUBVHAggregate> agg;
    mutable std::mutex singleRay;
};

}  // namespace pbrt

#endif  // NNBVH_INTEGRATION_PBRT_GPUBVH_H
