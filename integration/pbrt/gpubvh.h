// integration/pbrt/gpubvh.h -- the reference-side binding of the aggregate ("gpubvh" accelerator).
//
// What a maintainer of the reference adds as src/pbrt/cpu/gpubvh.h: a class with BVHAggregate's surface
// (/root/reference/src/pbrt/cpu/aggregates.h:26-60: Create, Bounds, Intersect, IntersectP) plus the batch
// calls, whose traversal runs in libnnbvh.so (include/nnbvh.hpp) and whose per-hit follow-up is the
// reference's own code:
//   Triangle::InteractionFromIntersection          shapes.h:884-1010 (public static)
//   Sphere::InteractionFromIntersection            shapes.h:238-280
//   SurfaceInteraction::SetIntersectionProperties  interaction.h:218-229
// so that Resolve() returns exactly what SimplePrimitive / GeometricPrimitive::Intersect would have
// (cpu/primitive.cpp:50-84, :100-110).
//
// To make it selectable from a scene file it also has to be listed in Primitive's TaggedPointer
// (cpu/primitive.h:33-35) and named in CreateAccelerator (cpu/aggregates.cpp:1163-1178):
//     else if (name == "gpubvh") accel = GPUBVHAggregate::Create(std::move(prims), parameters);
// Those two edits and three friend declarations (GeometricPrimitive, SimplePrimitive: cpu/primitive.h:58,
// :76; Triangle and Sphere: shapes.h) are the whole patch.  tests/test_integration_compile.py compiles
// this file against the reference's real headers (the private sections opened by a #define instead).
//
// Contract: SimplePrimitive / alpha-free GeometricPrimitive over Triangle or Sphere.
#ifndef NNBVH_INTEGRATION_PBRT_GPUBVH_H
#define NNBVH_INTEGRATION_PBRT_GPUBVH_H

#include <pbrt/cpu/aggregates.h>
#include <pbrt/cpu/primitive.h>
#include <pbrt/paramdict.h>
#include <pbrt/shapes.h>
#include <pbrt/util/error.h>

#include <cmath>
#include <memory>
#include <vector>

#include <nnbvh.hpp>

namespace pbrt {

class GPUBVHAggregate {
  public:
    GPUBVHAggregate(std::vector<Primitive> prims, int maxPrimsInNode, BVHAggregate::SplitMethod splitMethod,
                    int device = -1)
        : primitives(std::move(prims)) {
        // triangles first, then spheres: the order of nnbvh::GPUBVHAggregate's mixed constructor; `slot` maps
        // its primitive numbering back to this list
        std::vector<float> tris9;
        std::vector<nnbvh_sphere> spheres;
        std::vector<int> sphereSlots;
        for (size_t i = 0; i < primitives.size(); ++i) {
            Parts parts = PartsOf(primitives[i]);
            if (parts.tri) {
                const TriangleMesh *mesh = parts.tri->GetMesh();
                const int *v = &mesh->vertexIndices[3 * parts.tri->triIndex];
                for (int k = 0; k < 3; ++k) {
                    Point3f q = mesh->p[v[k]];
                    tris9.push_back(q.x), tris9.push_back(q.y), tris9.push_back(q.z);
                }
                slot.push_back((int)i);
            } else if (parts.sphere) {
                spheres.push_back(SphereRecord(*parts.sphere));
                sphereSlots.push_back((int)i);
            } else
                ErrorExit("gpubvh: only alpha-free triangle and sphere primitives are supported");
        }
        slot.insert(slot.end(), sphereSlots.begin(), sphereSlots.end());
        // BVHAggregate::SplitMethod { SAH, HLBVH, Middle, EqualCounts } (cpu/aggregates.h:33) has the
        // library's numbering
        auto sm = nnbvh::GPUBVHAggregate::SplitMethod((int)splitMethod);
        if (spheres.empty())
            agg = std::make_unique<nnbvh::GPUBVHAggregate>(tris9, maxPrimsInNode, sm, device);
        else
            agg = std::make_unique<nnbvh::GPUBVHAggregate>(tris9, spheres, maxPrimsInNode, sm, device);
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
        {
            std::lock_guard<std::mutex> lock(singleRay);
            agg->IntersectBatch(&r, 1, &h);
        }
        if (h.prim == NNBVH_MISS)
            return {};
        return Resolve(h, ray);
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

    // The batch surface: what an integrator that owns a ray batch calls
    const nnbvh::GPUBVHAggregate &Engine() const { return *agg; }

    // {prim, t, b0, b1, b2} -> what Primitive::Intersect would have returned for this ray
    ShapeIntersection Resolve(const nnbvh_hit &h, const Ray &ray) const {
        Parts parts = PartsOf(primitives[slot[agg->PrimitiveIndex(h.prim)]]);
        SurfaceInteraction intr;
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
            if (!gp->alpha)
                shape = gp->shape;
            out.material = gp->material;
            out.areaLight = gp->areaLight;
            out.mediumInterface = &gp->mediumInterface;
        }
        out.tri = shape.CastOrNullptr<Triangle>();
        out.sphere = shape.CastOrNullptr<Sphere>();
        return out;
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
    std::vector<int> slot;              // library primitive index (triangles, then spheres) -> index in it
    std::unique_ptr<nnbvh::GPUBVHAggregate> agg;
    mutable std::mutex singleRay;
};

}  // namespace pbrt

#endif  // NNBVH_INTEGRATION_PBRT_GPUBVH_H
