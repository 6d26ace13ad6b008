// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE, not product code.
//
// A thin extern "C" driver around the UNMODIFIED reference translation units
// (compiled where they lie under /root/reference/src by oracle/Makefile into
// oracle/_ref/libnnbvh_ref.so).  It runs the reference's own
//   BVHAggregate ctor / flattenBVH     (src/pbrt/cpu/aggregates.cpp:140-190, 505-522)
//   BVHAggregate::Intersect/IntersectP (src/pbrt/cpu/aggregates.cpp:529-624)
//   Triangle::Intersect/IntersectP     (src/pbrt/shapes.cpp:320-358)
//   IntersectTriangle                  (src/pbrt/shapes.cpp:172-273)
//   Bounds3f::IntersectP               (src/pbrt/util/vecmath.h:1573-1608)
//   Triangle::InteractionFromIntersection (src/pbrt/shapes.h:884-1010)
// so that the C restatement in oracle/nnbvh_oracle.cpp and the CUDA engine can
// be pinned against the real thing.  Nothing here is copied from the
// reference; it only #includes its headers at build time.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load the resulting library.

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

// BVHAggregate::nodes/primitives, SimplePrimitive::shape and
// Triangle::triIndex are private (cpu/aggregates.h:66-69, cpu/primitive.h:76-79,
// shapes.h:1186-1189); the reference tree is read-only, so the driver opens
// them for itself.  Layout is unaffected by access specifiers.
#define private public
#include <pbrt/pbrt.h>
#include <pbrt/cpu/aggregates.h>
#include <pbrt/textures.h>
#include <pbrt/util/hash.h>
#include <pbrt/cpu/primitive.h>
#include <pbrt/materials.h>
#include <pbrt/options.h>
#include <pbrt/shapes.h>
#include <pbrt/util/buffercache.h>
#include <pbrt/util/log.h>
#include <pbrt/util/loopsubdiv.h>
#include <pbrt/util/mesh.h>
#include <pbrt/util/parallel.h>
#include <pbrt/util/stats.h>
#include <pbrt/util/transform.h>
#include <pbrt/util/vecmath.h>
#undef private

namespace {

// Mirror of the 32-byte record defined in cpu/aggregates.cpp:128-137 (the
// definition lives in the .cpp, so the driver states the layout itself and
// static_asserts it).
struct alignas(32) NodeMirror {
    float bounds[6];
    int offset;
    uint16_t nPrimitives;
    uint8_t axis;
    uint8_t pad;
};
static_assert(sizeof(NodeMirror) == 32, "LinearBVHNode must be 32 bytes");

struct RefSphere {
    pbrt::Transform renderFromObject, objectFromRender;
    pbrt::Sphere *shape = nullptr;
};

// nnbvh_sphere (include/nnbvh.h): renderFromObject[12], objectFromRender[12], radius, zMin, zMax, phiMax.
pbrt::SquareMatrix<4> Matrix12(const float *m) {
    float a[4][4] = {{m[0], m[1], m[2], m[3]}, {m[4], m[5], m[6], m[7]}, {m[8], m[9], m[10], m[11]}, {0, 0, 0, 1}};
    return pbrt::SquareMatrix<4>(a);
}
RefSphere *MakeSphere(const float *s28) {
    auto *rs = new RefSphere;
    pbrt::SquareMatrix<4> r = Matrix12(s28), o = Matrix12(s28 + 12);
    rs->renderFromObject = pbrt::Transform(r, o);
    rs->objectFromRender = pbrt::Transform(o, r);
    rs->shape = new pbrt::Sphere(&rs->renderFromObject, &rs->objectFromRender, false, s28[24], s28[25],
                                 s28[26], s28[27]);
    return rs;
}
// SimplePrimitive::Intersect copies its Material handle into the interaction without looking at it
// (cpu/primitive.cpp:105-108): a fake pointer value carries the primitive's index through.
bool g_tagMaterials = true;  // nnbvh_ref_tag_materials(0): null materials instead, for the queue harnesses below
pbrt::Material TagMaterial(int index) {
    if (!g_tagMaterials)
        return pbrt::Material(nullptr);
    return pbrt::Material((pbrt::DiffuseMaterial *)(uintptr_t)(16 * (uintptr_t)(index + 1)));
}
int TagOf(pbrt::Material m) {
    return (int)((uintptr_t)m.ptr() / 16) - 1;
}

// An ObjectInstance prototype: its own mesh and BVHAggregate, in object space (scene.cpp:1541-1571).
struct RefObject {
    pbrt::TriangleMesh *mesh = nullptr;
    pbrt::BVHAggregate *agg = nullptr;
    int nTris = 0, nNodes = 0;
    std::vector<int> orderedToOrig, origToOrdered;
};
struct RefInstance {
    pbrt::Transform renderFromPrimitive;
    pbrt::TransformedPrimitive *prim = nullptr;
    int object = 0;
};

struct RefScene {
    std::vector<RefObject *> objects;
    std::vector<RefInstance *> instances;  // primitives nTris .. of the list (after the triangles)
    std::vector<RefSphere *> spheres;  // primitives nTris .. nTris + spheres.size() - 1 of the list
    pbrt::TriangleMesh *mesh = nullptr;
    pbrt::BVHAggregate *agg = nullptr;
    std::vector<pbrt::Primitive> primList;  // the list BVHAggregate was given, in the caller's order
    double buildSeconds = 0;                // wall time of the BVHAggregate constructor alone
    int nTris = 0;
    int nNodes = 0;
    std::vector<int> orderedToOrig;  // ordered prim index -> original triangle
    std::vector<int> origToOrdered;
};

bool g_inited = false;

int CountNodes(const NodeMirror *nodes) {
    // Pre-order layout: the node count is one past the largest index reachable.
    int maxIdx = 0;
    std::vector<int> stack{0};
    while (!stack.empty()) {
        int i = stack.back();
        stack.pop_back();
        maxIdx = std::max(maxIdx, i);
        if (nodes[i].nPrimitives == 0) {
            stack.push_back(i + 1);
            stack.push_back(nodes[i].offset);
        }
    }
    return maxIdx + 1;
}

inline pbrt::Ray MakeRay(const float *r) {
    return pbrt::Ray(pbrt::Point3f(r[0], r[1], r[2]), pbrt::Vector3f(r[4], r[5], r[6]), r[7]);
}

}  // namespace

extern "C" {

// nThreads <= 0 -> all cores (ParallelInit(-1) semantics, util/parallel.h:30).
int nnbvh_ref_init(int nThreads) {
    if (g_inited)
        return 0;
    pbrt::PBRTOptions opt;
    opt.nThreads = nThreads > 0 ? nThreads : 0;
    opt.quiet = true;
    pbrt::Options = new pbrt::PBRTOptions(opt);
    pbrt::InitLogging(pbrt::LogLevel::Error, "", false, false);
    pbrt::ParallelInit(nThreads > 0 ? nThreads : pbrt::AvailableCores());
    pbrt::Triangle::Init({});
    pbrt::InitBufferCaches();
    g_inited = true;
    return 0;
}

// Re-create the reference's thread pool with n workers (n<=0: all cores).
// One thread makes the >128K-primitive parallel build deterministic
// (cpu/aggregates.cpp:359-370 orders leaves by an atomic fetch_add).
int nnbvh_ref_set_threads(int nThreads) {
    if (!g_inited)
        return nnbvh_ref_init(nThreads);
    pbrt::ParallelCleanup();
    pbrt::ParallelInit(nThreads > 0 ? nThreads : pbrt::AvailableCores());
    return 0;
}

int nnbvh_ref_threads(void) {
    return pbrt::RunningThreads();
}

// splitMethod: 0 SAH, 1 HLBVH, 2 Middle, 3 EqualCounts (cpu/aggregates.h:31).
void *nnbvh_ref_scene_create_mixed(const float *verts, int nVerts, const int *indices, int nTris,
                                   const float *spheres28, int nSpheres, int maxPrimsInNode,
                                   int splitMethod);

void *nnbvh_ref_scene_create(const float *verts, int nVerts, const int *indices, int nTris,
                             int maxPrimsInNode, int splitMethod) {
    return nnbvh_ref_scene_create_mixed(verts, nVerts, indices, nTris, nullptr, 0, maxPrimsInNode,
                                        splitMethod);
}

// The primitive list is: triangles 0..nTris-1 (SimplePrimitive over Triangle), then the spheres
// (SimplePrimitive over Sphere).  nTris may be 0 when there are spheres.
void *nnbvh_ref_scene_create_mixed(const float *verts, int nVerts, const int *indices, int nTris,
                                   const float *spheres28, int nSpheres, int maxPrimsInNode,
                                   int splitMethod) {
    if (!g_inited)
        nnbvh_ref_init(0);
    auto *s = new RefScene;
    std::vector<int> idx(indices, indices + 3 * (size_t)nTris);
    std::vector<pbrt::Point3f> p(nVerts);
    for (int i = 0; i < nVerts; ++i)
        p[i] = pbrt::Point3f(verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]);
    std::vector<int> face(nTris);
    for (int i = 0; i < nTris; ++i)
        face[i] = i;  // faceIndex of a hit == original triangle id
    pbrt::Allocator alloc;
    s->mesh = alloc.new_object<pbrt::TriangleMesh>(pbrt::Transform(), false, idx, p,
                                                   std::vector<pbrt::Vector3f>(),
                                                   std::vector<pbrt::Normal3f>(),
                                                   std::vector<pbrt::Point2f>(), face, alloc);
    pstd::vector<pbrt::Shape> shapes = pbrt::Triangle::CreateTriangles(s->mesh, alloc);
    std::vector<pbrt::Primitive> prims;
    prims.reserve(nTris);
    for (pbrt::Shape sh : shapes)
        prims.push_back(pbrt::Primitive(
            alloc.new_object<pbrt::SimplePrimitive>(sh, pbrt::Material(nullptr))));
    for (int k = 0; k < nSpheres; ++k) {
        s->spheres.push_back(MakeSphere(spheres28 + 28 * (size_t)k));
        prims.push_back(pbrt::Primitive(alloc.new_object<pbrt::SimplePrimitive>(
            pbrt::Shape(s->spheres.back()->shape), TagMaterial(nTris + k))));
    }
    s->nTris = nTris;
    const int nPrims = nTris + nSpheres;
    s->primList = prims;
    auto t0 = std::chrono::steady_clock::now();
    s->agg = new pbrt::BVHAggregate(std::move(prims), maxPrimsInNode,
                                    (pbrt::BVHAggregate::SplitMethod)splitMethod);
    s->buildSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    s->nNodes = CountNodes(reinterpret_cast<const NodeMirror *>(s->agg->nodes));
    s->orderedToOrig.resize(nPrims);
    s->origToOrdered.resize(nPrims);
    for (int i = 0; i < nPrims; ++i) {
        const pbrt::SimplePrimitive *sp = s->agg->primitives[i].Cast<pbrt::SimplePrimitive>();
        int orig = sp->shape.Is<pbrt::Triangle>() ? sp->shape.Cast<pbrt::Triangle>()->triIndex
                                                  : TagOf(sp->material);
        s->orderedToOrig[i] = orig;
        s->origToOrdered[orig] = i;
    }
    return s;
}

// Wall time of BVHAggregate::BVHAggregate (build + flatten) for this scene, seconds.
double nnbvh_ref_scene_build_seconds(void *scene) {
    return ((RefScene *)scene)->buildSeconds;
}
// Triangles behind GeometricPrimitives with a constant alpha texture each (cpu/primitive.cpp:32-84):
// alpha[i] < 1 -> GeometricPrimitive(shape, no material, no light, no media, FloatConstantTexture(alpha[i])),
// else SimplePrimitive.  alpha = 0 primitives stay in the list (pbrt's scene loader would drop them).
void *nnbvh_ref_scene_create_alpha(const float *verts, int nVerts, const int *indices, int nTris,
                                   const float *alpha, int maxPrimsInNode, int splitMethod) {
    if (!g_inited)
        nnbvh_ref_init(0);
    auto *s = new RefScene;
    std::vector<int> idx(indices, indices + 3 * (size_t)nTris);
    std::vector<pbrt::Point3f> p(nVerts);
    for (int i = 0; i < nVerts; ++i)
        p[i] = pbrt::Point3f(verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]);
    std::vector<int> face(nTris);
    for (int i = 0; i < nTris; ++i)
        face[i] = i;
    pbrt::Allocator alloc;
    s->mesh = alloc.new_object<pbrt::TriangleMesh>(pbrt::Transform(), false, idx, p,
                                                   std::vector<pbrt::Vector3f>(),
                                                   std::vector<pbrt::Normal3f>(),
                                                   std::vector<pbrt::Point2f>(), face, alloc);
    pstd::vector<pbrt::Shape> shapes = pbrt::Triangle::CreateTriangles(s->mesh, alloc);
    std::vector<pbrt::Primitive> prims;
    prims.reserve(nTris);
    for (int i = 0; i < nTris; ++i) {
        if (alpha[i] < 1) {
            pbrt::FloatTexture tex(alloc.new_object<pbrt::FloatConstantTexture>(alpha[i]));
            prims.push_back(pbrt::Primitive(alloc.new_object<pbrt::GeometricPrimitive>(
                shapes[i], pbrt::Material(nullptr), pbrt::Light(nullptr), pbrt::MediumInterface(), tex)));
        } else
            prims.push_back(pbrt::Primitive(
                alloc.new_object<pbrt::SimplePrimitive>(shapes[i], pbrt::Material(nullptr))));
    }
    s->nTris = nTris;
    s->primList = prims;
    s->agg = new pbrt::BVHAggregate(std::move(prims), maxPrimsInNode,
                                    (pbrt::BVHAggregate::SplitMethod)splitMethod);
    s->nNodes = CountNodes(reinterpret_cast<const NodeMirror *>(s->agg->nodes));
    s->orderedToOrig.resize(nTris);
    s->origToOrdered.resize(nTris);
    for (int i = 0; i < nTris; ++i) {
        const pbrt::Primitive &pr = s->agg->primitives[i];
        pbrt::Shape sh = pr.Is<pbrt::SimplePrimitive>() ? pr.Cast<pbrt::SimplePrimitive>()->shape
                                                        : pr.Cast<pbrt::GeometricPrimitive>()->shape;
        int orig = sh.Cast<pbrt::Triangle>()->triIndex;
        s->orderedToOrig[i] = orig;
        s->origToOrdered[orig] = i;
    }
    return s;
}
// Origin of si->intr.SpawnRay(r.d) (cpu/primitive.cpp:63) after Triangle::Intersect on one triangle
// (p9), and tHit; returns 0 when the ray misses.
int nnbvh_ref_spawn_origin(const float *ray8, const float *p9, float *out4) {
    if (!g_inited)
        nnbvh_ref_init(0);
    std::vector<int> idx{0, 1, 2};
    std::vector<pbrt::Point3f> p{pbrt::Point3f(p9[0], p9[1], p9[2]), pbrt::Point3f(p9[3], p9[4], p9[5]),
                                 pbrt::Point3f(p9[6], p9[7], p9[8])};
    pbrt::Allocator alloc;
    pbrt::TriangleMesh *mesh = alloc.new_object<pbrt::TriangleMesh>(
        pbrt::Transform(), false, idx, p, std::vector<pbrt::Vector3f>(), std::vector<pbrt::Normal3f>(),
        std::vector<pbrt::Point2f>(), std::vector<int>(), alloc);
    pstd::vector<pbrt::Shape> shapes = pbrt::Triangle::CreateTriangles(mesh, alloc);
    pbrt::Ray ray = MakeRay(ray8);
    pstd::optional<pbrt::ShapeIntersection> si = shapes[0].Intersect(ray, ray8[3]);
    if (!si)
        return 0;
    pbrt::Ray next = si->intr.SpawnRay(ray.d);
    out4[0] = next.o.x, out4[1] = next.o.y, out4[2] = next.o.z, out4[3] = si->tHit;
    return 1;
}
// HashFloat(Point3f, Vector3f) as GeometricPrimitive::Intersect calls it (cpu/primitive.cpp:60)
float nnbvh_ref_hash_float(const float *o, const float *d) {
    return pbrt::HashFloat(pbrt::Point3f(o[0], o[1], o[2]), pbrt::Vector3f(d[0], d[1], d[2]));
}

int nnbvh_ref_scene_num_nodes(void *scene) {
    return ((RefScene *)scene)->nNodes;
}
int nnbvh_ref_scene_num_prims(void *scene) {
    RefScene *s = (RefScene *)scene;
    return s->nTris + (int)s->spheres.size() + (int)s->instances.size();
}

// nodesOut: nNodes*32 bytes, byte-identical LinearBVHNode records (pad zeroed).
// orderOut: nTris ints, ordered prim -> original triangle.
void nnbvh_ref_scene_export(void *scene, void *nodesOut, int *orderOut) {
    RefScene *s = (RefScene *)scene;
    const NodeMirror *src = reinterpret_cast<const NodeMirror *>(s->agg->nodes);
    NodeMirror *dst = (NodeMirror *)nodesOut;
    for (int i = 0; i < s->nNodes; ++i) {
        dst[i] = src[i];
        dst[i].pad = 0;
        if (dst[i].nPrimitives > 0)
            dst[i].axis = 0;  // axis is uninitialised for leaves in the reference
    }
    std::memcpy(orderOut, s->orderedToOrig.data(), sizeof(int) * s->orderedToOrig.size());
}

// rays: n * 8 floats {ox,oy,oz,tMax,dx,dy,dz,time}.  hits: n * 5 words
// {prim (ordered index, 0xFFFFFFFF miss), t, b0, b1, b2}.
// parallel != 0 -> ParallelFor over rays exactly as CPUAggregate::IntersectClosest
// does (wavefront/aggregate.cpp:42-57).
void nnbvh_ref_intersect(void *scene, const float *rays, int64_t n, uint32_t *hits,
                         int parallel) {
    RefScene *s = (RefScene *)scene;
    pbrt::Primitive top(s->agg);
    auto body = [&](int64_t i) {
        const float *r = rays + 8 * i;
        pbrt::Ray ray = MakeRay(r);
        pstd::optional<pbrt::ShapeIntersection> si = top.Intersect(ray, r[3]);
        uint32_t *h = hits + 5 * i;
        if (!si) {
            h[0] = 0xFFFFFFFFu;
            h[1] = h[2] = h[3] = h[4] = 0;
            return;
        }
        if (int tag = TagOf(si->intr.material); tag >= 0) {
            // a sphere: {prim, tHit, pObj}.  pObj does not survive in SurfaceInteraction either; the
            // accepted hit is the one BasicIntersect returns for tMax = Infinity (shapes.h:183-222).
            h[0] = (uint32_t)s->origToOrdered[tag];
            pstd::optional<pbrt::QuadricIntersection> qi =
                s->spheres[tag - s->nTris]->shape->BasicIntersect(ray, pbrt::Infinity);
            float t = si->tHit, q[3] = {0, 0, 0};
            if (qi)
                q[0] = qi->pObj.x, q[1] = qi->pObj.y, q[2] = qi->pObj.z;
            std::memcpy(&h[1], &t, 4);
            std::memcpy(&h[2], q, 12);
            return;
        }
        int orig = si->intr.faceIndex;
        h[0] = (uint32_t)s->origToOrdered[orig];
        // The barycentrics do not survive in SurfaceInteraction; recover them
        // from the reference's own IntersectTriangle on the winning triangle
        // (b and t do not depend on tMax, shapes.cpp:245-247).
        const int *v = &s->mesh->vertexIndices[3 * orig];
        pstd::optional<pbrt::TriangleIntersection> ti = pbrt::IntersectTriangle(
            ray, pbrt::Infinity, s->mesh->p[v[0]], s->mesh->p[v[1]], s->mesh->p[v[2]]);
        float t = si->tHit, b0 = 0, b1 = 0, b2 = 0;
        if (ti) {
            b0 = ti->b0;
            b1 = ti->b1;
            b2 = ti->b2;
        }
        std::memcpy(&h[1], &t, 4);
        std::memcpy(&h[2], &b0, 4);
        std::memcpy(&h[3], &b1, 4);
        std::memcpy(&h[4], &b2, 4);
    };
    if (parallel)
        pbrt::ParallelFor(0, n, body);
    else
        for (int64_t i = 0; i < n; ++i)
            body(i);
}

// Timing variant: the reference as-is, nothing but the call and a 1-byte result.
void nnbvh_ref_intersect_only(void *scene, const float *rays, int64_t n, uint8_t *hit) {
    RefScene *s = (RefScene *)scene;
    pbrt::Primitive top(s->agg);
    pbrt::ParallelFor(0, n, [&](int64_t i) {
        const float *r = rays + 8 * i;
        hit[i] = top.Intersect(MakeRay(r), r[3]).has_value();
    });
}

void nnbvh_ref_intersectP(void *scene, const float *rays, int64_t n, uint8_t *occluded,
                          int parallel) {
    RefScene *s = (RefScene *)scene;
    pbrt::Primitive top(s->agg);
    auto body = [&](int64_t i) {
        const float *r = rays + 8 * i;
        occluded[i] = top.IntersectP(MakeRay(r), r[3]) ? 1 : 0;
    };
    if (parallel)
        pbrt::ParallelFor(0, n, body);
    else
        for (int64_t i = 0; i < n; ++i)
            body(i);
}

// Direct known-answer hooks ---------------------------------------------------
int nnbvh_ref_triangle(const float *ray8, const float *p9, float *out4) {
    pstd::optional<pbrt::TriangleIntersection> ti = pbrt::IntersectTriangle(
        MakeRay(ray8), ray8[3], pbrt::Point3f(p9[0], p9[1], p9[2]),
        pbrt::Point3f(p9[3], p9[4], p9[5]), pbrt::Point3f(p9[6], p9[7], p9[8]));
    if (!ti)
        return 0;
    out4[0] = ti->t;
    out4[1] = ti->b0;
    out4[2] = ti->b1;
    out4[3] = ti->b2;
    return 1;
}

int nnbvh_ref_box(const float *ray8, const float *box6) {
    pbrt::Bounds3f b(pbrt::Point3f(box6[0], box6[1], box6[2]),
                     pbrt::Point3f(box6[3], box6[4], box6[5]));
    // Bounds3f(p1,p2) takes min/max of its arguments; callers pass pMin<=pMax.
    pbrt::Point3f o(ray8[0], ray8[1], ray8[2]);
    pbrt::Vector3f d(ray8[4], ray8[5], ray8[6]);
    pbrt::Vector3f invDir(1 / d.x, 1 / d.y, 1 / d.z);
    int dirIsNeg[3] = {int(invDir.x < 0), int(invDir.y < 0), int(invDir.z < 0)};
    return b.IntersectP(o, d, ray8[3], invDir, dirIsNeg) ? 1 : 0;
}

// ---- instances: TransformedPrimitive over per-object BVHAggregates (cpu/primitive.cpp:112-130) ----
// Primitive list: top-level triangles 0..nTris-1, then instance k as primitive nTris + k.
// objTris9: the objects' soups back to back, object j = triangles [objStart[j], objStart[j+1]).
// inst24: renderFromPrimitive[12], primitiveFromRender[12] per instance; instObject: its object.
void *nnbvh_ref_scene_create_instanced(const float *tris9, int nTris, const float *objTris9,
                                       const int *objStart, int nObjects, const float *inst24,
                                       const int *instObject, int nInstances, int maxPrimsInNode,
                                       int splitMethod) {
    if (!g_inited)
        nnbvh_ref_init(0);
    auto *s = new RefScene;
    pbrt::Allocator alloc;
    auto makeMesh = [&](const float *t9, int n) {
        std::vector<int> idx(3 * (size_t)n), face(n);
        std::vector<pbrt::Point3f> p(3 * (size_t)n);
        for (int i = 0; i < 3 * n; ++i) {
            idx[i] = i;
            p[i] = pbrt::Point3f(t9[3 * i], t9[3 * i + 1], t9[3 * i + 2]);
        }
        for (int i = 0; i < n; ++i)
            face[i] = i;
        return alloc.new_object<pbrt::TriangleMesh>(pbrt::Transform(), false, idx, p,
                                                    std::vector<pbrt::Vector3f>(),
                                                    std::vector<pbrt::Normal3f>(),
                                                    std::vector<pbrt::Point2f>(), face, alloc);
    };
    for (int j = 0; j < nObjects; ++j) {
        auto *o = new RefObject;
        o->nTris = objStart[j + 1] - objStart[j];
        o->mesh = makeMesh(objTris9 + 9 * (size_t)objStart[j], o->nTris);
        std::vector<pbrt::Primitive> prims;
        for (pbrt::Shape sh : pbrt::Triangle::CreateTriangles(o->mesh, alloc))
            prims.push_back(pbrt::Primitive(alloc.new_object<pbrt::SimplePrimitive>(sh, TagMaterial(j))));
        o->agg = new pbrt::BVHAggregate(std::move(prims), maxPrimsInNode,
                                        (pbrt::BVHAggregate::SplitMethod)splitMethod);
        o->nNodes = CountNodes(reinterpret_cast<const NodeMirror *>(o->agg->nodes));
        o->orderedToOrig.resize(o->nTris);
        o->origToOrdered.resize(o->nTris);
        for (int i = 0; i < o->nTris; ++i) {
            const pbrt::SimplePrimitive *sp = o->agg->primitives[i].Cast<pbrt::SimplePrimitive>();
            int orig = sp->shape.Cast<pbrt::Triangle>()->triIndex;
            o->orderedToOrig[i] = orig;
            o->origToOrdered[orig] = i;
        }
        s->objects.push_back(o);
    }
    s->mesh = makeMesh(tris9, nTris);
    std::vector<pbrt::Primitive> prims;
    for (pbrt::Shape sh : pbrt::Triangle::CreateTriangles(s->mesh, alloc))
        prims.push_back(pbrt::Primitive(alloc.new_object<pbrt::SimplePrimitive>(sh, pbrt::Material(nullptr))));
    std::map<const void *, int> instIndex;
    for (int k = 0; k < nInstances; ++k) {
        auto *in = new RefInstance;
        in->renderFromPrimitive = pbrt::Transform(Matrix12(inst24 + 24 * (size_t)k), Matrix12(inst24 + 24 * (size_t)k + 12));
        in->object = instObject[k];
        in->prim = new pbrt::TransformedPrimitive(pbrt::Primitive(s->objects[in->object]->agg), &in->renderFromPrimitive);
        instIndex[in->prim] = k;
        prims.push_back(pbrt::Primitive(in->prim));
        s->instances.push_back(in);
    }
    s->nTris = nTris;
    const int nPrims = nTris + nInstances;
    s->agg = new pbrt::BVHAggregate(std::move(prims), maxPrimsInNode,
                                    (pbrt::BVHAggregate::SplitMethod)splitMethod);
    s->nNodes = CountNodes(reinterpret_cast<const NodeMirror *>(s->agg->nodes));
    s->orderedToOrig.resize(nPrims);
    s->origToOrdered.resize(nPrims);
    for (int i = 0; i < nPrims; ++i) {
        pbrt::Primitive p = s->agg->primitives[i];
        int orig = p.Is<pbrt::TransformedPrimitive>()
                       ? nTris + instIndex[p.Cast<pbrt::TransformedPrimitive>()]
                       : p.Cast<pbrt::SimplePrimitive>()->shape.Cast<pbrt::Triangle>()->triIndex;
        s->orderedToOrig[i] = orig;
        s->origToOrdered[orig] = i;
    }
    return s;
}

int nnbvh_ref_scene_object_num_nodes(void *scene, int j) {
    return ((RefScene *)scene)->objects[j]->nNodes;
}
void nnbvh_ref_scene_export_object(void *scene, int j, void *nodesOut, int *orderOut) {
    RefObject *o = ((RefScene *)scene)->objects[j];
    const NodeMirror *src = reinterpret_cast<const NodeMirror *>(o->agg->nodes);
    NodeMirror *dst = (NodeMirror *)nodesOut;
    for (int i = 0; i < o->nNodes; ++i) {
        dst[i] = src[i];
        dst[i].pad = 0;
        if (dst[i].nPrimitives > 0)
            dst[i].axis = 0;
    }
    std::memcpy(orderOut, o->orderedToOrig.data(), sizeof(int) * o->nTris);
}

// Closest hits of an instanced scene: hits5 as nnbvh_ref_intersect (prim = ordered TOP-LEVEL
// primitive; for an instance t/b are the inner triangle's, in the instance's ray parametrisation,
// exactly what TransformedPrimitive::Intersect returns) and inner[i] = ordered primitive inside the
// object (0xFFFFFFFF for a top-level triangle or a miss).  SurfaceInteraction does not say WHICH
// instance was hit: it is found by asking each instance of the hit object on its own.
void nnbvh_ref_intersect_instanced(void *scene, const float *rays, int64_t n, uint32_t *hits,
                                   uint32_t *inner) {
    RefScene *s = (RefScene *)scene;
    pbrt::Primitive top(s->agg);
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        pbrt::Ray ray = MakeRay(r);
        pstd::optional<pbrt::ShapeIntersection> si = top.Intersect(ray, r[3]);
        uint32_t *h = hits + 5 * i;
        inner[i] = 0xFFFFFFFFu;
        h[0] = 0xFFFFFFFFu;
        h[1] = h[2] = h[3] = h[4] = 0;
        if (!si)
            continue;
        float t = si->tHit, b[3] = {0, 0, 0};
        int obj = TagOf(si->intr.material), face = si->intr.faceIndex;
        if (obj < 0) {
            h[0] = (uint32_t)s->origToOrdered[face];
            const int *v = &s->mesh->vertexIndices[3 * face];
            if (auto ti = pbrt::IntersectTriangle(ray, pbrt::Infinity, s->mesh->p[v[0]], s->mesh->p[v[1]], s->mesh->p[v[2]]))
                b[0] = ti->b0, b[1] = ti->b1, b[2] = ti->b2;
        } else {
            int found = -1;
            for (size_t k = 0; k < s->instances.size() && found < 0; ++k) {
                if (s->instances[k]->object != obj)
                    continue;
                auto sk = s->instances[k]->prim->Intersect(ray, pbrt::Infinity);
                if (sk && sk->tHit == t && sk->intr.faceIndex == face)
                    found = (int)k;
            }
            if (found < 0) {
                h[0] = 0xFFFFFFFEu;  // cannot happen; makes a test fail loudly
                continue;
            }
            RefObject *o = s->objects[obj];
            h[0] = (uint32_t)s->origToOrdered[s->nTris + found];
            inner[i] = (uint32_t)o->origToOrdered[face];
            pbrt::Ray local = s->instances[found]->renderFromPrimitive.ApplyInverse(ray, nullptr);
            const int *v = &o->mesh->vertexIndices[3 * face];
            if (auto ti = pbrt::IntersectTriangle(local, pbrt::Infinity, o->mesh->p[v[0]], o->mesh->p[v[1]], o->mesh->p[v[2]]))
                b[0] = ti->b0, b[1] = ti->b1, b[2] = ti->b2;
        }
        std::memcpy(&h[1], &t, 4);
        std::memcpy(&h[2], b, 12);
    }
}

// ---- general scenes: every primitive kind of the path in ONE list ------------------------------------------
// Primitive list: triangles 0..nTris-1 (GeometricPrimitive with FloatConstantTexture(alpha[i]) where alpha[i] < 1,
// else SimplePrimitive), spheres nTris..nTris+nSpheres-1 (SimplePrimitive, tagged), instances after them
// (TransformedPrimitive over per-object BVHAggregates whose triangles carry the object's tag).
constexpr int kObjectTagBase = 1 << 24;
// Scenes created while this is off carry null materials everywhere (EnqueueWorkAfterIntersection's `!material`
// branch, wavefront/intersect.h:101-109, must not meet a tag); nnbvh_ref_intersect* cannot name the primitive
// hit in such a scene, nnbvh_ref_intersect_surface and the queue harnesses can be used.
void nnbvh_ref_tag_materials(int on) { g_tagMaterials = on != 0; }

void *nnbvh_ref_scene_create_general(const float *tris9, int nTris, const float *alpha, const float *spheres28,
                                     int nSpheres, const float *objTris9, const int *objStart, int nObjects,
                                     const float *inst24, const int *instObject, int nInstances,
                                     int maxPrimsInNode, int splitMethod) {
    if (!g_inited)
        nnbvh_ref_init(0);
    auto *s = new RefScene;
    pbrt::Allocator alloc;
    auto makeMesh = [&](const float *t9, int n) {
        std::vector<int> idx(3 * (size_t)n), face(n);
        std::vector<pbrt::Point3f> p(3 * (size_t)n);
        for (int i = 0; i < 3 * n; ++i) {
            idx[i] = i;
            p[i] = pbrt::Point3f(t9[3 * i], t9[3 * i + 1], t9[3 * i + 2]);
        }
        for (int i = 0; i < n; ++i)
            face[i] = i;
        return alloc.new_object<pbrt::TriangleMesh>(pbrt::Transform(), false, idx, p, std::vector<pbrt::Vector3f>(),
                                                    std::vector<pbrt::Normal3f>(), std::vector<pbrt::Point2f>(), face,
                                                    alloc);
    };
    for (int j = 0; j < nObjects; ++j) {
        auto *o = new RefObject;
        o->nTris = objStart[j + 1] - objStart[j];
        o->mesh = makeMesh(objTris9 + 9 * (size_t)objStart[j], o->nTris);
        std::vector<pbrt::Primitive> prims;
        for (pbrt::Shape sh : pbrt::Triangle::CreateTriangles(o->mesh, alloc))
            prims.push_back(pbrt::Primitive(alloc.new_object<pbrt::SimplePrimitive>(sh, TagMaterial(kObjectTagBase + j))));
        o->agg = new pbrt::BVHAggregate(std::move(prims), maxPrimsInNode, (pbrt::BVHAggregate::SplitMethod)splitMethod);
        o->nNodes = CountNodes(reinterpret_cast<const NodeMirror *>(o->agg->nodes));
        o->orderedToOrig.resize(o->nTris);
        o->origToOrdered.resize(o->nTris);
        for (int i = 0; i < o->nTris; ++i) {
            int orig = o->agg->primitives[i].Cast<pbrt::SimplePrimitive>()->shape.Cast<pbrt::Triangle>()->triIndex;
            o->orderedToOrig[i] = orig;
            o->origToOrdered[orig] = i;
        }
        s->objects.push_back(o);
    }
    s->mesh = makeMesh(tris9, nTris);
    pstd::vector<pbrt::Shape> shapes = pbrt::Triangle::CreateTriangles(s->mesh, alloc);
    std::vector<pbrt::Primitive> prims;
    for (int i = 0; i < nTris; ++i) {
        if (alpha && alpha[i] < 1) {
            pbrt::FloatTexture tex(alloc.new_object<pbrt::FloatConstantTexture>(alpha[i]));
            prims.push_back(pbrt::Primitive(alloc.new_object<pbrt::GeometricPrimitive>(
                shapes[i], pbrt::Material(nullptr), pbrt::Light(nullptr), pbrt::MediumInterface(), tex)));
        } else
            prims.push_back(pbrt::Primitive(alloc.new_object<pbrt::SimplePrimitive>(shapes[i], pbrt::Material(nullptr))));
    }
    for (int k = 0; k < nSpheres; ++k) {
        s->spheres.push_back(MakeSphere(spheres28 + 28 * (size_t)k));
        prims.push_back(pbrt::Primitive(alloc.new_object<pbrt::SimplePrimitive>(pbrt::Shape(s->spheres.back()->shape),
                                                                                TagMaterial(nTris + k))));
    }
    std::map<const void *, int> instIndex;
    for (int k = 0; k < nInstances; ++k) {
        auto *in = new RefInstance;
        in->renderFromPrimitive = pbrt::Transform(Matrix12(inst24 + 24 * (size_t)k), Matrix12(inst24 + 24 * (size_t)k + 12));
        in->object = instObject[k];
        in->prim = new pbrt::TransformedPrimitive(pbrt::Primitive(s->objects[in->object]->agg), &in->renderFromPrimitive);
        instIndex[in->prim] = k;
        prims.push_back(pbrt::Primitive(in->prim));
        s->instances.push_back(in);
    }
    s->nTris = nTris;
    const int nPrims = nTris + nSpheres + nInstances;
    s->primList = prims;
    s->agg = new pbrt::BVHAggregate(std::move(prims), maxPrimsInNode, (pbrt::BVHAggregate::SplitMethod)splitMethod);
    s->nNodes = CountNodes(reinterpret_cast<const NodeMirror *>(s->agg->nodes));
    s->orderedToOrig.resize(nPrims);
    s->origToOrdered.resize(nPrims);
    for (int i = 0; i < nPrims; ++i) {
        pbrt::Primitive p = s->agg->primitives[i];
        int orig;
        if (p.Is<pbrt::TransformedPrimitive>())
            orig = nTris + nSpheres + instIndex[p.Cast<pbrt::TransformedPrimitive>()];
        else if (p.Is<pbrt::GeometricPrimitive>())
            orig = p.Cast<pbrt::GeometricPrimitive>()->shape.Cast<pbrt::Triangle>()->triIndex;
        else {
            const pbrt::SimplePrimitive *sp = p.Cast<pbrt::SimplePrimitive>();
            orig = sp->shape.Is<pbrt::Triangle>() ? sp->shape.Cast<pbrt::Triangle>()->triIndex : TagOf(sp->material);
        }
        s->orderedToOrig[i] = orig;
        if (orig < 0)  // a sphere of a scene made with nnbvh_ref_tag_materials(0): it cannot be named
            continue;
        s->origToOrdered[orig] = i;
    }
    return s;
}

// BVHAggregate::Intersect over a general scene: hits5 / inner as nnbvh_ref_intersect_instanced, sphere hits as
// nnbvh_ref_intersect ({prim, tHit, pObj}).
void nnbvh_ref_intersect_general(void *scene, const float *rays, int64_t n, uint32_t *hits, uint32_t *inner) {
    RefScene *s = (RefScene *)scene;
    pbrt::Primitive top(s->agg);
    const int nSph = (int)s->spheres.size();
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        pbrt::Ray ray = MakeRay(r);
        pstd::optional<pbrt::ShapeIntersection> si = top.Intersect(ray, r[3]);
        uint32_t *h = hits + 5 * i;
        inner[i] = 0xFFFFFFFFu;
        h[0] = 0xFFFFFFFFu;
        h[1] = h[2] = h[3] = h[4] = 0;
        if (!si)
            continue;
        float t = si->tHit, b[3] = {0, 0, 0};
        int tag = TagOf(si->intr.material), face = si->intr.faceIndex;
        if (tag < 0) {  // a top-level triangle
            h[0] = (uint32_t)s->origToOrdered[face];
            const int *v = &s->mesh->vertexIndices[3 * face];
            const pbrt::Point3f p0 = s->mesh->p[v[0]], p1 = s->mesh->p[v[1]], p2 = s->mesh->p[v[2]];
            // barycentrics as nnbvh_ref_intersect recovers them (b and t do not depend on tMax, shapes.cpp:245-247);
            // an alpha-rejected hit is re-traced against the same flat triangle, i.e. in practice skipped
            if (auto ti = pbrt::IntersectTriangle(ray, pbrt::Infinity, p0, p1, p2))
                b[0] = ti->b0, b[1] = ti->b1, b[2] = ti->b2;
        } else if (tag < kObjectTagBase) {  // a sphere
            h[0] = (uint32_t)s->origToOrdered[tag];
            pstd::optional<pbrt::QuadricIntersection> qi = s->spheres[tag - s->nTris]->shape->BasicIntersect(ray, pbrt::Infinity);
            if (qi)
                b[0] = qi->pObj.x, b[1] = qi->pObj.y, b[2] = qi->pObj.z;
        } else {  // a triangle of an object: find the instance
            const int obj = tag - kObjectTagBase;
            int found = -1;
            for (size_t k = 0; k < s->instances.size() && found < 0; ++k) {
                if (s->instances[k]->object != obj)
                    continue;
                auto sk = s->instances[k]->prim->Intersect(ray, pbrt::Infinity);
                if (sk && sk->tHit == t && sk->intr.faceIndex == face)
                    found = (int)k;
            }
            if (found < 0) {
                h[0] = 0xFFFFFFFEu;  // cannot happen; makes a test fail loudly
                continue;
            }
            RefObject *o = s->objects[obj];
            h[0] = (uint32_t)s->origToOrdered[s->nTris + nSph + found];
            inner[i] = (uint32_t)o->origToOrdered[face];
            pbrt::Ray local = s->instances[found]->renderFromPrimitive.ApplyInverse(ray, nullptr);
            const int *v = &o->mesh->vertexIndices[3 * face];
            if (auto ti = pbrt::IntersectTriangle(local, pbrt::Infinity, o->mesh->p[v[0]], o->mesh->p[v[1]], o->mesh->p[v[2]]))
                b[0] = ti->b0, b[1] = ti->b1, b[2] = ti->b2;
        }
        std::memcpy(&h[1], &t, 4);
        std::memcpy(&h[2], b, 12);
    }
}

// TransformedPrimitive::Bounds() of instance k (cpu/primitive.h:94).
void nnbvh_ref_instance_bounds(void *scene, int k, float *out6) {
    pbrt::Bounds3f b = ((RefScene *)scene)->instances[k]->prim->Bounds();
    out6[0] = b.pMin.x, out6[1] = b.pMin.y, out6[2] = b.pMin.z;
    out6[3] = b.pMax.x, out6[4] = b.pMax.y, out6[5] = b.pMax.z;
}

// Sphere::BasicIntersect (shapes.h:147-232) and Sphere::Bounds (shapes.cpp:37-40) on one sphere
// given as 28 floats (nnbvh_sphere).  out4 = {tHit, pObj.xyz}.
int nnbvh_ref_sphere(const float *ray8, const float *sphere28, float *out4) {
    RefSphere *rs = MakeSphere(sphere28);
    pstd::optional<pbrt::QuadricIntersection> qi = rs->shape->BasicIntersect(MakeRay(ray8), ray8[3]);
    int hit = qi.has_value();
    if (hit) {
        out4[0] = qi->tHit;
        out4[1] = qi->pObj.x;
        out4[2] = qi->pObj.y;
        out4[3] = qi->pObj.z;
    }
    delete rs->shape;
    delete rs;
    return hit;
}
void nnbvh_ref_sphere_bounds(const float *sphere28, float *out6) {
    RefSphere *rs = MakeSphere(sphere28);
    pbrt::Bounds3f b = rs->shape->Bounds();
    out6[0] = b.pMin.x, out6[1] = b.pMin.y, out6[2] = b.pMin.z;
    out6[3] = b.pMax.x, out6[4] = b.pMax.y, out6[5] = b.pMax.z;
    delete rs->shape;
    delete rs;
}

// 39 words of a SurfaceInteraction: pi.low xyz, pi.high xyz, n, uv, dpdu, dpdv, dndu, dndv,
// shading.{n, dpdu, dpdv, dndu, dndv}, faceIndex (int).
static void PutSurface39(const pbrt::SurfaceInteraction &si, float *o) {
    auto put3 = [&o](float x, float y, float z) { *o++ = x, *o++ = y, *o++ = z; };
    put3(si.pi.x.LowerBound(), si.pi.y.LowerBound(), si.pi.z.LowerBound());
    put3(si.pi.x.UpperBound(), si.pi.y.UpperBound(), si.pi.z.UpperBound());
    put3(si.n.x, si.n.y, si.n.z);
    *o++ = si.uv.x, *o++ = si.uv.y;
    put3(si.dpdu.x, si.dpdu.y, si.dpdu.z);
    put3(si.dpdv.x, si.dpdv.y, si.dpdv.z);
    put3(si.dndu.x, si.dndu.y, si.dndu.z);
    put3(si.dndv.x, si.dndv.y, si.dndv.z);
    put3(si.shading.n.x, si.shading.n.y, si.shading.n.z);
    put3(si.shading.dpdu.x, si.shading.dpdu.y, si.shading.dpdu.z);
    put3(si.shading.dpdv.x, si.shading.dpdv.y, si.shading.dpdv.z);
    put3(si.shading.dndu.x, si.shading.dndu.y, si.shading.dndu.z);
    put3(si.shading.dndv.x, si.shading.dndv.y, si.shading.dndv.z);
    int fi = si.faceIndex;
    std::memcpy(o, &fi, 4);
}

// Triangle::InteractionFromIntersection (shapes.h:884-1010) on a caller-supplied mesh: the host
// step that turns an engine hit record {prim,t,b0,b1,b2} back into a SurfaceInteraction
// (pins include/nnbvh_wavefront.hpp).  normals / tangents / uvs may be NULL.
// query i: triangle tri[i], barycentrics b[3i..], wo_time[4i..] = {wo.xyz, time}.
// out: 39 words per query = pi.low xyz, pi.high xyz, n, uv, dpdu, dpdv, dndu, dndv,
// shading.{n, dpdu, dpdv, dndu, dndv}, faceIndex (int).
int nnbvh_ref_triangle_interaction(const float *verts, int nVerts, const int *indices, int nTris,
                                   const float *normals, const float *tangents, const float *uvs,
                                   int reverseOrientation, const int *tri, const float *b,
                                   const float *wo_time, int64_t n, float *out) {
    if (!g_inited)
        nnbvh_ref_init(0);
    std::vector<int> idx(indices, indices + 3 * (size_t)nTris);
    std::vector<pbrt::Point3f> p(nVerts);
    std::vector<pbrt::Normal3f> nn;
    std::vector<pbrt::Vector3f> ss;
    std::vector<pbrt::Point2f> uv;
    for (int i = 0; i < nVerts; ++i)
        p[i] = pbrt::Point3f(verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]);
    if (normals)
        for (int i = 0; i < nVerts; ++i)
            nn.push_back(pbrt::Normal3f(normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]));
    if (tangents)
        for (int i = 0; i < nVerts; ++i)
            ss.push_back(pbrt::Vector3f(tangents[3 * i], tangents[3 * i + 1], tangents[3 * i + 2]));
    if (uvs)
        for (int i = 0; i < nVerts; ++i)
            uv.push_back(pbrt::Point2f(uvs[2 * i], uvs[2 * i + 1]));
    std::vector<int> face(nTris);
    for (int i = 0; i < nTris; ++i)
        face[i] = i;
    pbrt::Allocator alloc;
    pbrt::TriangleMesh *mesh = alloc.new_object<pbrt::TriangleMesh>(
        pbrt::Transform(), reverseOrientation != 0, idx, p, ss, nn, uv, face, alloc);
    for (int64_t i = 0; i < n; ++i) {
        pbrt::TriangleIntersection ti{b[3 * i], b[3 * i + 1], b[3 * i + 2], 1.f};
        const float *w = wo_time + 4 * i;
        pbrt::SurfaceInteraction si = pbrt::Triangle::InteractionFromIntersection(
            mesh, tri[i], ti, w[3], pbrt::Vector3f(w[0], w[1], w[2]));
        PutSurface39(si, out + 39 * i);
    }
    return 0;
}

// Sphere::Intersect (shapes.h:130-139) = BasicIntersect + InteractionFromIntersection (:238-280) on
// one sphere (28 floats, nnbvh_sphere) with the given reverseOrientation, for n rays
// {o, tMax, d, time}.  hit[i] = 0/1; where 1: tHit[i] and out[42 i ..] = the 39 words above + the
// interaction's wo.  Pins include/nnbvh_wavefront.hpp SphereInteractionFromIntersection.
int nnbvh_ref_sphere_interaction(const float *sphere28, int reverseOrientation, const float *rays8,
                                 int64_t n, int *hit, float *tHit, float *out) {
    if (!g_inited)
        nnbvh_ref_init(0);
    pbrt::SquareMatrix<4> r = Matrix12(sphere28), o = Matrix12(sphere28 + 12);
    pbrt::Transform renderFromObject(r, o), objectFromRender(o, r);
    pbrt::Sphere sphere(&renderFromObject, &objectFromRender, reverseOrientation != 0, sphere28[24],
                        sphere28[25], sphere28[26], sphere28[27]);
    for (int64_t i = 0; i < n; ++i) {
        const float *q = rays8 + 8 * i;
        pstd::optional<pbrt::ShapeIntersection> si = sphere.Intersect(MakeRay(q), q[3]);
        hit[i] = si.has_value();
        if (!si)
            continue;
        tHit[i] = si->tHit;
        PutSurface39(si->intr, out + 42 * i);
        out[42 * i + 39] = si->intr.wo.x, out[42 * i + 40] = si->intr.wo.y, out[42 * i + 41] = si->intr.wo.z;
    }
    return 0;
}

// BVHAggregate::Intersect on any scene of this driver (triangles, triangles + spheres, triangles +
// instances), returning the SurfaceInteraction itself: hit[i], tHit[i] and 42 words per ray (the 39
// of PutSurface39 + wo).  The end-to-end pin for include/nnbvh_wavefront.hpp.
void nnbvh_ref_intersect_surface(void *scene, const float *rays, int64_t n, int *hit, float *tHit,
                                 float *out42) {
    RefScene *s = (RefScene *)scene;
    pbrt::Primitive top(s->agg);
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        pstd::optional<pbrt::ShapeIntersection> si = top.Intersect(MakeRay(r), r[3]);
        hit[i] = si.has_value();
        if (!si)
            continue;
        tHit[i] = si->tHit;
        PutSurface39(si->intr, out42 + 42 * i);
        out42[42 * i + 39] = si->intr.wo.x, out42[42 * i + 40] = si->intr.wo.y, out42[42 * i + 41] = si->intr.wo.z;
    }
}

// Interaction(pi, n, uv, wo, time) (interaction.h:30-32) keeps Normalize(wo): what a triangle
// hit's SurfaceInteraction::wo holds for the wo = -ray.d it is given.
void nnbvh_ref_interaction_wo(const float *wo3, int64_t n, float *out3) {
    for (int64_t i = 0; i < n; ++i) {
        pbrt::Interaction it(pbrt::Point3fi(), pbrt::Normal3f(), pbrt::Point2f(),
                             pbrt::Vector3f(wo3[3 * i], wo3[3 * i + 1], wo3[3 * i + 2]), 0.f);
        out3[3 * i] = it.wo.x, out3[3 * i + 1] = it.wo.y, out3[3 * i + 2] = it.wo.z;
    }
}

// Loop subdivision through the reference (util/loopsubdiv.cpp) so that
// config 1's killeroo geometry matches what pbrt itself would trace.
// Returns a malloc'd mesh via out params; caller frees with nnbvh_ref_free.
int nnbvh_ref_loop_subdivide(int nLevels, const int *indices, int nIdx, const float *verts,
                             int nVerts, int **outIdx, int *outNIdx, float **outVerts,
                             int *outNVerts) {
    if (!g_inited)
        nnbvh_ref_init(0);
    std::vector<int> idx(indices, indices + nIdx);
    std::vector<pbrt::Point3f> p(nVerts);
    for (int i = 0; i < nVerts; ++i)
        p[i] = pbrt::Point3f(verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]);
    pbrt::Transform identity;
    pbrt::TriangleMesh *m =
        pbrt::LoopSubdivide(&identity, false, nLevels, idx, p, pbrt::Allocator());
    *outNIdx = 3 * m->nTriangles;
    *outNVerts = m->nVertices;
    *outIdx = (int *)std::malloc(sizeof(int) * *outNIdx);
    *outVerts = (float *)std::malloc(sizeof(float) * 3 * *outNVerts);
    std::memcpy(*outIdx, m->vertexIndices, sizeof(int) * *outNIdx);
    for (int i = 0; i < m->nVertices; ++i) {
        (*outVerts)[3 * i] = m->p[i].x;
        (*outVerts)[3 * i + 1] = m->p[i].y;
        (*outVerts)[3 * i + 2] = m->p[i].z;
    }
    return 0;
}

void nnbvh_ref_free(void *p) {
    std::free(p);
}

// Counter read-out: merges the calling thread's thread-local STAT counters
// (cpu/aggregates.cpp:27, shapes.cpp:148) into the global accumulator and
// parses PrintStats.  Only meaningful after single-threaded (parallel=0)
// calls made on this same thread.  Resets the counters.
int nnbvh_ref_stats(int64_t *nodesVisited, int64_t *triTests, int64_t *triHits) {
    pbrt::ReportThreadStats();
    char *buf = nullptr;
    size_t len = 0;
    FILE *f = open_memstream(&buf, &len);
    pbrt::PrintStats(f);
    std::fclose(f);
    pbrt::ClearStats();
    *nodesVisited = 0;
    *triTests = 0;
    *triHits = 0;
    std::istringstream in(std::string(buf ? buf : "", len));
    std::free(buf);
    std::string line;
    while (std::getline(in, line)) {
        if (line.find("Nodes visited") != std::string::npos) {
            size_t p = line.find("Nodes visited") + 13;
            *nodesVisited = std::strtoll(line.c_str() + p, nullptr, 10);
        } else if (line.find("Ray-Triangle intersection tests") != std::string::npos) {
            // "<hits> / <tests> (x.xx%)"
            size_t p = line.find("tests") + 5;
            char *end = nullptr;
            *triHits = std::strtoll(line.c_str() + p, &end, 10);
            while (*end && (*end == ' ' || *end == '/'))
                ++end;
            *triTests = std::strtoll(end, nullptr, 10);
        }
    }
    return 0;
}

}  // extern "C"

// ---- the reference-side binding, run on the reference's own primitives --------------------------------
// Compiled only into oracle/_ref/libnnbvh_ref_binding.so (which links libnnbvh.so): pbrt::GPUBVHAggregate of
// integration/pbrt/gpubvh.h is constructed over the SAME std::vector<Primitive> a BVHAggregate of this
// driver got, so its Intersect / IntersectP / Resolve can be compared with BVHAggregate's, interaction
// for interaction (tests/test_binding.py; needs a GPU).
#ifdef NNBVH_WITH_BINDING
#include "../integration/pbrt/gpubvh.h"
#include "../integration/pbrt/gpubvh_wavefront.h"

namespace {
thread_local std::string g_bindingError;
void PutSurface42(const pbrt::SurfaceInteraction &intr, float *o) {
    PutSurface39(intr, o);
    o[39] = intr.wo.x, o[40] = intr.wo.y, o[41] = intr.wo.z;
}
}  // namespace

extern "C" {
const char *nnbvh_binding_last_error(void) { return g_bindingError.c_str(); }

void *nnbvh_binding_create(void *scene, int maxPrimsInNode, int splitMethod) {
    RefScene *s = (RefScene *)scene;
    try {
        return new pbrt::GPUBVHAggregate(s->primList, maxPrimsInNode,
                                         (pbrt::BVHAggregate::SplitMethod)splitMethod);
    } catch (const nnbvh::Error &e) {
        g_bindingError = e.what();
        return nullptr;
    }
}
void nnbvh_binding_destroy(void *b) { delete (pbrt::GPUBVHAggregate *)b; }

// Primitive's single-ray surface: Intersect(ray, tMax) per ray
int nnbvh_binding_intersect_surface(void *binding, const float *rays, int64_t n, int *hit, float *tHit,
                                    float *out42) {
    const pbrt::GPUBVHAggregate *b = (const pbrt::GPUBVHAggregate *)binding;
    try {
        for (int64_t i = 0; i < n; ++i) {
            const float *r = rays + 8 * i;
            pstd::optional<pbrt::ShapeIntersection> si = b->Intersect(MakeRay(r), r[3]);
            hit[i] = si.has_value();
            if (!si)
                continue;
            tHit[i] = si->tHit;
            PutSurface42(si->intr, out42 + 42 * i);
        }
    } catch (const nnbvh::Error &e) {
        g_bindingError = e.what();
        return e.code;
    }
    return 0;
}
// The same from pbrt's ParallelFor workers at once: what an integrator that holds the binding as its
// Primitive does.  nnbvh_engine is not re-entrant; the binding's single-ray calls take a lock.
int nnbvh_binding_intersect_surface_parallel(void *binding, const float *rays, int64_t n, int *hit, float *tHit,
                                             float *out42) {
    const pbrt::GPUBVHAggregate *b = (const pbrt::GPUBVHAggregate *)binding;
    std::atomic<int> failed{0};
    pbrt::ParallelFor(0, n, [&](int64_t i) {
        try {
            const float *r = rays + 8 * i;
            pstd::optional<pbrt::ShapeIntersection> si = b->Intersect(MakeRay(r), r[3]);
            hit[i] = si.has_value();
            if (!si)
                return;
            tHit[i] = si->tHit;
            PutSurface42(si->intr, out42 + 42 * i);
        } catch (const nnbvh::Error &e) {
            failed = e.code;
        }
    });
    if (failed)
        g_bindingError = "a single-ray call failed in a worker thread";
    return failed;
}
// The batch surface: one IntersectBatch, then Resolve per hit
int nnbvh_binding_intersect_surface_batch(void *binding, const float *rays, int64_t n, int *hit,
                                          float *tHit, float *out42) {
    const pbrt::GPUBVHAggregate *b = (const pbrt::GPUBVHAggregate *)binding;
    try {
        std::vector<nnbvh_hit> hits((size_t)n);
        std::vector<uint32_t> inner((size_t)n, NNBVH_MISS);
        b->IntersectBatch(reinterpret_cast<const nnbvh_ray *>(rays), (uint64_t)n, hits.data(), inner.data());
        for (int64_t i = 0; i < n; ++i) {
            hit[i] = hits[i].prim != NNBVH_MISS;
            if (!hit[i])
                continue;
            pbrt::ShapeIntersection si = b->Resolve(hits[i], inner[i], MakeRay(rays + 8 * i));
            tHit[i] = si.tHit;
            PutSurface42(si.intr, out42 + 42 * i);
        }
    } catch (const nnbvh::Error &e) {
        g_bindingError = e.what();
        return e.code;
    }
    return 0;
}
int nnbvh_binding_intersect_p(void *binding, const float *rays, int64_t n, uint8_t *occluded) {
    const pbrt::GPUBVHAggregate *b = (const pbrt::GPUBVHAggregate *)binding;
    try {
        for (int64_t i = 0; i < n; ++i)
            occluded[i] = b->IntersectP(MakeRay(rays + 8 * i), rays[8 * i + 3]) ? 1 : 0;
    } catch (const nnbvh::Error &e) {
        g_bindingError = e.what();
        return e.code;
    }
    return 0;
}
// ---- the wavefront subclass, run with the reference's own queues ----
// One IntersectClosest over a real RayQueue filled by PushCameraRay (pixelIndex = ray index), either by
// pbrt::GPUBVHWavefrontAggregate (useBinding) or by CPUAggregate::IntersectClosest's loop
// (wavefront/aggregate.cpp:42-57) over the reference BVHAggregate.  The scene's materials are null, so
// every hit takes EnqueueWorkAfterIntersection's `!material` branch (wavefront/intersect.h:101-109): the
// ray intr.SpawnRay(r.ray.d) -- pi's error bounds, n -- goes to nextRayQ; misses go to escapedRayQ.
// kind[i]: 0 escaped, 1 continued, -1 lost; nextRays8[i] = the continued ray {o, 0, d, time}.
int nnbvh_binding_wavefront_closest(void *scene, int maxPrimsInNode, int splitMethod, int useBinding,
                                    const float *rays, int n, int8_t *kind, float *nextRays8) {
    RefScene *s = (RefScene *)scene;
    pbrt::Allocator alloc;
    pbrt::RayQueue rayQ(n, alloc), nextQ(n, alloc);
    pbrt::EscapedRayQueue escapedQ(n, alloc);
    pbrt::SampledWavelengths lambda = pbrt::SampledWavelengths::SampleUniform(0.5f);
    for (int i = 0; i < n; ++i)
        rayQ.PushCameraRay(MakeRay(rays + 8 * (size_t)i), lambda, i);
    try {
        if (useBinding) {
            pbrt::GPUBVHWavefrontAggregate agg(s->primList, maxPrimsInNode,
                                               (pbrt::BVHAggregate::SplitMethod)splitMethod, n);
            const pbrt::WavefrontAggregate &w = agg;  // through the reference's interface
            w.IntersectClosest(n, &rayQ, &escapedQ, nullptr, nullptr, nullptr, nullptr, &nextQ);
        } else {
            pbrt::Primitive aggregate(s->agg);
            const pbrt::RayQueue *rq = &rayQ;
            pbrt::EscapedRayQueue *eq = &escapedQ;
            pbrt::RayQueue *nq = &nextQ;
            pbrt::ParallelFor(0, rayQ.Size(), [=](int index) {
                const pbrt::RayWorkItem r = (*rq)[index];
                pstd::optional<pbrt::ShapeIntersection> si = aggregate.Intersect(r.ray, pbrt::Infinity);
                if (!si)
                    pbrt::EnqueueWorkAfterMiss(r, nullptr, eq);
                else
                    pbrt::EnqueueWorkAfterIntersection(r, r.ray.medium, si->tHit, si->intr, nullptr, nq, nullptr,
                                                       nullptr, nullptr);
            });
        }
    } catch (const nnbvh::Error &e) {
        g_bindingError = e.what();
        return e.code;
    }
    for (int i = 0; i < n; ++i)
        kind[i] = -1;
    for (int j = 0; j < escapedQ.Size(); ++j)
        kind[escapedQ.pixelIndex[j]] = 0;
    for (int j = 0; j < nextQ.Size(); ++j) {
        pbrt::RayWorkItem r = nextQ[j];
        kind[r.pixelIndex] = 1;
        float *o = nextRays8 + 8 * (size_t)r.pixelIndex;
        o[0] = r.ray.o.x, o[1] = r.ray.o.y, o[2] = r.ray.o.z, o[3] = 0;
        o[4] = r.ray.d.x, o[5] = r.ray.d.y, o[6] = r.ray.d.z, o[7] = r.ray.time;
    }
    return 0;
}

// One IntersectShadow over a real ShadowRayQueue: the radiance RecordShadowRayResult (wavefront/intersect.h:31-46)
// leaves in SOA<PixelSampleState>::L, first wavelength, per ray (Ld = i % 7 + 1 on the way in).
int nnbvh_binding_wavefront_shadow(void *scene, int maxPrimsInNode, int splitMethod, int useBinding,
                                   const float *rays, int n, float *L0) {
    RefScene *s = (RefScene *)scene;
    pbrt::Allocator alloc;
    pbrt::ShadowRayQueue shadowQ(n, alloc);
    pbrt::SOA<pbrt::PixelSampleState> pss(n, alloc);
    pbrt::SampledWavelengths lambda = pbrt::SampledWavelengths::SampleUniform(0.5f);
    for (int i = 0; i < n; ++i) {
        const float *r = rays + 8 * (size_t)i;
        pbrt::ShadowRayWorkItem w;
        w.ray = MakeRay(r);
        w.tMax = r[3];
        w.lambda = lambda;
        w.Ld = pbrt::SampledSpectrum(float(i % 7 + 1));
        w.r_u = w.r_l = pbrt::SampledSpectrum(1.f);
        w.pixelIndex = i;
        shadowQ.Push(w);
        pss.L[i] = pbrt::SampledSpectrum(0.f);
    }
    try {
        if (useBinding) {
            pbrt::GPUBVHWavefrontAggregate agg(s->primList, maxPrimsInNode,
                                               (pbrt::BVHAggregate::SplitMethod)splitMethod, n);
            const pbrt::WavefrontAggregate &w = agg;
            w.IntersectShadow(n, &shadowQ, &pss);
        } else {
            pbrt::Primitive aggregate(s->agg);
            pbrt::ShadowRayQueue *sq = &shadowQ;
            pbrt::SOA<pbrt::PixelSampleState> *ps = &pss;
            pbrt::ParallelFor(0, shadowQ.Size(), [=](int index) {  // aggregate.cpp:63-67
                const pbrt::ShadowRayWorkItem w = (*sq)[index];
                bool hit = aggregate.IntersectP(w.ray, w.tMax);
                pbrt::RecordShadowRayResult(w, ps, hit);
            });
        }
    } catch (const nnbvh::Error &e) {
        g_bindingError = e.what();
        return e.code;
    }
    for (int i = 0; i < n; ++i) {
        pbrt::SampledSpectrum L = pss.L[i];
        L0[i] = L[0];
    }
    return 0;
}

// One IntersectShadowTr over a real ShadowRayQueue.  Materials and media are null in these scenes, so every
// surface is a transparent interface: TraceTransmittance (wavefront/intersect.h:165-275) walks each ray from
// surface to surface with intr.SpawnRayTo(pLight) until nothing is left before the light, and adds Ld to
// PixelSampleState::L.  useBinding: pbrt::GPUBVHWavefrontAggregate (one GPU batch per round of segments);
// else CPUAggregate::IntersectShadowTr's loop (wavefront/aggregate.cpp:70-92) over the reference BVHAggregate.
// segments[i] = closest-hit queries the reference's walk of ray i made (CPU leg only; may be NULL).
int nnbvh_binding_wavefront_shadow_tr(void *scene, int maxPrimsInNode, int splitMethod, int useBinding,
                                      const float *rays, int n, float *L0, int *segments, int *gpuBatches) {
    RefScene *s = (RefScene *)scene;
    pbrt::Allocator alloc;
    pbrt::ShadowRayQueue shadowQ(n, alloc);
    pbrt::SOA<pbrt::PixelSampleState> pss(n, alloc);
    pbrt::SampledWavelengths lambda = pbrt::SampledWavelengths::SampleUniform(0.5f);
    for (int i = 0; i < n; ++i) {
        const float *r = rays + 8 * (size_t)i;
        pbrt::ShadowRayWorkItem w;
        w.ray = MakeRay(r);
        w.tMax = r[3];
        w.lambda = lambda;
        w.Ld = pbrt::SampledSpectrum(float(i % 7 + 1));
        w.r_u = w.r_l = pbrt::SampledSpectrum(1.f);
        w.pixelIndex = i;
        shadowQ.Push(w);
        pss.L[i] = pbrt::SampledSpectrum(0.f);
    }
    if (gpuBatches)
        *gpuBatches = 0;
    try {
        if (useBinding) {
            pbrt::GPUBVHWavefrontAggregate agg(s->primList, maxPrimsInNode,
                                               (pbrt::BVHAggregate::SplitMethod)splitMethod, n);
            const pbrt::WavefrontAggregate &w = agg;
            w.IntersectShadowTr(n, &shadowQ, &pss);
            if (gpuBatches)
                *gpuBatches = agg.LastShadowTrBatches();
        } else {
            pbrt::Primitive aggregate(s->agg);
            pbrt::ShadowRayQueue *sq = &shadowQ;
            pbrt::SOA<pbrt::PixelSampleState> *ps = &pss;
            pbrt::ParallelFor(0, shadowQ.Size(), [=](int index) {  // aggregate.cpp:72-91
                const pbrt::ShadowRayWorkItem w = (*sq)[index];
                pstd::optional<pbrt::ShapeIntersection> si;
                int asked = 0;
                pbrt::TraceTransmittance(
                    w, ps,
                    [&](pbrt::Ray ray, pbrt::Float tMax) -> pbrt::TransmittanceTraceResult {
                        ++asked;
                        si = aggregate.Intersect(ray, tMax);
                        if (!si)
                            return pbrt::TransmittanceTraceResult{false, pbrt::Point3f(), pbrt::Material()};
                        return pbrt::TransmittanceTraceResult{true, si->intr.p(), si->intr.material};
                    },
                    [&](pbrt::Point3f p) -> pbrt::Ray { return si->intr.SpawnRayTo(p); });
                if (segments)
                    segments[index] = asked;
            });
        }
    } catch (const nnbvh::Error &e) {
        g_bindingError = e.what();
        return e.code;
    }
    for (int i = 0; i < n; ++i) {
        pbrt::SampledSpectrum L = pss.L[i];
        L0[i] = L[0];
    }
    return 0;
}

// One IntersectOneRandom over a real SubsurfaceScatterQueue: probes p0 -> p1 (seg6[i] = p0.xyz, p1.xyz), the
// probe's material null like every surface's, so each surface between the two points is a candidate.
// Out per probe: reservoirPDF and, where it is non-zero, the chosen SubsurfaceInteraction's p() and n.
int nnbvh_binding_wavefront_one_random(void *scene, int maxPrimsInNode, int splitMethod, int useBinding,
                                       const float *seg6, int n, float *pdf, float *pn6, int *gpuBatches) {
    RefScene *s = (RefScene *)scene;
    pbrt::Allocator alloc;
    pbrt::SubsurfaceScatterQueue q(n, alloc);
    pbrt::SampledWavelengths lambda = pbrt::SampledWavelengths::SampleUniform(0.5f);
    for (int i = 0; i < n; ++i) {
        const float *p = seg6 + 6 * (size_t)i;
        q.Push(pbrt::Point3f(p[0], p[1], p[2]), pbrt::Point3f(p[3], p[4], p[5]), 0, pbrt::Material(),
               pbrt::TabulatedBSSRDF(), lambda, pbrt::SampledSpectrum(1.f), pbrt::SampledSpectrum(1.f),
               pbrt::MediumInterface(), 1.f, i);
    }
    if (gpuBatches)
        *gpuBatches = 0;
    try {
        if (useBinding) {
            pbrt::GPUBVHWavefrontAggregate agg(s->primList, maxPrimsInNode,
                                               (pbrt::BVHAggregate::SplitMethod)splitMethod, n);
            const pbrt::WavefrontAggregate &w = agg;
            w.IntersectOneRandom(n, &q);
            if (gpuBatches)
                *gpuBatches = agg.LastOneRandomBatches();
        } else {
            pbrt::Primitive aggregate(s->agg);
            pbrt::SubsurfaceScatterQueue *sq = &q;
            pbrt::ParallelFor(0, q.Size(), [=](int index) {  // aggregate.cpp:96-123
                const pbrt::SubsurfaceScatterWorkItem &w = (*sq)[index];
                uint64_t seed = pbrt::Hash(w.p0, w.p1);
                pbrt::WeightedReservoirSampler<pbrt::SubsurfaceInteraction> wrs(seed);
                pbrt::Interaction base(w.p0, 0.f, pbrt::Medium());
                while (true) {
                    pbrt::Ray r = base.SpawnRayTo(w.p1);
                    if (r.d == pbrt::Vector3f(0, 0, 0))
                        break;
                    pstd::optional<pbrt::ShapeIntersection> si = aggregate.Intersect(r, 1);
                    if (!si)
                        break;
                    base = si->intr;
                    if (si->intr.material == w.material)
                        wrs.Add(pbrt::SubsurfaceInteraction(si->intr), 1.f);
                }
                if (wrs.HasSample()) {
                    sq->reservoirPDF[index] = wrs.SampleProbability();
                    sq->ssi[index] = wrs.GetSample();
                } else
                    sq->reservoirPDF[index] = 0;
            });
        }
    } catch (const nnbvh::Error &e) {
        g_bindingError = e.what();
        return e.code;
    }
    for (int i = 0; i < n; ++i) {
        pdf[i] = q.reservoirPDF[i];
        float *o = pn6 + 6 * (size_t)i;
        for (int k = 0; k < 6; ++k)
            o[k] = 0;
        if (pdf[i] != 0) {
            pbrt::SubsurfaceInteraction ssi = q.ssi[i];
            pbrt::Point3f p = ssi.p();
            o[0] = p.x, o[1] = p.y, o[2] = p.z, o[3] = ssi.n.x, o[4] = ssi.n.y, o[5] = ssi.n.z;
        }
    }
    return 0;
}

void nnbvh_binding_bounds(void *binding, float *out6) {
    pbrt::Bounds3f b = ((const pbrt::GPUBVHAggregate *)binding)->Bounds();
    out6[0] = b.pMin.x, out6[1] = b.pMin.y, out6[2] = b.pMin.z;
    out6[3] = b.pMax.x, out6[4] = b.pMax.y, out6[5] = b.pMax.z;
}
}  // extern "C"
#endif  // NNBVH_WITH_BINDING
