// tests/cpp/wavefront_shim.cpp -- a C-callable consumer of include/nnbvh_wavefront.hpp for the tests.
//
// Plays the integrator's part around nnbvh::wavefront::GPUBVHWavefrontAggregate: owns SoA ray
// queues (the layout of the reference's SOA<RayWorkItem>, wavefront/workitems.soa) and a sink that
// records what EnqueueWorkAfterMiss / EnqueueWorkAfterIntersection / RecordShadowRayResult
// (wavefront/intersect.h:16-163) would be handed.  wf_interaction() exposes the host-side
// Triangle::InteractionFromIntersection restatement on its own (no GPU needed).
#include <cstring>
#include <mutex>

#include "nnbvh_wavefront.hpp"

using namespace nnbvh::wavefront;

namespace {
thread_local std::string g_what;

struct Mesh {
    std::vector<float> p, n, s, uv;
    std::vector<int> idx, face;
    TriangleMeshView View(bool reverseOrientation) const {
        TriangleMeshView v;
        v.vertexIndices = idx.data();
        v.p = p.data();
        v.n = n.empty() ? nullptr : n.data();
        v.s = s.empty() ? nullptr : s.data();
        v.uv = uv.empty() ? nullptr : uv.data();
        v.faceIndices = face.data();
        v.reverseOrientation = reverseOrientation;
        return v;
    }
};

Mesh MakeMesh(const float *verts, int nVerts, const int *indices, int nTris, const float *normals,
              const float *tangents, const float *uvs, bool reverseOrientation = false) {
    Mesh m;
    m.p.assign(verts, verts + 3 * (size_t)nVerts);
    m.idx.assign(indices, indices + 3 * (size_t)nTris);
    if (normals) m.n.assign(normals, normals + 3 * (size_t)nVerts);
    // TriangleMesh stores the normals negated when reverseOrientation is set (util/mesh.cpp:51-55)
    if (reverseOrientation)
        for (float &x : m.n) x = -x;
    if (tangents) m.s.assign(tangents, tangents + 3 * (size_t)nVerts);
    if (uvs) m.uv.assign(uvs, uvs + 2 * (size_t)nVerts);
    m.face.resize(nTris);
    for (int i = 0; i < nTris; ++i) m.face[i] = i;
    return m;
}

// 39 words, the order oracle/ref_driver.cpp nnbvh_ref_triangle_interaction writes
void Pack(const SurfaceHit &h, float *o) {
    auto put = [&o](const float *v, int k) { std::memcpy(o, v, 4 * k); o += k; };
    put(h.piLow, 3), put(h.piHigh, 3), put(h.n, 3), put(h.uv, 2), put(h.dpdu, 3), put(h.dpdv, 3);
    put(h.dndu, 3), put(h.dndv, 3), put(h.shading.n, 3), put(h.shading.dpdu, 3);
    put(h.shading.dpdv, 3), put(h.shading.dndu, 3), put(h.shading.dndv, 3);
    std::memcpy(o, &h.faceIndex, 4);
}

struct Scene {
    Mesh mesh;
    std::vector<Mesh> objectMeshes;
    std::unique_ptr<GPUBVHWavefrontAggregate> agg;
};

// what the reference's queues would receive, by ray index
struct Sink {
    int8_t *kind;     // 0 EnqueueWorkAfterMiss, 1 EnqueueWorkAfterIntersection
    uint32_t *prim;   // index in the constructor's primitive list
    float *tHit;
    float *surf39;
    float *woTime;    // 4 per ray
    uint8_t *shadow;  // RecordShadowRayResult(foundIntersection)
    std::atomic<int> calls{0};
    void EnqueueWorkAfterMiss(int i) { kind[i] = 0, ++calls; }
    void EnqueueWorkAfterIntersection(int i, uint32_t p, float t, const SurfaceHit &h) {
        kind[i] = 1, prim[i] = p, tHit[i] = t;
        Pack(h, surf39 + 39 * (size_t)i);
        std::memcpy(woTime + 4 * (size_t)i, h.wo, 12);
        woTime[4 * (size_t)i + 3] = h.time;
        ++calls;
    }
    void RecordShadowRayResult(int i, bool found) { shadow[i] = found, ++calls; }
};
// A triangle soup (9 floats per triangle) as an indexed mesh without normals / uv
Mesh SoupMesh(const float *tris9, int nTris) {
    std::vector<int> idx(3 * (size_t)nTris);
    for (size_t i = 0; i < idx.size(); ++i) idx[i] = (int)i;
    return MakeMesh(tris9, 3 * nTris, idx.data(), nTris, nullptr, nullptr, nullptr);
}

nnbvh_instance MakeInstance(const float *m24, int object) {
    nnbvh_instance in{};
    std::memcpy(in.renderFromPrimitive, m24, 48);
    std::memcpy(in.primitiveFromRender, m24 + 12, 48);
    in.object = (uint32_t)object;
    return in;
}
}  // namespace

extern "C" {

const char *wf_last_error(void) { return g_what.c_str(); }

int wf_interaction(const float *verts, int nVerts, const int *indices, int nTris,
                   const float *normals, const float *tangents, const float *uvs,
                   int reverseOrientation, const int *tri, const float *b, const float *woTime,
                   int64_t n, float *out39) {
    Mesh m = MakeMesh(verts, nVerts, indices, nTris, normals, tangents, uvs, reverseOrientation != 0);
    TriangleMeshView v = m.View(reverseOrientation != 0);
    for (int64_t i = 0; i < n; ++i) {
        SurfaceHit h;
        if (!InteractionFromIntersection(v, tri[i], b[3 * i], b[3 * i + 1], b[3 * i + 2],
                                         woTime[4 * i + 3], woTime + 4 * i, &h))
            return -1;
        Pack(h, out39 + 39 * i);
    }
    return 0;
}

// Sphere::InteractionFromIntersection on the host: query i is pObj[3i..] (what the engine returns
// for a sphere leaf) with woTime[4i..]; out42 = the 39 words + the interaction's wo.
int wf_sphere_interaction(const float *sphere28, int reverseOrientation, const float *pObj,
                          const float *woTime, int64_t n, float *out42) {
    nnbvh_sphere s;
    std::memcpy(&s, sphere28, sizeof s);
    SphereShape shape = MakeSphereShape(s, reverseOrientation != 0);
    for (int64_t i = 0; i < n; ++i) {
        SurfaceHit h;
        SphereInteractionFromIntersection(shape, pObj + 3 * i, woTime + 4 * i, woTime[4 * i + 3], &h);
        Pack(h, out42 + 42 * i);
        std::memcpy(out42 + 42 * i + 39, h.wo, 12);
    }
    return 0;
}

// nnbvh::SpherePhiMaxDegrees on n values
void wf_phimax_degrees(const float *radians, int64_t n, float *degrees) {
    for (int64_t i = 0; i < n; ++i) degrees[i] = nnbvh::SpherePhiMaxDegrees(radians[i]);
}

// The wo a triangle hit's interaction stores for the wo it is given
void wf_interaction_wo(const float *verts9, const float *wo3, int64_t n, float *out3) {
    int idx[3] = {0, 1, 2};
    TriangleMeshView v;
    v.vertexIndices = idx, v.p = verts9;
    for (int64_t i = 0; i < n; ++i) {
        SurfaceHit h;
        InteractionFromIntersection(v, 0, 0.25f, 0.25f, 0.5f, 0.f, wo3 + 3 * i, &h);
        std::memcpy(out3 + 3 * i, h.wo, 12);
    }
}

void *wf_create(const float *verts, int nVerts, const int *indices, int nTris, const float *normals,
                const float *uvs, int maxPrimsInNode, int splitMethod, int hostThreads) {
    try {
        auto *s = new Scene;
        s->mesh = MakeMesh(verts, nVerts, indices, nTris, normals, nullptr, uvs);
        std::vector<GPUBVHWavefrontAggregate::PrimitiveRef> prims(nTris);
        for (int i = 0; i < nTris; ++i) prims[i] = {0, i};
        s->agg.reset(new GPUBVHWavefrontAggregate(
            {s->mesh.View(false)}, prims, maxPrimsInNode,
            (nnbvh::GPUBVHAggregate::SplitMethod)splitMethod, -1, hostThreads));
        return s;
    } catch (const nnbvh::Error &e) {
        g_what = e.what();
        return nullptr;
    }
}

// The same scene with nSpheres `Shape "sphere"` primitives (28 floats each, nnbvh_sphere) after the
// triangles; reverse: one flag per sphere or NULL.
void *wf_create_mixed(const float *verts, int nVerts, const int *indices, int nTris, const float *normals,
                      const float *uvs, const float *spheres28, int nSpheres, const uint8_t *reverse,
                      int maxPrimsInNode, int splitMethod, int hostThreads) {
    try {
        auto *s = new Scene;
        s->mesh = MakeMesh(verts, nVerts, indices, nTris, normals, nullptr, uvs);
        std::vector<GPUBVHWavefrontAggregate::PrimitiveRef> prims(nTris);
        for (int i = 0; i < nTris; ++i) prims[i] = {0, i};
        std::vector<nnbvh_sphere> sph(nSpheres);
        std::memcpy(sph.data(), spheres28, sizeof(nnbvh_sphere) * (size_t)nSpheres);
        std::vector<uint8_t> rev;
        if (reverse) rev.assign(reverse, reverse + nSpheres);
        s->agg.reset(new GPUBVHWavefrontAggregate(
            {s->mesh.View(false)}, prims, sph, rev, maxPrimsInNode,
            (nnbvh::GPUBVHAggregate::SplitMethod)splitMethod, -1, hostThreads));
        return s;
    } catch (const nnbvh::Error &e) {
        g_what = e.what();
        return nullptr;
    }
}

// TransformedPrimitive::Intersect's interaction on the host: query i hits triangle tri[i] of the object's
// soup with barycentrics b[3i..], for the render-space ray direction / time dTime[4i..].
int wf_instance_interaction(const float *inst24, const float *soup9, int nTris, const int *tri,
                            const float *b, const float *dTime, int64_t n, float *out42) {
    Mesh m = SoupMesh(soup9, nTris);
    TriangleMeshView v = m.View(false);
    nnbvh_instance in = MakeInstance(inst24, 0);
    for (int64_t i = 0; i < n; ++i) {
        SurfaceHit h;
        if (!InstanceInteractionFromIntersection(in, v, tri[i], b[3 * i], b[3 * i + 1], b[3 * i + 2],
                                                 dTime[4 * i + 3], dTime + 4 * i, &h))
            return -1;
        Pack(h, out42 + 42 * i);
        std::memcpy(out42 + 42 * i + 39, h.wo, 12);
    }
    return 0;
}

// Top-level soup + objects (soups in object space, concatenated; objStart[j] .. objStart[j+1]) +
// instances (24 floats: renderFromPrimitive, primitiveFromRender; instObject[k]).
void *wf_create_instanced(const float *top9, int nTop, const float *obj9, const int *objStart, int nObjects,
                          const float *inst24, const int *instObject, int nInst, int maxPrimsInNode,
                          int splitMethod, int hostThreads) {
    try {
        auto *s = new Scene;
        using Ref = GPUBVHWavefrontAggregate::PrimitiveRef;
        s->mesh = SoupMesh(top9, nTop);
        std::vector<TriangleMeshView> views{s->mesh.View(false)};
        std::vector<Ref> prims(nTop);
        for (int i = 0; i < nTop; ++i) prims[i] = {0, i};
        std::vector<std::vector<Ref>> objects(nObjects);
        s->objectMeshes.reserve(nObjects);
        for (int j = 0; j < nObjects; ++j) {
            int n = objStart[j + 1] - objStart[j];
            s->objectMeshes.push_back(SoupMesh(obj9 + 9 * (size_t)objStart[j], n));
            views.push_back(s->objectMeshes.back().View(false));
            for (int i = 0; i < n; ++i) objects[j].push_back(Ref{j + 1, i});
        }
        std::vector<nnbvh_instance> inst(nInst);
        for (int k = 0; k < nInst; ++k) inst[k] = MakeInstance(inst24 + 24 * (size_t)k, instObject[k]);
        s->agg.reset(new GPUBVHWavefrontAggregate(views, prims, objects, inst, maxPrimsInNode,
                                                  (nnbvh::GPUBVHAggregate::SplitMethod)splitMethod, -1, hostThreads));
        return s;
    } catch (const nnbvh::Error &e) {
        g_what = e.what();
        return nullptr;
    }
}

void wf_destroy(void *scene) { delete (Scene *)scene; }

// planes: ox oy oz dx dy dz time (7 pointers; host or device memory).  Returns the number of sink
// calls (== n) or a negative nnbvh_status.
int wf_intersect_closest(void *scene, int maxRays, const float *const *planes, int n, int8_t *kind,
                         uint32_t *prim, float *tHit, float *surf39, float *woTime) {
    try {
        RayQueueView q{planes[0], planes[1], planes[2], planes[3], planes[4], planes[5], planes[6], n};
        Sink sink;
        sink.kind = kind, sink.prim = prim, sink.tHit = tHit, sink.surf39 = surf39, sink.woTime = woTime;
        ((Scene *)scene)->agg->IntersectClosest(maxRays, q, sink);
        return sink.calls;
    } catch (const nnbvh::Error &e) {
        g_what = e.what();
        return e.code;
    }
}

// planes: ox oy oz dx dy dz time tMax (8 pointers).
int wf_intersect_shadow(void *scene, int maxRays, const float *const *planes, int n, uint8_t *found,
                        int transmittance) {
    try {
        ShadowRayQueueView q{planes[0], planes[1], planes[2], planes[3], planes[4],
                             planes[5], planes[6], planes[7], n};
        Sink sink;
        sink.shadow = found;
        if (transmittance)
            ((Scene *)scene)->agg->IntersectShadowTr(maxRays, q, sink);
        else
            ((Scene *)scene)->agg->IntersectShadow(maxRays, q, sink);
        return sink.calls;
    } catch (const nnbvh::Error &e) {
        g_what = e.what();
        return e.code;
    }
}

}  // extern "C"
