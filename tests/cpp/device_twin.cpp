// tests/cpp/device_twin.cpp -- the DEVICE math headers compiled for the host (tests/test_device_twin.py).
//
// nn_bvh_b200/csrc/{exact_math,sphere,atan2_host,alpha_math}.cuh are included unchanged, with
// tests/cpp/cuda_stub/cuda_runtime.h supplying IEEE-exact host equivalents of the CUDA intrinsics.  The CPU
// suite then runs the device source -- the same statements the GPU executes -- on the reference's golden
// vectors and against the checker, so a slip in the device text is caught without a GPU.
#include <algorithm>

#include <vector>

#include "../../nn_bvh_b200/csrc/alpha.cuh"
#include "../../nn_bvh_b200/csrc/alpha_math.cuh"
#include "../../nn_bvh_b200/csrc/atan2_host.cuh"
#include "../../nn_bvh_b200/csrc/exact_math.cuh"
#include "../../nn_bvh_b200/csrc/sphere.cuh"

using namespace nnbvh;

namespace {
RayConst RayOf(const float *r) { return MakeRayConst(r[0], r[1], r[2], r[4], r[5], r[6]); }
// what MakeSphereDev (engine.cu) stores: Sphere::Sphere's clamped values (shapes.h:117-128)
SphereDev SphereOf(const float *s28) {
    SphereDev d;
    for (int k = 0; k < 12; ++k) d.m[k] = s28[12 + k];
    auto clamp = [](float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); };
    const float r = s28[24];
    d.radius = r;
    d.zMin = clamp(std::min(s28[25], s28[26]), -r, r);
    d.zMax = clamp(std::max(s28[25], s28[26]), -r, r);
    d.phiMax = (3.14159265358979323846f / 180) * clamp(s28[27], 0.f, 360.f);
    return d;
}

// What PackTrisKernel / BuildWideLayout (engine.cu) put on the device for the one-thread-per-ray kernels: three
// float4 per ordered primitive, the flags in the first one's .w; `special` >= 0 marks a sphere / instance slot.
std::vector<float4> Pack(const nnbvh_node *nodes, uint32_t nNodes, const float *tris9, uint32_t nPrims,
                         const int32_t *special, uint32_t specialFlag) {
    std::vector<uint8_t> last(nPrims, 0);
    for (uint32_t i = 0; i < nNodes; ++i)
        if (nodes[i].nPrimitives > 0) last[(uint32_t)nodes[i].offset + nodes[i].nPrimitives - 1] = 1;
    std::vector<float4> out(3 * (size_t)nPrims);
    for (uint32_t i = 0; i < nPrims; ++i) {
        if (special && special[i] >= 0) {
            out[3 * (size_t)i] = make_float4(__uint_as_float((uint32_t)special[i]), 0.f, 0.f,
                                             __uint_as_float(specialFlag | (last[i] ? kTriLast : 0u)));
            out[3 * (size_t)i + 1] = out[3 * (size_t)i + 2] = make_float4(0.f, 0.f, 0.f, 0.f);
            continue;
        }
        const float *v = tris9 + 9 * (size_t)i;
        uint32_t flags = 0;
        if (TriangleDegenerate(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7], v[8])) flags |= kTriDegenerate;
        if (last[i]) flags |= kTriLast;
        out[3 * (size_t)i] = make_float4(v[0], v[1], v[2], __uint_as_float(flags));
        out[3 * (size_t)i + 1] = make_float4(v[3], v[4], v[5], 0.f);
        out[3 * (size_t)i + 2] = make_float4(v[6], v[7], v[8], 0.f);
    }
    return out;
}

// Runs kernel(i) for every "thread" i < n, 128 to a block, as the engine launches the one-thread-per-ray kernels.
template <typename F>
void ForThreads(int64_t n, F kernel) {
    blockDim.x = 128;
    for (int64_t i = 0; i < n; ++i) {
        blockIdx.x = (unsigned)(i / 128), threadIdx.x = (unsigned)(i % 128);
        kernel();
    }
}
}  // namespace

extern "C" {

// BasicKernel / AlphaKernel over a whole tree, one simulated thread per ray.  spheres28: NULL or the sphere
// table with sphereOf[i] (-1 = triangle); alpha: NULL or one value per ordered primitive.  hits5: n x 5 words
// or, with occluded != NULL, the any-hit answers.
void dt_traverse(const void *nodes, uint32_t nNodes, const float *orderedTris9, uint32_t nPrims,
                 const int32_t *sphereOf, const float *spheres28, uint32_t nSpheres, const float *alpha,
                 const float *rays8, int64_t n, uint32_t *hits5, uint8_t *occluded) {
    const nnbvh_node *nd = (const nnbvh_node *)nodes;
    std::vector<float4> packed = Pack(nd, nNodes, orderedTris9, nPrims, spheres28 ? sphereOf : nullptr, kPrimSphere);
    std::vector<SphereDev> sph(nSpheres);
    for (uint32_t k = 0; k < nSpheres; ++k) sph[k] = SphereOf(spheres28 + 28 * (size_t)k);
    SceneDev sc{};
    sc.nodes = nd;
    sc.tris = packed.data();
    sc.nNodes = nNodes;
    sc.nPrims = nPrims;
    sc.spheres = spheres28 ? sph.data() : nullptr;
    RayInput in{};
    in.aos = reinterpret_cast<const float4 *>(rays8);
    nnbvh_hit *hits = reinterpret_cast<nnbvh_hit *>(hits5);
    const uint64_t un = (uint64_t)n;
    if (alpha) {
        if (occluded) ForThreads(n, [&] { AlphaKernel<true>(sc, alpha, in, un, hits, occluded); });
        else ForThreads(n, [&] { AlphaKernel<false>(sc, alpha, in, un, hits, occluded); });
    } else if (spheres28) {
        if (occluded) ForThreads(n, [&] { BasicKernel<true, false, true>(sc, in, un, hits, occluded, nullptr, nullptr); });
        else ForThreads(n, [&] { BasicKernel<false, false, true>(sc, in, un, hits, occluded, nullptr, nullptr); });
    } else {
        if (occluded) ForThreads(n, [&] { BasicKernel<true, false>(sc, in, un, hits, occluded, nullptr, nullptr); });
        else ForThreads(n, [&] { BasicKernel<false, false>(sc, in, un, hits, occluded, nullptr, nullptr); });
    }
}

// out4 = {t, b0, b1, b2}
int dt_triangle(const float *ray8, const float *p, float *out4) {
    if (TriangleDegenerate(p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8])) return 0;
    return TriangleTest(RayOf(ray8), ray8[3], p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8], &out4[0], &out4[1],
                        &out4[2], &out4[3])
               ? 1
               : 0;
}
int dt_box(const float *ray8, const float *box6) { return SlabTest(RayOf(ray8), box6, ray8[3]) ? 1 : 0; }
// out4 = {tHit, pObj}
int dt_sphere(const float *ray8, const float *sphere28, float *out4) {
    SphereDev sd = SphereOf(sphere28);
    SphereHit h;
    if (!SphereTest(&sd, ray8[0], ray8[1], ray8[2], ray8[4], ray8[5], ray8[6], ray8[3], &h)) return 0;
    out4[0] = h.t, out4[1] = h.px, out4[2] = h.py, out4[3] = h.pz;
    return 1;
}
void dt_atan2(const float *y, const float *x, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = Atan2fHost(y[i], x[i]);
}
float dt_hash_float(const float *o, const float *d) { return HashFloatOD(o[0], o[1], o[2], d[0], d[1], d[2]); }
// out4 = {origin of intr.SpawnRay(ray.d), tHit}
int dt_spawn_origin(const float *ray8, const float *p, float *out4) {
    float t, b0, b1, b2;
    if (!dt_triangle(ray8, p, out4)) return 0;
    t = out4[0], b0 = out4[1], b1 = out4[2], b2 = out4[3];
    SpawnOrigin(p, p + 3, p + 6, b0, b1, b2, ray8[4], ray8[5], ray8[6], out4);
    out4[3] = t;
    return 1;
}
// GeometricPrimitive::Intersect over one triangle with constant alpha; out4 = {t, b0, b1, b2}
int dt_alpha_triangle(const float *ray8, const float *p, float alpha, float *out4) {
    if (TriangleDegenerate(p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8])) return 0;
    if (!(alpha < 1)) return dt_triangle(ray8, p, out4);
    return AlphaTriangleTest(RayOf(ray8), ray8[3], p, p + 3, p + 6, alpha, &out4[0], &out4[1], &out4[2], &out4[3]) ? 1 : 0;
}
// Transform::ApplyInverse(Ray, &tMax) as InstanceRay does it; local8 = {o, tMax, d, time}
void dt_instance_ray(const float *primitiveFromRender12, const float *ray8, float *local8) {
    InstanceDev in{};
    for (int k = 0; k < 12; ++k) in.mInv[k] = primitiveFromRender12[k];
    float o[3], d[3], t;
    InstanceRay(&in, ray8[0], ray8[1], ray8[2], ray8[4], ray8[5], ray8[6], ray8[3], o, d, &t);
    local8[0] = o[0], local8[1] = o[1], local8[2] = o[2], local8[3] = t;
    local8[4] = d[0], local8[5] = d[1], local8[6] = d[2], local8[7] = ray8[7];
}

// BasicKernel<ANY, COUNT = true>: the reference's per-ray statistics (bvhNodesVisited, nTriTests) on a triangle scene
void dt_traverse_counters(const void *nodes, uint32_t nNodes, const float *orderedTris9, uint32_t nPrims,
                          const float *rays8, int64_t n, int anyHit, uint32_t *nodesVisited, uint32_t *triTests) {
    const nnbvh_node *nd = (const nnbvh_node *)nodes;
    std::vector<float4> packed = Pack(nd, nNodes, orderedTris9, nPrims, nullptr, 0);
    SceneDev sc{};
    sc.nodes = nd;
    sc.tris = packed.data();
    sc.nNodes = nNodes;
    sc.nPrims = nPrims;
    RayInput in{};
    in.aos = reinterpret_cast<const float4 *>(rays8);
    std::vector<nnbvh_hit> hits((size_t)n);
    std::vector<uint8_t> occ((size_t)n);
    const uint64_t un = (uint64_t)n;
    if (anyHit)
        ForThreads(n, [&] { BasicKernel<true, true>(sc, in, un, hits.data(), occ.data(), nodesVisited, triTests); });
    else
        ForThreads(n, [&] { BasicKernel<false, true>(sc, in, un, hits.data(), occ.data(), nodesVisited, triTests); });
}

// BasicKernel<ANY, false, false, INST = true> over an instanced scene (the layout nnbvh_upload_scene_instanced builds):
// inst24[k] = {renderFromPrimitive, primitiveFromRender}, instObject[k]; object j's tree is objNodes[objNodeBase[j] ..
// objNodeBase[j+1]) and its ordered triangles objTris9[objPrimBase[j] .. objPrimBase[j+1]).
void dt_traverse_instanced(const void *nodes, uint32_t nNodes, const float *orderedTris9, uint32_t nPrims,
                           const int32_t *instanceOf, const float *inst24, const int32_t *instObject, uint32_t nInst,
                           const void *objNodes, const int32_t *objNodeBase, const float *objTris9,
                           const int32_t *objPrimBase, uint32_t nObjects, const float *rays8, int64_t n,
                           uint32_t *hits5, uint32_t *inner, uint8_t *occluded) {
    const nnbvh_node *nd = (const nnbvh_node *)nodes, *on = (const nnbvh_node *)objNodes;
    std::vector<float4> packed = Pack(nd, nNodes, orderedTris9, nPrims, instanceOf, kPrimInstance);
    std::vector<float4> objPacked;
    for (uint32_t j = 0; j < nObjects; ++j) {
        const uint32_t nb = (uint32_t)objNodeBase[j], ne = (uint32_t)objNodeBase[j + 1];
        const uint32_t pb = (uint32_t)objPrimBase[j], pe = (uint32_t)objPrimBase[j + 1];
        std::vector<float4> one = Pack(on + nb, ne - nb, objTris9 + 9 * (size_t)pb, pe - pb, nullptr, 0);
        objPacked.insert(objPacked.end(), one.begin(), one.end());
    }
    std::vector<InstanceDev> inst(nInst);
    for (uint32_t k = 0; k < nInst; ++k) {
        for (int c = 0; c < 12; ++c) inst[k].mInv[c] = inst24[24 * (size_t)k + 12 + c];
        const int j = instObject[k];
        inst[k].nodeBase = (uint32_t)objNodeBase[j];
        inst[k].primBase = (uint32_t)objPrimBase[j];
        inst[k].nNodes = (uint32_t)(objNodeBase[j + 1] - objNodeBase[j]);
        inst[k].pad = 0;
    }
    SceneDev sc{};
    sc.nodes = nd;
    sc.tris = packed.data();
    sc.nNodes = nNodes;
    sc.nPrims = nPrims;
    sc.instances = inst.data();
    sc.objNodes = on;
    sc.objTris = objPacked.data();
    RayInput in{};
    in.aos = reinterpret_cast<const float4 *>(rays8);
    nnbvh_hit *hits = reinterpret_cast<nnbvh_hit *>(hits5);
    const uint64_t un = (uint64_t)n;
    if (occluded)
        ForThreads(n, [&] { BasicKernel<true, false, false, true>(sc, in, un, hits, occluded, nullptr, nullptr, nullptr); });
    else
        ForThreads(n, [&] { BasicKernel<false, false, false, true>(sc, in, un, hits, occluded, nullptr, nullptr, inner); });
}

}  // extern "C"
