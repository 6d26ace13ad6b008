// tests/cpp/host_mirror_test.cpp -- the C++ host mirror (include/nnbvh.hpp) against the oracle.
//
// Reads like a reference-side test of BVHAggregate: build over a triangle soup,
// Bounds(), single-ray Intersect/IntersectP, then a batch, each checked bit for
// bit against the CPU restatement (oracle/nnbvh_oracle.h -- tests may link it).
// Usage: host_mirror_test [--compile-only-check]
//   exit 0 = pass; prints the first mismatch otherwise.  Needs a CUDA device
//   (the pytest wrapper runs it under -m gpu; the CPU suite only compiles it and
//   checks that it fails loudly with NNBVH_ERR_CUDA when no device is present).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <random>
#include <vector>

#include "../../include/nnbvh.hpp"
#include "../../oracle/nnbvh_oracle.h"

using Split = nnbvh::GPUBVHAggregate::SplitMethod;

static int RunCase(Split split, int oracleSplit, const char *name) {
    const uint32_t T = 20000, N = 100000;
    std::mt19937 rng(1234);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    std::vector<float> tris(9 * (size_t)T);
    for (uint32_t i = 0; i < T; ++i) {
        float c[3] = {U(rng), U(rng), U(rng)};
        for (int v = 0; v < 3; ++v)
            for (int k = 0; k < 3; ++k) tris[9 * (size_t)i + 3 * v + k] = c[k] + 0.02f * (2 * U(rng) - 1);
    }
    try {
        nnbvh::GPUBVHAggregate agg(tris, 4, split);  // HLBVH: built on the device

        // the tree must be the reference's, byte for byte
        std::vector<nnbvh_oracle_node> onodes(2 * (size_t)T);
        std::vector<int32_t> oorder(T);
        int64_t m = nnbvh_oracle_build(tris.data(), (int32_t)T, 4, oracleSplit, onodes.data(), oorder.data());
        if (m != (int64_t)agg.Nodes().size() ||
            std::memcmp(onodes.data(), agg.Nodes().data(), sizeof(nnbvh_node) * (size_t)m) != 0) {
            std::printf("FAIL: tree differs from the oracle's (%lld vs %zu nodes)\n", (long long)m, agg.Nodes().size());
            return 1;
        }
        for (uint32_t i = 0; i < T; ++i)
            if ((uint32_t)oorder[i] != agg.PrimitiveIndex(i)) { std::printf("FAIL: leaf order differs at %u\n", i); return 1; }
        nnbvh::Bounds3f b = agg.Bounds();
        if (std::memcmp(&b, onodes[0].bounds, 24) != 0) { std::printf("FAIL: Bounds()\n"); return 1; }

        std::vector<nnbvh_ray> rays(N);
        std::normal_distribution<float> G(0.f, 1.f);
        for (uint32_t i = 0; i < N; ++i) {
            nnbvh_ray &r = rays[i];
            r.ox = -0.25f + 1.5f * U(rng); r.oy = -0.25f + 1.5f * U(rng); r.oz = -0.25f + 1.5f * U(rng);
            r.dx = G(rng); r.dy = G(rng); r.dz = G(rng);
            r.tMax = (i & 1) ? U(rng) : std::numeric_limits<float>::infinity();
            r.time = 0.f;
        }
        std::vector<nnbvh_hit> hits(N);
        std::vector<uint8_t> occ(N);
        agg.IntersectBatch(rays.data(), N, hits.data());
        agg.IntersectPBatch(rays.data(), N, occ.data());
        std::vector<uint32_t> want(5 * (size_t)N);
        std::vector<uint8_t> wocc(N);
        nnbvh_oracle_intersect(onodes.data(), agg.OrderedTriangles().data(), &rays[0].ox, N, want.data(), nullptr, nullptr, 0);
        nnbvh_oracle_intersectP(onodes.data(), agg.OrderedTriangles().data(), &rays[0].ox, N, wocc.data(), nullptr, nullptr, 0);
        static_assert(sizeof(nnbvh_hit) == 20, "hit record is 20 bytes");
        if (std::memcmp(hits.data(), want.data(), 20 * (size_t)N) != 0) { std::printf("FAIL: closest-hit records differ\n"); return 1; }
        if (std::memcmp(occ.data(), wocc.data(), N) != 0) { std::printf("FAIL: any-hit results differ\n"); return 1; }

        // a clone shares the uploaded scene and answers the same
        {
            auto twin = agg.Clone();
            std::vector<uint8_t> occ2(N);
            twin->IntersectPBatch(rays.data(), N, occ2.data());
            if (std::memcmp(occ2.data(), wocc.data(), N) != 0) { std::printf("FAIL: clone's any-hit results differ\n"); return 1; }
        }

        // single-ray surface
        uint32_t nHit = 0;
        for (uint32_t i = 0; i < 64; ++i) {
            nnbvh::Ray r{{rays[i].ox, rays[i].oy, rays[i].oz}, {rays[i].dx, rays[i].dy, rays[i].dz}, 0.f};
            auto h = agg.Intersect(r, rays[i].tMax);
            bool p = agg.IntersectP(r, rays[i].tMax);
            const uint32_t *w = &want[5 * (size_t)i];
            if (h.has_value() != (w[0] != NNBVH_MISS) || p != (wocc[i] != 0)) { std::printf("FAIL: single ray %u\n", i); return 1; }
            if (h) {
                ++nHit;
                if (h->prim != w[0] || std::memcmp(&h->tHit, &w[1], 16) != 0) { std::printf("FAIL: single-ray record %u\n", i); return 1; }
            }
        }
        std::printf("PASS %s: %u triangles, %zu nodes, %u rays bit-exact vs oracle (%u/64 single rays hit), kernel %.3f ms\n",
                    name, T, agg.Nodes().size(), N, nHit, agg.LastKernelMs());
        return 0;
    } catch (const nnbvh::Error &e) {
        std::printf("nnbvh::Error %d: %s\n", e.code, e.what());
        return e.code == NNBVH_ERR_CUDA ? 3 : 2;
    }
}

// Triangles and spheres in one primitive list (SimplePrimitive over Triangle / Sphere).
static int RunSphereCase() {
    const uint32_t T = 8000, S = 60, N = 60000;
    std::mt19937 rng(77);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    std::normal_distribution<float> G(0.f, 1.f);
    std::vector<float> tris(9 * (size_t)T);
    for (uint32_t i = 0; i < T; ++i) {
        float c[3] = {U(rng), U(rng), U(rng)};
        for (int v = 0; v < 3; ++v)
            for (int k = 0; k < 3; ++k) tris[9 * (size_t)i + 3 * v + k] = c[k] + 0.02f * (2 * U(rng) - 1);
    }
    std::vector<nnbvh_sphere> sph(S);
    for (nnbvh_sphere &s : sph) {  // a uniform scale and a translation: the inverse is exact enough to state by hand
        const float k = 0.5f + U(rng), c[3] = {U(rng), U(rng), U(rng)};
        const float r[12] = {k, 0, 0, c[0], 0, k, 0, c[1], 0, 0, k, c[2]};
        const float o[12] = {1 / k, 0, 0, -c[0] / k, 0, 1 / k, 0, -c[1] / k, 0, 0, 1 / k, -c[2] / k};
        std::memcpy(s.renderFromObject, r, sizeof r);
        std::memcpy(s.objectFromRender, o, sizeof o);
        s.radius = 0.02f + 0.05f * U(rng);
        s.zMin = (&s - sph.data()) % 3 ? -s.radius : -0.5f * s.radius;
        s.zMax = s.radius;
        s.phiMax = 360.f;
    }
    try {
        nnbvh::GPUBVHAggregate agg(tris, sph, 4, Split::SAH);
        const uint32_t P = T + S;
        std::vector<float> bounds(6 * (size_t)P);
        for (uint32_t i = 0; i < T; ++i)
            for (int k = 0; k < 3; ++k) {
                const float *p = &tris[9 * (size_t)i];
                bounds[6 * (size_t)i + k] = std::min(std::min(p[k], p[3 + k]), p[6 + k]);
                bounds[6 * (size_t)i + 3 + k] = std::max(std::max(p[k], p[3 + k]), p[6 + k]);
            }
        static_assert(sizeof(nnbvh_sphere) == 28 * sizeof(float), "nnbvh_sphere is 28 floats");
        for (uint32_t k = 0; k < S; ++k) nnbvh_oracle_sphere_bounds(sph[k].renderFromObject, &bounds[6 * (size_t)(T + k)]);
        std::vector<nnbvh_oracle_node> onodes(2 * (size_t)P);
        std::vector<int32_t> oorder(P), sphereOf(P);
        int64_t m = nnbvh_oracle_build_bounds(bounds.data(), (int32_t)P, 4, NNBVH_ORACLE_SAH, onodes.data(), oorder.data());
        if (m != (int64_t)agg.Nodes().size() ||
            std::memcmp(onodes.data(), agg.Nodes().data(), sizeof(nnbvh_node) * (size_t)m) != 0) {
            std::printf("FAIL: mixed tree differs from the oracle's\n");
            return 1;
        }
        for (uint32_t i = 0; i < P; ++i) {
            if ((uint32_t)oorder[i] != agg.PrimitiveIndex(i)) { std::printf("FAIL: mixed leaf order differs at %u\n", i); return 1; }
            sphereOf[i] = oorder[i] >= (int32_t)T ? oorder[i] - (int32_t)T : -1;
            if (agg.IsSphere(i) != (sphereOf[i] >= 0)) { std::printf("FAIL: IsSphere(%u)\n", i); return 1; }
        }
        std::vector<nnbvh_ray> rays(N);
        for (uint32_t i = 0; i < N; ++i) {
            nnbvh_ray &r = rays[i];
            r.ox = -0.25f + 1.5f * U(rng); r.oy = -0.25f + 1.5f * U(rng); r.oz = -0.25f + 1.5f * U(rng);
            r.dx = G(rng); r.dy = G(rng); r.dz = G(rng);
            r.tMax = (i & 1) ? U(rng) : std::numeric_limits<float>::infinity();
            r.time = 0.f;
        }
        std::vector<nnbvh_hit> hits(N);
        std::vector<uint8_t> occ(N), wocc(N);
        std::vector<uint32_t> want(5 * (size_t)N);
        agg.IntersectBatch(rays.data(), N, hits.data());
        agg.IntersectPBatch(rays.data(), N, occ.data());
        nnbvh_oracle_intersect_mixed(onodes.data(), agg.OrderedTriangles().data(), sphereOf.data(),
                                     sph[0].renderFromObject, &rays[0].ox, N, want.data(), nullptr, nullptr, 0);
        nnbvh_oracle_intersectP_mixed(onodes.data(), agg.OrderedTriangles().data(), sphereOf.data(),
                                      sph[0].renderFromObject, &rays[0].ox, N, wocc.data(), nullptr, nullptr, 0);
        if (std::memcmp(hits.data(), want.data(), 20 * (size_t)N) != 0) { std::printf("FAIL: mixed closest-hit records differ\n"); return 1; }
        if (std::memcmp(occ.data(), wocc.data(), N) != 0) { std::printf("FAIL: mixed any-hit results differ\n"); return 1; }
        uint32_t onSphere = 0;
        for (uint32_t i = 0; i < N; ++i) onSphere += hits[i].prim != NNBVH_MISS && agg.IsSphere(hits[i].prim);
        if (onSphere < 100) { std::printf("FAIL: only %u sphere hits\n", onSphere); return 1; }
        std::printf("PASS spheres: %u triangles + %u spheres, %u rays bit-exact vs oracle (%u sphere hits)\n", T, S, N, onSphere);
        return 0;
    } catch (const nnbvh::Error &e) {
        std::printf("nnbvh::Error %d: %s\n", e.code, e.what());
        return e.code == NNBVH_ERR_CUDA ? 3 : 2;
    }
}

// TransformedPrimitive instances: top-level triangles + instances of two objects.
static int RunInstanceCase() {
    const uint32_t T = 5000, K = 80, N = 60000;
    std::mt19937 rng(5);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    std::normal_distribution<float> G(0.f, 1.f);
    auto soup = [&](uint32_t n, float lo, float hi, float s) {
        std::vector<float> t(9 * (size_t)n);
        for (uint32_t i = 0; i < n; ++i) {
            float c[3] = {lo + (hi - lo) * U(rng), lo + (hi - lo) * U(rng), lo + (hi - lo) * U(rng)};
            for (int v = 0; v < 3; ++v)
                for (int k = 0; k < 3; ++k) t[9 * (size_t)i + 3 * v + k] = c[k] + s * (2 * U(rng) - 1);
        }
        return t;
    };
    std::vector<float> tris = soup(T, 0.f, 1.f, 0.02f);
    std::vector<std::vector<float>> objs = {soup(600, -0.2f, 0.2f, 0.03f), soup(1, -0.1f, 0.1f, 0.1f)};
    std::vector<nnbvh_instance> inst(K);
    for (uint32_t k = 0; k < K; ++k) {  // rotation about z, uniform scale, translation; inverse written out by hand
        const float a = 6.2831853f * U(rng), c = std::cos(a), s = std::sin(a), sc = 0.2f + 0.3f * U(rng);
        const float t[3] = {U(rng), U(rng), U(rng)};
        const float r[12] = {sc * c, -sc * s, 0, t[0], sc * s, sc * c, 0, t[1], 0, 0, sc, t[2]};
        const float ic = c / sc, is = s / sc;
        const float o[12] = {ic, is, 0, -(ic * t[0] + is * t[1]), -is, ic, 0, -(-is * t[0] + ic * t[1]), 0, 0, 1 / sc, -t[2] / sc};
        std::memcpy(inst[k].renderFromPrimitive, r, sizeof r);
        std::memcpy(inst[k].primitiveFromRender, o, sizeof o);
        inst[k].object = k % 2;
        inst[k].reserved[0] = inst[k].reserved[1] = inst[k].reserved[2] = 0;
    }
    try {
        nnbvh::GPUBVHAggregate agg(tris, objs, inst, 4, Split::SAH);
        // the oracle's view of the same trees
        std::vector<nnbvh_oracle_node> objNodes;
        std::vector<float> objTris;
        std::vector<int32_t> nodeBase, primBase, instObject(K);
        for (const auto &o : agg.Objects()) {
            nodeBase.push_back((int32_t)objNodes.size());
            primBase.push_back((int32_t)(objTris.size() / 9));
            const nnbvh_oracle_node *nd = reinterpret_cast<const nnbvh_oracle_node *>(o.nodes.data());
            objNodes.insert(objNodes.end(), nd, nd + o.nodes.size());
            objTris.insert(objTris.end(), o.orderedTris9.begin(), o.orderedTris9.end());
        }
        std::vector<float> inst24(24 * (size_t)K);
        for (uint32_t k = 0; k < K; ++k) {
            std::memcpy(&inst24[24 * (size_t)k], inst[k].renderFromPrimitive, 96);
            instObject[k] = (int32_t)inst[k].object;
        }
        std::vector<nnbvh_ray> rays(N);
        for (uint32_t i = 0; i < N; ++i) {
            nnbvh_ray &r = rays[i];
            r.ox = -0.25f + 1.5f * U(rng); r.oy = -0.25f + 1.5f * U(rng); r.oz = -0.25f + 1.5f * U(rng);
            r.dx = G(rng); r.dy = G(rng); r.dz = G(rng);
            r.tMax = (i & 1) ? U(rng) : std::numeric_limits<float>::infinity();
            r.time = 0.f;
        }
        std::vector<nnbvh_hit> hits(N);
        std::vector<uint32_t> inner(N), want(5 * (size_t)N), winner(N);
        std::vector<uint8_t> occ(N), wocc(N);
        agg.IntersectBatchInstanced(rays.data(), N, hits.data(), inner.data());
        agg.IntersectPBatch(rays.data(), N, occ.data());
        const nnbvh_oracle_node *top = reinterpret_cast<const nnbvh_oracle_node *>(agg.Nodes().data());
        nnbvh_oracle_intersect_instanced(top, agg.OrderedTriangles().data(), agg.OrderedInstanceIndex().data(), inst24.data(),
                                         instObject.data(), objNodes.data(), nodeBase.data(), objTris.data(), primBase.data(),
                                         &rays[0].ox, N, want.data(), winner.data(), nullptr, nullptr, nullptr, 0);
        nnbvh_oracle_intersect_instanced(top, agg.OrderedTriangles().data(), agg.OrderedInstanceIndex().data(), inst24.data(),
                                         instObject.data(), objNodes.data(), nodeBase.data(), objTris.data(), primBase.data(),
                                         &rays[0].ox, N, nullptr, nullptr, wocc.data(), nullptr, nullptr, 0);
        if (std::memcmp(hits.data(), want.data(), 20 * (size_t)N) != 0) { std::printf("FAIL: instanced closest-hit records differ\n"); return 1; }
        if (std::memcmp(inner.data(), winner.data(), 4 * (size_t)N) != 0) { std::printf("FAIL: inner primitives differ\n"); return 1; }
        if (std::memcmp(occ.data(), wocc.data(), N) != 0) { std::printf("FAIL: instanced any-hit results differ\n"); return 1; }
        uint32_t onInst = 0;
        for (uint32_t i = 0; i < N; ++i) onInst += inner[i] != NNBVH_MISS;
        if (onInst < 500) { std::printf("FAIL: only %u instance hits\n", onInst); return 1; }
        std::printf("PASS instances: %u triangles + %u instances of 2 objects, %u rays bit-exact vs oracle (%u instance hits)\n", T, K, N, onInst);
        return 0;
    } catch (const nnbvh::Error &e) {
        std::printf("nnbvh::Error %d: %s\n", e.code, e.what());
        return e.code == NNBVH_ERR_CUDA ? 3 : 2;
    }
}

int main(int, char **) {
    int rc = RunCase(Split::SAH, NNBVH_ORACLE_SAH, "sah");
    if (rc) return rc;
    rc = RunCase(Split::HLBVH, NNBVH_ORACLE_HLBVH, "hlbvh");
    if (rc) return rc;
    rc = RunSphereCase();
    if (rc) return rc;
    return RunInstanceCase();
}
