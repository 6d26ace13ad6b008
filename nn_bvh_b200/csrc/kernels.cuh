// nn_bvh_b200/csrc/kernels.cuh -- traversal kernels (sm_100a).
//
// Two families, same results bit for bit:
//   BasicKernel : one thread per ray over the reference's own 32-byte nodes, in
//                 the statement order of cpu/aggregates.cpp:529-624.  Carries the
//                 reference's per-ray counters; it is the in-library yardstick.
//   WideKernel  : the product path.  Persistent warps pull rays from a global
//                 cursor, traverse 64-byte two-child nodes (one fetch = two box
//                 tests, leaves never fetched), keep the top of the tree in
//                 shared memory (staged with a TMA bulk copy), and carry the
//                 far child's entry distance on the stack so a popped node needs
//                 no re-fetch to re-test against the shrunken tMax.
//
// Why the WideKernel visits exactly what the reference visits: the reference
// tests a node's box when it is popped/descended into; the only input of that
// test that changes over time is raytMax, and it enters only the final
// `tMin < raytMax` (util/vecmath.h:1607).  So the static part is evaluated when
// the parent is fetched, tMin is kept, and the comparison is made at the moment
// the reference would make it.  Near/far order, leaf order and the in-leaf
// sequential tMax update are the reference's (aggregates.cpp:547-569).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "alpha_math.cuh"
#include "exact_math.cuh"
#include "internal.h"
#include "sphere.cuh"

namespace nnbvh {

struct SceneDev {
    const nnbvh_node *nodes;  // 32-byte reference nodes (BasicKernel)
    const WideNode *wide;     // 64-byte nodes (WideKernel)
    const float4 *tris;       // 3 float4 per ordered primitive
    uint32_t nNodes, nWide, nPrims;
    // Root handling: the reference tests the root's own box first.
    float rootBox[6];
    int32_t rootRef;          // WideNode index (>=0) or leaf ref (<0) of the root
    uint32_t topNodes;        // WideNodes [0, topNodes) are the BFS top of the tree
    const SphereDev *spheres; // sphere table (scenes uploaded with nnbvh_upload_scene_mixed), else NULL
    // instanced scenes (nnbvh_upload_scene_instanced): instance table and the objects' trees / triangles
    const InstanceDev *instances;
    const nnbvh_node *objNodes;
    const float4 *objTris;
    // general scenes (nnbvh_upload_scene_general): constant alpha per ordered top-level primitive, >= 1 = none
    const float *alpha;
    // instanced scenes on the StepKernel: per instance the object's root box, root ref and first triangle in the
    // COMBINED arrays (the objects' WideNodes / triangles follow the top level's in `wide` / `tris`)
    const InstanceStep *instStep;
};

struct RayInput {
    const float4 *aos;  // 2 float4 per ray, or NULL for SoA
    const float *ox, *oy, *oz, *dx, *dy, *dz, *tMax;
    // one-thread-per-ray kernels: optional ray order (binning.cuh); thread j takes ray perm[j] when the
    // device flag *permOn (or a NULL permOn) says the batch was sorted.  Results go to the ray's own slot.
    const uint32_t *perm, *permOn;
};

__device__ __forceinline__ uint64_t OrderedRay(const RayInput &in, uint64_t j) {
    return (in.perm && (!in.permOn || __ldg(in.permOn))) ? (uint64_t)__ldg(in.perm + j) : j;
}

struct RayRegs {
    float ox, oy, oz, tMax, dx, dy, dz;
};

__device__ __forceinline__ RayRegs LoadRay(const RayInput &in, uint64_t i) {
    RayRegs r;
    if (in.aos) {
        float4 a = __ldg(in.aos + 2 * i), b = __ldg(in.aos + 2 * i + 1);
        r.ox = a.x; r.oy = a.y; r.oz = a.z; r.tMax = a.w;
        r.dx = b.x; r.dy = b.y; r.dz = b.z;
    } else {
        r.ox = __ldg(in.ox + i); r.oy = __ldg(in.oy + i); r.oz = __ldg(in.oz + i);
        r.dx = __ldg(in.dx + i); r.dy = __ldg(in.dy + i); r.dz = __ldg(in.dz + i);
        r.tMax = in.tMax ? __ldg(in.tMax + i) : __int_as_float(0x7f800000);
    }
    return r;
}

struct HitRegs {
    uint32_t prim;
    float t, b0, b1, b2;
    uint32_t inner = NNBVH_MISS;  // instanced scenes: ordered primitive inside the hit object
};

__device__ __forceinline__ void StoreHit(nnbvh_hit *hits, uint64_t i, const HitRegs &h) {
    uint32_t *o = reinterpret_cast<uint32_t *>(hits + i);
    o[0] = h.prim;
    o[1] = __float_as_uint(h.t);
    o[2] = __float_as_uint(h.b0);
    o[3] = __float_as_uint(h.b1);
    o[4] = __float_as_uint(h.b2);
}

// Tests the primitives of one leaf in order, each against the CURRENT tMax
// (aggregates.cpp:547-555).  Returns true as soon as ANY finds a hit.
template <bool ANY, bool COUNT>
__device__ __noinline__ bool InnerTraverse(const SceneDev &sc, uint32_t instance, const RayConst &outer, float tMaxOuter,
                              HitRegs &hit, uint32_t prim, float &tMax, uint32_t &nodesVisited,
                              uint32_t &triTests);

template <bool ANY, bool COUNT, bool SPHERES, bool INST = false, bool ALPHA = false>
__device__ __forceinline__ bool LeafTest(const float4 *__restrict__ tris, uint32_t first,
                                         const RayConst &rc, float &tMax, HitRegs &hit,
                                         uint32_t &triTests, const SphereDev *__restrict__ spheres,
                                         const SceneDev *sc = nullptr, uint32_t *nodesVisited = nullptr) {
    uint32_t prim = first;
    while (true) {
        const float4 *tp = tris + 3 * (size_t)prim;
        float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
        uint32_t flags = __float_as_uint(v0.w);
        if (INST && (flags & kPrimInstance)) {
            // TransformedPrimitive::Intersect / IntersectP (cpu/primitive.cpp:112-130)
            uint32_t nv = 0;
            const bool got = InnerTraverse<ANY, COUNT>(*sc, __float_as_uint(v0.x), rc, tMax, hit, prim, tMax, nv, triTests);
            if (COUNT) *nodesVisited += nv;
            if (ANY && got) return true;
            if (flags & kTriLast) break;
            ++prim;
            continue;
        }
        if (SPHERES && (flags & kPrimSphere)) {
            // Sphere::Intersect / IntersectP (shapes.h:137-145, 234-236); nTriTests counts triangles only
            SphereHit sh;
            if (SphereTest(spheres + __float_as_uint(v0.x), rc.ox, rc.oy, rc.oz, rc.dx, rc.dy, rc.dz,
                           tMax, &sh)) {
                if (ANY) return true;
                hit.prim = prim;
                hit.inner = NNBVH_MISS;
                hit.t = sh.t;
                hit.b0 = sh.px;
                hit.b1 = sh.py;
                hit.b2 = sh.pz;
                tMax = sh.t;
            }
            if (flags & kTriLast) break;
            ++prim;
            continue;
        }
        if (COUNT) ++triTests;
        if (!(flags & kTriDegenerate)) {
            float t, b0, b1, b2;
            bool got;
            if (ALPHA) {
                // GeometricPrimitive::Intersect with a constant alpha texture (cpu/primitive.cpp:50-84; alpha.cuh)
                const float a = __ldg(sc->alpha + prim);
                const float p0[3] = {v0.x, v0.y, v0.z}, p1[3] = {v1.x, v1.y, v1.z}, p2[3] = {v2.x, v2.y, v2.z};
                got = a < 1 ? AlphaTriangleTest(rc, tMax, p0, p1, p2, a, &t, &b0, &b1, &b2)
                            : TriangleTest(rc, tMax, v0.x, v0.y, v0.z, v1.x, v1.y, v1.z, v2.x, v2.y, v2.z, &t, &b0, &b1, &b2);
            } else {
                got = TriangleTest(rc, tMax, v0.x, v0.y, v0.z, v1.x, v1.y, v1.z, v2.x, v2.y, v2.z, &t, &b0, &b1, &b2);
            }
            if (got) {
                if (ANY) return true;
                hit.prim = prim;
                hit.inner = NNBVH_MISS;
                hit.t = t;
                hit.b0 = b0;
                hit.b1 = b1;
                hit.b2 = b2;
                tMax = t;
            }
        }
        if (flags & kTriLast) break;
        ++prim;
    }
    return false;
}

// BVHAggregate::Intersect / IntersectP (cpu/aggregates.cpp:529-624) over 32-byte reference nodes, in
// the reference's statement order.  primBase: added to leaf offsets (objects' triangles lie back to
// back).  Returns true when ANY found a hit.
template <bool ANY, bool COUNT, bool SPHERES, bool INST, bool ALPHA = false>
__device__ __forceinline__ bool TraverseRef(const nnbvh_node *__restrict__ nodes, const float4 *__restrict__ tris,
                                            uint32_t primBase, const SceneDev &sc, const RayConst &rc, float &tMax,
                                            HitRegs &hit, uint32_t &nodesVisited, uint32_t &triTests) {
    int stack[kMaxStack];
    int sp = 0, cur = 0;
    const int neg[3] = {rc.negx, rc.negy, rc.negz};
    while (true) {
        if (COUNT) ++nodesVisited;
        const float4 *np = reinterpret_cast<const float4 *>(nodes + cur);
        float4 a = __ldg(np), b = __ldg(np + 1);
        float box[6] = {a.x, a.y, a.z, a.w, b.x, b.y};
        int offset = __float_as_int(b.z);
        uint32_t meta = __float_as_uint(b.w);
        uint32_t nPrims = meta & 0xffffu, axis = (meta >> 16) & 0xffu;
        if (SlabTest(rc, box, tMax)) {
            if (nPrims > 0) {
                if (LeafTest<ANY, COUNT, SPHERES, INST, ALPHA>(tris + 3 * (size_t)primBase, (uint32_t)offset, rc, tMax, hit,
                                                        triTests, sc.spheres, &sc, &nodesVisited))
                    return true;
                if (sp == 0) break;
                cur = stack[--sp];
            } else if (neg[axis]) {
                stack[sp++] = cur + 1;
                cur = offset;
            } else {
                stack[sp++] = offset;
                cur = cur + 1;
            }
        } else {
            if (sp == 0) break;
            cur = stack[--sp];
        }
    }
    return false;
}

template <bool ANY, bool COUNT>
__device__ __noinline__ bool InnerTraverse(const SceneDev &sc, uint32_t instance, const RayConst &outer,
                                           float tMaxOuter, HitRegs &hit, uint32_t prim, float &tMax,
                                           uint32_t &nodesVisited, uint32_t &triTests) {
    const InstanceDev *in = sc.instances + instance;
    float o[3], d[3], tLocal;
    InstanceRay(in, outer.ox, outer.oy, outer.oz, outer.dx, outer.dy, outer.dz, tMaxOuter, o, d, &tLocal);
    const RayConst rc = MakeRayConst(o[0], o[1], o[2], d[0], d[1], d[2]);
    const uint32_t nodeBase = __ldg(&in->nodeBase), primBase = __ldg(&in->primBase);
    HitRegs h{NNBVH_MISS, 0.f, 0.f, 0.f, 0.f};
    const bool any = TraverseRef<ANY, COUNT, false, false>(sc.objNodes + nodeBase, sc.objTris, primBase, sc, rc,
                                                           tLocal, h, nodesVisited, triTests);
    if (ANY) return any;
    if (h.prim == NNBVH_MISS) return false;
    hit.prim = prim;          // the top-level slot of the instance
    hit.inner = h.prim;       // ordered primitive inside the object (LeafTest numbers them from primBase 0)
    hit.t = h.t;              // in the instance ray's parametrisation, as the reference returns it
    hit.b0 = h.b0;
    hit.b1 = h.b1;
    hit.b2 = h.b2;
    tMax = h.t;               // BVHAggregate::Intersect: tMax = si->tHit (aggregates.cpp:552)
    return true;
}

// ---------------------------------------------------------------------------
// BasicKernel: cpu/aggregates.cpp:529-624 per thread.
// ---------------------------------------------------------------------------
// SPHERES, INST, ALPHA all true = the general instantiation: any mix of triangles (with or without a constant
// alpha), spheres and instances in ONE primitive list (nnbvh_upload_scene_general).
template <bool ANY, bool COUNT, bool SPHERES = false, bool INST = false, bool ALPHA = false>
__global__ void __launch_bounds__(128)
BasicKernel(SceneDev sc, RayInput in, uint64_t n, nnbvh_hit *__restrict__ hits,
            uint8_t *__restrict__ occluded, uint32_t *__restrict__ nodesVisitedOut,
            uint32_t *__restrict__ triTestsOut, uint32_t *__restrict__ innerOut = nullptr) {
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint64_t i = OrderedRay(in, j);
    RayRegs rr = LoadRay(in, i);
    RayConst rc = MakeRayConst(rr.ox, rr.oy, rr.oz, rr.dx, rr.dy, rr.dz);
    float tMax = rr.tMax;
    HitRegs hit{NNBVH_MISS, 0.f, 0.f, 0.f, 0.f};
    bool any = false;
    uint32_t nodesVisited = 0, triTests = 0;
    if (sc.nNodes)
        any = TraverseRef<ANY, COUNT, SPHERES, INST, ALPHA>(sc.nodes, sc.tris, 0u, sc, rc, tMax, hit, nodesVisited, triTests);
    if (ANY) occluded[i] = any ? 1 : 0;
    else {
        StoreHit(hits, i, hit);
        if (INST && innerOut) innerOut[i] = hit.inner;
    }
    if (COUNT) {
        nodesVisitedOut[i] = nodesVisited;
        triTestsOut[i] = triTests;
    }
}

// Everything below uses warp intrinsics, shared memory and PTX: device compilation only (the host twin of
// tests/cpp/device_twin.cpp stops here).
#ifdef __CUDACC__
// ---------------------------------------------------------------------------
// WideKernel
// ---------------------------------------------------------------------------

// TMA 1-D bulk copy global -> shared, completion on an mbarrier
// (cp.async.bulk; SASS UBLKCP).  Used once per CTA to stage the top of the tree.
__device__ __forceinline__ void MbarInit(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(bar)),
                 "r"(count));
}
__device__ __forceinline__ void MbarExpectTx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void BulkG2S(void *dstSmem, const void *srcGmem, uint32_t bytes,
                                        uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"((uint32_t)__cvta_generic_to_shared(dstSmem)),
        "l"(srcGmem), "r"(bytes), "r"((uint32_t)__cvta_generic_to_shared(bar))
        : "memory");
}
__device__ __forceinline__ void MbarWait(uint64_t *bar, uint32_t phase) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"((uint32_t)__cvta_generic_to_shared(bar)),
        "r"(phase)
        : "memory");
}

struct WideParams {
    SceneDev sc;
    RayInput in;
    uint64_t n;
    nnbvh_hit *hits;
    uint8_t *occluded;
    unsigned long long *cursor;  // global ray cursor (zeroed before launch)
    const uint32_t *perm;        // optional ray order (binning.cuh); results go to slot perm[j]
    uint32_t smemNodes;          // WideNodes staged in shared memory (<= sc.topNodes)
    uint32_t refillBelow;        // refill finished lanes when fewer than this many are live
    uint32_t leafQuorum;         // run the leaf phase once this many lanes wait at a leaf
};

struct NodeRegs {
    float4 q0, q1, q2;
    int4 q3;
};

__device__ __forceinline__ NodeRegs LoadWide(const WideNode *__restrict__ g,
                                             const WideNode *smem, uint32_t smemNodes, int idx) {
    NodeRegs r;
    if ((uint32_t)idx < smemNodes) {
        const float4 *p = reinterpret_cast<const float4 *>(smem + idx);
        r.q0 = p[0]; r.q1 = p[1]; r.q2 = p[2];
        r.q3 = *reinterpret_cast<const int4 *>(p + 3);
    } else {
        const float4 *p = reinterpret_cast<const float4 *>(g + idx);
        r.q0 = __ldg(p); r.q1 = __ldg(p + 1); r.q2 = __ldg(p + 2);
        r.q3 = __ldg(reinterpret_cast<const int4 *>(p + 3));
    }
    return r;
}

// MINB = CTAs per SM the register budget is sized for (256 threads per CTA).
template <bool ANY, int MINB>
__global__ void __launch_bounds__(256, MINB) WideKernel(WideParams p) {
    extern __shared__ __align__(128) unsigned char smemRaw[];
    WideNode *top = reinterpret_cast<WideNode *>(smemRaw);
    __shared__ uint64_t bar;

    // ---- stage the top of the tree (TMA bulk copy, one elected thread) ----
    if (p.smemNodes) {
        if (threadIdx.x == 0) {
            MbarInit(&bar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t bytes = p.smemNodes * (uint32_t)sizeof(WideNode);
            MbarExpectTx(&bar, bytes);
            const unsigned char *src = reinterpret_cast<const unsigned char *>(p.sc.wide);
            for (uint32_t off = 0; off < bytes; off += 16384) {
                uint32_t len = min(16384u, bytes - off);
                BulkG2S(smemRaw + off, src + off, len, &bar);
            }
        }
        MbarWait(&bar, 0);
    }

    const SceneDev &sc = p.sc;
    const unsigned lane = threadIdx.x & 31;
    const float4 *__restrict__ tris = sc.tris;

    // per-lane traversal state
    uint64_t rayIdx = 0;
    bool alive = false;
    RayConst rc;
    float tMax = 0.f;
    HitRegs hit{NNBVH_MISS, 0.f, 0.f, 0.f, 0.f};
    int cur = 0, sp = 0;
    int stackRef[kMaxStack];
    float stackT[kMaxStack];
    bool exhausted = false;  // the global cursor ran past n (warp-uniform)
    uint32_t dummy = 0;

    // Phase-separated loop.  An if/else "node or leaf" body runs both bodies
    // almost every iteration (some lane of 32 is nearly always at a leaf), which
    // measured at 8.7 active lanes per instruction.  Instead: lanes that reach a
    // leaf PARK (no speculation -- the reference's order is kept exactly) while the
    // others keep descending, until `leafQuorum` lanes are parked or nobody is
    // descending; then the leaf body runs once for all parked lanes.
    const unsigned leafQuorum = p.leafQuorum;
    while (true) {
        // ---- refill: finished lanes take the next rays from the global cursor ----
        unsigned liveMask = __ballot_sync(0xffffffffu, alive);
        if (!exhausted && (unsigned)__popc(liveMask) < p.refillBelow) {
            unsigned want = ~liveMask;
            unsigned nWant = __popc(want);
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(p.cursor, (unsigned long long)nWant);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base + nWant >= p.n) exhausted = true;
            if (!alive) {
                uint64_t idx = base + __popc(want & ((1u << lane) - 1));
                if (idx < p.n) {
                    if (p.perm) idx = __ldg(p.perm + idx);
                    rayIdx = idx;
                    RayRegs rr = LoadRay(p.in, idx);
                    rc = MakeRayConst(rr.ox, rr.oy, rr.oz, rr.dx, rr.dy, rr.dz);
                    tMax = rr.tMax;
                    hit.prim = NNBVH_MISS;
                    hit.t = hit.b0 = hit.b1 = hit.b2 = 0.f;
                    sp = 0;
                    // Root: the reference tests the root's own box first
                    // (aggregates.cpp:543); nodes == nullptr -> miss (:531).
                    if (sc.nNodes && SlabTest(rc, sc.rootBox, tMax)) {
                        cur = sc.rootRef;
                        alive = true;
                    } else if (ANY) p.occluded[idx] = 0;
                    else StoreHit(p.hits, idx, hit);
                }
            }
            liveMask = __ballot_sync(0xffffffffu, alive);
        }
        if (liveMask == 0) {
            if (exhausted) break;
            continue;
        }

        // ---- phase A: interior nodes ----
        while (true) {
            const bool atNode = alive && cur >= 0;
            const unsigned nodeMask = __ballot_sync(0xffffffffu, atNode);
            const unsigned leafMask = __ballot_sync(0xffffffffu, alive && cur < 0);
            if (nodeMask == 0 || (unsigned)__popc(leafMask) >= leafQuorum) break;
            if (atNode) {
                NodeRegs nd = LoadWide(sc.wide, top, p.smemNodes, cur);
                const float lo0x = nd.q0.x, lo0y = nd.q0.y, lo0z = nd.q0.z;
                const float hi0x = nd.q0.w, hi0y = nd.q1.x, hi0z = nd.q1.y;
                const float lo1x = nd.q1.z, lo1y = nd.q1.w, lo1z = nd.q2.x;
                const float hi1x = nd.q2.y, hi1y = nd.q2.z, hi1z = nd.q2.w;
                float t0, t1;
                const bool ok0 = SlabStatic(rc, rc.negx ? hi0x : lo0x, rc.negx ? lo0x : hi0x,
                                            rc.negy ? hi0y : lo0y, rc.negy ? lo0y : hi0y,
                                            rc.negz ? hi0z : lo0z, rc.negz ? lo0z : hi0z, &t0);
                const bool ok1 = SlabStatic(rc, rc.negx ? hi1x : lo1x, rc.negx ? lo1x : hi1x,
                                            rc.negy ? hi1y : lo1y, rc.negy ? lo1y : hi1y,
                                            rc.negz ? hi1z : lo1z, rc.negz ? lo1z : hi1z, &t1);
                const int axis = nd.q3.z;
                const bool secondFirst = axis == 0 ? rc.negx : (axis == 1 ? rc.negy : rc.negz);
                // "static" = every part of the box test that does not involve tMax.
                const bool stNear = secondFirst ? ok1 : ok0, stFar = secondFirst ? ok0 : ok1;
                const int refNear = secondFirst ? nd.q3.y : nd.q3.x;
                const int refFar = secondFirst ? nd.q3.x : nd.q3.y;
                const float tNear = secondFirst ? t1 : t0, tFar = secondFirst ? t0 : t1;
                if (stNear && tNear < tMax) {
                    // The far child is stacked whenever its static test passes, NOT only
                    // when tFar < tMax now: the reference's tMax can GROW by an ulp (a
                    // later primitive is accepted in scaled space, tScaled <= tMax*det,
                    // with a rounded t above tMax; shapes.cpp:239-247), so a box that
                    // fails `tMin < tMax` now may pass when the reference pops it.
                    if (stFar) {
                        stackRef[sp] = refFar;
                        stackT[sp] = tFar;
                        ++sp;
                    }
                    cur = refNear;
                } else if (stFar && tFar < tMax) {
                    cur = refFar;
                } else {
                    // pop until an entry whose box beats the current tMax (:570-574)
                    while (true) {
                        if (sp == 0) {
                            alive = false;
                            if (ANY) p.occluded[rayIdx] = 0;
                            else StoreHit(p.hits, rayIdx, hit);
                            break;
                        }
                        --sp;
                        if (stackT[sp] < tMax) {
                            cur = stackRef[sp];
                            break;
                        }
                    }
                }
            }
        }

        // ---- phase B: leaves (all parked lanes together) ----
        if (alive && cur < 0) {
            if (LeafTest<ANY, false, false>(tris, (uint32_t)cur & 0x7fffffffu, rc, tMax, hit, dummy, nullptr)) {
                alive = false;  // ANY: first accepted primitive ends the ray (:598-601)
                p.occluded[rayIdx] = 1;
            } else {
                while (true) {  // :556-558
                    if (sp == 0) {
                        alive = false;
                        if (ANY) p.occluded[rayIdx] = 0;
                        else StoreHit(p.hits, rayIdx, hit);
                        break;
                    }
                    --sp;
                    if (stackT[sp] < tMax) {
                        cur = stackRef[sp];
                        break;
                    }
                }
            }
        }
    }
}

#endif  // __CUDACC__

}  // namespace nnbvh
