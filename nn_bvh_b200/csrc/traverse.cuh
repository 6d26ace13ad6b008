// nn_bvh_b200/csrc/traverse.cuh -- StepKernel: the product traversal kernel (sm_100a).
//
// Same results, bit for bit, as BasicKernel / the reference (cpu/aggregates.cpp:529-624),
// organised for the resource that actually binds on B200 once the scene sits in L1/L2:
// warp issue slots and L1 wavefronts (profiles/r01a_*: 70 % issue utilisation at 13 of 32
// lanes active, L1 at 75 %).
//
//  * Persistent warps pull rays from a global cursor; finished lanes are refilled.
//  * Every lane carries `cur`, the node or leaf it is AT.  Both boxes of a node are tested when the node
//    is fetched; `tEntry < tMax` -- the only part of Bounds3f::IntersectP that depends on the shrinking
//    tMax (util/vecmath.h:1607) -- is evaluated at the moment the lane steps onto a child: at once for
//    the child it descends to, at the pop for the stacked one (the entry carries its tEntry), which is
//    the moment the reference tests it.  tMax only changes while the lane is parked at a leaf, so a lane
//    at a node never has to look at it again.
//  * One step per lane per iteration, straight-line: a lane whose node has no passing child pops (a
//    branch; entries that fail the test against the current tMax are dropped inside that pop).
//  * A 64-byte node is fetched with two 256-bit loads (LDG.E.ENL2.256): half the L1
//    wavefronts of four LDG.128.
//  * Fast slab (EXACT=false): for rays whose invDir is finite and non-zero and whose origin
//    is finite -- no 0*inf NaN can arise -- the compare/select cascade of vecmath.h:1587-1605
//    equals  max3(tx0,ty0,tz0) <= min3(tx1',ty1',tz1')  &&  min3 > 0  with the entry distance
//    max3 (proof in DESIGN.md section 5.1; needs lo <= hi and finite boxes, validated at
//    upload).  Rays that do not qualify are appended to a deferred list and traced by the
//    EXACT=true instantiation, which runs the reference's cascade verbatim (SlabStatic).
//  * Leaves: lanes park at a leaf until `leafQuorum` lanes wait (or nobody is at an interior
//    node); then every parked lane tests ONE triangle.  A lane stays parked for the next
//    triangle of its leaf, so the triangle body always starts with many lanes.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "exact_math.cuh"
#include "internal.h"
#include "kernels.cuh"

#ifndef NNBVH_PREFETCH_AHEAD
#define NNBVH_PREFETCH_AHEAD 0  // refill: L2-prefetch the rays this many permutation entries ahead of the cursor (0 = off)
#endif
#ifndef NNBVH_PREFETCH_TRI
#define NNBVH_PREFETCH_TRI 0  // 1: L1-prefetch the first triangle when a lane steps onto a leaf; 2: both of its lines
#endif
#ifndef NNBVH_ANY_ORDER
#define NNBVH_ANY_ORDER 1
#endif
#ifndef NNBVH_ANY_PRED_POP
#define NNBVH_ANY_PRED_POP 0  // 1: any hit pops with a predicated load too -- measured 1-2 % slower than the branch (r02zk)
#endif
#ifndef NNBVH_CLOSEST_PRED_POP
#define NNBVH_CLOSEST_PRED_POP 0  // 1: closest hit pops with a predicated ld.local in the step's straight line -- measured 1-4 % slower than the branch (r02zl)
#endif
#ifndef NNBVH_L1_NOALLOC
#define NNBVH_L1_NOALLOC 0  // 1: triangle loads bypass L1 allocation (ld.global.nc.L1::no_allocate); 2: the refill's ray loads too
#endif
#ifndef NNBVH_STACK_CACHE
#define NNBVH_STACK_CACHE 0  // 1: closest hit (no instances) keeps the top stack entry in two registers
#endif
#ifndef NNBVH_SPC
#define NNBVH_SPC 4  // node steps per look at the leaf quorum (unrolled)
#endif

// popOne takes the address of the top stack entry before the first push has been written (it is not
// dereferenced then): nvcc's "used before its value is set" does not apply
#pragma nv_diag_suppress 549

namespace nnbvh {

// A 16-byte read-only load that does not allocate in the L1 (data with no reuse inside the SM: a triangle is tested
// once per visit, a ray is read once) -- the L1 is shared with the lanes' stacks and the nodes.
__device__ __forceinline__ float4 LdgNoAllocate(const float4 *p) {
    float4 v;
    asm("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
        : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

struct StepParams {
    SceneDev sc;
    RayInput in;
    uint64_t n;                   // rays (or capacity of `perm` when nPtr is used)
    const uint32_t *nPtr;         // if set: number of entries of `perm` to process, on the device
    nnbvh_hit *hits;
    uint8_t *occluded;            // any hit: ZEROED by the caller before the launch; the kernel only writes the 1s
    unsigned long long *cursor;   // zeroed before launch
    const uint32_t *perm;         // optional ray order; results go to slot perm[j]
    const uint32_t *permOn;       // optional device flag: 0 = ignore perm (binning decided not to sort)
    uint32_t *deferred;           // EXACT=false: indices of rays needing the exact slab test
    uint32_t *nDeferred;
    uint32_t refillBelow, leafQuorum;
    uint32_t sphereQuorum;        // SPHERES: run the sphere tests when this many lanes wait at sphere slots
    unsigned long long *stats;    // STATS instantiation only: 8 lane-occupancy counters
    uint32_t *inner;              // INST, closest hit: optional plane, ordered primitive inside the hit object
};

struct F8 {
    float v[8];
};

// 256-bit read-only global load (PTX ISA 8.8, sm_100+): SASS LDG.E.ENL2.256.CONSTANT.
__device__ __forceinline__ F8 Ldg256(const void *p) {
    F8 r;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]),
                   "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7])
                 : "l"(p));
    return r;
}

// The per-ray constants the step loop keeps in registers (no d: it is folded into inv / S).
struct StepRay {
    float ox, oy, oz;
    float ix, iy, iz;
    float Sx, Sy, Sz;
    int kz;         // MaxComponentIndex(Abs(d)), shapes.cpp:186
    unsigned neg;   // bit k = dirIsNeg[k], cpu/aggregates.cpp:535
};

__device__ __forceinline__ RayConst ToRayConst(const StepRay &s) {
    RayConst r;
    r.ox = s.ox; r.oy = s.oy; r.oz = s.oz;
    r.dx = r.dy = r.dz = 0.f;  // unused by SlabStatic / TriangleTest
    r.ix = s.ix; r.iy = s.iy; r.iz = s.iz;
    r.negx = s.neg & 1; r.negy = (s.neg >> 1) & 1; r.negz = (s.neg >> 2) & 1;
    r.kz = s.kz;
    r.kx = s.kz == 2 ? 0 : s.kz + 1;
    r.ky = r.kx == 2 ? 0 : r.kx + 1;
    r.Sx = s.Sx; r.Sy = s.Sy; r.Sz = s.Sz;
    return r;
}

constexpr float kNegInf = -__builtin_huge_valf();

template <bool ANY> struct StackWord { typedef unsigned long long type; };
template <> struct StackWord<true> { typedef unsigned type; };

// INST: the top-level list holds TransformedPrimitive slots (kPrimInstance).  The objects' WideNodes and triangles
// live in the SAME arrays as the top level (refs are absolute), so entering an instance is: transform the ray
// (InstanceRay = Transform::ApplyInverse(Ray, &tMax)), test the object's root box, push a sentinel holding where
// the outer traversal continues (and its tMax), and carry on with `cur` = the object's root; leaving it -- when the
// stack is back at the sentinel -- re-derives the outer ray's constants from the ray in memory.  A ray whose
// TRANSFORMED direction does not qualify for the fast slab is handed to the exact launch as a whole.
constexpr int kPopMark = 0x7fffffff;  // sentinel: "the instance was the last primitive of its leaf: pop"

// ALPHA: sc.alpha holds a constant alpha per TOP-LEVEL ordered primitive (>= 1: none); a triangle with alpha < 1 is
// tested as GeometricPrimitive::Intersect does (alpha_math.cuh) with the ray as the caller gave it (read again:
// the lane keeps no direction).  SPHERES + INST + ALPHA together = the general instantiation.
template <bool ANY, bool EXACT, int MINB, bool STATS = false, bool SPHERES = false, bool INST = false, bool ALPHA = false>
__global__ void __launch_bounds__(256, MINB) StepKernel(StepParams p) {
    // STATS: [0] phase-A iterations, [1] lanes at a node, [2] lanes parked at a leaf, [3] lanes without
    // a ray, [4] phase-B passes, [5] lanes parked there, [6] lanes that ran a triangle test, [7] refills
    unsigned long long st[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const SceneDev &sc = p.sc;
    const unsigned lane = threadIdx.x & 31;
    const float4 *__restrict__ tris = sc.tris;
    const unsigned char *__restrict__ wideBase = reinterpret_cast<const unsigned char *>(sc.wide);
    const uint64_t nRays = p.nPtr ? (uint64_t)__ldg(p.nPtr) : p.n;
    const uint32_t *__restrict__ perm = (p.perm && (!p.permOn || __ldg(p.permOn))) ? p.perm : nullptr;

    uint32_t rayIdx = 0;
    StepRay ry{};
    float tMax = 0.f;
    // Closest hit: no lane carries its hit record.  Every accepted hit is written to the ray's output slot at once
    // (a later, closer one overwrites it: 1.2 writes per ray on bathroom) and tMax IS the record's t; what stays in
    // registers is one bit, kHasHit of ry.neg, so that a ray that ends without a hit stores the miss record.
    // Five registers less in the node loop (at 48 per thread the compiler kept the record in local memory there).
    constexpr unsigned kHasHit = 0x80000000u;
    const HitRegs kMissRecord{NNBVH_MISS, 0.f, 0.f, 0.f, 0.f};
    // kDead: the lane has no ray.  A lane is at an interior node when cur >= 0 and parked at a leaf slot when
    // cur < kDead (bit 31 + the ordered primitive, which is never 0x7fffffff): one compare each, no flag to carry.
    constexpr int kDead = -1;
    int cur = kDead, sp = 0;
    // NNBVH_STACK_CACHE: the newest stack entry lives in (topRef, topT) and only goes to local memory when a second one
    // is pushed on top of it; a pop that follows a push touches no memory.  (The stack is a third of the L1's sector
    // traffic)  Closest hit without instances only: the instance code addresses stack[] directly.
    constexpr bool CACHE = NNBVH_STACK_CACHE && !NNBVH_CLOSEST_PRED_POP && !ANY && !INST;
    constexpr int kNone = -1;  // no ref has this value (kDead)
    int topRef = kNone;
    float topT = 0.f;
    // stack entry = {ref, tEntry} packed into 64 bits (one STL.64 / LDL.64 per push / pop).
    // ANY: IntersectP never shrinks tMax (aggregates.cpp:581-624), so `tMin < raytMax` of a child can be
    // settled for good when its parent is fetched: only children that pass it are entered or stacked, the
    // entry is the bare 32-bit ref and nothing is re-validated at a pop.
    typename StackWord<ANY>::type stack[kMaxStack];
    bool exhausted = false;
    // SPHERES: a lane that steps onto a sphere slot waits (sphWait) until `sphereQuorum` lanes of its warp
    // wait at spheres or no lane can do anything else: Sphere::BasicIntersect is ~20x a triangle test, and
    // run for one or two lanes at a time it was most of the kernel (crown's environment shell: every
    // escaping ray tests it once).  Waiting changes nothing a ray can observe: its own state is frozen.
    bool sphWait = false;
    // INST: >= 0 while the lane is inside an instance (the instance's top-level slot); stack entries below spBase
    // belong to the outer traversal, stack[spBase - 1] is the sentinel
    int instSlot = -1, spBase = 0;
    uint32_t instPrimBase = 0;
    bool instHit = false;

    const unsigned leafQuorum = p.leafQuorum;

    auto fastSlabSafe = [](const RayConst &rc) {  // finite, non-zero invDir and finite origin <=> the fast slab is exact
        const float big = __int_as_float(0x7f800000);
        const bool okInv = fabsf(rc.ix) < big && fabsf(rc.iy) < big && fabsf(rc.iz) < big &&
                           rc.ix != 0.f && rc.iy != 0.f && rc.iz != 0.f;
        const bool okOrg = fabsf(rc.ox) < big && fabsf(rc.oy) < big && fabsf(rc.oz) < big;
        return okInv && okOrg;
    };
    auto setRay = [&](const RayConst &rc) {
        ry.ox = rc.ox; ry.oy = rc.oy; ry.oz = rc.oz;
        ry.ix = rc.ix; ry.iy = rc.iy; ry.iz = rc.iz;
        ry.Sx = rc.Sx; ry.Sy = rc.Sy; ry.Sz = rc.Sz;
        ry.kz = rc.kz;
        ry.neg = (ry.neg & kHasHit) | (unsigned)rc.negx | ((unsigned)rc.negy << 1) | ((unsigned)rc.negz << 2);
    };
    auto finish = [&]() {  // the ray is done: a hit's record is already out
        cur = kDead;
        if (ANY) {
            // occluded[] is zeroed before the launch (StepParams::occluded): an unoccluded ray just ends
        } else if (!(ry.neg & kHasHit)) {
            StoreHit(p.hits, rayIdx, kMissRecord);
            if (INST && p.inner) p.inner[rayIdx] = NNBVH_MISS;
        }
    };
    auto accept = [&](uint32_t prim, uint32_t inner, float t, float b0, float b1, float b2) {
        StoreHit(p.hits, rayIdx, HitRegs{prim, t, b0, b1, b2});
        if (INST && p.inner) p.inner[rayIdx] = inner;
        ry.neg |= kHasHit;
        tMax = t;
    };
    // The pop of both phases (aggregates.cpp:556-558, :570-574) for the lanes that `want` it: a branch.  Closest
    // hit: entries whose box test fails against the current tMax (`tEntry < tMax` false) are discarded inside the
    // pop -- the reference pops those one by one with the same outcome.  The same pop written as ONE predicated
    // ld.local in the step's straight line (NNBVH_CLOSEST_PRED_POP / NNBVH_ANY_PRED_POP = 1; stale entries and the
    // ray's end in a rare arm) issues 5 % fewer instructions and measured 1-4 % SLOWER for both kinds of ray
    // (profiles/r02_experiments.txt: the load's latency then sits in every lane's step), so it stays off.
    auto popOne = [&](const bool want) {
        if (ANY) {
#if NNBVH_ANY_PRED_POP
            const bool pop = want && sp > 0;
            cur = (want && !pop) ? kDead : cur;  // unoccluded: occluded[] was zeroed before the launch
            const size_t top = __cvta_generic_to_local(&stack[sp - 1]);  // not dereferenced unless pop
            asm volatile("{\n .reg .pred q;\n setp.ne.s32 q, %3, 0;\n @q ld.local.b32 %0, [%2];\n @q add.s32 %1, %1, -1;\n}"
                         : "+r"(cur), "+r"(sp) : "l"(top), "r"((int)pop) : "memory");
#else
            if (want) {
                if (sp > 0) {
                    --sp;
                    cur = (int)(unsigned)stack[sp];
                } else {
                    cur = kDead;  // unoccluded: occluded[] was zeroed before the launch
                }
            }
#endif
            return;
        }
#if !NNBVH_CLOSEST_PRED_POP
        if (CACHE) {
            if (want) {
                bool got = false;
                while (true) {
                    float t;
                    if (topRef != kNone) {
                        cur = topRef;
                        t = topT;
                        topRef = kNone;
                    } else if (sp > 0) {
                        --sp;
                        const unsigned long long e = (unsigned long long)stack[sp];
                        cur = (int)(unsigned)e;
                        t = __uint_as_float((unsigned)(e >> 32));
                    } else
                        break;
                    if (t < tMax) {
                        got = true;
                        break;
                    }
                }
                if (!got) finish();
            }
            return;
        }
        if (want) {
            bool got = false;
            while (sp > 0) {
                --sp;
                const unsigned long long e = (unsigned long long)stack[sp];
                cur = (int)(unsigned)e;
                if (__uint_as_float((unsigned)(e >> 32)) < tMax) {
                    got = true;
                    break;
                }
            }
            if (!got) finish();
        }
        return;
#endif
        const bool pop = want && sp > 0;
        float tEntry;  // written by the load; only looked at under `pop`
        // `if (pop) { e = stack[--sp]; ... }` compiles to a branch around the load; as a predicated ld.local the
        // registers simply keep their values in the other lanes.  entry = {ref (low word), tEntry bits (high word)}
        const size_t top = __cvta_generic_to_local(&stack[sp - 1]);  // not dereferenced unless pop
        asm volatile("{\n .reg .pred q;\n setp.ne.s32 q, %4, 0;\n @q ld.local.v2.b32 {%0, %1}, [%3];\n"
                     " @q add.s32 %2, %2, -1;\n}"
                     : "+r"(cur), "=f"(tEntry), "+r"(sp) : "l"(top), "r"((int)pop) : "memory");
        if ((want && !pop) || (pop && !(tEntry < tMax))) {
            bool got = false;
            while (sp > 0) {
                --sp;
                const unsigned long long e = (unsigned long long)stack[sp];
                cur = (int)(unsigned)e;
                if (__uint_as_float((unsigned)(e >> 32)) < tMax) {
                    got = true;
                    break;
                }
            }
            if (!got) finish();
        }
    };
    // INST: the pop of both phases.  Entries above spBase are the current level's; with none left inside an instance
    // the sentinel is taken: TransformedPrimitive::Intersect has returned, BVHAggregate::Intersect sets
    // tMax = tHit when it brought a hit (aggregates.cpp:549-553) and goes on with the leaf's next primitive.
    auto popInst = [&]() {
        while (true) {
            bool got = false;
            if (ANY) {
                if (sp > spBase) {
                    --sp;
                    cur = (int)(unsigned)stack[sp];
                    got = true;
                }
            } else {
                while (sp > spBase) {
                    --sp;
                    const unsigned long long e = (unsigned long long)stack[sp];
                    cur = (int)(unsigned)e;
                    if (__uint_as_float((unsigned)(e >> 32)) < tMax) {
                        got = true;
                        break;
                    }
                }
            }
            if (got) return;
            if (instSlot < 0) {
                finish();
                return;
            }
            --sp;  // the sentinel
            const unsigned long long e = (unsigned long long)stack[sp];
            const int outerNext = (int)(unsigned)e;
            const RayRegs rr = LoadRay(p.in, rayIdx);
            setRay(MakeRayConst(rr.ox, rr.oy, rr.oz, rr.dx, rr.dy, rr.dz));
            if (ANY) tMax = rr.tMax;
            else tMax = instHit ? tMax : __uint_as_float((unsigned)(e >> 32));  // the hit's t is the inner tMax
            instSlot = -1;
            spBase = 0;
            instHit = false;
            if (outerNext != kPopMark) {
                cur = outerNext;  // the next primitive of the outer leaf: no box test again
                return;
            }
        }
    };

    while (true) {
        // ---- refill: finished lanes take the next rays from the global cursor ----
        unsigned liveMask = __ballot_sync(0xffffffffu, cur != kDead);
        if (!exhausted && (unsigned)__popc(liveMask) < p.refillBelow) {
            unsigned want = ~liveMask;
            unsigned nWant = __popc(want);
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(p.cursor, (unsigned long long)nWant);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base + nWant >= nRays) exhausted = true;
            if (STATS) st[7] += 1;
            if (NNBVH_PREFETCH_AHEAD > 0 && perm && p.in.aos) {
                // Through the sorted permutation a refill reads rays from random places of the caller's array:
                // DRAM latency with the whole warp waiting.  Every refill touches the rays a later refill (of any
                // warp) will take, NNBVH_PREFETCH_AHEAD entries ahead of the cursor, so that they wait in the L2.
                const uint64_t ja = base + NNBVH_PREFETCH_AHEAD + lane;
                if (ja < nRays) {
                    const uint32_t ia = __ldg(perm + ja);
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(p.in.aos + 2 * (size_t)ia));
                }
            }
            if (cur == kDead) {
                uint64_t j = base + __popc(want & ((1u << lane) - 1));
                if (j < nRays) {
                    uint32_t idx = perm ? __ldg(perm + j) : (uint32_t)j;
#if NNBVH_L1_NOALLOC >= 2
                    RayRegs rr;
                    if (p.in.aos) {
                        const float4 a = LdgNoAllocate(p.in.aos + 2 * (size_t)idx), b = LdgNoAllocate(p.in.aos + 2 * (size_t)idx + 1);
                        rr.ox = a.x; rr.oy = a.y; rr.oz = a.z; rr.tMax = a.w;
                        rr.dx = b.x; rr.dy = b.y; rr.dz = b.z;
                    } else
                        rr = LoadRay(p.in, idx);
#else
                    RayRegs rr = LoadRay(p.in, idx);
#endif
                    RayConst rc = MakeRayConst(rr.ox, rr.oy, rr.oz, rr.dx, rr.dy, rr.dz);
                    bool defer = false;
                    if (!EXACT) {
                        // finite, non-zero invDir and finite origin <=> the fast slab is exact
                        const float big = __int_as_float(0x7f800000);
                        bool okInv = fabsf(rc.ix) < big && fabsf(rc.iy) < big && fabsf(rc.iz) < big &&
                                     rc.ix != 0.f && rc.iy != 0.f && rc.iz != 0.f;
                        bool okOrg = fabsf(rc.ox) < big && fabsf(rc.oy) < big && fabsf(rc.oz) < big;
                        defer = !(okInv && okOrg);
                    }
                    if (defer) {
                        uint32_t pos = atomicAdd(p.nDeferred, 1u);
                        p.deferred[pos] = idx;
                    } else {
                        rayIdx = idx;
                        ry.ox = rc.ox; ry.oy = rc.oy; ry.oz = rc.oz;
                        ry.ix = rc.ix; ry.iy = rc.iy; ry.iz = rc.iz;
                        ry.Sx = rc.Sx; ry.Sy = rc.Sy; ry.Sz = rc.Sz;
                        ry.kz = rc.kz;
                        ry.neg = (unsigned)rc.negx | ((unsigned)rc.negy << 1) | ((unsigned)rc.negz << 2);
                        tMax = rr.tMax;
                        if (INST) {
                            instSlot = -1;
                            spBase = 0;
                            instHit = false;
                        }
                        sp = 0;
                        topRef = kNone;
                        // Root: the reference tests the root's own box first (aggregates.cpp:543);
                        // nodes == nullptr -> miss (:531).  Exact cascade, once per ray.
                        float tRoot = 0.f;
                        bool rootOk = false;
                        if (sc.nNodes) {
                            const float *b = sc.rootBox;
                            rootOk = SlabStatic(rc, rc.negx ? b[3] : b[0], rc.negx ? b[0] : b[3],
                                                rc.negy ? b[4] : b[1], rc.negy ? b[1] : b[4],
                                                rc.negz ? b[5] : b[2], rc.negz ? b[2] : b[5], &tRoot);
                        }
                        if (rootOk && tRoot < tMax) {
                            cur = sc.rootRef;
                        } else if (ANY) {
                        } else {
                            StoreHit(p.hits, idx, kMissRecord);
                            if (INST && p.inner) p.inner[idx] = NNBVH_MISS;
                        }
                    }
                }
            }
            liveMask = __ballot_sync(0xffffffffu, cur != kDead);
        }
        if (liveMask == 0) {
            if (exhausted) break;
            continue;
        }

        // ---- phase A: interior nodes, one step per lane per iteration ----
        while (true) {
            const bool atNode = cur >= 0;
            const unsigned nodeMask = __ballot_sync(0xffffffffu, atNode);
            const unsigned leafMask = __ballot_sync(0xffffffffu, cur < kDead && !(SPHERES && sphWait));
            if (nodeMask == 0 || (unsigned)__popc(leafMask) >= leafQuorum) break;
            // NNBVH_SPC node steps between two looks at the leaf quorum: the two ballots, the popc
            // and the compares were 7 % of the issued instructions at one step per look.  A lane that
            // reaches a leaf (or loses its ray) inside the group sits out the rest of it.
#pragma unroll
            for (int u = 0; u < NNBVH_SPC; ++u) {
            if (STATS) {  // per node step, not per look
                const unsigned nm = __ballot_sync(0xffffffffu, cur >= 0);
                const unsigned lm = __ballot_sync(0xffffffffu, cur < kDead);
                st[0] += 1;
                st[1] += __popc(nm);
                st[2] += __popc(lm);
                st[3] += 32 - __popc(nm | lm);
            }
            if (cur >= 0) {
                // `descend`: the lane moves to a child of `cur`; otherwise it pops.  The child selection is
                // written with selects and ONE predicated push, not as the reference's if / else-if chain:
                // lanes of a warp take different arms at almost every node, and as branches the arms ran
                // one after the other with 3-10 lanes each -- a third of the node step's issue slots
                // (profiles/r02c, 0x2190-0x24c0 of StepKernel<0,0,5>).
                bool descend = false;
                {
                    const unsigned char *np = wideBase + (size_t)(unsigned)cur * sizeof(WideNode);
                    const F8 a = Ldg256(np), b = Ldg256(np + 32);
                    // WideNode: lo0 xyz, hi0 xyz, lo1 xy | lo1 z, hi1 xyz, ref0, ref1, axis, 1 << axis
                    const int ref0 = __float_as_int(b.v[4]), ref1 = __float_as_int(b.v[5]);
                    const unsigned axisBit = __float_as_uint(b.v[7]);  // 1 << axis
                    // aggregates.cpp:561.  IntersectP's answer does not depend on the visit order (no tMax update,
                    // any accepted primitive ends the ray): NNBVH_ANY_ORDER 0 = child 0 first, 2 = the child whose
                    // box is entered first (smaller tEntry), 1 = the reference's order (default).
#if NNBVH_ANY_ORDER == 0
                    const bool secondFirst = ANY ? false : (ry.neg & axisBit) != 0;
#else
                    const bool secondFirst = (ry.neg & axisBit) != 0;
#endif
                    const int refNear = secondFirst ? ref1 : ref0, refFar = secondFirst ? ref0 : ref1;
                    float tNear, tFar;
                    bool stNear, stFar;
                    if (EXACT) {
                        const bool nx = ry.neg & 1, ny = ry.neg & 2, nz = ry.neg & 4;
                        const float n0x = nx ? a.v[3] : a.v[0], f0x = nx ? a.v[0] : a.v[3];
                        const float n0y = ny ? a.v[4] : a.v[1], f0y = ny ? a.v[1] : a.v[4];
                        const float n0z = nz ? a.v[5] : a.v[2], f0z = nz ? a.v[2] : a.v[5];
                        const float n1x = nx ? b.v[1] : a.v[6], f1x = nx ? a.v[6] : b.v[1];
                        const float n1y = ny ? b.v[2] : a.v[7], f1y = ny ? a.v[7] : b.v[2];
                        const float n1z = nz ? b.v[3] : b.v[0], f1z = nz ? b.v[0] : b.v[3];
                        RayConst rc = ToRayConst(ry);
                        float t0, t1;
                        const bool ok0 = SlabStatic(rc, n0x, f0x, n0y, f0y, n0z, f0z, &t0);
                        const bool ok1 = SlabStatic(rc, n1x, f1x, n1y, f1y, n1z, f1z, &t1);
                        stNear = secondFirst ? ok1 : ok0;
                        stFar = secondFirst ? ok0 : ok1;
                        tNear = secondFirst ? t1 : t0;
                        tFar = secondFirst ? t0 : t1;
                    } else {
                        // Both planes of every slab, then min / max: with lo <= hi, finite non-zero invDir and
                        // a finite origin (the fast path's preconditions) rounding is monotone, so the near
                        // plane's distance -- bounds[dirIsNeg] in the reference -- IS the smaller of the two,
                        // whatever the sign of invDir; and the far distances are padded after the min3
                        // (min_k fl(t_k * c) = fl((min_k t_k) * c) for c > 0): one multiply per box, not three.
                        const float pad = SlabPad();
                        const float lx0 = Mul(Sub(a.v[0], ry.ox), ry.ix), hx0 = Mul(Sub(a.v[3], ry.ox), ry.ix);
                        const float ly0 = Mul(Sub(a.v[1], ry.oy), ry.iy), hy0 = Mul(Sub(a.v[4], ry.oy), ry.iy);
                        const float lz0 = Mul(Sub(a.v[2], ry.oz), ry.iz), hz0 = Mul(Sub(a.v[5], ry.oz), ry.iz);
                        const float lx1 = Mul(Sub(a.v[6], ry.ox), ry.ix), hx1 = Mul(Sub(b.v[1], ry.ox), ry.ix);
                        const float ly1 = Mul(Sub(a.v[7], ry.oy), ry.iy), hy1 = Mul(Sub(b.v[2], ry.oy), ry.iy);
                        const float lz1 = Mul(Sub(b.v[0], ry.oz), ry.iz), hz1 = Mul(Sub(b.v[3], ry.oz), ry.iz);
                        const float t0 = fmaxf(fmaxf(fminf(lx0, hx0), fminf(ly0, hy0)), fminf(lz0, hz0));
                        const float t1 = fmaxf(fmaxf(fminf(lx1, hx1), fminf(ly1, hy1)), fminf(lz1, hz1));
                        const float x0 = Mul(fminf(fminf(fmaxf(lx0, hx0), fmaxf(ly0, hy0)), fmaxf(lz0, hz0)), pad);
                        const float x1 = Mul(fminf(fminf(fmaxf(lx1, hx1), fmaxf(ly1, hy1)), fmaxf(lz1, hz1)), pad);
                        // order the VALUES, then compare: the predicates never live in registers
                        tNear = secondFirst ? t1 : t0;
                        tFar = secondFirst ? t0 : t1;
                        const float xNear = secondFirst ? x1 : x0, xFar = secondFirst ? x0 : x1;
                        stNear = (tNear <= xNear) && (xNear > 0.f);
                        stFar = (tFar <= xFar) && (xFar > 0.f);
                    }
                    const bool goNear = stNear && (tNear < tMax);
                    const bool goFar = stFar && (tFar < tMax);
                    // The far child is stacked whenever its tMax-independent test passes: tMax
                    // can GROW by an ulp before the pop (shapes.cpp:239-247), see DESIGN.md.
                    const bool push = goNear && (ANY ? goFar : stFar);
                    if (CACHE) {
                        if (push) {
                            if (topRef != kNone) {
                                stack[sp] = ((unsigned long long)__float_as_uint(topT) << 32) | (unsigned)topRef;
                                ++sp;
                            }
                            topRef = refFar;
                            topT = tFar;
                        }
                    } else if (push) {
                        if (ANY) stack[sp] = (unsigned)refFar;
                        else stack[sp] = ((unsigned long long)__float_as_uint(tFar) << 32) | (unsigned)refFar;
                        ++sp;
                    }
                    // near child if it passes (tNear < tMax already established: next step passes), else the
                    // far child if it does (aggregates.cpp:560-569); with neither, the values are dead
                    cur = goNear ? refNear : (goFar ? refFar : cur);
                    descend = goNear || goFar;
                }
                if (INST && !descend) {
                    popInst();
                } else {
                    popOne(!descend);
                }
#if NNBVH_PREFETCH_TRI
                // the lane has just stepped onto a leaf: it will wait there for the quorum, and its first triangle
                // can travel to the L1 meanwhile
                if (cur < kDead) {
                    const float4 *tp = tris + 3 * (size_t)((uint32_t)cur & 0x7fffffffu);
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(tp));
#if NNBVH_PREFETCH_TRI > 1
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(tp + 2));
#endif
                }
#endif
            }
            }
        }

        // ---- phase B: every parked lane tests ONE triangle of its leaf ----
        if (STATS) {
            const unsigned parked = __ballot_sync(0xffffffffu, cur < kDead);
            const unsigned testing = parked;
            st[4] += 1;
            st[5] += __popc(parked);
            st[6] += __popc(testing);
        }
        bool doSphere = false;
        if (SPHERES) {
            const unsigned sphMask = __ballot_sync(0xffffffffu, sphWait);
            if (sphMask) {
                const unsigned other = __ballot_sync(0xffffffffu, cur != kDead && !sphWait);
                doSphere = (unsigned)__popc(sphMask) >= p.sphereQuorum || other == 0;
            }
        }
        if (cur < kDead && (!(SPHERES && sphWait) || doSphere)) {
            bool needPop = true;
            {  // the leaf node's own `tMin < raytMax` was settled when the lane stepped onto it (:547-555)
                const uint32_t prim = (uint32_t)cur & 0x7fffffffu;
                const float4 *tp = tris + 3 * (size_t)prim;
#if NNBVH_L1_NOALLOC >= 1
                const float4 v0 = LdgNoAllocate(tp), v1 = LdgNoAllocate(tp + 1), v2 = LdgNoAllocate(tp + 2);
#else
                const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
#endif
                const uint32_t flags = __float_as_uint(v0.w);
                bool found = false;
                if (INST && (flags & kPrimInstance)) {
                    // TransformedPrimitive::Intersect / IntersectP (cpu/primitive.cpp:112-130): the ray in the
                    // instance's space, the object's aggregate traversed there with the shortened tMax
                    const uint32_t k = __float_as_uint(v0.x);
                    const RayRegs rr = LoadRay(p.in, rayIdx);
                    float o[3], d[3], tLocal;
                    InstanceRay(sc.instances + k, rr.ox, rr.oy, rr.oz, rr.dx, rr.dy, rr.dz, tMax, o, d, &tLocal);
                    const RayConst rc = MakeRayConst(o[0], o[1], o[2], d[0], d[1], d[2]);
                    if (!EXACT && !fastSlabSafe(rc)) {
                        // this instance's ray needs the verbatim slab cascade: the exact launch traces the ray again
                        const uint32_t pos = atomicAdd(p.nDeferred, 1u);
                        p.deferred[pos] = rayIdx;
                        cur = kDead;
                        needPop = false;
                    } else {
                        const float4 *is = reinterpret_cast<const float4 *>(sc.instStep + k);
                        const float4 b0 = __ldg(is), b1 = __ldg(is + 1);  // rootBox[6], rootRef, primBase
                        float tRoot = 0.f;
                        const bool rootOk = SlabStatic(rc, rc.negx ? b0.w : b0.x, rc.negx ? b0.x : b0.w,
                                                       rc.negy ? b1.x : b0.y, rc.negy ? b0.y : b1.x,
                                                       rc.negz ? b1.y : b0.z, rc.negz ? b0.z : b1.y, &tRoot);
                        if (rootOk && tRoot < tLocal) {
                            const int outerNext = (flags & kTriLast) ? kPopMark : cur + 1;
                            if (ANY) stack[sp] = (unsigned)outerNext;
                            else stack[sp] = ((unsigned long long)__float_as_uint(tMax) << 32) | (unsigned)outerNext;
                            ++sp;
                            spBase = sp;
                            instSlot = (int)prim;
                            instPrimBase = __float_as_uint(b1.w);
                            instHit = false;
                            setRay(rc);
                            tMax = tLocal;
                            cur = __float_as_int(b1.z);
                            needPop = false;
                        } else if (!(flags & kTriLast)) {
                            cur = cur + 1;  // the instance is missed as a whole: next primitive of the leaf
                            needPop = false;
                        }
                    }
                } else if (SPHERES && (flags & kPrimSphere) && !sphWait) {
                    sphWait = true;  // first arrival: wait for company (see above); nothing else moves
                    needPop = false;
                } else if (SPHERES && (flags & kPrimSphere)) {
                    sphWait = false;
                    // Sphere::Intersect / IntersectP (shapes.h:137-145, 234-236).  The lane keeps no
                    // direction (it is folded into invDir and the shear): the ray is read again.
                    const RayRegs rr = LoadRay(p.in, rayIdx);
                    SphereHit sh;
                    if (SphereTest(sc.spheres + __float_as_uint(v0.x), rr.ox, rr.oy, rr.oz, rr.dx, rr.dy,
                                   rr.dz, tMax, &sh)) {
                        found = true;
                        if (!ANY) accept(prim, NNBVH_MISS, sh.t, sh.px, sh.py, sh.pz);
                    }
                } else if (!(flags & kTriDegenerate)) {
                    RayConst rc = ToRayConst(ry);
                    float t, b0, b1, b2;
                    bool got;
                    const float a = (ALPHA && (!INST || instSlot < 0)) ? __ldg(sc.alpha + prim) : 1.f;
                    if (ALPHA && a < 1) {
                        const RayRegs rr = LoadRay(p.in, rayIdx);  // HashFloat(r.o, r.d) and the re-trace want d
                        rc.dx = rr.dx; rc.dy = rr.dy; rc.dz = rr.dz;
                        const float q0[3] = {v0.x, v0.y, v0.z}, q1[3] = {v1.x, v1.y, v1.z}, q2[3] = {v2.x, v2.y, v2.z};
                        got = AlphaTriangleTest(rc, tMax, q0, q1, q2, a, &t, &b0, &b1, &b2);
                    } else {
                        got = TriangleTest(rc, tMax, v0.x, v0.y, v0.z, v1.x, v1.y, v1.z, v2.x, v2.y, v2.z,
                                           &t, &b0, &b1, &b2);
                    }
                    if (got) {
                        found = true;
                        if (ANY) {
                        } else if (INST && instSlot >= 0) {  // inside an instance: its top-level slot + the primitive inside the object
                            accept((uint32_t)instSlot, prim - instPrimBase, t, b0, b1, b2);
                            instHit = true;
                        } else {
                            accept(prim, NNBVH_MISS, t, b0, b1, b2);
                        }
                    }
                }
                if (INST && (flags & kPrimInstance)) {
                    // handled above
                } else if (SPHERES && sphWait) {
                    // parked at the sphere slot: `cur` stays
                } else if (ANY && found) {  // first accepted primitive ends the ray (:598-601)
                    cur = kDead;
                    needPop = false;
                    p.occluded[rayIdx] = 1;
                } else if (!(flags & kTriLast)) {
                    cur = cur + 1;  // next primitive of the same leaf (bit 31 stays set)
                    needPop = false;
                }
            }
            if (INST && needPop) {
                popInst();
            } else {  // :556-558; culled entries are skipped as in phase A
                popOne(needPop);
            }
        }
    }
    if (STATS && lane == 0)
        for (int k = 0; k < 8; ++k) atomicAdd(p.stats + k, st[k]);
}

}  // namespace nnbvh
