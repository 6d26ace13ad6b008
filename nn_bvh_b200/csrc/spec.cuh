// nn_bvh_b200/csrc/spec.cuh -- SpecKernel: StepKernel + one postponed leaf per lane (sm_100a).
//
// profiles/r01b_*: in StepKernel only 14.7 of 32 lanes are at an interior node in an average
// node iteration; most of the others are PARKED at a leaf waiting for the leaf quorum.  Here a
// lane that reaches a leaf postpones it (`pend`), pops its next stack entry and keeps
// traversing; it only blocks when it reaches a second leaf.  Every lane with a postponed leaf
// tests one triangle of it whenever the leaf phase runs.
//
// Exactness (results stay bit-identical to cpu/aggregates.cpp:529-624):
//  * Leaves are tested in the reference's order: `pend` is always the earliest untested leaf,
//    `cur` (if a leaf) the next one, and a lane never passes a second leaf.
//  * While a leaf is postponed the lane decides with a STALE tMax = S; the reference would have
//    used T, the value after that leaf's tests.  T <= S except that an accepted hit can come
//    out a few ulps ABOVE the tMax it was accepted against (acceptance is tScaled <= tMax*det
//    in scaled space, shapes.cpp:239-247).  The lane therefore keeps tLim >= every T it may
//    still see (tLim = S + |S|*2^-18, shrinking with tMax, == tMax when nothing is postponed):
//      - entry distance t <  tMax        : take it; it is re-validated against the then-current
//                                          tMax when the lane steps onto it (child boxes lie
//                                          inside parent boxes -- validated at upload -- so by
//                                          monotonicity of rounding no descendant of a node that
//                                          the reference would have culled can pass that test);
//      - t >= tLim                        : reject; the reference rejects too for every T <= tLim;
//      - tMax <= t < tLim                 : undecidable now -> the lane stalls at this node until
//                                          its postponed leaf is resolved;
//      - an accepted hit with t > tLim    : the assumption broke (only possible with denormal
//                                          products) -> the ray is appended to the deferred list
//                                          and retraced by the non-speculative exact pass.
//  * SPEC=false (the deferred / exact pass, or trees without box containment): a lane with a
//    postponed leaf does not step, so every decision is made with the reference's own tMax.
#pragma once

#include "traverse.cuh"

namespace nnbvh {

constexpr int kCurNoMore = -1;  // stack exhausted, only `pend` left (prim 0x7fffffff cannot exist)
constexpr int kCurDead = -2;    // no ray in this lane
constexpr uint32_t kNoIdx = 0xffffffffu;

__device__ __forceinline__ float SpecLimit(float tMax) {
    return fmaf(fabsf(tMax), 3.814697265625e-06f /* 2^-18 */, tMax);
}

template <bool ANY, bool EXACT, bool SPEC, int MINB>
__global__ void __launch_bounds__(256, MINB) SpecKernel(StepParams p) {
    const SceneDev &sc = p.sc;
    const unsigned lane = threadIdx.x & 31;
    const float4 *__restrict__ tris = sc.tris;
    const unsigned char *__restrict__ wideBase = reinterpret_cast<const unsigned char *>(sc.wide);
    const uint64_t nRays = p.nPtr ? (uint64_t)__ldg(p.nPtr) : p.n;

    uint32_t rayIdx = kNoIdx;  // != kNoIdx while a result is still to be stored
    StepRay ry{};
    float tMax = 0.f, tLim = 0.f;
    HitRegs hit{NNBVH_MISS, 0.f, 0.f, 0.f, 0.f};  // ANY: hit.prim = 0/1 occluded
    int cur = kCurDead, pend = 0;                 // pend < 0: postponed leaf ref
    float curT = 0.f;
    bool stall = false;
    unsigned long long stack[kMaxStack];  // {tEntry, ref} packed: one STL.64 / LDL.64
    int sp = 0;
    bool exhausted = false;
    const unsigned leafQuorum = p.leafQuorum;

// pop one entry (aggregates.cpp:556-558, 570-574); an empty stack ends the ray unless a leaf is
// still postponed
#define NNBVH_POP()                                                     \
    do {                                                                \
        if (sp == 0) cur = (pend < 0) ? kCurNoMore : kCurDead;          \
        else {                                                          \
            unsigned long long e_ = stack[--sp];                        \
            cur = (int)(unsigned)e_;                                    \
            curT = __uint_as_float((unsigned)(e_ >> 32));               \
        }                                                               \
    } while (0)

    while (true) {
        // ---- results of finished rays (kept out of the hot loops) ----
        if (cur == kCurDead && rayIdx != kNoIdx) {
            if (ANY) p.occluded[rayIdx] = (uint8_t)hit.prim;
            else StoreHit(p.hits, rayIdx, hit);
            rayIdx = kNoIdx;
        }
        // ---- refill: finished lanes take the next rays from the global cursor ----
        unsigned liveMask = __ballot_sync(0xffffffffu, cur != kCurDead);
        if (!exhausted && (unsigned)__popc(liveMask) < p.refillBelow) {
            unsigned want = ~liveMask;
            unsigned nWant = __popc(want);
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(p.cursor, (unsigned long long)nWant);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base + nWant >= nRays) exhausted = true;
            if (cur == kCurDead) {
                uint64_t j = base + __popc(want & ((1u << lane) - 1));
                if (j < nRays) {
                    uint32_t idx = p.perm ? __ldg(p.perm + j) : (uint32_t)j;
                    RayRegs rr = LoadRay(p.in, idx);
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
                        ry.ox = rc.ox; ry.oy = rc.oy; ry.oz = rc.oz;
                        ry.ix = rc.ix; ry.iy = rc.iy; ry.iz = rc.iz;
                        ry.Sx = rc.Sx; ry.Sy = rc.Sy; ry.Sz = rc.Sz;
                        ry.kz = rc.kz;
                        ry.neg = (unsigned)rc.negx | ((unsigned)rc.negy << 1) | ((unsigned)rc.negz << 2);
                        tMax = rr.tMax;
                        tLim = tMax;
                        hit.prim = ANY ? 0u : NNBVH_MISS;
                        hit.t = hit.b0 = hit.b1 = hit.b2 = 0.f;
                        sp = 0;
                        pend = 0;
                        stall = false;
                        rayIdx = idx;
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
                            curT = tRoot;
                        } else {  // miss: stored at the top of the next outer iteration
                            cur = kCurDead;
                        }
                    }
                }
            }
            // rays that missed the root (or were deferred) are not live
            if (cur == kCurDead && rayIdx != kNoIdx) {
                if (ANY) p.occluded[rayIdx] = (uint8_t)hit.prim;
                else StoreHit(p.hits, rayIdx, hit);
                rayIdx = kNoIdx;
            }
            liveMask = __ballot_sync(0xffffffffu, cur != kCurDead);
        }
        if (liveMask == 0) {
            if (exhausted) break;
            continue;
        }
#include "spec_body.inl"
    }
#undef NNBVH_POP
}

}  // namespace nnbvh
