// nn_bvh_b200/csrc/hlbvh_emit.h -- LBVH treelet emission shared by the host and the device HLBVH
// builders (compiled by g++ in build_host.cpp and by nvcc in hlbvh.cu).
//
// Replaces BVHAggregate::emitLBVH (/root/reference/src/pbrt/cpu/aggregates.cpp:451-503), which is a
// recursion over a pointer tree with a shared atomic leaf counter.  Here one treelet is emitted
// iteratively, straight into pre-order slots of its private node region:
//   * the reference creates a node BEFORE its children (`node = buildNodes++`, :463/:491) and
//     flattenBVH (:505-522) lays nodes out in that same order, so the slot counter IS the
//     final pre-order index inside the treelet;
//   * single-threaded, the reference hands out leaf slots in treelet order and, inside a
//     treelet, left to right: a leaf's firstPrimOffset is simply the position of its first
//     primitive in the Morton-sorted array.  No counter is needed and treelets are independent.
// Interior `offset`s are treelet-relative; the assembler rebases them.
#pragma once

#include <stdint.h>

#include "../../include/nnbvh.h"

#ifdef __CUDACC__
#define NNBVH_HD __host__ __device__ __forceinline__
#else
#define NNBVH_HD inline
#endif

namespace nnbvh {

constexpr uint32_t kTreeletMask = 0x3ffc0000u;  // top 12 of the 30 Morton bits, aggregates.cpp:416
constexpr int kTreeletFirstBit = 29 - 12;       // :436

// util/math.h:99-119
NNBVH_HD uint32_t LeftShift3(uint32_t x) {
    if (x == (1u << 10)) --x;
    x = (x | (x << 16)) & 0x30000ffu;
    x = (x | (x << 8)) & 0x300f00fu;
    x = (x | (x << 4)) & 0x30c30c3u;
    x = (x | (x << 2)) & 0x9249249u;
    return x;
}
NNBVH_HD uint32_t EncodeMorton3(float x, float y, float z) {
    return (LeftShift3((uint32_t)z) << 2) | (LeftShift3((uint32_t)y) << 1) | LeftShift3((uint32_t)x);
}

NNBVH_HD float MinF(float a, float b) { return b < a ? b : a; }  // std::min / std::max semantics
NNBVH_HD float MaxF(float a, float b) { return a < b ? b : a; }

// Emits the treelet over sorted positions [start, start+n) into out[0 .. return value).
// codes / prim: Morton-sorted codes and primitive indices; box6: per-primitive bounds
// (pMin.xyz, pMax.xyz) indexed by primitive.  *overflow is set when a leaf would hold
// >= 65536 primitives (the reference CHECKs that in flattenBVH, :511).
NNBVH_HD uint32_t EmitTreelet(const uint32_t *codes, const uint32_t *prim, const float *box6,
                              uint32_t start, uint32_t n, int maxPrimsInNode, nnbvh_node *out,
                              int *overflow) {
    struct Frame {
        uint32_t a, n, slot, split;
        int bit, stage;
    };
    Frame st[24];  // one frame per Morton bit below the treelet prefix, plus slack
    int sp = 0;
    uint32_t count = 0;
    st[0] = Frame{start, n, 0u, 0u, kTreeletFirstBit, 0};
    while (sp >= 0) {
        Frame &f = st[sp];
        if (f.stage == 0) {
            // :458 leaf test first, then :476-480 skip bits on which the whole range agrees
            while (!(f.bit == -1 || (int)f.n < maxPrimsInNode) &&
                   ((codes[f.a] ^ codes[f.a + f.n - 1]) & (1u << f.bit)) == 0)
                --f.bit;
            f.slot = count++;
            nnbvh_node &nd = out[f.slot];
            if (f.bit == -1 || (int)f.n < maxPrimsInNode) {
                float lo[3], hi[3];
                for (int k = 0; k < 3; ++k) {  // Bounds3f(): max / lowest, vecmath.h:1259-1264
                    lo[k] = 3.402823466e+38f;
                    hi[k] = -3.402823466e+38f;
                }
                for (uint32_t i = 0; i < f.n; ++i) {
                    const float *b = box6 + 6 * (size_t)prim[f.a + i];
                    for (int k = 0; k < 3; ++k) {
                        lo[k] = MinF(lo[k], b[k]);
                        hi[k] = MaxF(hi[k], b[3 + k]);
                    }
                }
                for (int k = 0; k < 3; ++k) {
                    nd.bounds[k] = lo[k];
                    nd.bounds[3 + k] = hi[k];
                }
                nd.offset = (int32_t)f.a;
                nd.nPrimitives = (uint16_t)(f.n < 65535u ? f.n : 65535u);
                nd.axis = 0;
                nd.pad = 0;
                if (f.n >= 65536u) *overflow = 1;
                --sp;
                continue;
            }
            // split point: first position whose bit differs from the first element's (:483-487;
            // FindInterval, util/math.h:507-520, over a monotone predicate)
            const uint32_t mask = 1u << f.bit, first = codes[f.a] & mask;
            uint32_t lo = 0, hi = f.n - 1;  // codes[a+lo] agrees, codes[a+hi] differs
            while (hi - lo > 1) {
                uint32_t mid = (lo + hi) >> 1;
                if ((codes[f.a + mid] & mask) == first) lo = mid;
                else hi = mid;
            }
            f.split = hi;
            f.stage = 1;
            const Frame child{f.a, hi, 0u, 0u, f.bit - 1, 0};
            st[++sp] = child;
            continue;
        }
        nnbvh_node &nd = out[f.slot];
        if (f.stage == 1) {
            // first child done: it occupies slots [slot+1, count); the second starts at `count`
            nd.offset = (int32_t)count;
            f.stage = 2;
            const Frame child{f.a + f.split, f.n - f.split, 0u, 0u, f.bit - 1, 0};
            st[++sp] = child;
            continue;
        }
        // both children done: InitInterior, :113-120
        const nnbvh_node &c0 = out[f.slot + 1], &c1 = out[nd.offset];
        for (int k = 0; k < 3; ++k) {
            nd.bounds[k] = MinF(c0.bounds[k], c1.bounds[k]);
            nd.bounds[3 + k] = MaxF(c0.bounds[3 + k], c1.bounds[3 + k]);
        }
        nd.nPrimitives = 0;
        nd.axis = (uint8_t)(f.bit % 3);
        nd.pad = 0;
        --sp;
    }
    return count;
}

}  // namespace nnbvh
