// nn_bvh_b200/csrc/internal.h -- shared between the host and device halves of libnnbvh.
#pragma once

#include <stdint.h>

#include "../../include/nnbvh.h"

namespace nnbvh {

// Records a printf-style message for nnbvh_last_error() (thread-local) and
// returns `code`, so call sites read `return SetError(...)`.
int SetError(int code, const char *fmt, ...);

// ---- device layout (private to the library) --------------------------------
//
// WideNode: one per INTERIOR node of the reference tree, holding BOTH children's
// boxes, so one 64-byte fetch (4 x LDG.128, two 32-byte sectors) serves two box
// tests and leaf nodes are never fetched at all.  Children keep the
// reference's identity: child 0 is the node at idx+1, child 1 the node at
// secondChildOffset, `axis` is the parent's split axis (near/far choice,
// cpu/aggregates.cpp:560-569).
//
// ref encoding: >= 0  -> interior child: index of ITS WideNode
//               <  0  -> leaf child: first ordered primitive = ref & 0x7fffffff;
//                        the leaf ends at the first triangle whose flag word has
//                        kTriLast set.
struct alignas(64) WideNode {
    float lo0[3], hi0[3];  // child 0 box
    float lo1[3], hi1[3];  // child 1 box
    int32_t ref0, ref1;
    int32_t axis;
    int32_t axisBit;  // 1 << axis: StepKernel's near/far choice is one AND with the ray's sign bits
};
static_assert(sizeof(WideNode) == 64, "WideNode must be 64 bytes");

// PackedTri: 48 bytes = 3 x float4 {p0.xyz, flags}, {p1.xyz, 0}, {p2.xyz, 0}.
constexpr uint32_t kTriDegenerate = 1u;  // shapes.cpp:176 reject, evaluated at upload
constexpr uint32_t kTriLast = 2u;        // last primitive of its leaf
constexpr uint32_t kPrimSphere = 4u;     // the slot is a sphere: .x of the first float4 = index into the sphere table
constexpr uint32_t kPrimInstance = 8u;   // the slot is a TransformedPrimitive: .x = index into the instance table
constexpr int kMaxStack = 64;            // cpu/aggregates.cpp:538


// ---- HLBVH assembly, shared by the host builder and the device builder (build_host.cpp) ----
struct TreeletRef {
    uint32_t start, n;  // range of the Morton-sorted primitive array
    uint32_t slabAt;    // its nodes are slab[slabAt .. slabAt + nNodes), pre-order, interior
    uint32_t nNodes;    // offsets relative to slabAt
};
// buildUpperSAH (cpu/aggregates.cpp:626-723) over the treelet roots, then the final pre-order
// layout of flattenBVH (:505-522) written straight into nodesOut.  treelets[] is permuted.
int AssembleHLBVH(const nnbvh_node *slab, TreeletRef *treelets, uint32_t nTreelets,
                  nnbvh_node *nodesOut, uint32_t *nNodesOut);

}  // namespace nnbvh
