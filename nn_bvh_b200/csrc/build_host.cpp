// nn_bvh_b200/csrc/build_host.cpp
//
// Host-side tree construction for the engine: nnbvh_build_bvh() produces the
// same flattened tree, byte for byte and in the same leaf order, as the
// reference's BVHAggregate constructor running on one thread
// (/root/reference/src/pbrt/cpu/aggregates.cpp:140-387 buildRecursive,
// :505-522 flattenBVH), for a triangle soup.
//
// Design notes (not a transliteration):
//  * The reference recursion permutes ONE array in place and hands each child
//    a contiguous sub-span, and leaves claim their slots in orderedPrims in DFS
//    order.  So a leaf's firstPrimOffset is simply the start of its span and
//    the ordered primitive list is the final permutation of the array: no
//    shared counter is needed, and disjoint spans can be built by different
//    threads with a bit-identical result (the reference's own parallel build is
//    NOT deterministic, :359-370).
//  * Nodes are emitted straight into pre-order position: a subtree over n
//    primitives gets a private slab of 2n-1 slots, is built there in pre-order,
//    and slabs are compacted afterwards (offsets fixed up) -- no pointer tree.
//  * std::partition / std::nth_element are the same libstdc++ algorithms the
//    reference calls, so the permutations (and therefore leaf order) match.
//
// Float arithmetic here decides the tree shape, so this file must be compiled
// with -ffp-contract=off (the build script does).

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

#include "../../include/nnbvh.h"
#include "internal.h"

namespace nnbvh {
namespace {

struct Aabb {
    float lo[3], hi[3];
    Aabb() {  // Bounds3 default ctor, util/vecmath.h:1259-1264
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::numeric_limits<float>::max();
            hi[k] = std::numeric_limits<float>::lowest();
        }
    }
    void Grow(const Aabb &b) {  // Union(b1,b2), vecmath.h:1487
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::min(lo[k], b.lo[k]);
            hi[k] = std::max(hi[k], b.hi[k]);
        }
    }
    void Grow(const float p[3]) {  // Union(b,p), vecmath.h:1479
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::min(lo[k], p[k]);
            hi[k] = std::max(hi[k], p[k]);
        }
    }
    float Area() const {  // SurfaceArea, vecmath.h:1296-1299
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return 2 * (dx * dy + dx * dz + dy * dz);
    }
    int WidestAxis() const {  // MaxDimension, vecmath.h:1307-1315
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        if (dx > dy && dx > dz) return 0;
        return dy > dz ? 1 : 2;
    }
    float Rel(float p, int k) const {  // Offset()[k], vecmath.h:1321-1330
        float o = p - lo[k];
        if (hi[k] > lo[k]) o /= hi[k] - lo[k];
        return o;
    }
};

struct Item {  // BVHPrimitive, aggregates.cpp:91-100
    size_t index;
    Aabb box;
    float Mid(int k) const { return .5f * box.lo[k] + .5f * box.hi[k]; }
};

constexpr int kBuckets = 12;

inline int BucketOf(const Aabb &cb, const Item &it, int dim) {
    int b = kBuckets * cb.Rel(it.Mid(dim), dim);
    return b == kBuckets ? kBuckets - 1 : b;
}

struct Ctx {
    Item *items;
    nnbvh_node *slab;  // 2*nTris-1 slots; subtree over span [s,s+n) owns slots [2s, 2s+2n-1)
    int maxLeaf;
    int split;
    std::atomic<bool> *overflow;
};

void WriteBox(nnbvh_node &nd, const Aabb &b) {
    for (int k = 0; k < 3; ++k) {
        nd.bounds[k] = b.lo[k];
        nd.bounds[3 + k] = b.hi[k];
    }
}

// Chooses the split of span [s, s+n) exactly as buildRecursive does and
// permutes it; returns mid (0 => make a leaf).
size_t SplitSpan(const Ctx &c, size_t s, size_t n, const Aabb &bounds, int *axisOut) {
    Item *a = c.items + s;
    if (bounds.Area() == 0 || n == 1) return 0;  // :205
    Aabb cb;
    for (size_t i = 0; i < n; ++i) {
        float m[3] = {a[i].Mid(0), a[i].Mid(1), a[i].Mid(2)};
        cb.Grow(m);
    }
    int dim = cb.WidestAxis();
    *axisOut = dim;
    if (cb.hi[dim] == cb.lo[dim]) return 0;  // :225
    auto byMid = [dim](const Item &x, const Item &y) { return x.Mid(dim) < y.Mid(dim); };
    auto median = [&]() {
        std::nth_element(a, a + n / 2, a + n, byMid);
        return n / 2;
    };
    if (c.split == NNBVH_SPLIT_EQUAL_COUNTS) return median();
    if (c.split == NNBVH_SPLIT_MIDDLE) {  // :240-253, falls through to EqualCounts
        float pmid = (cb.lo[dim] + cb.hi[dim]) / 2;
        Item *m = std::partition(a, a + n, [=](const Item &x) { return x.Mid(dim) < pmid; });
        if (m != a && m != a + n) return size_t(m - a);
        return median();
    }
    // SAH, :265-355
    if (n <= 2) return median();
    int count[kBuckets] = {};
    Aabb bb[kBuckets];
    for (size_t i = 0; i < n; ++i) {
        int b = BucketOf(cb, a[i], dim);
        ++count[b];
        bb[b].Grow(a[i].box);
    }
    float cost[kBuckets - 1] = {};
    {
        int below = 0;
        Aabb acc;
        for (int i = 0; i < kBuckets - 1; ++i) {
            acc.Grow(bb[i]);
            below += count[i];
            cost[i] += below * acc.Area();
        }
        int above = 0;
        Aabb acc2;
        for (int i = kBuckets - 1; i >= 1; --i) {
            acc2.Grow(bb[i]);
            above += count[i];
            cost[i - 1] += above * acc2.Area();
        }
    }
    int best = -1;
    float bestCost = std::numeric_limits<float>::infinity();
    for (int i = 0; i < kBuckets - 1; ++i)
        if (cost[i] < bestCost) {
            bestCost = cost[i];
            best = i;
        }
    float leafCost = n;
    bestCost = 1.f / 2.f + bestCost / bounds.Area();
    if (n > (size_t)c.maxLeaf || bestCost < leafCost) {
        Item *m = std::partition(a, a + n, [&](const Item &x) { return BucketOf(cb, x, dim) <= best; });
        return size_t(m - a);
    }
    return 0;
}

// Builds the subtree over span [s, s+n) in pre-order into c.slab starting at
// slot `at`; returns the number of nodes written.  Offsets are slab-absolute.
// `par` = remaining levels at which children may run on a fresh thread.
uint32_t BuildSpan(const Ctx &c, size_t s, size_t n, uint32_t at, int par) {
    nnbvh_node &nd = c.slab[at];
    std::memset(&nd, 0, sizeof(nd));
    Aabb bounds;
    for (size_t i = 0; i < n; ++i) bounds.Grow(c.items[s + i].box);
    int axis = 0;
    size_t mid = SplitSpan(c, s, n, bounds, &axis);
    if (mid == 0) {
        // SplitSpan returns 0 only for "make a leaf": the median split has
        // mid >= 1, Middle guards it, and a SAH partition with an empty side
        // sends the reference itself into unbounded recursion.
        WriteBox(nd, bounds);
        nd.offset = (int32_t)s;
        nd.nPrimitives = (uint16_t)std::min<size_t>(n, 65535);
        if (n >= 65536) c.overflow->store(true);  // CHECK_LT(nPrimitives, 65536), :511
        return 1;
    }
    uint32_t nl, nr;
    if (par > 0 && n > 64 * 1024) {
        // Right child into a private region sized for the worst case, moved
        // down next to the left subtree afterwards.
        uint32_t rightAt = at + 1 + (uint32_t)(2 * mid - 1);
        uint32_t nrT = 0;
        std::thread t([&] { nrT = BuildSpan(c, s + mid, n - mid, rightAt, par - 1); });
        nl = BuildSpan(c, s, mid, at + 1, par - 1);
        t.join();
        nr = nrT;
        uint32_t dst = at + 1 + nl;
        if (dst != rightAt) {
            int32_t shift = (int32_t)(rightAt - dst);
            std::memmove(&c.slab[dst], &c.slab[rightAt], sizeof(nnbvh_node) * nr);
            for (uint32_t i = dst; i < dst + nr; ++i)
                if (c.slab[i].nPrimitives == 0) c.slab[i].offset -= shift;
        }
    } else {
        nl = BuildSpan(c, s, mid, at + 1, 0);
        nr = BuildSpan(c, s + mid, n - mid, at + 1 + nl, 0);
    }
    // InitInterior, :113-120: bounds = Union(c0, c1)
    const nnbvh_node &L = c.slab[at + 1], &R = c.slab[at + 1 + nl];
    for (int k = 0; k < 3; ++k) {
        nd.bounds[k] = std::min(L.bounds[k], R.bounds[k]);
        nd.bounds[3 + k] = std::max(L.bounds[3 + k], R.bounds[3 + k]);
    }
    nd.axis = (uint8_t)axis;
    nd.nPrimitives = 0;
    nd.offset = (int32_t)(at + 1 + nl);
    return 1 + nl + nr;
}

}  // namespace
}  // namespace nnbvh

using namespace nnbvh;

extern "C" int nnbvh_build_bvh(const float *triVerts9, uint32_t nTris, int32_t maxPrimsInNode,
                               int32_t splitMethod, int32_t nThreads, nnbvh_node *nodesOut,
                               uint32_t *nNodesOut, uint32_t *orderOut) {
    if (!triVerts9 || !nodesOut || !nNodesOut || !orderOut || nTris == 0)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_bvh: null argument or empty soup "
                                           "(the reference CHECKs !primitives.empty())");
    if (splitMethod != NNBVH_SPLIT_SAH && splitMethod != NNBVH_SPLIT_MIDDLE &&
        splitMethod != NNBVH_SPLIT_EQUAL_COUNTS)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_bvh: split method %d not supported on "
                                           "the host (HLBVH is built on the device)", splitMethod);
    if (nTris > (1u << 30))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_bvh: too many triangles");
    std::vector<Item> items(nTris);
    for (uint32_t i = 0; i < nTris; ++i) {
        // Triangle::Bounds = Union(Bounds3f(p0,p1), p2), shapes.cpp:294-301
        const float *p = triVerts9 + 9 * (size_t)i;
        Aabb b;
        for (int k = 0; k < 3; ++k) {
            b.lo[k] = std::min(p[k], p[3 + k]);
            b.hi[k] = std::max(p[k], p[3 + k]);
        }
        b.Grow(p + 6);
        items[i].index = i;
        items[i].box = b;
    }
    Ctx c;
    c.items = items.data();
    c.slab = nodesOut;
    c.maxLeaf = std::min(255, (int)maxPrimsInNode);  // aggregates.cpp:142
    c.split = splitMethod;
    std::atomic<bool> overflow{false};
    c.overflow = &overflow;
    int nt = nThreads > 0 ? nThreads : (int)std::thread::hardware_concurrency();
    int par = 0;
    while ((1 << par) < nt) ++par;
    uint32_t total = BuildSpan(c, 0, nTris, 0, par);
    if (overflow.load())
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_bvh: a leaf holds >= 65536 primitives "
                                           "(the reference CHECKs nPrimitives < 65536)");
    for (uint32_t i = 0; i < nTris; ++i) orderOut[i] = (uint32_t)items[i].index;
    *nNodesOut = total;
    return NNBVH_OK;
}

extern "C" int nnbvh_tree_stack_depth(const nnbvh_node *nodes, uint32_t nNodes,
                                      uint32_t *depthOut) {
    if (!depthOut) return SetError(NNBVH_ERR_INVALID, "nnbvh_tree_stack_depth: null argument");
    *depthOut = 0;
    if (nNodes == 0) return NNBVH_OK;
    if (!nodes) return SetError(NNBVH_ERR_INVALID, "nnbvh_tree_stack_depth: null nodes");
    // Every interior node on the current root-to-node path holds exactly one
    // pending sibling on the stack, so the need is the maximum interior depth.
    std::vector<std::pair<uint32_t, uint32_t>> st;
    st.emplace_back(0u, 0u);
    uint32_t worst = 0;
    while (!st.empty()) {
        auto [i, d] = st.back();
        st.pop_back();
        if (i >= nNodes) return SetError(NNBVH_ERR_INVALID, "tree: node index %u out of range", i);
        if (nodes[i].nPrimitives == 0) {
            worst = std::max(worst, d + 1);
            st.emplace_back(i + 1, d + 1);
            st.emplace_back((uint32_t)nodes[i].offset, d + 1);
        }
    }
    *depthOut = worst;
    return NNBVH_OK;
}
