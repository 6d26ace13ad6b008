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
//    NOT deterministic, :359-370).  Large children are forked while threads are
//    spare; a large span's own passes run on a Team (below) with results that
//    equal the serial ones bit for bit.
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
#include "hlbvh_emit.h"

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
constexpr size_t kForkMin = 32 * 1024;  // smallest span worth a thread of its own

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
    std::atomic<int> *spare;  // threads that may still be started (nThreads - 1 at the beginning)
};

void WriteBox(nnbvh_node &nd, const Aabb &b) {
    for (int k = 0; k < 3; ++k) {
        nd.bounds[k] = b.lo[k];
        nd.bounds[3 + k] = b.hi[k];
    }
}

// ---- several threads on ONE large span ----------------------------------------------------------------
// Near the root a span is too large to leave to one thread (and SAH's first splits are often lopsided, so
// forking the children does not help yet).  Its passes are reductions whose serial results can be
// reproduced exactly from per-chunk partial results combined in chunk order: std::min / std::max keep the
// LEFTMOST of equal values (that covers -0 / +0) and never take a NaN, which is associative; counts are
// integers.  The partition is libstdc++'s two-pointer scheme (bits/stl_algo.h __partition, bidirectional):
// with m = the number of elements satisfying the predicate, the i-th failing element of [0, m) from the left
// is exchanged with the i-th satisfying element of [m, n) from the right -- positions that can be listed
// per chunk and exchanged independently.  So the permutation, hence the tree, stays bit-identical.
class Team {
  public:
    Team(const Ctx &c, int want) : c_(c) {
        while (want > 0) {
            int cur = c.spare->load();
            int take = std::min(cur, want);
            if (take <= 0) break;
            if (c.spare->compare_exchange_weak(cur, cur - take)) {
                extra_ = take;
                break;
            }
        }
    }
    ~Team() { Release(); }
    void Release() {
        if (extra_) c_.spare->fetch_add(extra_);
        extra_ = 0;
    }
    int Size() const { return extra_ + 1; }
    template <typename F>
    void Run(F body) const {  // body(t) for t in [0, Size())
        std::vector<std::thread> th;
        for (int t = 1; t <= extra_; ++t) th.emplace_back([&body, t] { body(t); });
        body(0);
        for (std::thread &x : th) x.join();
    }

  private:
    const Ctx &c_;
    int extra_ = 0;
};
constexpr size_t kTeamMin = 128 * 1024;  // smallest span that gets a team

inline void Chunk(size_t n, int t, int T, size_t *b, size_t *e) {
    *b = n * (size_t)t / (size_t)T;
    *e = n * (size_t)(t + 1) / (size_t)T;
}

// Union of the boxes and of the centroids of a[0..n)
void TeamBounds(const Team &team, const Item *a, size_t n, Aabb *bounds, Aabb *cb) {
    const int T = team.Size();
    std::vector<Aabb> pb(T), pc(T);
    team.Run([&](int t) {
        size_t b, e;
        Chunk(n, t, T, &b, &e);
        Aabb bb, cc;
        for (size_t i = b; i < e; ++i) {
            bb.Grow(a[i].box);
            float m[3] = {a[i].Mid(0), a[i].Mid(1), a[i].Mid(2)};
            cc.Grow(m);
        }
        pb[t] = bb, pc[t] = cc;
    });
    for (int t = 0; t < T; ++t) bounds->Grow(pb[t]), cb->Grow(pc[t]);
}

// std::partition(a, a + n, pred) with the same resulting permutation; returns the partition point.
template <typename Pred>
size_t TeamPartition(const Team &team, Item *a, size_t n, Pred pred) {
    const int T = team.Size();
    std::vector<uint8_t> flag(n);
    std::vector<size_t> cnt(T);
    team.Run([&](int t) {
        size_t b, e, k = 0;
        Chunk(n, t, T, &b, &e);
        for (size_t i = b; i < e; ++i) k += (flag[i] = pred(a[i]) ? 1 : 0);
        cnt[t] = k;
    });
    size_t m = 0;
    for (int t = 0; t < T; ++t) m += cnt[t];
    std::vector<std::vector<uint32_t>> fl(T), tr(T);  // failing in [0, m), satisfying in [m, n): ascending
    team.Run([&](int t) {
        size_t b, e;
        Chunk(n, t, T, &b, &e);
        for (size_t i = b; i < std::min(e, m); ++i)
            if (!flag[i]) fl[t].push_back((uint32_t)i);
        for (size_t i = std::max(b, m); i < e; ++i)
            if (flag[i]) tr[t].push_back((uint32_t)i);
    });
    std::vector<size_t> fo(T + 1, 0), to(T + 1, 0);
    for (int t = 0; t < T; ++t) fo[t + 1] = fo[t] + fl[t].size(), to[t + 1] = to[t] + tr[t].size();
    const size_t k = fo[T];  // == to[T]
    std::vector<uint32_t> R(k);
    team.Run([&](int t) { std::copy(tr[t].begin(), tr[t].end(), R.begin() + to[t]); });
    team.Run([&](int t) {  // the i-th failing element from the left <-> the i-th satisfying one from the right
        for (size_t j = 0; j < fl[t].size(); ++j) std::swap(a[fl[t][j]], a[R[k - 1 - (fo[t] + j)]]);
    });
    return m;
}

// Chooses the split of span [s, s+n) exactly as buildRecursive does and
// permutes it; returns mid (0 => make a leaf).  team / cbPre: several threads on this span, with the
// centroid bounds already known.
size_t SplitSpan(const Ctx &c, size_t s, size_t n, const Aabb &bounds, int *axisOut, const Team *team = nullptr,
                 const Aabb *cbPre = nullptr) {
    Item *a = c.items + s;
    if (bounds.Area() == 0 || n == 1) return 0;  // :205
    Aabb cb;
    if (cbPre)
        cb = *cbPre;
    else
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
        auto below = [=](const Item &x) { return x.Mid(dim) < pmid; };
        Item *m = team ? a + TeamPartition(*team, a, n, below) : std::partition(a, a + n, below);
        if (m != a && m != a + n) return size_t(m - a);
        return median();
    }
    // SAH, :265-355
    if (n <= 2) return median();
    int count[kBuckets] = {};
    Aabb bb[kBuckets];
    if (team) {
        const int T = team->Size();
        struct Part {
            int count[kBuckets] = {};
            Aabb bb[kBuckets];
        };
        std::vector<Part> part(T);
        team->Run([&](int t) {
            size_t b0, e0;
            Chunk(n, t, T, &b0, &e0);
            Part p;
            for (size_t i = b0; i < e0; ++i) {
                int b = BucketOf(cb, a[i], dim);
                ++p.count[b];
                p.bb[b].Grow(a[i].box);
            }
            part[t] = p;
        });
        for (int t = 0; t < T; ++t)
            for (int b = 0; b < kBuckets; ++b) count[b] += part[t].count[b], bb[b].Grow(part[t].bb[b]);
    } else
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
        auto inLeft = [&](const Item &x) { return BucketOf(cb, x, dim) <= best; };
        Item *m = team ? a + TeamPartition(*team, a, n, inLeft) : std::partition(a, a + n, inLeft);
        return size_t(m - a);
    }
    return 0;
}

// Builds the subtree over span [s, s+n) in pre-order into c.slab starting at
// slot `at`; returns the number of nodes written.  Offsets are slab-absolute.
// `par`: children of this span may run on a fresh thread (while c.spare lasts).
uint32_t BuildSpan(const Ctx &c, size_t s, size_t n, uint32_t at, bool par) {
    nnbvh_node &nd = c.slab[at];
    std::memset(&nd, 0, sizeof(nd));
    Aabb bounds;
    int axis = 0;
    size_t mid;
    {
        Team team(c, par && n >= kTeamMin ? 15 : 0);  // what is spare right now (the passes are memory-bound)
        if (team.Size() > 1) {
            Aabb cb;
            TeamBounds(team, c.items + s, n, &bounds, &cb);
            mid = SplitSpan(c, s, n, bounds, &axis, &team, &cb);
        } else {
            for (size_t i = 0; i < n; ++i) bounds.Grow(c.items[s + i].box);
            mid = SplitSpan(c, s, n, bounds, &axis);
        }
    }  // the slots go back before the children start
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
    // A fork pays only when BOTH children are large: SAH splits near the root are often lopsided (a few
    // thousand primitives against the rest), and then the large child simply keeps the whole thread budget.
    const bool bigL = mid > kForkMin, bigR = n - mid > kForkMin;
    if (par && bigL && bigR && c.spare->fetch_sub(1) > 0) {
        // Right child into a private region sized for the worst case, moved
        // down next to the left subtree afterwards.
        uint32_t rightAt = at + 1 + (uint32_t)(2 * mid - 1);
        uint32_t nrT = 0;
        std::thread t([&] {
            nrT = BuildSpan(c, s + mid, n - mid, rightAt, true);
            c.spare->fetch_add(1);  // the slot is free again as soon as this subtree is done
        });
        nl = BuildSpan(c, s, mid, at + 1, true);
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
        if (par && bigL && bigR) c.spare->fetch_add(1);  // no slot was free: undo the probe
        nl = BuildSpan(c, s, mid, at + 1, par && bigL);
        nr = BuildSpan(c, s + mid, n - mid, at + 1 + nl, par && bigR);
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


// ---------------------------------------------------------------------------
// HLBVH (aggregates.cpp:389-503, 626-723).  Morton codes -> stable sort -> independent treelets
// (hlbvh_emit.h) -> upper SAH over the treelet roots, emitted straight into pre-order.
// ---------------------------------------------------------------------------
namespace nnbvh {
namespace {

struct UpperCtx {
    const nnbvh_node *slab;
    TreeletRef *tr;
    nnbvh_node *out;
    int error;  // 1: the reference's CHECK_NE(centroidBounds...) / CHECK_GT(mid, start) would fire
};

Aabb BoxOf(const nnbvh_node &nd) {
    Aabb b;
    for (int k = 0; k < 3; ++k) {
        b.lo[k] = nd.bounds[k];
        b.hi[k] = nd.bounds[3 + k];
    }
    return b;
}

// Writes the subtree over treelets [s, e) at out[at...]; returns the number of nodes written.
uint32_t EmitUpper(UpperCtx &c, uint32_t s, uint32_t e, uint32_t at) {
    if (c.error) return 0;
    if (e - s == 1) {  // :633-634: the treelet itself; rebase its interior offsets
        const TreeletRef &t = c.tr[s];
        std::memcpy(c.out + at, c.slab + t.slabAt, sizeof(nnbvh_node) * (size_t)t.nNodes);
        for (uint32_t i = at; i < at + t.nNodes; ++i)
            if (c.out[i].nPrimitives == 0) c.out[i].offset += (int32_t)at;
        return t.nNodes;
    }
    Aabb bounds, cb;
    for (uint32_t i = s; i < e; ++i) {
        Aabb b = BoxOf(c.slab[c.tr[i].slabAt]);
        bounds.Grow(b);
        float m[3];
        for (int k = 0; k < 3; ++k) m[k] = (b.lo[k] + b.hi[k]) * 0.5f;  // :645-646
        cb.Grow(m);
    }
    const int dim = cb.WidestAxis();
    if (cb.hi[dim] == cb.lo[dim]) {  // CHECK_NE, :652
        c.error = 1;
        return 0;
    }
    auto bucketOf = [&](const TreeletRef &t) {
        const nnbvh_node &r = c.slab[t.slabAt];
        float centroid = (r.bounds[dim] + r.bounds[3 + dim]) * 0.5f;
        int b = kBuckets * ((centroid - cb.lo[dim]) / (cb.hi[dim] - cb.lo[dim]));
        return b == kBuckets ? kBuckets - 1 : b;
    };
    int count[kBuckets] = {};
    Aabb bb[kBuckets];
    for (uint32_t i = s; i < e; ++i) {
        int b = bucketOf(c.tr[i]);
        ++count[b];
        bb[b].Grow(BoxOf(c.slab[c.tr[i].slabAt]));
    }
    float best = 0;
    int bestBucket = -1;
    for (int i = 0; i < kBuckets - 1; ++i) {  // :679-692, O(buckets^2) there; same sums here
        Aabb b0, b1;
        int c0 = 0, c1 = 0;
        for (int j = 0; j <= i; ++j) {
            b0.Grow(bb[j]);
            c0 += count[j];
        }
        for (int j = i + 1; j < kBuckets; ++j) {
            b1.Grow(bb[j]);
            c1 += count[j];
        }
        float cost = .125f + (c0 * b0.Area() + c1 * b1.Area()) / bounds.Area();
        if (bestBucket < 0 || cost < best) {
            best = cost;
            bestBucket = i;
        }
    }
    TreeletRef *m = std::partition(c.tr + s, c.tr + e,
                                   [&](const TreeletRef &t) { return bucketOf(t) <= bestBucket; });
    const uint32_t mid = (uint32_t)(m - c.tr);
    if (!(mid > s && mid < e)) {  // CHECK_GT / CHECK_LT, :716-717
        c.error = 1;
        return 0;
    }
    const uint32_t nl = EmitUpper(c, s, mid, at + 1);
    const uint32_t nr = EmitUpper(c, mid, e, at + 1 + nl);
    if (c.error) return 0;
    nnbvh_node &nd = c.out[at];
    std::memset(&nd, 0, sizeof(nd));
    const nnbvh_node &L = c.out[at + 1], &R = c.out[at + 1 + nl];
    for (int k = 0; k < 3; ++k) {
        nd.bounds[k] = std::min(L.bounds[k], R.bounds[k]);
        nd.bounds[3 + k] = std::max(L.bounds[3 + k], R.bounds[3 + k]);
    }
    nd.axis = (uint8_t)dim;
    nd.offset = (int32_t)(at + 1 + nl);
    return 1 + nl + nr;
}

}  // namespace

int AssembleHLBVH(const nnbvh_node *slab, TreeletRef *treelets, uint32_t nTreelets,
                  nnbvh_node *nodesOut, uint32_t *nNodesOut) {
    UpperCtx c{slab, treelets, nodesOut, 0};
    uint32_t total = EmitUpper(c, 0, nTreelets, 0);
    if (c.error)
        return SetError(NNBVH_ERR_INVALID, "hlbvh: treelet roots cannot be split (the reference "
                                           "CHECKs centroidBounds / mid in buildUpperSAH, "
                                           "cpu/aggregates.cpp:652,716)");
    *nNodesOut = total;
    return NNBVH_OK;
}

namespace {

// Triangle::Bounds = Union(Bounds3f(p0,p1), p2), shapes.cpp:294-301
void TriBox(const float *p, float *b6) {
    for (int k = 0; k < 3; ++k) {
        b6[k] = std::min(std::min(p[k], p[3 + k]), p[6 + k]);
        b6[3 + k] = std::max(std::max(p[k], p[3 + k]), p[6 + k]);
    }
}

// box: 6 floats per primitive (Primitive::Bounds()), original order
int BuildHLBVHHost(const std::vector<float> &box, uint32_t nTris, int maxLeaf, int nThreads,
                   nnbvh_node *nodesOut, uint32_t *nNodesOut, uint32_t *orderOut) {
    Aabb cb;
    for (uint32_t i = 0; i < nTris; ++i) {
        const float *b = &box[6 * (size_t)i];
        float m[3] = {.5f * b[0] + .5f * b[3], .5f * b[1] + .5f * b[4], .5f * b[2] + .5f * b[5]};
        cb.Grow(m);
    }
    std::vector<uint32_t> code(nTris), prim(nTris), code2(nTris), prim2(nTris);
    for (uint32_t i = 0; i < nTris; ++i) {
        const float *b = &box[6 * (size_t)i];
        float o[3];
        for (int k = 0; k < 3; ++k) o[k] = cb.Rel(.5f * b[k] + .5f * b[3 + k], k) * 1024;  // :404-408
        code[i] = EncodeMorton3(o[0], o[1], o[2]);
        prim[i] = i;
    }
    // stable LSD radix sort on the 30 code bits, 3 passes of 10 (the reference does 5 of 6, :41-82;
    // every stable sort by the same key yields the same permutation)
    for (int pass = 0; pass < 3; ++pass) {
        const int shift = 10 * pass;
        std::vector<uint32_t> cnt(1025, 0);
        for (uint32_t i = 0; i < nTris; ++i) ++cnt[((code[i] >> shift) & 1023u) + 1];
        for (int d = 0; d < 1024; ++d) cnt[d + 1] += cnt[d];
        for (uint32_t i = 0; i < nTris; ++i) {
            uint32_t at = cnt[(code[i] >> shift) & 1023u]++;
            code2[at] = code[i];
            prim2[at] = prim[i];
        }
        code.swap(code2);
        prim.swap(prim2);
    }
    std::vector<TreeletRef> tr;
    for (uint32_t start = 0, end = 1; end <= nTris; ++end)
        if (end == nTris || ((code[start] ^ code[end]) & kTreeletMask)) {
            tr.push_back(TreeletRef{start, end - start, 2 * start, 0});
            start = end;
        }
    std::vector<nnbvh_node> slab(2 * (size_t)nTris);
    std::atomic<uint32_t> next{0};
    std::atomic<int> overflow{0};
    auto work = [&] {
        for (uint32_t t; (t = next.fetch_add(1)) < tr.size();) {
            int ov = 0;
            tr[t].nNodes = EmitTreelet(code.data(), prim.data(), box.data(), tr[t].start, tr[t].n,
                                       maxLeaf, slab.data() + tr[t].slabAt, &ov);
            if (ov) overflow.store(1);
        }
    };
    int nt = std::max(1, std::min(nThreads, 64));
    std::vector<std::thread> pool;
    for (int k = 1; k < nt; ++k) pool.emplace_back(work);
    work();
    for (std::thread &t : pool) t.join();
    if (overflow.load())
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_bvh: a leaf holds >= 65536 primitives "
                                           "(the reference CHECKs nPrimitives < 65536)");
    for (uint32_t i = 0; i < nTris; ++i) orderOut[i] = prim[i];
    return AssembleHLBVH(slab.data(), tr.data(), (uint32_t)tr.size(), nodesOut, nNodesOut);
}

}  // namespace
}  // namespace nnbvh

using namespace nnbvh;

// body(begin, end) over [0, n) on up to nt threads (independent elements: no ordering involved)
template <typename F>
static void ForRanges(size_t n, int nt, F body) {
    nt = (int)std::min<size_t>(std::max(1, std::min(nt, 64)), n / (64 * 1024) + 1);
    std::vector<std::thread> th;
    for (int t = 1; t < nt; ++t) th.emplace_back([=] { body(n * (size_t)t / nt, n * (size_t)(t + 1) / nt); });
    body(0, n / nt);
    for (std::thread &x : th) x.join();
}

static int BuildFromBoxes(const std::vector<float> &box, uint32_t nPrims, int32_t maxPrimsInNode,
                          int32_t splitMethod, int32_t nThreads, nnbvh_node *nodesOut,
                          uint32_t *nNodesOut, uint32_t *orderOut) {
    if (splitMethod < NNBVH_SPLIT_SAH || splitMethod > NNBVH_SPLIT_EQUAL_COUNTS)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_bvh: unknown split method %d", splitMethod);
    if (nPrims > (1u << 30))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_bvh: too many primitives");
    if (splitMethod == NNBVH_SPLIT_HLBVH) {
        int nt = nThreads > 0 ? nThreads : (int)std::thread::hardware_concurrency();
        return BuildHLBVHHost(box, nPrims, std::min(255, (int)maxPrimsInNode), nt, nodesOut,
                              nNodesOut, orderOut);
    }
    const int nt = nThreads > 0 ? nThreads : (int)std::thread::hardware_concurrency();
    std::vector<Item> items(nPrims);
    ForRanges(nPrims, nt, [&](size_t b0, size_t e0) {
        for (size_t i = b0; i < e0; ++i) {
            const float *b6 = &box[6 * i];
            Aabb b;
            for (int k = 0; k < 3; ++k) {
                b.lo[k] = b6[k];
                b.hi[k] = b6[3 + k];
            }
            items[i].index = i;
            items[i].box = b;
        }
    });
    Ctx c;
    c.items = items.data();
    c.slab = nodesOut;
    c.maxLeaf = std::min(255, (int)maxPrimsInNode);  // aggregates.cpp:142
    c.split = splitMethod;
    std::atomic<bool> overflow{false};
    c.overflow = &overflow;
    std::atomic<int> spare{std::max(0, nt - 1)};
    c.spare = &spare;
    uint32_t total = BuildSpan(c, 0, nPrims, 0, nt > 1);
    if (overflow.load())
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_bvh: a leaf holds >= 65536 primitives "
                                           "(the reference CHECKs nPrimitives < 65536)");
    for (uint32_t i = 0; i < nPrims; ++i) orderOut[i] = (uint32_t)items[i].index;
    *nNodesOut = total;
    return NNBVH_OK;
}

extern "C" int nnbvh_build_bvh(const float *triVerts9, uint32_t nTris, int32_t maxPrimsInNode,
                               int32_t splitMethod, int32_t nThreads, nnbvh_node *nodesOut,
                               uint32_t *nNodesOut, uint32_t *orderOut) {
    if (!triVerts9 || !nodesOut || !nNodesOut || !orderOut || nTris == 0)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_bvh: null argument or empty soup "
                                           "(the reference CHECKs !primitives.empty())");
    if (nTris > (1u << 30))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_bvh: too many triangles");
    std::vector<float> box(6 * (size_t)nTris);
    ForRanges(nTris, nThreads > 0 ? nThreads : (int)std::thread::hardware_concurrency(), [&](size_t b0, size_t e0) {
        for (size_t i = b0; i < e0; ++i) TriBox(triVerts9 + 9 * i, &box[6 * i]);
    });
    return BuildFromBoxes(box, nTris, maxPrimsInNode, splitMethod, nThreads, nodesOut, nNodesOut, orderOut);
}

extern "C" int nnbvh_build_bvh_bounds(const float *bounds6, uint32_t nPrims, int32_t maxPrimsInNode,
                                      int32_t splitMethod, int32_t nThreads, nnbvh_node *nodesOut,
                                      uint32_t *nNodesOut, uint32_t *orderOut) {
    if (!bounds6 || !nodesOut || !nNodesOut || !orderOut || nPrims == 0)
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_bvh_bounds: null argument or no primitives "
                                           "(the reference CHECKs !primitives.empty())");
    if (nPrims > (1u << 30))
        return SetError(NNBVH_ERR_INVALID, "nnbvh_build_bvh_bounds: too many primitives");
    std::vector<float> box(bounds6, bounds6 + 6 * (size_t)nPrims);
    return BuildFromBoxes(box, nPrims, maxPrimsInNode, splitMethod, nThreads, nodesOut, nNodesOut, orderOut);
}

// Sphere::Bounds (shapes.cpp:37-40) = renderFromObject(Bounds3f((-r,-r,zMin), (r,r,zMax))):
// Transform::operator()(Bounds3f) unions the eight transformed corners (util/transform.cpp:134-139),
// each through Transform::operator()(Point3f) (util/transform.h:310-319, w == 1 for an affine matrix).
extern "C" int nnbvh_sphere_bounds(const nnbvh_sphere *s, float *bounds6) {
    if (!s || !bounds6) return SetError(NNBVH_ERR_INVALID, "nnbvh_sphere_bounds: NULL argument");
    const float r = s->radius;
    // the constructor's clamp, shapes.h:124-125 (Clamp: util/math.h, `val < low ? low : val > high ? high : val`)
    auto clamp = [](float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); };
    const float zMin = clamp(std::min(s->zMin, s->zMax), -r, r), zMax = clamp(std::max(s->zMin, s->zMax), -r, r);
    const float lo[3] = {-r, -r, zMin}, hi[3] = {r, r, zMax};
    // Bounds3f(p1, p2) takes Min / Max of its arguments (vecmath.h:1268-1270)
    const float bl[3] = {std::min(lo[0], hi[0]), std::min(lo[1], hi[1]), std::min(lo[2], hi[2])};
    const float bh[3] = {std::max(lo[0], hi[0]), std::max(lo[1], hi[1]), std::max(lo[2], hi[2])};
    float mn[3], mx[3];
    for (int k = 0; k < 3; ++k) {  // Bounds3(): pMin = max float, pMax = lowest (vecmath.h:1259-1264)
        mn[k] = std::numeric_limits<float>::max();
        mx[k] = std::numeric_limits<float>::lowest();
    }
    const float *m = s->renderFromObject;
    for (int c = 0; c < 8; ++c) {  // Corner(c), vecmath.h:1284-1288
        const float x = (c & 1) ? bh[0] : bl[0], y = (c & 2) ? bh[1] : bl[1], z = (c & 4) ? bh[2] : bl[2];
        const float q[3] = {m[0] * x + m[1] * y + m[2] * z + m[3], m[4] * x + m[5] * y + m[6] * z + m[7],
                            m[8] * x + m[9] * y + m[10] * z + m[11]};
        for (int k = 0; k < 3; ++k) {
            mn[k] = std::min(mn[k], q[k]);
            mx[k] = std::max(mx[k], q[k]);
        }
    }
    for (int k = 0; k < 3; ++k) bounds6[k] = mn[k], bounds6[3 + k] = mx[k];
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
