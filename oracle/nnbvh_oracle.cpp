// oracle/nnbvh_oracle.cpp -- TEST INFRASTRUCTURE (the checker), not product code.
//
// CPU restatement of the reference's hot path, operation for operation in
// IEEE binary32 (build with -ffp-contract=off; fused multiply-adds only where
// the reference source says FMA()).  Each function cites the reference
// file:line it follows (paths relative to /root/reference/src/pbrt).
//
// Parity status: PINNED against the unmodified reference built into
// oracle/_ref/libnnbvh_ref.so (tests/test_oracle.py::test_oracle_vs_reference_live) and against the
// golden vectors under tests/golden/ generated from that library.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may use this file.

#include "nnbvh_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#include <atomic>
#include <thread>

namespace {

// ---- util/float.h:43,195-197 ------------------------------------------------
constexpr float kMachineEpsilon = std::numeric_limits<float>::epsilon() * 0.5;
constexpr float Gamma(int n) {
    return (n * kMachineEpsilon) / (1 - n * kMachineEpsilon);
}
constexpr float kInfinity = std::numeric_limits<float>::infinity();

// ---- util/math.h:569-575 (float instantiation; FMA = std::fma, float.h:100) --
inline float DifferenceOfProducts(float a, float b, float c, float d) {
    float cd = c * d;
    float differenceOfProducts = std::fma(a, b, -cd);
    float error = std::fma(-c, d, cd);
    return differenceOfProducts + error;
}

struct V3 {
    float x, y, z;
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};

// util/vecmath.h:447-451 -- std::max over an initializer list (first largest).
inline float MaxComponentValue(float a, float b, float c) {
    return std::max({a, b, c});
}

// ---- util/vecmath.h:1254-1336 -----------------------------------------------
struct Box {
    float mn[3], mx[3];
    Box() {
        for (int i = 0; i < 3; ++i) {
            mn[i] = std::numeric_limits<float>::max();
            mx[i] = std::numeric_limits<float>::lowest();
        }
    }
    float SurfaceArea() const {
        float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
        return 2 * (dx * dy + dx * dz + dy * dz);
    }
    int MaxDimension() const {
        float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
        if (dx > dy && dx > dz)
            return 0;
        else if (dy > dz)
            return 1;
        else
            return 2;
    }
    // Bounds3::Offset (vecmath.h:1321-1330), one component.
    float Offset(const float p[3], int dim) const {
        float o = p[dim] - mn[dim];
        if (mx[dim] > mn[dim])
            o /= mx[dim] - mn[dim];
        return o;
    }
};
inline Box Union(const Box &a, const Box &b) {  // vecmath.h:1487-1492
    Box r;
    for (int i = 0; i < 3; ++i) {
        r.mn[i] = std::min(a.mn[i], b.mn[i]);
        r.mx[i] = std::max(a.mx[i], b.mx[i]);
    }
    return r;
}
inline Box Union(const Box &a, const float p[3]) {  // vecmath.h:1479-1484
    Box r;
    for (int i = 0; i < 3; ++i) {
        r.mn[i] = std::min(a.mn[i], p[i]);
        r.mx[i] = std::max(a.mx[i], p[i]);
    }
    return r;
}

// ---- cpu/aggregates.cpp:91-100 ----------------------------------------------
struct BuildPrim {
    size_t primitiveIndex;
    Box bounds;
    void Centroid(float c[3]) const {
        for (int i = 0; i < 3; ++i)
            c[i] = .5f * bounds.mn[i] + .5f * bounds.mx[i];
    }
    float Centroid(int dim) const { return .5f * bounds.mn[dim] + .5f * bounds.mx[dim]; }
};

struct BuildNode {  // cpu/aggregates.cpp:103-126
    Box bounds;
    BuildNode *children[2];
    int splitAxis, firstPrimOffset, nPrimitives;
};

struct Builder {
    int maxPrimsInNode;
    int splitMethod;
    std::vector<BuildNode *> pool;
    int totalNodes = 0;
    int orderedPrimsOffset = 0;
    int32_t *orderOut;

    ~Builder() {
        for (BuildNode *n : pool)
            delete n;
    }

    BuildNode *MakeLeaf(BuildNode *node, BuildPrim *prims, size_t n, const Box &bounds) {
        // aggregates.cpp:207-215 / 227-234 / 342-351
        int firstPrimOffset = orderedPrimsOffset;
        orderedPrimsOffset += (int)n;
        for (size_t i = 0; i < n; ++i)
            orderOut[firstPrimOffset + i] = (int32_t)prims[i].primitiveIndex;
        node->firstPrimOffset = firstPrimOffset;
        node->nPrimitives = (int)n;
        node->bounds = bounds;
        node->children[0] = node->children[1] = nullptr;
        return node;
    }

    // cpu/aggregates.cpp:192-387 (sequential recursion order == the
    // reference's order when it runs on one thread).
    BuildNode *BuildRecursive(BuildPrim *prims, size_t n) {
        BuildNode *node = new BuildNode;
        pool.push_back(node);
        ++totalNodes;
        Box bounds;
        for (size_t i = 0; i < n; ++i)
            bounds = Union(bounds, prims[i].bounds);

        if (bounds.SurfaceArea() == 0 || n == 1)
            return MakeLeaf(node, prims, n, bounds);

        Box centroidBounds;
        for (size_t i = 0; i < n; ++i) {
            float c[3];
            prims[i].Centroid(c);
            centroidBounds = Union(centroidBounds, c);
        }
        int dim = centroidBounds.MaxDimension();
        if (centroidBounds.mx[dim] == centroidBounds.mn[dim])
            return MakeLeaf(node, prims, n, bounds);

        size_t mid = n / 2;
        auto equalCounts = [&]() {
            mid = n / 2;
            std::nth_element(prims, prims + mid, prims + n,
                             [dim](const BuildPrim &a, const BuildPrim &b) {
                                 return a.Centroid(dim) < b.Centroid(dim);
                             });
        };
        switch (splitMethod) {
        case NNBVH_ORACLE_MIDDLE: {  // :240-253
            float pmid = (centroidBounds.mn[dim] + centroidBounds.mx[dim]) / 2;
            BuildPrim *midIter =
                std::partition(prims, prims + n, [dim, pmid](const BuildPrim &pi) {
                    return pi.Centroid(dim) < pmid;
                });
            mid = midIter - prims;
            if (midIter != prims && midIter != prims + n)
                break;
            equalCounts();  // falls through to EqualCounts in the reference
            break;
        }
        case NNBVH_ORACLE_EQUAL_COUNTS:  // :254-264
            equalCounts();
            break;
        case NNBVH_ORACLE_SAH:
        default: {  // :265-355
            if (n <= 2) {
                equalCounts();
            } else {
                constexpr int nBuckets = 12;
                struct Bucket {
                    int count = 0;
                    Box bounds;
                } buckets[nBuckets];
                for (size_t i = 0; i < n; ++i) {
                    float c[3];
                    prims[i].Centroid(c);
                    int b = nBuckets * centroidBounds.Offset(c, dim);
                    if (b == nBuckets)
                        b = nBuckets - 1;
                    buckets[b].count++;
                    buckets[b].bounds = Union(buckets[b].bounds, prims[i].bounds);
                }
                constexpr int nSplits = nBuckets - 1;
                float costs[nSplits] = {};
                int countBelow = 0;
                Box boundBelow;
                for (int i = 0; i < nSplits; ++i) {
                    boundBelow = Union(boundBelow, buckets[i].bounds);
                    countBelow += buckets[i].count;
                    costs[i] += countBelow * boundBelow.SurfaceArea();
                }
                int countAbove = 0;
                Box boundAbove;
                for (int i = nSplits; i >= 1; --i) {
                    boundAbove = Union(boundAbove, buckets[i].bounds);
                    countAbove += buckets[i].count;
                    costs[i - 1] += countAbove * boundAbove.SurfaceArea();
                }
                int minCostSplitBucket = -1;
                float minCost = kInfinity;
                for (int i = 0; i < nSplits; ++i) {
                    if (costs[i] < minCost) {
                        minCost = costs[i];
                        minCostSplitBucket = i;
                    }
                }
                float leafCost = n;
                minCost = 1.f / 2.f + minCost / bounds.SurfaceArea();
                if (n > (size_t)maxPrimsInNode || minCost < leafCost) {
                    BuildPrim *midIter =
                        std::partition(prims, prims + n, [=](const BuildPrim &bp) {
                            float c[3];
                            bp.Centroid(c);
                            int b = nBuckets * centroidBounds.Offset(c, dim);
                            if (b == nBuckets)
                                b = nBuckets - 1;
                            return b <= minCostSplitBucket;
                        });
                    mid = midIter - prims;
                } else {
                    return MakeLeaf(node, prims, n, bounds);
                }
            }
            break;
        }
        }
        BuildNode *c0 = BuildRecursive(prims, mid);
        BuildNode *c1 = BuildRecursive(prims + mid, n - mid);
        // InitInterior, :113-120
        node->children[0] = c0;
        node->children[1] = c1;
        node->bounds = Union(c0->bounds, c1->bounds);
        node->splitAxis = dim;
        node->nPrimitives = 0;
        return node;
    }


    // ---- HLBVH: cpu/aggregates.cpp:29-82 (RadixSort), 389-449 (buildHLBVH), 451-503 (emitLBVH),
    //      626-723 (buildUpperSAH); util/math.h:99-119 (LeftShift3 / EncodeMorton3) ----------
    struct MortonPrim {
        int primitiveIndex;
        uint32_t mortonCode;
    };
    static uint32_t LeftShift3(uint32_t x) {
        if (x == (1 << 10)) --x;
        x = (x | (x << 16)) & 0x30000ffu;
        x = (x | (x << 8)) & 0x300f00fu;
        x = (x | (x << 4)) & 0x30c30c3u;
        x = (x | (x << 2)) & 0x9249249u;
        return x;
    }
    static uint32_t EncodeMorton3(float x, float y, float z) {
        return (LeftShift3((uint32_t)z) << 2) | (LeftShift3((uint32_t)y) << 1) | LeftShift3((uint32_t)x);
    }
    static void RadixSort(std::vector<MortonPrim> *v) {
        std::vector<MortonPrim> temp(v->size());
        const int bitsPerPass = 6, nBits = 30, nPasses = nBits / bitsPerPass;
        for (int pass = 0; pass < nPasses; ++pass) {
            int lowBit = pass * bitsPerPass;
            std::vector<MortonPrim> &in = (pass & 1) ? temp : *v;
            std::vector<MortonPrim> &out = (pass & 1) ? *v : temp;
            const int nBuckets = 1 << bitsPerPass, bitMask = nBuckets - 1;
            int bucketCount[1 << 6] = {0};
            for (const MortonPrim &mp : in) ++bucketCount[(mp.mortonCode >> lowBit) & bitMask];
            int outIndex[1 << 6];
            outIndex[0] = 0;
            for (int i = 1; i < nBuckets; ++i) outIndex[i] = outIndex[i - 1] + bucketCount[i - 1];
            for (const MortonPrim &mp : in) out[outIndex[(mp.mortonCode >> lowBit) & bitMask]++] = mp;
        }
        if (nPasses & 1) std::swap(*v, temp);
    }

    BuildNode *NewNode() {
        BuildNode *n = new BuildNode();
        pool.push_back(n);
        ++totalNodes;
        return n;
    }

    BuildNode *EmitLBVH(const std::vector<BuildPrim> &prims, MortonPrim *mp, int n, int bitIndex) {
        if (bitIndex == -1 || n < maxPrimsInNode) {
            BuildNode *node = NewNode();
            Box bounds;
            int first = orderedPrimsOffset;
            orderedPrimsOffset += n;
            for (int i = 0; i < n; ++i) {
                orderOut[first + i] = mp[i].primitiveIndex;
                bounds = Union(bounds, prims[mp[i].primitiveIndex].bounds);
            }
            node->firstPrimOffset = first;
            node->nPrimitives = n;
            node->bounds = bounds;
            node->children[0] = node->children[1] = nullptr;
            return node;
        }
        uint32_t mask = 1u << bitIndex;
        if ((mp[0].mortonCode & mask) == (mp[n - 1].mortonCode & mask))
            return EmitLBVH(prims, mp, n, bitIndex - 1);
        // FindInterval (util/math.h:518-534): last index in [0, n-2] whose predicate holds
        int size = n - 2, first = 1;
        while (size > 0) {
            int half = size >> 1, middle = first + half;
            bool pred = (mp[0].mortonCode & mask) == (mp[middle].mortonCode & mask);
            first = pred ? middle + 1 : first;
            size = pred ? size - (half + 1) : half;
        }
        int splitOffset = std::min(std::max(first - 1, 0), n - 2) + 1;
        BuildNode *node = NewNode();   // created BEFORE its children (pre-order slot in the treelet)
        BuildNode *c0 = EmitLBVH(prims, mp, splitOffset, bitIndex - 1);
        BuildNode *c1 = EmitLBVH(prims, mp + splitOffset, n - splitOffset, bitIndex - 1);
        node->children[0] = c0;
        node->children[1] = c1;
        node->bounds = Union(c0->bounds, c1->bounds);
        node->splitAxis = bitIndex % 3;
        node->nPrimitives = 0;
        return node;
    }

    static float SurfaceArea(const Box &b) {
        float dx = b.mx[0] - b.mn[0], dy = b.mx[1] - b.mn[1], dz = b.mx[2] - b.mn[2];
        return 2 * (dx * dy + dx * dz + dy * dz);
    }

    BuildNode *BuildUpperSAH(std::vector<BuildNode *> &roots, int start, int end) {
        int nNodes = end - start;
        if (nNodes == 1) return roots[start];
        BuildNode *node = NewNode();
        Box bounds, cb;
        for (int i = start; i < end; ++i) bounds = Union(bounds, roots[i]->bounds);
        for (int i = start; i < end; ++i) {
            float c[3];
            for (int k = 0; k < 3; ++k) c[k] = (roots[i]->bounds.mn[k] + roots[i]->bounds.mx[k]) * 0.5f;
            cb = Union(cb, c);
        }
        float ex = cb.mx[0] - cb.mn[0], ey = cb.mx[1] - cb.mn[1], ez = cb.mx[2] - cb.mn[2];
        int dim = (ex > ey && ex > ez) ? 0 : (ey > ez ? 1 : 2);   // MaxDimension, vecmath.h:1307-1315
        const int nBuckets = 12;
        auto bucketOf = [&](const BuildNode *nd) {
            float centroid = (nd->bounds.mn[dim] + nd->bounds.mx[dim]) * 0.5f;
            int b = nBuckets * ((centroid - cb.mn[dim]) / (cb.mx[dim] - cb.mn[dim]));
            return b == nBuckets ? nBuckets - 1 : b;
        };
        int count[12] = {0};
        Box bb[12];
        for (int i = start; i < end; ++i) {
            int b = bucketOf(roots[i]);
            count[b]++;
            bb[b] = Union(bb[b], roots[i]->bounds);
        }
        float cost[11];
        for (int i = 0; i < nBuckets - 1; ++i) {
            Box b0, b1;
            int c0 = 0, c1 = 0;
            for (int j = 0; j <= i; ++j) { b0 = Union(b0, bb[j]); c0 += count[j]; }
            for (int j = i + 1; j < nBuckets; ++j) { b1 = Union(b1, bb[j]); c1 += count[j]; }
            cost[i] = .125f + (c0 * SurfaceArea(b0) + c1 * SurfaceArea(b1)) / SurfaceArea(bounds);
        }
        float minCost = cost[0];
        int minBucket = 0;
        for (int i = 1; i < nBuckets - 1; ++i)
            if (cost[i] < minCost) { minCost = cost[i]; minBucket = i; }
        BuildNode **pmid = std::partition(&roots[start], &roots[end - 1] + 1,
                                          [&](const BuildNode *nd) { return bucketOf(nd) <= minBucket; });
        int mid = (int)(pmid - &roots[0]);
        BuildNode *c0 = BuildUpperSAH(roots, start, mid);
        BuildNode *c1 = BuildUpperSAH(roots, mid, end);
        node->children[0] = c0;
        node->children[1] = c1;
        node->bounds = Union(c0->bounds, c1->bounds);
        node->splitAxis = dim;
        node->nPrimitives = 0;
        return node;
    }

    BuildNode *BuildHLBVH(const std::vector<BuildPrim> &prims) {
        Box bounds;
        for (const BuildPrim &p : prims) {
            float c[3];
            p.Centroid(c);
            bounds = Union(bounds, c);
        }
        std::vector<MortonPrim> mp(prims.size());
        for (size_t i = 0; i < prims.size(); ++i) {
            float c[3], o[3];
            prims[i].Centroid(c);
            for (int k = 0; k < 3; ++k) {   // Bounds3::Offset, vecmath.h:1323-1330
                o[k] = c[k] - bounds.mn[k];
                if (bounds.mx[k] > bounds.mn[k]) o[k] /= bounds.mx[k] - bounds.mn[k];
                o[k] = o[k] * 1024;
            }
            mp[i].primitiveIndex = (int)prims[i].primitiveIndex;
            mp[i].mortonCode = EncodeMorton3(o[0], o[1], o[2]);
        }
        RadixSort(&mp);
        std::vector<BuildNode *> roots;
        const uint32_t mask = 0x3ffc0000u;   // 0b00111111111111000000000000000000
        for (size_t start = 0, end = 1; end <= mp.size(); ++end) {
            if (end == mp.size() || ((mp[start].mortonCode & mask) != (mp[end].mortonCode & mask))) {
                roots.push_back(EmitLBVH(prims, &mp[start], (int)(end - start), 29 - 12));
                start = end;
            }
        }
        return BuildUpperSAH(roots, 0, (int)roots.size());
    }

    // cpu/aggregates.cpp:505-522
    int Flatten(const BuildNode *node, nnbvh_oracle_node *nodes, int *offset) {
        nnbvh_oracle_node *ln = &nodes[*offset];
        std::memset(ln, 0, sizeof(*ln));
        for (int i = 0; i < 3; ++i) {
            ln->bounds[i] = node->bounds.mn[i];
            ln->bounds[3 + i] = node->bounds.mx[i];
        }
        int nodeOffset = (*offset)++;
        if (node->nPrimitives > 0) {
            ln->offset = node->firstPrimOffset;
            ln->nPrimitives = (uint16_t)node->nPrimitives;
        } else {
            ln->axis = (uint8_t)node->splitAxis;
            ln->nPrimitives = 0;
            Flatten(node->children[0], nodes, offset);
            ln->offset = Flatten(node->children[1], nodes, offset);
        }
        return nodeOffset;
    }
};

// ---- per-ray constants --------------------------------------------------------
struct RayCtx {
    float ox, oy, oz, dx, dy, dz;
    float invx, invy, invz;
    int neg[3];
};

inline RayCtx MakeRayCtx(const float *r) {
    RayCtx c;
    c.ox = r[0];
    c.oy = r[1];
    c.oz = r[2];
    c.dx = r[4];
    c.dy = r[5];
    c.dz = r[6];
    // cpu/aggregates.cpp:534-535
    c.invx = 1 / c.dx;
    c.invy = 1 / c.dy;
    c.invz = 1 / c.dz;
    c.neg[0] = int(c.invx < 0);
    c.neg[1] = int(c.invy < 0);
    c.neg[2] = int(c.invz < 0);
    return c;
}

// ---- util/vecmath.h:1573-1608 -------------------------------------------------
inline bool BoxIntersectP(const float *b, const RayCtx &r, float raytMax) {
    // bounds[i] is pMin for i==0, pMax for i==1 (vecmath.h:1272-1276)
    float tMin = (b[3 * r.neg[0] + 0] - r.ox) * r.invx;
    float tMax = (b[3 * (1 - r.neg[0]) + 0] - r.ox) * r.invx;
    float tyMin = (b[3 * r.neg[1] + 1] - r.oy) * r.invy;
    float tyMax = (b[3 * (1 - r.neg[1]) + 1] - r.oy) * r.invy;
    tMax *= 1 + 2 * Gamma(3);
    tyMax *= 1 + 2 * Gamma(3);
    if (tMin > tyMax || tyMin > tMax)
        return false;
    if (tyMin > tMin)
        tMin = tyMin;
    if (tyMax < tMax)
        tMax = tyMax;
    float tzMin = (b[3 * r.neg[2] + 2] - r.oz) * r.invz;
    float tzMax = (b[3 * (1 - r.neg[2]) + 2] - r.oz) * r.invz;
    tzMax *= 1 + 2 * Gamma(3);
    if (tMin > tzMax || tzMin > tMax)
        return false;
    if (tzMin > tMin)
        tMin = tzMin;
    if (tzMax < tMax)
        tMax = tzMax;
    return (tMin < raytMax) && (tMax > 0);
}

// ---- shapes.cpp:172-273 ----------------------------------------------------------
inline bool TriangleIntersect(const RayCtx &r, float tMax, const float *p, float *tOut,
                              float *b0Out, float *b1Out, float *b2Out) {
    V3 p0{p[0], p[1], p[2]}, p1{p[3], p[4], p[5]}, p2{p[6], p[7], p[8]};
    // :176  LengthSquared(Cross(p2 - p0, p1 - p0)) == 0
    {
        V3 v{p2.x - p0.x, p2.y - p0.y, p2.z - p0.z};
        V3 w{p1.x - p0.x, p1.y - p0.y, p1.z - p0.z};
        float cx = DifferenceOfProducts(v.y, w.z, v.z, w.y);
        float cy = DifferenceOfProducts(v.z, w.x, v.x, w.z);
        float cz = DifferenceOfProducts(v.x, w.y, v.y, w.x);
        if (cx * cx + cy * cy + cz * cz == 0)
            return false;
    }
    // :181-183
    V3 p0t{p0.x - r.ox, p0.y - r.oy, p0.z - r.oz};
    V3 p1t{p1.x - r.ox, p1.y - r.oy, p1.z - r.oz};
    V3 p2t{p2.x - r.ox, p2.y - r.oy, p2.z - r.oz};
    // :186-196
    V3 d0{r.dx, r.dy, r.dz};
    float ax = std::abs(d0.x), ay = std::abs(d0.y), az = std::abs(d0.z);
    int kz = (ax > ay) ? ((ax > az) ? 0 : 2) : ((ay > az) ? 1 : 2);
    int kx = kz + 1;
    if (kx == 3)
        kx = 0;
    int ky = kx + 1;
    if (ky == 3)
        ky = 0;
    V3 d{d0[kx], d0[ky], d0[kz]};
    p0t = V3{p0t[kx], p0t[ky], p0t[kz]};
    p1t = V3{p1t[kx], p1t[ky], p1t[kz]};
    p2t = V3{p2t[kx], p2t[ky], p2t[kz]};
    // :199-207
    float Sx = -d.x / d.z;
    float Sy = -d.y / d.z;
    float Sz = 1 / d.z;
    p0t.x += Sx * p0t.z;
    p0t.y += Sy * p0t.z;
    p1t.x += Sx * p1t.z;
    p1t.y += Sy * p1t.z;
    p2t.x += Sx * p2t.z;
    p2t.y += Sy * p2t.z;
    // :210-212
    float e0 = DifferenceOfProducts(p1t.x, p2t.y, p1t.y, p2t.x);
    float e1 = DifferenceOfProducts(p2t.x, p0t.y, p2t.y, p0t.x);
    float e2 = DifferenceOfProducts(p0t.x, p1t.y, p0t.y, p1t.x);
    // :215-225
    if (e0 == 0.0f || e1 == 0.0f || e2 == 0.0f) {
        double p2txp1ty = (double)p2t.x * (double)p1t.y;
        double p2typ1tx = (double)p2t.y * (double)p1t.x;
        e0 = (float)(p2typ1tx - p2txp1ty);
        double p0txp2ty = (double)p0t.x * (double)p2t.y;
        double p0typ2tx = (double)p0t.y * (double)p2t.x;
        e1 = (float)(p0typ2tx - p0txp2ty);
        double p1txp0ty = (double)p1t.x * (double)p0t.y;
        double p1typ0tx = (double)p1t.y * (double)p0t.x;
        e2 = (float)(p1typ0tx - p1txp0ty);
    }
    // :228-232
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0))
        return false;
    float det = e0 + e1 + e2;
    if (det == 0)
        return false;
    // :235-242
    p0t.z *= Sz;
    p1t.z *= Sz;
    p2t.z *= Sz;
    float tScaled = e0 * p0t.z + e1 * p1t.z + e2 * p2t.z;
    if (det < 0 && (tScaled >= 0 || tScaled < tMax * det))
        return false;
    else if (det > 0 && (tScaled <= 0 || tScaled > tMax * det))
        return false;
    // :245-247
    float invDet = 1 / det;
    float b0 = e0 * invDet, b1 = e1 * invDet, b2 = e2 * invDet;
    float t = tScaled * invDet;
    // :250-269
    float maxZt = MaxComponentValue(std::abs(p0t.z), std::abs(p1t.z), std::abs(p2t.z));
    float deltaZ = Gamma(3) * maxZt;
    float maxXt = MaxComponentValue(std::abs(p0t.x), std::abs(p1t.x), std::abs(p2t.x));
    float maxYt = MaxComponentValue(std::abs(p0t.y), std::abs(p1t.y), std::abs(p2t.y));
    float deltaX = Gamma(5) * (maxXt + maxZt);
    float deltaY = Gamma(5) * (maxYt + maxZt);
    float deltaE = 2 * (Gamma(2) * maxXt * maxYt + deltaY * maxXt + deltaX * maxYt);
    float maxE = MaxComponentValue(std::abs(e0), std::abs(e1), std::abs(e2));
    float deltaT =
        3 * (Gamma(3) * maxE * maxZt + deltaE * maxZt + deltaZ * maxE) * std::abs(invDet);
    if (t <= deltaT)
        return false;
    *tOut = t;
    *b0Out = b0;
    *b1Out = b1;
    *b2Out = b2;
    return true;
}

// ---- Sphere::BasicIntersect, shapes.h:147-232 ---------------------------------------------------
// Interval as the reference's HOST build evaluates it: outward rounding is one float step past
// the round-to-nearest result (util/float.h:199-320, the non-GPU branches).
inline float NextFloatUp(float v) {  // util/float.h:164-177
    if (std::isinf(v) && v > 0.f)
        return v;
    if (v == -0.f)
        v = 0.f;
    uint32_t ui;
    std::memcpy(&ui, &v, 4);
    if (v >= 0)
        ++ui;
    else
        --ui;
    std::memcpy(&v, &ui, 4);
    return v;
}
inline float NextFloatDown(float v) {  // util/float.h:179-193
    if (std::isinf(v) && v < 0.f)
        return v;
    if (v == 0.f)
        v = -0.f;
    uint32_t ui;
    std::memcpy(&ui, &v, 4);
    if (v > 0)
        --ui;
    else
        ++ui;
    std::memcpy(&v, &ui, 4);
    return v;
}
inline float AddRoundUp(float a, float b) { return NextFloatUp(a + b); }
inline float AddRoundDown(float a, float b) { return NextFloatDown(a + b); }
inline float SubRoundUp(float a, float b) { return AddRoundUp(a, -b); }
inline float SubRoundDown(float a, float b) { return AddRoundDown(a, -b); }
inline float MulRoundUp(float a, float b) { return NextFloatUp(a * b); }
inline float MulRoundDown(float a, float b) { return NextFloatDown(a * b); }
inline float DivRoundUp(float a, float b) { return NextFloatUp(a / b); }
inline float DivRoundDown(float a, float b) { return NextFloatDown(a / b); }
inline float SqrtRoundUp(float a) { return NextFloatUp(std::sqrt(a)); }
inline float SqrtRoundDown(float a) { return std::max<float>(0, NextFloatDown(std::sqrt(a))); }

struct Interval {  // util/math.h:819-935
    float low, high;
    Interval() : low(0), high(0) {}
    explicit Interval(float v) : low(v), high(v) {}
    Interval(float l, float h) : low(std::min(l, h)), high(std::max(l, h)) {}
    static Interval FromValueAndError(float v, float err) {
        Interval i;
        if (err == 0)
            i.low = i.high = v;
        else {
            i.low = SubRoundDown(v, err);
            i.high = AddRoundUp(v, err);
        }
        return i;
    }
    float Midpoint() const { return (low + high) / 2; }
    Interval operator+(Interval i) const { return {AddRoundDown(low, i.low), AddRoundUp(high, i.high)}; }
    Interval operator-(Interval i) const { return {SubRoundDown(low, i.high), SubRoundUp(high, i.low)}; }
    Interval operator*(Interval i) const {
        float lp[4] = {MulRoundDown(low, i.low), MulRoundDown(high, i.low), MulRoundDown(low, i.high),
                       MulRoundDown(high, i.high)};
        float hp[4] = {MulRoundUp(low, i.low), MulRoundUp(high, i.low), MulRoundUp(low, i.high),
                       MulRoundUp(high, i.high)};
        return {std::min({lp[0], lp[1], lp[2], lp[3]}), std::max({hp[0], hp[1], hp[2], hp[3]})};
    }
    Interval operator/(Interval i) const {  // :961-973
        if (0 >= i.low && 0 <= i.high)
            return Interval(-kInfinity, kInfinity);
        float lq[4] = {DivRoundDown(low, i.low), DivRoundDown(high, i.low), DivRoundDown(low, i.high),
                       DivRoundDown(high, i.high)};
        float hq[4] = {DivRoundUp(low, i.low), DivRoundUp(high, i.low), DivRoundUp(low, i.high),
                       DivRoundUp(high, i.high)};
        return {std::min({lq[0], lq[1], lq[2], lq[3]}), std::max({hq[0], hq[1], hq[2], hq[3]})};
    }
    bool operator==(Interval i) const { return low == i.low && high == i.high; }
};
inline Interval Sqr(Interval i) {  // :975-982
    float alow = std::abs(i.low), ahigh = std::abs(i.high);
    if (alow > ahigh)
        std::swap(alow, ahigh);
    if (0 >= i.low && 0 <= i.high)
        return Interval(0, MulRoundUp(ahigh, ahigh));
    return Interval(MulRoundDown(alow, alow), MulRoundUp(ahigh, ahigh));
}
inline Interval operator*(float f, Interval i) {  // :995-1000
    if (f > 0)
        return Interval(MulRoundDown(f, i.low), MulRoundUp(f, i.high));
    return Interval(MulRoundDown(f, i.high), MulRoundUp(f, i.low));
}
inline Interval Sqrt(Interval i) { return {SqrtRoundDown(i.low), SqrtRoundUp(i.high)}; }  // :1070-1072

// nnbvh_sphere as 28 floats: renderFromObject[12], objectFromRender[12], radius, zMin, zMax, phiMax.
struct SphereCtx {
    const float *m;  // objectFromRender rows 0..2 (row 3 = 0 0 0 1)
    float radius, zMin, zMax, phiMax;
};
inline float Clamp(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }
inline SphereCtx MakeSphereCtx(const float *s28) {  // Sphere::Sphere, shapes.h:117-128
    SphereCtx c;
    c.m = s28 + 12;
    c.radius = s28[24];
    c.zMin = Clamp(std::min(s28[25], s28[26]), -c.radius, c.radius);
    c.zMax = Clamp(std::max(s28[25], s28[26]), -c.radius, c.radius);
    c.phiMax = (3.14159265358979323846f / 180) * Clamp(s28[27], 0, 360);  // Radians(), util/math.h:261-263
    return c;
}

// phi comes from std::atan2 as in the reference (shapes.h:196-198): this checker links the same libm.
// For phiMax = 360 `phi > phiMax` (shapes.h:203) is never true -- Radians(360) rounds to the float 2*Pi
// and phi <= 2*Pi by construction.
inline bool SphereBasicIntersect(const SphereCtx &sp, const float *ray8, float tMax, float *tOut, float *pObj) {
    const float *m = sp.m;
    const float g3 = Gamma(3);
    Interval oi[3], di[3];
    {   // Transform::operator()(Point3fi) on an exact point, util/transform.h:133-176
        float x = Interval(ray8[0]).Midpoint(), y = Interval(ray8[1]).Midpoint(), z = Interval(ray8[2]).Midpoint();
        for (int k = 0; k < 3; ++k) {
            const float *r = m + 4 * k;
            float v = (r[0] * x + r[1] * y) + (r[2] * z + r[3]);
            float e = g3 * (std::abs(r[0] * x) + std::abs(r[1] * y) + std::abs(r[2] * z) + std::abs(r[3]));
            oi[k] = Interval::FromValueAndError(v, e);
        }
    }
    {   // Transform::operator()(Vector3fi) on an exact vector, util/transform.h:272-306
        float x = Interval(ray8[4]).Midpoint(), y = Interval(ray8[5]).Midpoint(), z = Interval(ray8[6]).Midpoint();
        for (int k = 0; k < 3; ++k) {
            const float *r = m + 4 * k;
            float e = g3 * (std::abs(r[0] * x) + std::abs(r[1] * y) + std::abs(r[2] * z));
            float v = r[0] * x + r[1] * y + r[2] * z;
            di[k] = Interval::FromValueAndError(v, e);
        }
    }
    const Interval radius(sp.radius);
    Interval a = Sqr(di[0]) + Sqr(di[1]) + Sqr(di[2]);
    Interval b = 2 * (di[0] * oi[0] + di[1] * oi[1] + di[2] * oi[2]);
    Interval c = Sqr(oi[0]) + Sqr(oi[1]) + Sqr(oi[2]) - Sqr(radius);
    Interval f = b / (2 * a);
    Interval v[3] = {oi[0] - f * di[0], oi[1] - f * di[1], oi[2] - f * di[2]};
    Interval length = Sqrt(Sqr(v[0]) + Sqr(v[1]) + Sqr(v[2]));
    Interval discrim = 4 * a * (radius + length) * (radius - length);
    if (discrim.low < 0)
        return false;
    Interval rootDiscrim = Sqrt(discrim);
    Interval q;
    if (b.Midpoint() < 0)
        q = -.5f * (b - rootDiscrim);
    else
        q = -.5f * (b + rootDiscrim);
    Interval t0 = q / a, t1 = c / q;
    if (t0.low > t1.low)
        std::swap(t0, t1);
    if (t0.high > tMax || t1.low <= 0)
        return false;
    Interval tShapeHit = t0;
    if (tShapeHit.low <= 0) {
        tShapeHit = t1;
        if (tShapeHit.high > tMax)
            return false;
    }
    auto point = [&](Interval t, float *p) {  // shapes.h:189-195
        float tm = t.Midpoint();
        for (int k = 0; k < 3; ++k)
            p[k] = oi[k].Midpoint() + tm * di[k].Midpoint();
        float d[3] = {p[0] - 0.f, p[1] - 0.f, p[2] - 0.f};
        float s = sp.radius / std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        p[0] *= s, p[1] *= s, p[2] *= s;
        if (p[0] == 0 && p[1] == 0)
            p[0] = 1e-5f * sp.radius;
    };
    auto clipped = [&](const float *p) {
        float phi = std::atan2(p[1], p[0]);
        if (phi < 0)
            phi += 2 * 3.14159265358979323846f;
        return (sp.zMin > -sp.radius && p[2] < sp.zMin) || (sp.zMax < sp.radius && p[2] > sp.zMax) ||
               phi > sp.phiMax;
    };
    float pHit[3];
    point(tShapeHit, pHit);
    if (clipped(pHit)) {
        if (tShapeHit == t1)
            return false;
        if (t1.high > tMax)
            return false;
        tShapeHit = t1;
        point(tShapeHit, pHit);
        if (clipped(pHit))
            return false;
    }
    *tOut = tShapeHit.Midpoint();
    pObj[0] = pHit[0], pObj[1] = pHit[1], pObj[2] = pHit[2];
    return true;
}

// Sphere::Bounds, shapes.cpp:37-40 with util/transform.cpp:134-139 and util/transform.h:310-319
inline void SphereBounds(const float *s28, float *out6) {
    SphereCtx sp = MakeSphereCtx(s28);
    const float *m = s28;  // renderFromObject
    Box bt;                // Bounds3(): pMin = max, pMax = lowest
    const float r = sp.radius;
    float lo[3] = {std::min(-r, r), std::min(-r, r), std::min(sp.zMin, sp.zMax)};
    float hi[3] = {std::max(-r, r), std::max(-r, r), std::max(sp.zMin, sp.zMax)};
    for (int c = 0; c < 8; ++c) {
        float x = (c & 1) ? hi[0] : lo[0], y = (c & 2) ? hi[1] : lo[1], z = (c & 4) ? hi[2] : lo[2];
        float p[3] = {m[0] * x + m[1] * y + m[2] * z + m[3], m[4] * x + m[5] * y + m[6] * z + m[7],
                      m[8] * x + m[9] * y + m[10] * z + m[11]};
        bt = Union(bt, p);
    }
    for (int k = 0; k < 3; ++k)
        out6[k] = bt.mn[k], out6[3 + k] = bt.mx[k];
}

// ---- GeometricPrimitive with an alpha texture (cpu/primitive.cpp:50-84) ------------------------------
// Only constant alpha values are restated (FloatConstantTexture, textures.h:269-283): image and procedural
// textures are outside this path.
// MurmurHash64A, util/hash.h:19-70 (the tail switch is unused here: 24 bytes)
inline uint64_t MurmurHash64A(const unsigned char *key, size_t len, uint64_t seed) {
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    const int r = 47;
    uint64_t h = seed ^ (len * m);
    const unsigned char *end = key + 8 * (len / 8);
    while (key != end) {
        uint64_t k;
        std::memcpy(&k, key, sizeof(uint64_t));
        key += 8;
        k *= m;
        k ^= k >> r;
        k *= m;
        h ^= k;
        h *= m;
    }
    h ^= h >> r;
    h *= m;
    h ^= h >> r;
    return h;
}
// HashFloat(r.o, r.d), util/hash.h:98-112: Point3f and Vector3f copied back to back, seed 0
inline float HashFloatOD(const float *o, const float *d) {
    uint64_t buf[3];
    std::memcpy((char *)buf, o, 12);
    std::memcpy((char *)buf + 12, d, 12);
    return uint32_t(MurmurHash64A((const unsigned char *)buf, 24, 0)) * 0x1p-32f;
}
// The origin of si->intr.SpawnRay(r.d) for a hit (b0,b1,b2) on triangle p (no shading normals):
// Triangle::InteractionFromIntersection's pHit / pError / n (shapes.h:924-941), Point3fi(pHit, pError)
// (util/math.h:829-838), OffsetRayOrigin (ray.h:75-92).  The orientation of n does not matter: Abs(n)
// scales the offset and Dot(w, n) < 0 flips it back.
inline void SpawnOrigin(const float *p, float b0, float b1, float b2, const float *w, float *po) {
    const float *p0 = p, *p1 = p + 3, *p2 = p + 6;
    float dp02[3], dp12[3];
    for (int k = 0; k < 3; ++k)
        dp02[k] = p0[k] - p2[k], dp12[k] = p1[k] - p2[k];
    float n[3] = {DifferenceOfProducts(dp02[1], dp12[2], dp02[2], dp12[1]),
                  DifferenceOfProducts(dp02[2], dp12[0], dp02[0], dp12[2]),
                  DifferenceOfProducts(dp02[0], dp12[1], dp02[1], dp12[0])};
    float len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
    for (int k = 0; k < 3; ++k)
        n[k] = n[k] / len;
    const float g7 = Gamma(7);
    float mid[3], err[3];
    for (int k = 0; k < 3; ++k) {
        float pHit = b0 * p0[k] + b1 * p1[k] + b2 * p2[k];
        float pErr = g7 * (std::abs(b0 * p0[k]) + std::abs(b1 * p1[k]) + std::abs(b2 * p2[k]));
        Interval pi = Interval::FromValueAndError(pHit, pErr);
        mid[k] = pi.Midpoint();           // Point3f(pi)
        err[k] = (pi.high - pi.low) / 2;  // pi.Error()
    }
    auto sumOfProducts = [](float a, float b, float c, float d) {  // util/math.h:577-583
        float cd = c * d;
        float sum = std::fma(a, b, cd);
        float error = std::fma(c, d, -cd);
        return sum + error;
    };
    // Dot(Normal3f, Vector3f), vecmath.h:1057-1069
    float d = std::fma(std::abs(n[0]), err[0], sumOfProducts(std::abs(n[1]), err[1], std::abs(n[2]), err[2]));
    float offset[3] = {d * n[0], d * n[1], d * n[2]};
    if (std::fma(w[0], n[0], sumOfProducts(w[1], n[1], w[2], n[2])) < 0)
        for (int k = 0; k < 3; ++k)
            offset[k] = -offset[k];
    for (int k = 0; k < 3; ++k) {
        po[k] = mid[k] + offset[k];
        if (offset[k] > 0)
            po[k] = NextFloatUp(po[k]);
        else if (offset[k] < 0)
            po[k] = NextFloatDown(po[k]);
    }
}
inline RayCtx MakeRayCtx(const float *o, const float *d) {
    float r8[8] = {o[0], o[1], o[2], 0, d[0], d[1], d[2], 0};
    return MakeRayCtx(r8);
}
// GeometricPrimitive::Intersect over a Triangle with constant alpha a.
inline bool GeometricAlphaIntersect(const float *o, const float *d, float tMax, const float *p, float a,
                                    float *tOut, float *b0, float *b1, float *b2, int depth = 0) {
    RayCtx r = MakeRayCtx(o, d);
    float t, c0, c1, c2;
    if (!TriangleIntersect(r, tMax, p, &t, &c0, &c1, &c2))
        return false;
    if (a < 1) {
        float u = (a <= 0) ? 1.f : HashFloatOD(o, d);
        if (u > a) {  // ignore this intersection and trace a new ray (against this primitive only)
            if (depth > 64)
                return false;
            float on[3], tn;
            SpawnOrigin(p, c0, c1, c2, d, on);
            if (!GeometricAlphaIntersect(on, d, tMax - t, p, a, &tn, b0, b1, b2, depth + 1))
                return false;
            *tOut = tn + t;  // siNext->tHit += si->tHit
            return true;
        }
    }
    *tOut = t, *b0 = c0, *b1 = c1, *b2 = c2;
    return true;
}

// ---- TransformedPrimitive (cpu/primitive.cpp:112-130) ------------------------------------------
// Ray Transform::ApplyInverse(const Ray &r, Float *tMax), util/transform.h:416-430, with
// Point3fi Transform::ApplyInverse(const Point3fi &) on an exact point (util/transform.cpp:263-303) and
// Vector3<T> Transform::ApplyInverse(Vector3<T>) (util/transform.h:398-406).  mInv: rows 0..2 of the
// inverse matrix (row 3 = 0 0 0 1, so wp == 1).
inline void InstanceRay(const float *mInv, const float *ray8, float *local8) {
    const float g3 = Gamma(3);
    float x = Interval(ray8[0]).Midpoint(), y = Interval(ray8[1]).Midpoint(), z = Interval(ray8[2]).Midpoint();
    Interval o[3];
    for (int k = 0; k < 3; ++k) {
        const float *r = mInv + 4 * k;
        float v = (r[0] * x + r[1] * y) + (r[2] * z + r[3]);
        float e = g3 * (std::abs(r[0] * x) + std::abs(r[1] * y) + std::abs(r[2] * z));
        o[k] = Interval::FromValueAndError(v, e);
    }
    float dx = ray8[4], dy = ray8[5], dz = ray8[6], d[3];
    for (int k = 0; k < 3; ++k) {
        const float *r = mInv + 4 * k;
        d[k] = r[0] * dx + r[1] * dy + r[2] * dz;
    }
    float tMax = ray8[3];
    float lengthSquared = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
    if (lengthSquared > 0) {
        float oErr[3] = {(o[0].high - o[0].low) / 2, (o[1].high - o[1].low) / 2, (o[2].high - o[2].low) / 2};
        float dt = (std::abs(d[0]) * oErr[0] + std::abs(d[1]) * oErr[1] + std::abs(d[2]) * oErr[2]) / lengthSquared;
        for (int k = 0; k < 3; ++k)
            o[k] = o[k] + Interval(dt * d[k]);  // Point3fi += Vector3f: Interval += Float (math.h:916-917)
        tMax -= dt;
    }
    local8[0] = o[0].Midpoint(), local8[1] = o[1].Midpoint(), local8[2] = o[2].Midpoint();
    local8[3] = tMax;
    local8[4] = d[0], local8[5] = d[1], local8[6] = d[2];
    local8[7] = ray8[7];
}

// Objects' trees and ordered triangles back to back; instance k = (inst24 + 24k, instObject[k]).
struct InstancedScene {
    const int32_t *instanceOf;  // per ordered top-level primitive: -1 triangle, else instance index
    const float *inst24;        // renderFromPrimitive[12], primitiveFromRender[12]
    const int32_t *instObject;
    const nnbvh_oracle_node *objNodes;
    const int32_t *objNodeBase;  // first node of object j
    const float *objTris;        // ordered, 9 floats each
    const int32_t *objPrimBase;  // first ordered primitive of object j
};

template <bool ANY>
inline bool Traverse(const nnbvh_oracle_node *nodes, const float *tris, const float *ray8,
                     uint32_t *hit5, uint32_t *nodesVisitedOut, uint32_t *triTestsOut,
                     const int32_t *sphereOf = nullptr, const float *spheres28 = nullptr,
                     const InstancedScene *inst = nullptr, uint32_t *innerOut = nullptr,
                     const float *alphaOf = nullptr) {
    RayCtx r = MakeRayCtx(ray8);
    float tMax = ray8[3];
    uint32_t hitPrim = 0xFFFFFFFFu, hitInner = 0xFFFFFFFFu;
    float hitT = 0, hb0 = 0, hb1 = 0, hb2 = 0;
    bool any = false;
    uint32_t nodesVisited = 0, triTests = 0;
    if (nodes) {
        int toVisitOffset = 0, currentNodeIndex = 0;
        int nodesToVisit[64];
        while (true) {
            ++nodesVisited;
            const nnbvh_oracle_node *node = &nodes[currentNodeIndex];
            if (BoxIntersectP(node->bounds, r, tMax)) {
                if (node->nPrimitives > 0) {
                    for (int i = 0; i < node->nPrimitives; ++i) {
                        float t, b0, b1, b2;
                        uint32_t prim = (uint32_t)(node->offset + i);
                        if (inst && inst->instanceOf[prim] >= 0) {
                            // TransformedPrimitive::Intersect / IntersectP: the object's own BVHAggregate is
                            // traversed with the transformed ray; its visits and tests land in the same
                            // counters (one STAT_COUNTER each, aggregates.cpp:27, shapes.cpp:148)
                            const int k = inst->instanceOf[prim], j = inst->instObject[k];
                            float local8[8];
                            float r8[8];
                            std::memcpy(r8, ray8, sizeof r8);
                            r8[3] = tMax;
                            InstanceRay(inst->inst24 + 24 * (size_t)k + 12, r8, local8);
                            uint32_t h5[5], nv = 0, tt = 0;
                            bool got = Traverse<ANY>(inst->objNodes + inst->objNodeBase[j],
                                                     inst->objTris + 9 * (size_t)inst->objPrimBase[j], local8, h5,
                                                     &nv, &tt);
                            nodesVisited += nv;
                            triTests += tt;
                            if (got) {
                                if (ANY) {
                                    any = true;
                                    goto done;
                                }
                                hitPrim = prim;
                                hitInner = h5[0];
                                std::memcpy(&hitT, &h5[1], 4);
                                std::memcpy(&hb0, &h5[2], 4);
                                std::memcpy(&hb1, &h5[3], 4);
                                std::memcpy(&hb2, &h5[4], 4);
                                tMax = hitT;
                            }
                            continue;
                        }
                        if (sphereOf && sphereOf[prim] >= 0) {
                            // Sphere::Intersect / IntersectP (shapes.h:137-145, 234-236): no nTriTests
                            float pObj[3];
                            if (SphereBasicIntersect(MakeSphereCtx(spheres28 + 28 * (size_t)sphereOf[prim]),
                                                     ray8, tMax, &t, pObj)) {
                                if (ANY) {
                                    any = true;
                                    goto done;
                                }
                                hitPrim = prim;
                                hitT = t;
                                hb0 = pObj[0];
                                hb1 = pObj[1];
                                hb2 = pObj[2];
                                tMax = t;
                            }
                            continue;
                        }
                        ++triTests;
                        if (alphaOf && alphaOf[prim] < 1
                                ? GeometricAlphaIntersect(ray8, ray8 + 4, tMax, tris + 9 * (size_t)prim, alphaOf[prim],
                                                          &t, &b0, &b1, &b2)
                                : TriangleIntersect(r, tMax, tris + 9 * (size_t)prim, &t, &b0, &b1, &b2)) {
                            if (ANY) {
                                any = true;
                                goto done;
                            }
                            hitPrim = prim;
                            hitInner = 0xFFFFFFFFu;
                            hitT = t;
                            hb0 = b0;
                            hb1 = b1;
                            hb2 = b2;
                            tMax = t;
                        }
                    }
                    if (toVisitOffset == 0)
                        break;
                    currentNodeIndex = nodesToVisit[--toVisitOffset];
                } else {
                    if (r.neg[node->axis]) {
                        nodesToVisit[toVisitOffset++] = currentNodeIndex + 1;
                        currentNodeIndex = node->offset;
                    } else {
                        nodesToVisit[toVisitOffset++] = node->offset;
                        currentNodeIndex = currentNodeIndex + 1;
                    }
                }
            } else {
                if (toVisitOffset == 0)
                    break;
                currentNodeIndex = nodesToVisit[--toVisitOffset];
            }
        }
    }
done:
    if (nodesVisitedOut)
        *nodesVisitedOut = nodesVisited;
    if (triTestsOut)
        *triTestsOut = triTests;
    if (ANY)
        return any;
    if (innerOut)
        *innerOut = hitInner;
    hit5[0] = hitPrim;
    std::memcpy(&hit5[1], &hitT, 4);
    std::memcpy(&hit5[2], &hb0, 4);
    std::memcpy(&hit5[3], &hb1, 4);
    std::memcpy(&hit5[4], &hb2, 4);
    return hitPrim != 0xFFFFFFFFu;
}

// Dynamic chunked loop over [0,n) on nThreads std::threads (<=0: all cores).
template <typename F>
void ParallelRays(int64_t n, int nThreads, F body) {
    int nt = nThreads > 0 ? nThreads : (int)std::thread::hardware_concurrency();
    if (nt < 1)
        nt = 1;
    const int64_t chunk = 4096;
    if (nt == 1 || n <= chunk) {
        for (int64_t i = 0; i < n; ++i)
            body(i);
        return;
    }
    std::atomic<int64_t> next{0};
    auto worker = [&]() {
        for (;;) {
            int64_t b = next.fetch_add(chunk);
            if (b >= n)
                return;
            int64_t e = std::min(n, b + chunk);
            for (int64_t i = b; i < e; ++i)
                body(i);
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nt; ++t)
        th.emplace_back(worker);
    worker();
    for (auto &t : th)
        t.join();
}

}  // namespace

extern "C" {

int64_t nnbvh_oracle_build_bounds(const float *bounds6, int32_t nPrims, int32_t maxPrimsInNode,
                                  int32_t splitMethod, nnbvh_oracle_node *nodesOut, int32_t *orderOut) {
    if (nPrims <= 0)
        return -1;
    Builder b;
    b.maxPrimsInNode = std::min(255, maxPrimsInNode);  // aggregates.cpp:142
    b.splitMethod = splitMethod;
    b.orderOut = orderOut;
    std::vector<BuildPrim> prims((size_t)nPrims);
    for (int32_t i = 0; i < nPrims; ++i) {
        Box bb;
        for (int k = 0; k < 3; ++k) {
            bb.mn[k] = bounds6[6 * (size_t)i + k];
            bb.mx[k] = bounds6[6 * (size_t)i + 3 + k];
        }
        prims[i].primitiveIndex = (size_t)i;
        prims[i].bounds = bb;
    }
    BuildNode *root = splitMethod == NNBVH_ORACLE_HLBVH ? b.BuildHLBVH(prims)
                                                        : b.BuildRecursive(prims.data(), prims.size());
    int offset = 0;
    b.Flatten(root, nodesOut, &offset);
    return offset;
}

void nnbvh_oracle_intersect_mixed(const nnbvh_oracle_node *nodes, const float *orderedTris9,
                                  const int32_t *orderedSphereIndex, const float *spheres28,
                                  const float *rays8, int64_t n, uint32_t *hits5, uint32_t *nodesVisited,
                                  uint32_t *triTests, int32_t nThreads) {
    ParallelRays(n, nThreads, [&](int64_t i) {
        Traverse<false>(nodes, orderedTris9, rays8 + 8 * i, hits5 + 5 * i,
                        nodesVisited ? nodesVisited + i : nullptr, triTests ? triTests + i : nullptr,
                        orderedSphereIndex, spheres28);
    });
}

void nnbvh_oracle_intersectP_mixed(const nnbvh_oracle_node *nodes, const float *orderedTris9,
                                   const int32_t *orderedSphereIndex, const float *spheres28,
                                   const float *rays8, int64_t n, uint8_t *occluded, uint32_t *nodesVisited,
                                   uint32_t *triTests, int32_t nThreads) {
    ParallelRays(n, nThreads, [&](int64_t i) {
        occluded[i] = Traverse<true>(nodes, orderedTris9, rays8 + 8 * i, nullptr,
                                     nodesVisited ? nodesVisited + i : nullptr,
                                     triTests ? triTests + i : nullptr, orderedSphereIndex, spheres28)
                          ? 1
                          : 0;
    });
}

// Triangles behind GeometricPrimitives with a constant alpha each (orderedAlpha[i] >= 1: none).  Closest hits
// or, with occluded != NULL, IntersectP (which is Intersect(...).has_value() for such primitives).
void nnbvh_oracle_intersect_alpha(const nnbvh_oracle_node *nodes, const float *orderedTris9,
                                  const float *orderedAlpha, const float *rays8, int64_t n, uint32_t *hits5,
                                  uint8_t *occluded, int32_t nThreads) {
    ParallelRays(n, nThreads, [&](int64_t i) {
        if (occluded)
            occluded[i] = Traverse<true>(nodes, orderedTris9, rays8 + 8 * i, nullptr, nullptr, nullptr, nullptr,
                                         nullptr, nullptr, nullptr, orderedAlpha)
                              ? 1
                              : 0;
        else
            Traverse<false>(nodes, orderedTris9, rays8 + 8 * i, hits5 + 5 * i, nullptr, nullptr, nullptr, nullptr,
                            nullptr, nullptr, orderedAlpha);
    });
}
// Origin of the ray GeometricPrimitive::Intersect re-traces after an ignored hit, and tHit; 0 = miss.
int32_t nnbvh_oracle_spawn_origin(const float *ray8, const float *p9, float *out4) {
    RayCtx r = MakeRayCtx(ray8);
    float t, b0, b1, b2;
    if (!TriangleIntersect(r, ray8[3], p9, &t, &b0, &b1, &b2))
        return 0;
    SpawnOrigin(p9, b0, b1, b2, ray8 + 4, out4);
    out4[3] = t;
    return 1;
}
// GeometricPrimitive::Intersect over one triangle with a constant alpha; out4 = {tHit, b0, b1, b2}; 0 = no hit.
int32_t nnbvh_oracle_alpha_triangle(const float *ray8, const float *p9, float alpha, float *out4) {
    if (alpha < 1)
        return GeometricAlphaIntersect(ray8, ray8 + 4, ray8[3], p9, alpha, &out4[0], &out4[1], &out4[2], &out4[3]) ? 1 : 0;
    RayCtx r = MakeRayCtx(ray8);
    return TriangleIntersect(r, ray8[3], p9, &out4[0], &out4[1], &out4[2], &out4[3]) ? 1 : 0;
}
// Transform::ApplyInverse(Ray, &tMax) (util/transform.h:416-430) with mInv = primitiveFromRender; local8 = {o, tMax, d, time}
void nnbvh_oracle_instance_ray(const float *primitiveFromRender12, const float *ray8, float *local8) {
    InstanceRay(primitiveFromRender12, ray8, local8);
}
float nnbvh_oracle_hash_float(const float *o3, const float *d3) {
    return HashFloatOD(o3, d3);
}

void nnbvh_oracle_intersect_instanced(const nnbvh_oracle_node *nodes, const float *orderedTris9,
                                      const int32_t *instanceOf, const float *inst24,
                                      const int32_t *instObject, const nnbvh_oracle_node *objNodes,
                                      const int32_t *objNodeBase, const float *objTris,
                                      const int32_t *objPrimBase, const float *rays8, int64_t n,
                                      uint32_t *hits5, uint32_t *inner, uint8_t *occluded,
                                      uint32_t *nodesVisited, uint32_t *triTests, int32_t nThreads) {
    InstancedScene sc{instanceOf, inst24, instObject, objNodes, objNodeBase, objTris, objPrimBase};
    ParallelRays(n, nThreads, [&](int64_t i) {
        if (occluded)
            occluded[i] = Traverse<true>(nodes, orderedTris9, rays8 + 8 * i, nullptr,
                                         nodesVisited ? nodesVisited + i : nullptr,
                                         triTests ? triTests + i : nullptr, nullptr, nullptr, &sc, nullptr)
                              ? 1
                              : 0;
        else
            Traverse<false>(nodes, orderedTris9, rays8 + 8 * i, hits5 + 5 * i,
                            nodesVisited ? nodesVisited + i : nullptr, triTests ? triTests + i : nullptr,
                            nullptr, nullptr, &sc, inner ? inner + i : nullptr);
    });
}

// TransformedPrimitive::Bounds (cpu/primitive.h:94): renderFromPrimitive(objectBounds), eight corners.
void nnbvh_oracle_instance_bounds(const float *renderFromPrimitive12, const float *objectBounds6, float *out6) {
    const float *m = renderFromPrimitive12;
    Box bt;
    for (int c = 0; c < 8; ++c) {
        float x = objectBounds6[(c & 1) ? 3 : 0], y = objectBounds6[(c & 2) ? 4 : 1], z = objectBounds6[(c & 4) ? 5 : 2];
        float p[3] = {m[0] * x + m[1] * y + m[2] * z + m[3], m[4] * x + m[5] * y + m[6] * z + m[7],
                      m[8] * x + m[9] * y + m[10] * z + m[11]};
        bt = Union(bt, p);
    }
    for (int k = 0; k < 3; ++k)
        out6[k] = bt.mn[k], out6[3 + k] = bt.mx[k];
}

int32_t nnbvh_oracle_sphere(const float *ray8, const float *sphere28, float *out4) {
    return SphereBasicIntersect(MakeSphereCtx(sphere28), ray8, ray8[3], &out4[0], &out4[1]) ? 1 : 0;
}

// std::atan2(float, float) of the libm this checker (and oracle/_ref) links: what Sphere::BasicIntersect
// compares with phiMax (shapes.h:196-203).
void nnbvh_oracle_atan2f(const float *y, const float *x, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i)
        out[i] = std::atan2(y[i], x[i]);
}

void nnbvh_oracle_sphere_bounds(const float *sphere28, float *out6) {
    SphereBounds(sphere28, out6);
}

int64_t nnbvh_oracle_build(const float *triVerts9, int32_t nTris, int32_t maxPrimsInNode,
                           int32_t splitMethod, nnbvh_oracle_node *nodesOut,
                           int32_t *orderOut) {
    if (nTris <= 0)
        return -1;
    Builder b;
    b.maxPrimsInNode = std::min(255, maxPrimsInNode);  // aggregates.cpp:142
    b.splitMethod = splitMethod;
    b.orderOut = orderOut;
    std::vector<BuildPrim> prims((size_t)nTris);
    for (int32_t i = 0; i < nTris; ++i) {
        // Triangle::Bounds = Union(Bounds3f(p0, p1), p2)  (shapes.cpp:294-301)
        const float *p = triVerts9 + 9 * (size_t)i;
        Box bb;
        for (int k = 0; k < 3; ++k) {
            bb.mn[k] = std::min(p[k], p[3 + k]);
            bb.mx[k] = std::max(p[k], p[3 + k]);
        }
        bb = Union(bb, p + 6);
        prims[i].primitiveIndex = (size_t)i;
        prims[i].bounds = bb;
    }
    BuildNode *root = splitMethod == NNBVH_ORACLE_HLBVH ? b.BuildHLBVH(prims)
                                                        : b.BuildRecursive(prims.data(), prims.size());
    int offset = 0;
    b.Flatten(root, nodesOut, &offset);
    return offset;
}

void nnbvh_oracle_intersect(const nnbvh_oracle_node *nodes, const float *orderedTris9,
                            const float *rays8, int64_t n, uint32_t *hits5,
                            uint32_t *nodesVisited, uint32_t *triTests, int32_t nThreads) {
    ParallelRays(n, nThreads, [&](int64_t i) {
        Traverse<false>(nodes, orderedTris9, rays8 + 8 * i, hits5 + 5 * i,
                        nodesVisited ? nodesVisited + i : nullptr,
                        triTests ? triTests + i : nullptr);
    });
}

void nnbvh_oracle_intersectP(const nnbvh_oracle_node *nodes, const float *orderedTris9,
                             const float *rays8, int64_t n, uint8_t *occluded,
                             uint32_t *nodesVisited, uint32_t *triTests, int32_t nThreads) {
    ParallelRays(n, nThreads, [&](int64_t i) {
        occluded[i] = Traverse<true>(nodes, orderedTris9, rays8 + 8 * i, nullptr,
                                     nodesVisited ? nodesVisited + i : nullptr,
                                     triTests ? triTests + i : nullptr)
                          ? 1
                          : 0;
    });
}

int32_t nnbvh_oracle_triangle(const float *ray8, const float *p9, float *out4) {
    RayCtx r = MakeRayCtx(ray8);
    return TriangleIntersect(r, ray8[3], p9, &out4[0], &out4[1], &out4[2], &out4[3]) ? 1 : 0;
}

int32_t nnbvh_oracle_box(const float *ray8, const float *box6) {
    RayCtx r = MakeRayCtx(ray8);
    return BoxIntersectP(box6, r, ray8[3]) ? 1 : 0;
}

}  // extern "C"
