// nn_bvh_b200/csrc/hlbvh.cuh -- HLBVH construction stages on the device (sm_100a).
//
// Replaces the data-parallel stages of BVHAggregate::buildHLBVH
// (/root/reference/src/pbrt/cpu/aggregates.cpp:389-449):
//   :392-394  centroid bounds            -> PrimBoundsKernel (warp min/max + ordered-uint atomics)
//   :398-409  Morton codes               -> MortonKernel
//   :412      RadixSort (:41-82)         -> RadixCountKernel / RadixScatterKernel + the generic tile scan of
//                                           binning.cuh; 4 stable passes of 8 bits over the 30 code bits
//   :415-431  treelet boundaries         -> TreeletHeadKernel + TreeletListKernel
//   :434-449  emitLBVH per treelet       -> one thread per NODE: LbvhTopologyKernel, a prefix sum,
//                                           LbvhNodesKernel, LbvhBoundsKernel (see below)
// buildUpperSAH over the <= 4096 treelet roots and the final pre-order layout stay on the host
// (AssembleHLBVH, build_host.cpp): a few thousand nodes of sequential float decisions.
//
// Every float operation that reaches the output is a min, a max, or one of the explicitly
// rounded intrinsics below, so the device tree is bit-identical to the host builder's.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "hlbvh_emit.h"

namespace nnbvh {

// Order-preserving float <-> uint map for atomicMin / atomicMax on floats.
__device__ __forceinline__ uint32_t OrderedBits(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ uint32_t OrderedBitsInv(uint32_t k) {
    return (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
}

// box6[i] = Triangle::Bounds (shapes.cpp:294-301); cb[0..2] / cb[3..5] = ordered-uint min / max of
// the box centroids (.5f*pMin + .5f*pMax, aggregates.cpp:97).  cb must hold 3x 0xffffffff, 3x 0.
__global__ void __launch_bounds__(256)
PrimBoundsKernel(const float *__restrict__ tris9, uint32_t n, float *__restrict__ box6,
                 uint32_t *__restrict__ cb) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3], hi[3];
    if (i < n) {
        const float *p = tris9 + 9 * (size_t)i;
        float v[9];
        for (int k = 0; k < 9; ++k) v[k] = p[k];
        for (int k = 0; k < 3; ++k) {
            float a = MinF(MinF(v[k], v[3 + k]), v[6 + k]);
            float b = MaxF(MaxF(v[k], v[3 + k]), v[6 + k]);
            box6[6 * (size_t)i + k] = a;
            box6[6 * (size_t)i + 3 + k] = b;
            float c = __fadd_rn(__fmul_rn(.5f, a), __fmul_rn(.5f, b));
            lo[k] = hi[k] = c;
        }
    } else {
        for (int k = 0; k < 3; ++k) {
            lo[k] = 3.402823466e+38f;
            hi[k] = -3.402823466e+38f;
        }
    }
    for (int k = 0; k < 3; ++k) {
        for (int o = 16; o; o >>= 1) {
            lo[k] = MinF(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = MaxF(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
        for (int k = 0; k < 3; ++k) {
            atomicMin(cb + k, OrderedBits(lo[k]));
            atomicMax(cb + 3 + k, OrderedBits(hi[k]));
        }
    }
}

// code[i] = EncodeMorton3(bounds.Offset(centroid) * 1024), aggregates.cpp:404-408; prim[i] = i.
__global__ void __launch_bounds__(256)
MortonKernel(const float *__restrict__ box6, uint32_t n, const uint32_t *__restrict__ cb,
             uint32_t *__restrict__ code, uint32_t *__restrict__ prim) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float o[3];
    for (int k = 0; k < 3; ++k) {
        float lo = __uint_as_float(OrderedBitsInv(cb[k])), hi = __uint_as_float(OrderedBitsInv(cb[3 + k]));
        float c = __fadd_rn(__fmul_rn(.5f, box6[6 * (size_t)i + k]), __fmul_rn(.5f, box6[6 * (size_t)i + 3 + k]));
        float r = __fsub_rn(c, lo);                          // Bounds3::Offset, vecmath.h:1321-1330
        if (hi > lo) r = __fdiv_rn(r, __fsub_rn(hi, lo));
        o[k] = __fmul_rn(r, 1024.f);
    }
    code[i] = EncodeMorton3(o[0], o[1], o[2]);
    prim[i] = i;
}

// ---- stable LSD radix sort, BITS (<= 8) bits per pass ------------------------------------------
// A block owns a tile of 2048 consecutive keys; warp w owns the tile's keys [256w, 256w+256) in 8
// rounds of 32 consecutive keys.
//   RadixCountKernel: hist[digit * nBlocks + block] = the tile's digit counts.  An exclusive scan
//     of hist (digit-major) then holds, per (digit, block), the first output position.
//   RadixScatterKernel: the rank of a key among the tile's keys with the same digit is (count in
//     earlier warps) + (count in earlier rounds of this warp) + (earlier lanes of this round), i.e.
//     input order, so the pass is stable.  Pairs are first placed in shared memory in digit order
//     and written out from there, so consecutive threads write consecutive addresses (runs of
//     tile/2^BITS pairs per digit) instead of 32 scattered 4-byte stores per instruction.
// Also the sort of the ray binning pre-pass (binning.cuh), which passes `enable`.
constexpr int kSortTile = 2048;

template <int BITS>
__global__ void __launch_bounds__(256)
RadixCountKernel(const uint32_t *__restrict__ keysIn, uint32_t n, int shift,
                 uint32_t *__restrict__ hist, const uint32_t *__restrict__ enable = nullptr) {
    constexpr uint32_t R = 1u << BITS;
    __shared__ uint32_t cnt[8][R];
    if (enable && *enable == 0u) return;  // uniform: the binning pre-pass decided not to sort
    const unsigned tid = threadIdx.x, w = tid >> 5;
    for (uint32_t k = tid; k < 8 * R; k += 256) (&cnt[0][0])[k] = 0;
    __syncthreads();
    const uint64_t base = (uint64_t)blockIdx.x * kSortTile;
    uint32_t key[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) {  // all eight loads in flight before the first shared atomic
        const uint64_t i = base + r * 256u + tid;  // order is irrelevant for counting: coalesced
        key[r] = i < n ? keysIn[i] : 0u;
    }
#pragma unroll
    for (int r = 0; r < 8; ++r)
        if (base + r * 256u + tid < n) atomicAdd(&cnt[w][(key[r] >> shift) & (R - 1)], 1u);
    __syncthreads();
    for (uint32_t d = tid; d < R; d += 256) {
        uint32_t t = 0;
        for (int k = 0; k < 8; ++k) t += cnt[k][d];
        hist[(size_t)d * gridDim.x + blockIdx.x] = t;
    }
}

template <int BITS>
__global__ void __launch_bounds__(256, 5)
RadixScatterKernel(const uint32_t *__restrict__ keysIn, const uint32_t *__restrict__ valsIn,
                   uint32_t n, int shift, const uint32_t *__restrict__ hist,
                   uint32_t *__restrict__ keysOut, uint32_t *__restrict__ valsOut,
                   const uint32_t *__restrict__ enable = nullptr) {
    constexpr uint32_t R = 1u << BITS;
    static_assert(R <= 256, "one thread per digit in the prefix stages");
    __shared__ uint32_t cnt[8][R];
    __shared__ uint32_t digitStart[R];  // tile-local exclusive prefix of the digit totals
    __shared__ uint32_t outBase[R];     // global position of the tile's first pair with this digit - digitStart
    __shared__ uint32_t warpTot[8];
    __shared__ uint32_t sKey[kSortTile], sVal[kSortTile];
    if (enable && *enable == 0u) return;
    const unsigned tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    for (uint32_t k = tid; k < 8 * R; k += 256) (&cnt[0][0])[k] = 0;
    __syncthreads();
    const uint64_t tile = (uint64_t)blockIdx.x * kSortTile, base = tile + w * 256u;
    uint32_t key[8], rank[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const uint64_t i = base + r * 32u + lane;
        key[r] = i < n ? keysIn[i] : 0u;
    }
    // the eight MATCH instructions are independent of each other and of the counters: issue them
    // together, then run the short dependent chain over the shared counters
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const bool valid = base + r * 32u + lane < n;
        rank[r] = __match_any_sync(0xffffffffu, valid ? ((key[r] >> shift) & (R - 1)) : R);
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const bool valid = base + r * 32u + lane < n;
        const uint32_t digit = (key[r] >> shift) & (R - 1);
        const unsigned peers = rank[r];
        const uint32_t before = valid ? cnt[w][digit] : 0u;
        rank[r] = before + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
        if (valid && (unsigned)(__ffs(peers) - 1) == lane) cnt[w][digit] = before + __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // thread d: exclusive prefix of digit d's counts over the 8 warps, then over the digits
    uint32_t total = 0;
    if (tid < R) {
        for (int k = 0; k < 8; ++k) {
            const uint32_t t = cnt[k][tid];
            cnt[k][tid] = total;
            total += t;
        }
    }
    uint32_t inc = total;
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (unsigned)o) inc += t;
    }
    if (lane == 31) warpTot[w] = inc;
    __syncthreads();
    if (tid < R) {
        uint32_t pre = 0;
        for (unsigned k = 0; k < w; ++k) pre += warpTot[k];
        const uint32_t start = pre + inc - total;
        digitStart[tid] = start;
        outBase[tid] = hist[(size_t)tid * gridDim.x + blockIdx.x] - start;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        if (base + r * 32u + lane < n) {
            const uint32_t d = (key[r] >> shift) & (R - 1);
            const uint32_t at = digitStart[d] + cnt[w][d] + rank[r];
            sKey[at] = key[r];
            const uint64_t i = base + r * 32u + lane;
            sVal[at] = valsIn ? valsIn[i] : (uint32_t)i;  // no valsIn: first pass, identity
        }
    }
    __syncthreads();
    const uint32_t count = (uint32_t)(n - tile < (uint64_t)kSortTile ? n - tile : (uint64_t)kSortTile);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const uint32_t j = r * 256u + tid;
        if (j < count) {
            const uint32_t k = sKey[j];
            const uint32_t pos = outBase[(k >> shift) & (R - 1)] + j;
            if (keysOut) keysOut[pos] = k;  // the last pass of a sort that only wants the permutation
            valsOut[pos] = sVal[j];
        }
    }
}

// ---- treelets -------------------------------------------------------------------------------
constexpr int kMaxTreelets = 4096;  // 12 prefix bits

// head[p] = first sorted position whose code has 12-bit prefix p (0xffffffff: no such code).
__global__ void __launch_bounds__(256)
TreeletHeadKernel(const uint32_t *__restrict__ code, uint32_t n, uint32_t *__restrict__ head) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t c = code[i];
    if (i == 0 || ((c ^ code[i - 1]) & kTreeletMask)) head[c >> 18] = i;
}

// One warp: compacts head[] into starts[0..count) in prefix (= Morton) order.
__global__ void __launch_bounds__(32)
TreeletListKernel(const uint32_t *__restrict__ head, uint32_t *__restrict__ starts,
                  uint32_t *__restrict__ count) {
    uint32_t at = 0;
    for (int b = 0; b < kMaxTreelets; b += 32) {
        uint32_t h = head[b + threadIdx.x];
        unsigned m = __ballot_sync(0xffffffffu, h != 0xffffffffu);
        if (h != 0xffffffffu) starts[at + __popc(m & ((1u << threadIdx.x) - 1u))] = h;
        at += __popc(m);
    }
    if (threadIdx.x == 0) *count = at;
}

// ---- treelet emission, one thread per NODE ---------------------------------------------------
// emitLBVH (aggregates.cpp:451-503) is a recursion per treelet.  The tree it produces is the
// radix tree of the sorted codes below the treelet prefix, pruned where a range holds fewer
// than maxPrimsInNode primitives or a single code, and laid out in pre-order (flattenBVH,
// :505-522).  All of that is computable per node:
//   * the interior node that splits between sorted positions i and i+1 on bit d (the highest
//     bit in which code[i] and code[i+1] differ) covers exactly the keys that share
//     code[i] >> (d+1): [l, r] by two searches in the sorted array.  It exists iff r-l+1 >=
//     maxPrimsInNode;
//   * node A precedes node X in pre-order iff A starts left of X's first key (A is then an
//     ancestor of X or entirely left of it) or A is an ancestor starting at the same key.
//     With S(p) = number of nodes starting before position p (a prefix sum) and the nodes that
//     start at one position forming a chain of strictly decreasing split bits that ends in a
//     leaf:  index(X) = S(l) + #{chain nodes at l with a higher split bit}.
//     chainMask[l] collects those split bits (atomicOr), bit 31 says "a leaf starts here";
//   * the first child sits at index+1, the second child (first key i+1, top of its chain) at
//     S(i+1).  S is taken over the whole array, so S(treelet start) is the treelet's offset in
//     the dense node array and no compaction pass is needed;
//   * bounds are min/max unions: leaves first, then up the parent links, the second child to
//     arrive at a node doing the union (order does not change a min or a max).
constexpr uint32_t kLeafBit = 1u << 31;
constexpr uint32_t kChainBits = 0x3ffffu;  // split bits 0..17
constexpr uint32_t kNoNode = 0xffffffffu;

// First index l <= i with code[l] >= key, given code[i] >= key (gallop left, then bisect).
__device__ __forceinline__ uint32_t FirstAtLeast(const uint32_t *__restrict__ code, uint32_t i,
                                                 uint32_t key) {
    int hi = (int)i, lo, step = 1;
    for (;;) {
        lo = hi - step;
        if (lo < 0) {
            lo = -1;
            break;
        }
        if (code[lo] < key) break;
        hi = lo;
        step <<= 1;
    }
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;  // lo >= -1, hi >= lo + 2: never negative
        if (code[mid] >= key) hi = mid;
        else lo = mid;
    }
    return (uint32_t)hi;
}
// Last index r >= i with code[r] <= key, given code[i] <= key (gallop right, then bisect).
__device__ __forceinline__ uint32_t LastAtMost(const uint32_t *__restrict__ code, uint32_t n,
                                               uint32_t i, uint32_t key) {
    uint32_t lo = i, hi, step = 1;
    for (;;) {
        hi = lo + step;
        if (hi >= n) {
            hi = n;
            break;
        }
        if (code[hi] > key) break;
        lo = hi;
        step <<= 1;
    }
    while (hi - lo > 1) {
        uint32_t mid = lo + ((hi - lo) >> 1);
        if (code[mid] <= key) lo = mid;
        else hi = mid;
    }
    return lo;
}

// Thread i: the candidate interior node splitting between i and i+1, and (treelet heads only)
// the treelet whose root is a leaf.  chainMask must be zero on entry.
__global__ void __launch_bounds__(256)
LbvhTopologyKernel(const uint32_t *__restrict__ code, uint32_t n, uint32_t maxPrims,
                   uint32_t *__restrict__ chainMask, uint32_t *__restrict__ leafN,
                   uint32_t *__restrict__ rangeL, int *__restrict__ overflow) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t c = code[i];
    if (i == 0 || ((c ^ code[i - 1]) & kTreeletMask)) {  // emitLBVH's entry test on the treelet
        const uint32_t e = LastAtMost(code, n, i, c | kChainBits);
        const uint32_t cnt = e - i + 1;
        if (cnt < maxPrims || code[e] == c) {
            atomicOr(chainMask + i, kLeafBit);
            leafN[i] = cnt;
            if (cnt >= 65536u) *overflow = 1;
        }
    }
    rangeL[i] = kNoNode;
    if (i + 1 >= n) return;
    const uint32_t c1 = code[i + 1], x = c ^ c1;
    if (x == 0 || (x & kTreeletMask)) return;
    const int d = 31 - __clz(x);
    const uint32_t low = (2u << d) - 1u;
    const uint32_t l = FirstAtLeast(code, i, c & ~low), r = LastAtMost(code, n, i + 1, c | low);
    if (r - l + 1 < maxPrims) return;  // :458: its whole range is one leaf (of an ancestor's making)
    rangeL[i] = l;
    atomicOr(chainMask + l, 1u << d);
    const uint32_t nl = i - l + 1, nr = r - i;
    if (nl < maxPrims || code[l] == c) {
        atomicOr(chainMask + l, kLeafBit);
        leafN[l] = nl;
        if (nl >= 65536u) *overflow = 1;
    }
    if (nr < maxPrims || code[r] == c1) {
        atomicOr(chainMask + i + 1, kLeafBit);
        leafN[i + 1] = nr;
        if (nr >= 65536u) *overflow = 1;
    }
}

// nodesAt[p] = number of nodes that start at sorted position p (p < n); nodesAt[n] = 0.
// An exclusive scan turns it into S.
__global__ void __launch_bounds__(256)
LbvhCountKernel(const uint32_t *__restrict__ chainMask, uint32_t n, uint32_t *__restrict__ nodesAt) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    const uint32_t m = i < n ? chainMask[i] : 0u;
    nodesAt[i] = __popc(m & kChainBits) + (m >> 31);
}

// Writes every node's topology fields and the parent links.  Interior offsets are relative to
// the treelet's first node, as AssembleHLBVH expects.
__global__ void __launch_bounds__(256)
LbvhNodesKernel(const uint32_t *__restrict__ code, uint32_t n, const uint32_t *__restrict__ chainMask,
                const uint32_t *__restrict__ leafN, const uint32_t *__restrict__ rangeL,
                const uint32_t *__restrict__ S, const uint32_t *__restrict__ head,
                nnbvh_node *__restrict__ nodes, uint32_t *__restrict__ parent) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t c = code[i], m = chainMask[i];
    if (i == 0 || ((c ^ code[i - 1]) & kTreeletMask)) parent[S[i]] = kNoNode;  // a treelet root
    if (m & kLeafBit) {
        nnbvh_node &nd = nodes[S[i] + __popc(m & kChainBits)];
        nd.offset = (int32_t)i;
        nd.nPrimitives = (uint16_t)leafN[i];
        nd.axis = 0;
        nd.pad = 0;
    }
    const uint32_t l = rangeL[i];
    if (l == kNoNode) return;
    const int d = 31 - __clz(c ^ code[i + 1]);
    const uint32_t pos = S[l] + __popc((chainMask[l] & kChainBits) >> (d + 1));
    const uint32_t second = S[i + 1];
    nnbvh_node &nd = nodes[pos];
    nd.offset = (int32_t)(second - S[head[c >> 18]]);
    nd.nPrimitives = 0;
    nd.axis = (uint8_t)(d % 3);
    nd.pad = 0;
    parent[pos + 1] = pos;
    parent[second] = pos;
}

// Leaf bounds, then interior bounds up the parent links.  visit must be zero on entry.
__global__ void __launch_bounds__(256)
LbvhBoundsKernel(const uint32_t *__restrict__ code, const uint32_t *__restrict__ prim,
                 const float *__restrict__ box6, uint32_t n, const uint32_t *__restrict__ chainMask,
                 const uint32_t *__restrict__ leafN, const uint32_t *__restrict__ S,
                 const uint32_t *__restrict__ head, const uint32_t *__restrict__ parent,
                 uint32_t *__restrict__ visit, nnbvh_node *nodes) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t m = chainMask[i];
    if (!(m & kLeafBit)) return;
    float lo[3], hi[3];
    for (int k = 0; k < 3; ++k) {
        lo[k] = 3.402823466e+38f;
        hi[k] = -3.402823466e+38f;
    }
    const uint32_t cnt = leafN[i];
    for (uint32_t j = 0; j < cnt; ++j) {
        const float *b = box6 + 6 * (size_t)prim[i + j];
        for (int k = 0; k < 3; ++k) {
            lo[k] = MinF(lo[k], b[k]);
            hi[k] = MaxF(hi[k], b[3 + k]);
        }
    }
    const uint32_t base = S[head[code[i] >> 18]];
    uint32_t cur = S[i] + __popc(m & kChainBits);
    for (;;) {
        float *nb = nodes[cur].bounds;
        for (int k = 0; k < 3; ++k) {
            nb[k] = lo[k];
            nb[3 + k] = hi[k];
        }
        const uint32_t par = parent[cur];
        if (par == kNoNode) break;
        __threadfence();
        if (atomicAdd(visit + par, 1u) == 0u) break;  // the sibling subtree is still open
        // Union(children[0], children[1]) with the FIRST child as std::min / std::max's first
        // argument (InitInterior, :113-120): only the sign of a zero can tell the difference
        const bool first = cur == par + 1;
        const float *sb = nodes[first ? base + (uint32_t)nodes[par].offset : par + 1].bounds;
        for (int k = 0; k < 3; ++k) {
            const float sl = __ldcg(sb + k), sh = __ldcg(sb + 3 + k);
            lo[k] = first ? MinF(lo[k], sl) : MinF(sl, lo[k]);
            hi[k] = first ? MaxF(hi[k], sh) : MaxF(sh, hi[k]);
        }
        cur = par;
    }
}

// offs[k] = S[starts[k]] for k < count, offs[count] = S[n] (the total node count).
__global__ void __launch_bounds__(256)
TreeletOffsetsKernel(const uint32_t *__restrict__ starts, const uint32_t *__restrict__ count,
                     const uint32_t *__restrict__ S, uint32_t n, uint32_t *__restrict__ offs) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x, nT = *count;
    if (k < nT) offs[k] = S[starts[k]];
    else if (k == nT) offs[k] = S[n];
}

}  // namespace nnbvh
