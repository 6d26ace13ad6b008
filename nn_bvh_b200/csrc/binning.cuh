// nn_bvh_b200/csrc/binning.cuh -- ray binning pre-pass (sm_100a).
//
// The reference traces rays one at a time in whatever order the integrator
// produces them (cpu/integrators.cpp:296-313); results are independent of that
// order.  On the GPU the order decides which rays share a warp, i.e. how many
// lanes execute the same instruction and how many distinct cache lines one node
// fetch touches.  This pre-pass computes a permutation that groups rays by the
// grid cell of their origin (Morton order, inside the scene's root box) and by a
// coarse octahedral direction cell; the traversal kernel then pulls rays through
// the permutation and writes each result back to the ray's ORIGINAL slot, so the
// caller sees exactly the reference's semantics: hits[i] belongs to rays[i].
//
// Keys (BinKeyKernel) -> stable LSD radix sort of (key, ray index), 8 bits per pass
// (RadixCount/RadixScatterKernel in hlbvh.cuh, the HLBVH builder's sort) -> the permutation.  A
// device-side flag skips the sort for batches that arrive coherent (camera rays):
// BinSampleKernel measures the number of distinct keys per warp of consecutive rays
// on a sample of the batch; every later kernel of the pre-pass returns at once when
// the flag is 0, and the traversal kernel then ignores the permutation.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "kernels.cuh"

namespace nnbvh {

struct BinParams {
    float lo[3];
    float scale[3];     // cells per unit length (0 for a flat axis)
    uint32_t cellBits;  // grid is (1<<cellBits)^3
    uint32_t dirBits;   // octahedral direction grid is (1<<dirBits)^2
};

__device__ __forceinline__ uint32_t Part1By2(uint32_t x) {  // 10 bits -> every third bit
    x &= 0x3ffu;
    x = (x | (x << 16)) & 0x030000ffu;
    x = (x | (x << 8)) & 0x0300f00fu;
    x = (x | (x << 4)) & 0x030c30c3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

__device__ __forceinline__ uint32_t RayKey(const BinParams &bp, const RayRegs &r) {
    const int g = (1 << bp.cellBits) - 1;
    // NaN / inf origins land in cell 0 or g; any value is fine for a sort key
    int cx = (int)fminf(fmaxf((r.ox - bp.lo[0]) * bp.scale[0], 0.f), (float)g);
    int cy = (int)fminf(fmaxf((r.oy - bp.lo[1]) * bp.scale[1], 0.f), (float)g);
    int cz = (int)fminf(fmaxf((r.oz - bp.lo[2]) * bp.scale[2], 0.f), (float)g);
    uint32_t cell = Part1By2((uint32_t)cx) | (Part1By2((uint32_t)cy) << 1) | (Part1By2((uint32_t)cz) << 2);
    // octahedral map of the direction to [0,1)^2
    float ax = fabsf(r.dx), ay = fabsf(r.dy), az = fabsf(r.dz);
    float s = ax + ay + az;
    float inv = s > 0.f ? 1.f / s : 0.f;
    float u = r.dx * inv, v = r.dy * inv;
    if (r.dz < 0.f) {
        float uu = (1.f - fabsf(v)) * (u >= 0.f ? 1.f : -1.f);
        float vv = (1.f - fabsf(u)) * (v >= 0.f ? 1.f : -1.f);
        u = uu;
        v = vv;
    }
    const int dg = (1 << bp.dirBits) - 1;
    int du = (int)fminf(fmaxf((u * 0.5f + 0.5f) * (float)(dg + 1), 0.f), (float)dg);
    int dv = (int)fminf(fmaxf((v * 0.5f + 0.5f) * (float)(dg + 1), 0.f), (float)dg);
    uint32_t dir = ((uint32_t)dv << bp.dirBits) | (uint32_t)du;
    return (cell << (2 * bp.dirBits)) | dir;
}

// How incoherent the batch already is: kBinSamples warps' worth of consecutive rays, spread evenly
// over the batch; each counts the different keys among its 32 rays (1 = all in one cell, 32 =
// all different).  stats[0] accumulates the counts, stats[1] the number of warps sampled.
constexpr int kBinSamples = 256;

__global__ void __launch_bounds__(32)
BinSampleKernel(BinParams bp, RayInput in, uint64_t n, unsigned long long *__restrict__ stats) {
    const uint64_t warps = (n + 31) / 32;
    const uint64_t w = warps * blockIdx.x / kBinSamples;  // distinct for warps >= kBinSamples
    if (blockIdx.x > 0 && w == warps * (blockIdx.x - 1) / kBinSamples) return;
    const uint64_t i = w * 32 + threadIdx.x;
    const bool valid = i < n;
    uint32_t key = 0xffffffffu;
    if (valid) key = RayKey(bp, LoadRay(in, i));
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    const unsigned leaders =
        __ballot_sync(0xffffffffu, valid && (unsigned)(__ffs(peers) - 1) == threadIdx.x);
    if (threadIdx.x == 0) {
        atomicAdd(stats, (unsigned long long)__popc(leaders));
        atomicAdd(stats + 1, 1ull);
    }
}

// One thread: sort only batches whose warps are incoherent as they come.  *flag = 1 when the
// average sampled warp holds more than `threshold` distinct keys (or when forced).
__global__ void BinDecideKernel(const unsigned long long *__restrict__ stats, uint32_t threshold,
                                uint32_t force, uint32_t *__restrict__ flag) {
    *flag = (force || stats[0] > stats[1] * threshold) ? 1u : 0u;
}

__global__ void __launch_bounds__(256)
BinKeyKernel(BinParams bp, RayInput in, uint64_t n, uint32_t *__restrict__ keys,
             const uint32_t *__restrict__ enable) {
    if (*enable == 0u) return;
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keys[i] = RayKey(bp, LoadRay(in, i));
}

// Exclusive scan of hist[0..nBins) in place, two launches over tiles of kScanTile entries.
constexpr int kScanTile = 4096;  // 1024 threads x 4

__global__ void __launch_bounds__(1024)
BinTileSumKernel(const uint32_t *__restrict__ hist, uint32_t nBins, uint32_t *__restrict__ tileSums,
                 const uint32_t *__restrict__ enable = nullptr) {
    __shared__ uint32_t warpSum[32];
    if (enable && *enable == 0u) return;
    uint32_t base = blockIdx.x * kScanTile + threadIdx.x * 4;
    uint32_t s = 0;
    if (base + 3 < nBins) {
        uint4 v = *reinterpret_cast<const uint4 *>(hist + base);
        s = v.x + v.y + v.z + v.w;
    } else {
        for (int k = 0; k < 4; ++k)
            if (base + k < nBins) s += hist[base + k];
    }
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) warpSum[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t t = warpSum[threadIdx.x];
        for (int o = 16; o; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (threadIdx.x == 0) tileSums[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(1024)
BinTileScanKernel(uint32_t *__restrict__ hist, uint32_t nBins, const uint32_t *__restrict__ tileSums,
                  const uint32_t *__restrict__ enable = nullptr) {
    __shared__ uint32_t warpSum[32];
    __shared__ uint32_t tileBase;
    if (enable && *enable == 0u) return;
    // prefix of the tiles before this one (<= a few hundred values)
    uint32_t pre = 0;
    for (uint32_t t = threadIdx.x; t < blockIdx.x; t += blockDim.x) pre += tileSums[t];
    for (int o = 16; o; o >>= 1) pre += __shfl_xor_sync(0xffffffffu, pre, o);
    if ((threadIdx.x & 31) == 0) warpSum[threadIdx.x >> 5] = pre;
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t t = warpSum[threadIdx.x];
        for (int o = 16; o; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (threadIdx.x == 0) tileBase = t;
    }
    __syncthreads();
    uint32_t base = blockIdx.x * kScanTile + threadIdx.x * 4;
    uint32_t v[4] = {0, 0, 0, 0};
    for (int k = 0; k < 4; ++k)
        if (base + k < nBins) v[k] = hist[base + k];
    uint32_t s = v[0] + v[1] + v[2] + v[3];
    // inclusive warp scan of s
    uint32_t inc = s;
    const unsigned lane = threadIdx.x & 31;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (unsigned)o) inc += t;
    }
    __syncthreads();  // warpSum reuse
    if (lane == 31) warpSum[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t t = warpSum[threadIdx.x], ti = t;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t u = __shfl_up_sync(0xffffffffu, ti, o);
            if (threadIdx.x >= (unsigned)o) ti += u;
        }
        warpSum[threadIdx.x] = ti - t;  // exclusive
    }
    __syncthreads();
    uint32_t ex = tileBase + warpSum[threadIdx.x >> 5] + (inc - s);
    for (int k = 0; k < 4; ++k) {
        if (base + k < nBins) hist[base + k] = ex;
        ex += v[k];
    }
}

}  // namespace nnbvh
