// tests/cpp/cuda_stub/cuda_runtime.h -- stands in for <cuda_runtime.h> when the device math headers
// (nn_bvh_b200/csrc/{exact_math,sphere,atan2_host,alpha_math}.cuh) are compiled for the HOST by
// tests/cpp/device_twin.cpp.  Every intrinsic those headers use gets an IEEE-754 equivalent: the _rn
// operations are plain float / double operations (compiled with -ffp-contract=off, round-to-nearest),
// fused multiply-add is std::fma, the bit casts are memcpy.  What the twin then verifies is the device
// SOURCE -- statement order, constants, branches -- against the checker; that the GPU's _rn instructions
// are IEEE-exact is what the -m gpu tests add.
#pragma once

#include <cmath>
#include <cstdint>
#include <cstring>

#define __device__
#define __host__
#define __global__
#define __forceinline__ inline
#define __noinline__ __attribute__((noinline))
#define __launch_bounds__(...)

struct float4 {
    float x, y, z, w;
};
inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }

inline float __fadd_rn(float a, float b) { return a + b; }
inline float __fsub_rn(float a, float b) { return a - b; }
inline float __fmul_rn(float a, float b) { return a * b; }
inline float __fdiv_rn(float a, float b) { return a / b; }
inline float __fsqrt_rn(float a) { return std::sqrt(a); }
inline float __fmaf_rn(float a, float b, float c) { return std::fma(a, b, c); }
inline double __dmul_rn(double a, double b) { return a * b; }
inline double __dsub_rn(double a, double b) { return a - b; }
inline double __dadd_rn(double a, double b) { return a + b; }
inline float __double2float_rn(double a) { return (float)a; }
inline float __uint2float_rn(uint32_t a) { return (float)a; }
inline uint32_t __float_as_uint(float f) {
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return u;
}
inline int __float_as_int(float f) {
    int u;
    std::memcpy(&u, &f, 4);
    return u;
}
inline float __uint_as_float(uint32_t u) {
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}
inline float __int_as_float(int u) {
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}
template <typename T>
inline T __ldg(const T *p) {
    return *p;
}

// One "thread" at a time: the twin sets these and calls a __global__ function as an ordinary function.
struct TwinDim3 {
    unsigned x = 0, y = 0, z = 0;
};
inline thread_local TwinDim3 blockIdx, threadIdx, blockDim;
