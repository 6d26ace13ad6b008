// nn_bvh_b200/csrc/atan2_host.cuh -- std::atan2(float, float) as the reference's HOST build computes it.
//
// Sphere::BasicIntersect clips a partial sphere by `phi > phiMax` with phi = std::atan2(pHit.y, pHit.x)
// (/root/reference/src/pbrt/shapes.h:196-203, :218-222).  That value is whatever the host's libm
// returns; it is not correctly rounded, so the device has to follow the same algorithm to take the same
// side of the comparison.  The reference is built against glibc (2.39 in this image), whose atan2f /
// atanf up to 2.40 are the fdlibm single-precision routines: argument reduction to one of four
// intervals, an 11-term odd/even polynomial, hi + lo constants.  This file follows that published
// algorithm step by step in round-to-nearest float arithmetic without contraction.  The host twin
// of this function was compared with glibc's atan2f on 10^9 inputs (arbitrary bit patterns and
// pairs in [-1,1]^2) without a difference; tests/test_spheres.py repeats the comparison on the device.
// A reference linked against another libm (glibc >= 2.41's correctly rounded atan2f, musl, ...) may
// take the other side for a hit within an ulp of the phiMax boundary.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "exact_math.cuh"

namespace nnbvh {

__device__ __forceinline__ float AtanfHost(float x) {
    // atan(0.5), atan(1), atan(1.5), atan(inf) as hi + lo
    const float hi0 = 4.6364760399e-01f, hi1 = 7.8539812565e-01f, hi2 = 9.8279368877e-01f, hi3 = 1.5707962513e+00f;
    const float lo0 = 5.0121582440e-09f, lo1 = 3.7748947079e-08f, lo2 = 3.4473217170e-08f, lo3 = 7.5497894159e-08f;
    const int32_t hx = __float_as_int(x), ix = hx & 0x7fffffff;
    if (ix >= 0x4c000000) {  // |x| >= 2^25
        if (ix > 0x7f800000) return Add(x, x);
        return hx > 0 ? Add(hi3, lo3) : Sub(-hi3, lo3);
    }
    int id;
    if (ix < 0x3ee00000) {  // |x| < 0.4375
        if (ix < 0x31000000) return x;  // |x| < 2^-29
        id = -1;
    } else {
        x = fabsf(x);
        if (ix < 0x3f980000) {      // |x| < 1.1875
            if (ix < 0x3f300000) {  // 7/16 <= |x| < 11/16
                id = 0;
                x = Div(Sub(Mul(2.0f, x), 1.0f), Add(2.0f, x));
            } else {                // 11/16 <= |x| < 19/16
                id = 1;
                x = Div(Sub(x, 1.0f), Add(x, 1.0f));
            }
        } else if (ix < 0x401c0000) {  // |x| < 2.4375
            id = 2;
            x = Div(Sub(x, 1.5f), Add(1.0f, Mul(1.5f, x)));
        } else {
            id = 3;
            x = Div(-1.0f, x);
        }
    }
    const float z = Mul(x, x), w = Mul(z, z);
    // the sum over aT[i] z^(i+1), split into its odd and even halves
    const float a0 = 3.3333334327e-01f, a1 = -2.0000000298e-01f, a2 = 1.4285714924e-01f, a3 = -1.1111110449e-01f,
                a4 = 9.0908870101e-02f, a5 = -7.6918758452e-02f, a6 = 6.6610731184e-02f, a7 = -5.8335702866e-02f,
                a8 = 4.9768779427e-02f, a9 = -3.6531571299e-02f, a10 = 1.6285819933e-02f;
    const float s1 = Mul(z, Add(a0, Mul(w, Add(a2, Mul(w, Add(a4, Mul(w, Add(a6, Mul(w, Add(a8, Mul(w, a10)))))))))));
    const float s2 = Mul(w, Add(a1, Mul(w, Add(a3, Mul(w, Add(a5, Mul(w, Add(a7, Mul(w, a9)))))))));
    if (id < 0) return Sub(x, Mul(x, Add(s1, s2)));
    const float hi = id == 0 ? hi0 : (id == 1 ? hi1 : (id == 2 ? hi2 : hi3));
    const float lo = id == 0 ? lo0 : (id == 1 ? lo1 : (id == 2 ? lo2 : lo3));
    const float r = Sub(hi, Sub(Sub(Mul(x, Add(s1, s2)), lo), x));
    return hx < 0 ? -r : r;
}

__device__ __forceinline__ float Atan2fHost(float y, float x) {
    const float tiny = 1.0e-30f, pi_o_4 = 7.8539818525e-01f, pi_o_2 = 1.5707963705e+00f, pi = 3.1415927410e+00f,
                pi_lo = -8.7422776573e-08f;
    const int32_t hx = __float_as_int(x), ix = hx & 0x7fffffff;
    const int32_t hy = __float_as_int(y), iy = hy & 0x7fffffff;
    if (ix > 0x7f800000 || iy > 0x7f800000) return Add(x, y);
    if (hx == 0x3f800000) return AtanfHost(y);
    const int m = ((hy >> 31) & 1) | ((hx >> 30) & 2);  // 2 * sign(x) + sign(y)
    if (iy == 0) return m < 2 ? y : (m == 2 ? Add(pi, tiny) : Sub(-pi, tiny));
    if (ix == 0) return hy < 0 ? Sub(-pi_o_2, tiny) : Add(pi_o_2, tiny);
    if (ix == 0x7f800000) {
        if (iy == 0x7f800000) {
            if (m == 0) return Add(pi_o_4, tiny);
            if (m == 1) return Sub(-pi_o_4, tiny);
            if (m == 2) return Add(Mul(3.0f, pi_o_4), tiny);
            return Sub(Mul(-3.0f, pi_o_4), tiny);
        }
        if (m == 0) return 0.0f;
        if (m == 1) return -0.0f;
        return m == 2 ? Add(pi, tiny) : Sub(-pi, tiny);
    }
    if (iy == 0x7f800000) return hy < 0 ? Sub(-pi_o_2, tiny) : Add(pi_o_2, tiny);
    const int k = (iy - ix) >> 23;
    float z;
    if (k > 60) z = Add(pi_o_2, Mul(0.5f, pi_lo));
    else if (hx < 0 && k < -60) z = 0.0f;
    else z = AtanfHost(fabsf(Div(y, x)));
    if (m == 0) return z;
    if (m == 1) return __int_as_float(__float_as_int(z) ^ (int)0x80000000);
    if (m == 2) return Sub(pi, Sub(z, pi_lo));
    return Sub(Sub(z, pi_lo), pi);
}

}  // namespace nnbvh
