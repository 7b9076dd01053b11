// x / s for one divisor s and many x, correctly rounded (bit-identical to IEEE division), without the reciprocal
// approximation, range check and slow-path call that __fdiv_rn spends on every element.
//
// r = RN(1/s) once (correctly rounded reciprocal).  q0 = RN(x r) is within 1.5 ulp of x/s; one residual step
// (e = x - s q exactly by FMA, q += e r) brings it within half an ulp plus a second-order term, i.e. to a faithful
// rounding; by Markstein's theorem a further residual step from a faithful q with the correctly rounded
// reciprocal yields the correctly rounded quotient.  The argument needs every intermediate to stay normal:
// guaranteed when s is in [2^-60, 2^60] and x is normal with a biased exponent in [64, 190]; anything else
// (zeros, subnormals, huge values) takes the IEEE routine.  scripts/div_check.cu compares the two exhaustively
// (all 2^32 x) for sample divisors.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace stereo {

struct ConstDivisor {
    float s, r;
    bool ok;
};

__device__ __forceinline__ ConstDivisor make_divisor(float s)
{
    ConstDivisor c;
    c.s = s;
    c.r = __frcp_rn(s);
    const float a = fabsf(s);
    c.ok = a >= 0x1p-60f && a <= 0x1p60f;
    return c;
}

__device__ __forceinline__ float div_exact(float x, const ConstDivisor &c)
{
    float q = __fmul_rn(x, c.r);
    float e = __fmaf_rn(-c.s, q, x);
    q = __fmaf_rn(e, c.r, q);
    e = __fmaf_rn(-c.s, q, x);
    q = __fmaf_rn(e, c.r, q);
    const uint32_t ex = (__float_as_uint(x) >> 23) & 0xffu;
    if (ex - 64u > 126u || !c.ok) q = __fdiv_rn(x, c.s);
    return q;
}

}  // namespace stereo
