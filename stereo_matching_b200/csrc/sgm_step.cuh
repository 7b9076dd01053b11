// One scanline step of adcensus.sgm2 (smooth/nyu.lua:59-60) for the KPL consecutive disparities a lane owns.
// Shared by the shared-memory ring kernels of the four original paths (sgm_ring.cu) and the diagonal-path kernels
// of the 8-path variant (post.cu) so that both do the original's arithmetic operation for operation.
#pragma once
#include "common.cuh"

namespace stereo {

constexpr float kSgmBig = 3.0e38f;        // "no such disparity": never the minimum, kSgmBig + P == kSgmBig

// The three penalty cases of the original; "P1 / alpha1" applies to the d-1 term on the downward path (SD 2)
// and to the d+1 term on the upward one (SD 3).
template <int SD>
struct SgmPenalties {
    float tau, P2a, P2b, P2c, P1a_m, P1a_p, P1b_m, P1b_p, P1c_m, P1c_p;
    __device__ explicit SgmPenalties(const SgmParams &p)
    {
        tau = p.tau_so;
        const float P1a = p.pi1;
        P2a = p.pi2;
        const float P1b = __fdiv_rn(p.pi1, __fmul_rn(p.q1, p.q2));
        P2b = __fdiv_rn(p.pi2, __fmul_rn(p.q1, p.q2));
        const float P1c = __fdiv_rn(p.pi1, p.q1);
        P2c = __fdiv_rn(p.pi2, p.q1);
        P1a_m = SD == 2 ? __fdiv_rn(P1a, p.alpha1) : P1a;
        P1a_p = SD == 3 ? __fdiv_rn(P1a, p.alpha1) : P1a;
        P1b_m = SD == 2 ? __fdiv_rn(P1b, p.alpha1) : P1b;
        P1b_p = SD == 3 ? __fdiv_rn(P1b, p.alpha1) : P1b;
        P1c_m = SD == 2 ? __fdiv_rn(P1c, p.alpha1) : P1c;
        P1c_p = SD == 3 ? __fdiv_rn(P1c, p.alpha1) : P1c;
    }
};

// prev: path costs of the previous step (kSgmBig where invalid); c: this step's costs; i1c / i1p: the other view's
// guide values at the matched pixels of this / the previous step; D1: guide difference of the reference view along
// the path; xx0: matched x of the lane's first disparity.  Writes val (kSgmBig where invalid).
template <int SD, int KPL>
__device__ __forceinline__ void sgm_step(const SgmPenalties<SD> &pen, float D1, const float (&c)[KPL],
                                         const float (&prev)[KPL], const float (&i1c)[KPL], const float (&i1p)[KPL],
                                         const bool (&valid)[KPL], int xx0, int direction, int w, bool has_below,
                                         bool has_above, float (&val)[KPL])
{
    // x step of the path: 0 ->, 1 <-, 2 down, 3 up, and the diagonals 4 down-right, 5 down-left, 6 up-right, 7 up-left
    constexpr int dx = (SD == 0 || SD == 4 || SD == 6) ? 1 : ((SD == 1 || SD == 5 || SD == 7) ? -1 : 0);
    float pmin = prev[0];
#pragma unroll
    for (int k = 1; k < KPL; ++k) pmin = fminf(pmin, prev[k]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) pmin = fminf(pmin, __shfl_xor_sync(0xffffffffu, pmin, o));
    float below = __shfl_up_sync(0xffffffffu, prev[KPL - 1], 1);      // prev[d0 - 1]
    float above = __shfl_down_sync(0xffffffffu, prev[0], 1);          // prev[d0 + KPL]
    below = has_below ? below : kSgmBig;
    above = has_above ? above : kSgmBig;
    // D1 is the same for the whole scanline step: pick the candidate penalty set once
    const bool d1_lt = D1 < pen.tau, d1_gt = D1 > pen.tau;
    const float P2x = d1_lt ? pen.P2a : pen.P2b, P1mx = d1_lt ? pen.P1a_m : pen.P1b_m, P1px = d1_lt ? pen.P1a_p : pen.P1b_p;
    const float pmin_p2x = __fadd_rn(pmin, P2x), pmin_p2c = __fadd_rn(pmin, pen.P2c);
#pragma unroll
    for (int k = 0; k < KPL; ++k) {
        const int xx = xx0 + k * direction;
        // the matched pixel of the previous step is (xx - dx, y - dy); its guide value is i1p[k]
        const bool inside = (unsigned)xx < (unsigned)w && (dx == 0 || (unsigned)(xx - dx) < (unsigned)w);
        const float D2 = inside ? fabsf(i1c[k] - i1p[k]) : 10.f;
        const bool both = d1_lt ? (D2 < pen.tau) : (d1_gt && D2 > pen.tau);
        const float P1m = both ? P1mx : pen.P1c_m, P1p = both ? P1px : pen.P1c_p;
        float m = fminf(prev[k], both ? pmin_p2x : pmin_p2c);
        const float pm = k > 0 ? prev[k > 0 ? k - 1 : 0] : below;
        const float pp = k < KPL - 1 ? prev[k < KPL - 1 ? k + 1 : 0] : above;
        m = fminf(m, __fadd_rn(pm, P1m));
        m = fminf(m, __fadd_rn(pp, P1p));
        const float v = __fadd_rn(__fadd_rn(c[k], m), -pmin);
        val[k] = valid[k] ? v : kSgmBig;
    }
}

// The same step with the penalty case looked up instead of recomputed: `want` classifies D1 of this scanline step
// (1: D1 < tau, 2: D1 > tau, 3: neither) and code[k] classifies D2 of disparity k (1: D2 < tau, 2: D2 > tau, 0:
// neither), both produced once per guide pair by sgm_prep_kernel with the original's comparisons, so that
// "both" = (D1 < tau && D2 < tau) || (D1 > tau && D2 > tau) is code[k] == want.  Invalid disparities carry
// c = prev = kSgmBig, which the additions leave unchanged (ulp(3e38) = 2e31), so no per-element select is needed.
template <int SD, int KPL>
__device__ __forceinline__ void sgm_step_coded(const SgmPenalties<SD> &pen, int want, const float (&c)[KPL],
                                               const float (&prev)[KPL], const int (&code)[KPL], bool has_below,
                                               bool has_above, float (&val)[KPL])
{
    float pmin = prev[0];
#pragma unroll
    for (int k = 1; k < KPL; ++k) pmin = fminf(pmin, prev[k]);
    // minimum over the warp in one instruction: redux.sync works on integers, and float order is integer order once
    // the magnitude bits of negative values are flipped (no NaNs here: costs are finite, kSgmBig included)
    {
        int key = __float_as_int(pmin);
        key ^= (key >> 31) & 0x7fffffff;
        key = __reduce_min_sync(0xffffffffu, key);
        key ^= (key >> 31) & 0x7fffffff;
        pmin = __int_as_float(key);
    }
    float below = __shfl_up_sync(0xffffffffu, prev[KPL - 1], 1);      // prev[d0 - 1]
    float above = __shfl_down_sync(0xffffffffu, prev[0], 1);          // prev[d0 + KPL]
    below = has_below ? below : kSgmBig;
    above = has_above ? above : kSgmBig;
    const bool d1_lt = want == 1;
    const float P2x = d1_lt ? pen.P2a : pen.P2b, P1mx = d1_lt ? pen.P1a_m : pen.P1b_m, P1px = d1_lt ? pen.P1a_p : pen.P1b_p;
    const float pmin_p2x = __fadd_rn(pmin, P2x), pmin_p2c = __fadd_rn(pmin, pen.P2c);
#pragma unroll
    for (int k = 0; k < KPL; ++k) {
        const bool both = code[k] == want;
        // on the horizontal paths no alpha1 is applied: P1 is the same for d - 1 and d + 1, one select less
        const float P1m = both ? P1mx : pen.P1c_m, P1p = SD < 2 ? P1m : (both ? P1px : pen.P1c_p);
        float m = fminf(prev[k], both ? pmin_p2x : pmin_p2c);
        const float pm = k > 0 ? prev[k > 0 ? k - 1 : 0] : below;
        const float pp = k < KPL - 1 ? prev[k < KPL - 1 ? k + 1 : 0] : above;
        m = fminf(m, __fadd_rn(pm, P1m));
        m = fminf(m, __fadd_rn(pp, P1p));
        val[k] = __fadd_rn(__fadd_rn(c[k], m), -pmin);
    }
}

}  // namespace stereo
