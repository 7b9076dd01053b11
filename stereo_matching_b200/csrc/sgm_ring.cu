// adcensus.sgm2 (smooth/nyu.lua:59-60) with the volumes streamed through shared memory by bulk copies.
//
// The register-prefetch kernels in post.cu (one warp per scanline, every lane loading its own 32 bytes of each
// step) move 2.0-2.6 TB/s: each warp asks DRAM for one 912-byte row per step, 1 500-3 700 independent streams at a
// time.  Here a CTA of NW scanline warps plus one producer warp pulls LARGE contiguous granules into a ring of
// shared-memory stages with cp.async.bulk (completion on an mbarrier), so that a request covers
//   horizontal paths: G consecutive steps of one scanline  (G x D floats, contiguous along x),
//   vertical paths:   one step of NW adjacent scanlines    (NW x D floats, contiguous along x),
// The penalty case of every (pixel, disparity) depends on the two guide images only: sgm_prep_kernel classifies
// D1 and D2 against tau once per call into byte images (`want` per pixel and path, `code` per matched pixel and
// path, padded so that no range check is left), and a step reads 8 bytes per lane one step ahead instead of
// gathering and comparing guide values.  The path-cost arithmetic is the original's, operation for operation
// (sgm_step_coded() in sgm_step.cuh): results are bit-identical to the register-prefetch kernels.
// D % 4 != 0 or volumes that are not 16-byte aligned take the same kernels: the bulk copies then fetch the
// 16-byte aligned superset of each run and the lanes read 4-byte words.
// Measured and dropped (profiles/r02_post_chain_ab.md): -(x / std) of post.lua:146,152 applied to the landed stages
// by two more warps per CTA instead of by an elementwise pass - the extra hand-over per stage cost the shallow
// pipelines of these kernels four times what the pass costs.
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "sgm_step.cuh"
#include "tc_common.cuh"

namespace stereo {
using namespace tc;

namespace {

constexpr int kMaxStages = 8;
constexpr int kBarBytes = 2 * 8 * kMaxStages;      // full and empty barriers

// Byte images of sgm_prep_kernel: want[path][h][w], code[path][h][pitch] with pitch = w + 2 * pad
struct SgmAux {
    const uint8_t *want, *code;
    int pad, pitch;
    // the last pass scales its sum by `final_scale` (0.25: nyu.lua:61,64 `/ 4`; 1: more paths follow) and, when
    // `wta` is set, writes the arg min over D of every pixel's result (the nyu_sgm stage image, post.lua:159)
    float final_scale;
    uint8_t *wta;
};

// Shared-memory layout of one stage: per array (costs, running sums) a row of SLOTS, one per bulk copy:
//   horizontal: slot = scanline of the CTA, holding G consecutive steps   (G * D floats)
//   vertical:   slot = step of the stage, holding the NW scanlines        (NW * D floats)
// Aligned volumes (D % 4 == 0, 16-byte aligned pointers): a slot is exactly its floats, read with 16-byte loads.
// Otherwise (e.g. the reference's default D = 201) the bulk copy fetches the 16-byte aligned superset of the run
// into a slot 32 bytes larger, the run starts `phase` = (address mod 16) bytes into the slot, and the lanes read
// 4-byte words.  The superset reaches at most 12 bytes before / after the run, inside the volume's allocation
// granule.
__host__ __device__ inline uint32_t ring_slot_bytes(bool aligned, uint32_t run_floats)
{
    return aligned ? run_floats * 4 : ((run_floats * 4 + 15) & ~15u) + 32;
}

template <int SD, int NACC, int KPL, int NW, int DIR, bool ALIGNED>
__device__ __forceinline__ void sgm_ring_scan(const SgmAux &aux, const float *__restrict__ cvol, const float *acc_a,
                                              const float *acc_b, float *outp, int h, int w, int D, const SgmParams &p,
                                              int last, int G, int S, int group)
{
    constexpr int dx = SD == 0 ? 1 : (SD == 1 ? -1 : 0);
    constexpr int dy = SD == 2 ? 1 : (SD == 3 ? -1 : 0);
    constexpr bool kH = dy == 0;
    constexpr int kArrays = 1 + NACC;
    extern __shared__ __align__(128) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int steps = kH ? w : h;
    const int n_lines = kH ? h : w;
    const int line0 = group * NW;
    const int lines_here = min(NW, n_lines - line0);
    const int n_stages = (steps + G - 1) / G;
    const uint32_t slot_bytes = ring_slot_bytes(ALIGNED, (uint32_t)(kH ? G : NW) * D);
    const uint32_t array_bytes = (uint32_t)(kH ? NW : G) * slot_bytes;  // one array of one stage
    const uint32_t stage_bytes = kArrays * array_bytes;
    const uint32_t s_data = smem_u32(smem);
    const uint32_t s_bar = s_data + (uint32_t)S * stage_bytes;
    const uint32_t bar_full = s_bar, bar_empty = s_bar + 8 * kMaxStages;
    const float *src[3] = {cvol, acc_a, acc_b};
    // unaligned volumes: results leave through a per-warp staging row (two of them, used alternately) so that a
    // warp store is a 128-byte run instead of 32 words 32 bytes apart
    const int Dp = (D + 3) & ~3;
    float *stg_rows = reinterpret_cast<float *>(smem + (size_t)S * stage_bytes + kBarBytes) + (size_t)warp * 2 * Dp;

    if (threadIdx.x == 0) {
        for (int i = 0; i < S; ++i) {
            mbar_init(bar_full + 8 * i, 1);
            mbar_init(bar_empty + 8 * i, (uint32_t)lines_here);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // global element offset of the run that fills slot `k` of stage `t`
    auto run_start = [&](int t, int k, int g_here) -> size_t {
        if (kH) {       // scanline line0 + k, g_here consecutive pixels; the right-to-left path walks them backwards
            const int x_lo = SD == 0 ? t * G : w - t * G - g_here;
            return ((size_t)(line0 + k) * w + x_lo) * D;
        }
        const int y = SD == 2 ? t * G + k : h - 1 - (t * G + k);      // step k of the stage: one image row
        return ((size_t)y * w + line0) * D;
    };

    if (warp == NW) {
        // ---- producer: one lane issues the bulk copies of every stage ----
        if (lane != 0) return;
        for (int t = 0; t < n_stages; ++t) {
            const int s = t % S;
            mbar_wait(bar_empty + 8 * s, ((t / S) & 1) ^ 1);
            const int g_here = min(G, steps - t * G);
            const int n_slots = kH ? lines_here : g_here;
            const uint32_t run_bytes = (uint32_t)(kH ? g_here : lines_here) * D * 4;
            const uint32_t dst0 = s_data + (uint32_t)s * stage_bytes;
            uint32_t total = 0;
            for (int pass = 0; pass < 2; ++pass) {          // pass 0 adds up the bytes, pass 1 issues the copies
                if (pass == 1) mbar_expect_tx(bar_full + 8 * s, total);
                for (int a = 0; a < kArrays; ++a)
                    for (int k = 0; k < n_slots; ++k) {
                        const uintptr_t from = (uintptr_t)(src[a] + run_start(t, k, g_here));
                        const uintptr_t lo = ALIGNED ? from : from & ~(uintptr_t)15;
                        const uint32_t len = ALIGNED ? run_bytes : (uint32_t)(((from + run_bytes + 15) & ~(uintptr_t)15) - lo);
                        if (pass == 0) total += len;
                        else bulk_g2s(dst0 + a * array_bytes + (uint32_t)k * slot_bytes, (const void *)lo, len, bar_full + 8 * s);
                    }
            }
        }
        return;
    }
    if (warp >= lines_here) return;

    // ---- consumers: one warp per scanline, KPL consecutive disparities per lane ----
    const int line = line0 + warp;
    const int x = kH ? (SD == 0 ? 0 : w - 1) : line;
    const int y = kH ? line : (SD == 2 ? 0 : h - 1);
    const int d0 = lane * KPL;
    const SgmPenalties<SD> pen(p);
    bool valid4[KPL / 4];
#pragma unroll
    for (int j = 0; j < KPL / 4; ++j) valid4[j] = d0 + 4 * j < D;      // whole float4s when D % 4 == 0
    const bool has_below = lane > 0, has_above = d0 + KPL < D;

    // classified guide differences of the NEXT step (bytes; one step ahead of their use)
    const uint8_t *cptr = aux.code + ((size_t)SD * h + y) * aux.pitch + aux.pad + x + d0 * DIR;
    const uint8_t *wptr = aux.want + ((size_t)SD * h + y) * w + x;
    const ptrdiff_t cstep = (ptrdiff_t)dy * aux.pitch + dx, wstep = (ptrdiff_t)dy * w + dx;
    int code_n[KPL], want_n = 3;
#pragma unroll
    for (int k = 0; k < KPL; ++k) code_n[k] = 0;

    float prev[KPL];
#pragma unroll
    for (int k = 0; k < KPL; ++k) prev[k] = kSgmBig;
    float *optr = outp + ((size_t)y * w + x) * D + (ALIGNED ? d0 : 0);
    const ptrdiff_t vstride = ((ptrdiff_t)dy * w + dx) * D;

    // unaligned volumes: running slot-relative byte offsets (see the gather below).  Horizontal paths move one pixel
    // (D floats) inside the run per step; vertical paths change the run, whose phase moves by (w * D * 4) mod 16.
    uint32_t o_run[3] = {0, 0, 0}, ph_run[3] = {0, 0, 0};
    const uint32_t o_step = kH ? (uint32_t)(dx * D * 4) : ((uint32_t)(dy * w * D * 4) & 15u);
    int step = 0;
    for (int t = 0; t < n_stages; ++t) {
        const int s = t % S;
        const int g_here = min(G, steps - t * G);
        mbar_wait(bar_full + 8 * s, (t / S) & 1);
        const uint8_t *stage = smem + (size_t)s * stage_bytes;
        if (!ALIGNED) {
            // byte offset of this lane's first float inside its slot at the first step of the stage: the phase of
            // the run (address mod 16) plus the run-relative position
            const size_t e0 = run_start(t, kH ? warp : 0, g_here);
            const int inner0 = kH ? (SD == 0 ? 0 : g_here - 1) : warp;
#pragma unroll
            for (int q = 0; q < kArrays; ++q) {
                ph_run[q] = (uint32_t)((uintptr_t)(src[q] + e0) & 15);
                o_run[q] = ph_run[q] + ((uint32_t)inner0 * D + d0) * 4;
            }
        }
        for (int i = 0; i < g_here; ++i, ++step) {
            int code[KPL];
#pragma unroll
            for (int k = 0; k < KPL; ++k) code[k] = code_n[k];
            const int want = want_n;
            if (step + 1 < steps) {
                cptr += cstep;
                wptr += wstep;
#pragma unroll
                for (int k = 0; k < KPL; ++k) code_n[k] = __ldg(cptr + k * DIR);
                want_n = __ldg(wptr);
            }
            // costs (and running sums) of this step from the stage
            const int slot = kH ? warp : i;
            const int inner = kH ? (SD == 0 ? i : g_here - 1 - i) : warp;        // run-relative pixel
            const uint8_t *base = stage + (uint32_t)slot * slot_bytes + ((uint32_t)inner * D + d0) * 4;
            float c[KPL], a[KPL], b[KPL];      // a, b: running sums (used when NACC >= 1, 2)
            if (ALIGNED) {
#pragma unroll
                for (int j = 0; j < KPL / 4; ++j) {
                    if (valid4[j]) {
                        const float4 v = *reinterpret_cast<const float4 *>(base + 16 * j);
                        c[4 * j] = v.x; c[4 * j + 1] = v.y; c[4 * j + 2] = v.z; c[4 * j + 3] = v.w;
                        if (NACC >= 1) {
                            const float4 u = *reinterpret_cast<const float4 *>(base + array_bytes + 16 * j);
                            a[4 * j] = u.x; a[4 * j + 1] = u.y; a[4 * j + 2] = u.z; a[4 * j + 3] = u.w;
                        }
                        if (NACC >= 2) {
                            const float4 u = *reinterpret_cast<const float4 *>(base + 2 * array_bytes + 16 * j);
                            b[4 * j] = u.x; b[4 * j + 1] = u.y; b[4 * j + 2] = u.z; b[4 * j + 3] = u.w;
                        }
                    } else {
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            c[4 * j + e] = kSgmBig;           // stays kSgmBig through the step: "no such disparity"
                            if (NACC >= 1) a[4 * j + e] = 0.f;
                            if (NACC >= 2) b[4 * j + e] = 0.f;
                        }
                    }
                }
            } else {
                // the run sits `phase` bytes into its slot (per array: the volumes need not share an alignment), so a
                // lane's KPL floats start (phase + row offset) mod 16 bytes into a 16-byte word - the same shift for
                // every lane.  Read the aligned words that cover them and pick by the (warp-uniform) shift.
                auto gather = [&](int q, float (&dst)[KPL], float fill) {
                    const uint32_t o = o_run[q];
                    const uint8_t *al = stage + q * array_bytes + (uint32_t)slot * slot_bytes + (o & ~15u);
                    float tq[KPL + 4];
#pragma unroll
                    for (int j = 0; j < KPL / 4 + 1; ++j) {
                        const float4 v = *reinterpret_cast<const float4 *>(al + 16 * j);
                        tq[4 * j] = v.x; tq[4 * j + 1] = v.y; tq[4 * j + 2] = v.z; tq[4 * j + 3] = v.w;
                    }
                    const uint32_t sh = (o & 15u) >> 2;
                    if (sh == 0) {
#pragma unroll
                        for (int k = 0; k < KPL; ++k) dst[k] = tq[k];
                    } else if (sh == 1) {
#pragma unroll
                        for (int k = 0; k < KPL; ++k) dst[k] = tq[k + 1];
                    } else if (sh == 2) {
#pragma unroll
                        for (int k = 0; k < KPL; ++k) dst[k] = tq[k + 2];
                    } else {
#pragma unroll
                        for (int k = 0; k < KPL; ++k) dst[k] = tq[k + 3];
                    }
                    if (d0 + KPL > D) {                      // only the lanes at the end of the disparity range
#pragma unroll
                        for (int k = 0; k < KPL; ++k) dst[k] = d0 + k < D ? dst[k] : fill;
                    }
                    // next step: the neighbouring pixel of the run (horizontal) / the next image row (vertical)
                    if (kH) {
                        o_run[q] = o + o_step;
                    } else {
                        const uint32_t ph = (ph_run[q] + o_step) & 15u;
                        o_run[q] = o - ph_run[q] + ph;
                        ph_run[q] = ph;
                    }
                };
                gather(0, c, kSgmBig);
                if constexpr (NACC >= 1) gather(1, a, 0.f);
                if constexpr (NACC >= 2) gather(2, b, 0.f);
            }
            float val[KPL];
            if (step == 0) {
#pragma unroll
                for (int k = 0; k < KPL; ++k) val[k] = c[k];
            } else {
                sgm_step_coded<SD, KPL>(pen, want, c, prev, code, has_below, has_above, val);
            }
            float best = kSgmBig;         // arg min over this lane's disparities (lowest index wins ties)
            int best_k = 0;
#pragma unroll
            for (int j = 0; j < KPL / 4; ++j) {
                if (valid4[j]) {
                    float r[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        float acc = val[4 * j + e];
                        if (NACC == 1) acc = __fadd_rn(a[4 * j + e], acc);
                        if (NACC == 2) acc = __fadd_rn(__fadd_rn(a[4 * j + e], b[4 * j + e]), acc);
                        r[e] = last ? __fmul_rn(acc, aux.final_scale) : acc;       // x / 4 == x * 0.25 exactly
                        if (last && (ALIGNED || d0 + 4 * j + e < D) && r[e] < best) {
                            best = r[e];
                            best_k = 4 * j + e;
                        }
                    }
                    if (ALIGNED) {
                        reinterpret_cast<float4 *>(optr)[j] = make_float4(r[0], r[1], r[2], r[3]);
                    } else {
                        *reinterpret_cast<float4 *>(stg_rows + (step & 1) * Dp + d0 + 4 * j) = make_float4(r[0], r[1], r[2], r[3]);
                    }
                }
            }
            if (last && aux.wta != nullptr) {
                // warp-wide arg min: one redux on order-preserving keys (-0 counted as +0), lowest lane = lowest d
                int key = __float_as_int(__fadd_rn(best, 0.f));
                key ^= (key >> 31) & 0x7fffffff;
                const int mn = __reduce_min_sync(0xffffffffu, key);
                const unsigned winners = __ballot_sync(0xffffffffu, key == mn);
                if (lane == __ffs(winners) - 1) {
                    const int px = kH ? (SD == 0 ? step : w - 1 - step) : line;
                    const int py = kH ? line : (SD == 2 ? step : h - 1 - step);
                    aux.wta[(size_t)py * w + px] = (uint8_t)(d0 + best_k);
                }
            }
            if (!ALIGNED) {
                __syncwarp();
                const float *row = stg_rows + (step & 1) * Dp;
                for (int e = lane; e < D; e += 32) optr[e] = row[e];
            }
#pragma unroll
            for (int k = 0; k < KPL; ++k) prev[k] = val[k];
            optr += vstride;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_empty + 8 * s);
    }
}

// Classifies the guide differences of the four paths (0: left->right, 1: right->left, 2: down, 3: up) against tau
// with the original's comparisons: D1 = |x0(p) - x0(p - r)|, D2 = |x1(q) - x1(q - r)| with q = (xx, y) the matched
// pixel, 10 where q or q - r (horizontally) is outside the image.  Rows / columns whose predecessor lies outside
// (the first step of a path) are never read.
__global__ void sgm_prep_kernel(const float *__restrict__ x0, const float *__restrict__ x1, int h, int w, float tau,
                                int pad, int pitch, uint8_t *__restrict__ want, uint8_t *__restrict__ code)
{
    const size_t total = (size_t)h * pitch;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(idx / pitch), xx = (int)(idx % pitch) - pad;
#pragma unroll
        for (int sd = 0; sd < 4; ++sd) {
            const int dx = sd == 0 ? 1 : (sd == 1 ? -1 : 0), dy = sd == 2 ? 1 : (sd == 3 ? -1 : 0);
            const int yp = y - dy, xp = xx - dx;
            const bool row_ok = (unsigned)yp < (unsigned)h;
            const bool inside = (unsigned)xx < (unsigned)w && (unsigned)xp < (unsigned)w && row_ok;
            const float D2 = inside ? fabsf(x1[(size_t)y * w + xx] - x1[(size_t)yp * w + xp]) : 10.f;
            code[((size_t)sd * h + y) * pitch + xx + pad] = D2 < tau ? 1 : (D2 > tau ? 2 : 0);
            if ((unsigned)xx < (unsigned)w) {
                const float D1 = inside ? fabsf(x0[(size_t)y * w + xx] - x0[(size_t)yp * w + xp]) : 0.f;
                want[((size_t)sd * h + y) * w + xx] = D1 < tau ? 1 : (D1 > tau ? 2 : 3);
            }
        }
    }
}

// both horizontal paths in one grid: blockIdx.y = 0 writes a (left->right), 1 writes b (right->left)
template <int KPL, int NW, int DIR, bool ALIGNED>
__global__ __launch_bounds__((NW + 1) * 32) void sgm_ring_h_kernel(SgmAux aux, const float *__restrict__ cvol,
                                                                   float *out_a, float *out_b, int h, int w, int D,
                                                                   SgmParams p, int G, int S)
{
    if (blockIdx.y == 0)
        sgm_ring_scan<0, 0, KPL, NW, DIR, ALIGNED>(aux, cvol, nullptr, nullptr, out_a, h, w, D, p, 0, G, S, blockIdx.x);
    else
        sgm_ring_scan<1, 0, KPL, NW, DIR, ALIGNED>(aux, cvol, nullptr, nullptr, out_b, h, w, D, p, 0, G, S, blockIdx.x);
}

// SD 2: a <- (a + b) + Lr (downward); SD 3: a <- (a + Lr) / 4 (upward)
template <int SD, int KPL, int NW, int DIR, bool ALIGNED>
__global__ __launch_bounds__((NW + 1) * 32) void sgm_ring_v_kernel(SgmAux aux, const float *__restrict__ cvol,
                                                                   float *out_a, const float *out_b, int h, int w, int D,
                                                                   SgmParams p, int G, int S)
{
    if (SD == 2)
        sgm_ring_scan<2, 2, KPL, NW, DIR, ALIGNED>(aux, cvol, out_a, out_b, out_a, h, w, D, p, 0, G, S, blockIdx.x);
    else
        sgm_ring_scan<3, 1, KPL, NW, DIR, ALIGNED>(aux, cvol, out_a, nullptr, out_a, h, w, D, p, 1, G, S, blockIdx.x);
}

struct RingShape {
    int G, S;
    size_t smem;
};

// stages of about `stage_target` bytes, as many as fit into `budget` (2..kMaxStages)
RingShape ring_shape(int arrays, int nw, int D, bool horizontal, bool aligned, size_t stage_target, size_t budget)
{
    auto stage = [&](int G) {
        return horizontal ? (size_t)arrays * nw * ring_slot_bytes(aligned, (uint32_t)G * D)
                          : (size_t)arrays * G * ring_slot_bytes(aligned, (uint32_t)nw * D);
    };
    int G = horizontal ? 16 : 4;
    while (G > 1 && stage(G) > stage_target) --G;
    int S = (int)std::min<size_t>(kMaxStages, budget / stage(G));
    while (S < 2 && G > 1) {
        --G;
        S = (int)std::min<size_t>(kMaxStages, budget / stage(G));
    }
    const size_t staging = aligned ? 0 : (size_t)nw * 2 * ((D + 3) & ~3) * 4;      // per-warp output rows
    return {G, S, (size_t)S * stage(G) + kBarBytes + staging};
}

template <int KPL, int NW, int DIR, bool ALIGNED>
int launch_ring_t(stereo_ctx *ctx, const float *x0, const float *x1, const float *cvol, float *out_a, float *out_b,
                  int h, int w, int D, const SgmParams &p, const SgmExtras &ex)
{
    // stage size / ring budget in KB per kernel; STEREO_SGM_RING="h_stage,h_budget,v2_stage,v2_budget,v3_stage,v3_budget"
    // overrides them (tuning sweeps, scripts/sgm_ring_sweep.py)
    static int cfg[6] = {32, 96, 24, 72, 16, 72};
    static const bool cfg_read = [] {
        if (const char *e = getenv("STEREO_SGM_RING")) sscanf(e, "%d,%d,%d,%d,%d,%d", cfg, cfg + 1, cfg + 2, cfg + 3, cfg + 4, cfg + 5);
        return true;
    }();
    (void)cfg_read;
    const RingShape sh_h = ring_shape(1, NW, D, true, ALIGNED, (size_t)cfg[0] * 1024, (size_t)cfg[1] * 1024);
    const RingShape sh_v2 = ring_shape(3, NW, D, false, ALIGNED, (size_t)cfg[2] * 1024, (size_t)cfg[3] * 1024);
    const RingShape sh_v3 = ring_shape(2, NW, D, false, ALIGNED, (size_t)cfg[4] * 1024, (size_t)cfg[5] * 1024);
    if (sh_h.S < 2 || sh_v2.S < 2 || sh_v3.S < 2) {
        set_error("sgm ring: disp_range %d does not fit the shared-memory stages", D);
        return STEREO_E_LIMIT;
    }
    // classified guide differences, one set per side (the two sides of the post chain run on two streams)
    SgmAux aux;
    aux.pad = 32 * KPL + 4;
    aux.pitch = w + 2 * aux.pad;
    DevBuf &buf = ctx->sgm_aux[DIR > 0];
    const size_t want_bytes = (size_t)4 * h * w, code_bytes = (size_t)4 * h * aux.pitch;
    STEREO_TRY(buf.reserve(want_bytes + code_bytes));
    uint8_t *want = buf.as<uint8_t>(), *code = want + want_bytes;
    aux.want = want;
    aux.code = code;
    aux.final_scale = ex.final_scale;
    aux.wta = ex.wta;
    sgm_prep_kernel<<<std::min<size_t>(((size_t)h * aux.pitch + 255) / 256, 148 * 8), 256, 0, ctx->stream>>>(
        x0, x1, h, w, p.tau_so, aux.pad, aux.pitch, want, code);

    auto kh = sgm_ring_h_kernel<KPL, NW, DIR, ALIGNED>;
    auto kv2 = sgm_ring_v_kernel<2, KPL, NW, DIR, ALIGNED>;
    auto kv3 = sgm_ring_v_kernel<3, KPL, NW, DIR, ALIGNED>;
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(kh, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh_h.smem));
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(kv2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh_v2.smem));
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(kv3, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh_v3.smem));
    const int threads = (NW + 1) * 32;
    kh<<<dim3((h + NW - 1) / NW, 2), threads, sh_h.smem, ctx->stream>>>(aux, cvol, out_a, out_b, h, w, D, p, sh_h.G, sh_h.S);
    kv2<<<(w + NW - 1) / NW, threads, sh_v2.smem, ctx->stream>>>(aux, cvol, out_a, out_b, h, w, D, p, sh_v2.G, sh_v2.S);
    kv3<<<(w + NW - 1) / NW, threads, sh_v3.smem, ctx->stream>>>(aux, cvol, out_a, out_b, h, w, D, p, sh_v3.G, sh_v3.S);
    ctx->launches += 4;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

template <int DIR, bool ALIGNED>
int launch_ring_dir(stereo_ctx *ctx, const float *x0, const float *x1, const float *cvol, float *out, float *scratch,
                    int h, int w, int D, const SgmParams &p, const SgmExtras &ex)
{
    if (D <= 128) return launch_ring_t<4, 4, DIR, ALIGNED>(ctx, x0, x1, cvol, out, scratch, h, w, D, p, ex);
    if (D <= 256) return launch_ring_t<8, 4, DIR, ALIGNED>(ctx, x0, x1, cvol, out, scratch, h, w, D, p, ex);
    return launch_ring_t<16, 4, DIR, ALIGNED>(ctx, x0, x1, cvol, out, scratch, h, w, D, p, ex);
}

}  // namespace

int launch_sgm_ring(stereo_ctx *ctx, const float *x0, const float *x1, const float *cvol, float *out, float *scratch,
                    int h, int w, int D, const SgmParams &p, int direction, const SgmExtras &ex)
{
    // bulk copies and 16-byte accesses need every [x][d] row to start 16-byte aligned
    const bool aligned = D % 4 == 0 && ((uintptr_t)cvol % 16 == 0) && ((uintptr_t)out % 16 == 0) &&
                         ((uintptr_t)scratch % 16 == 0);
    if (aligned)
        return direction > 0 ? launch_ring_dir<1, true>(ctx, x0, x1, cvol, out, scratch, h, w, D, p, ex)
                             : launch_ring_dir<-1, true>(ctx, x0, x1, cvol, out, scratch, h, w, D, p, ex);
    return direction > 0 ? launch_ring_dir<1, false>(ctx, x0, x1, cvol, out, scratch, h, w, D, p, ex)
                         : launch_ring_dir<-1, false>(ctx, x0, x1, cvol, out, scratch, h, w, D, p, ex);
}

}  // namespace stereo
