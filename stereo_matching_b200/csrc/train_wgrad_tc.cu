// Weight gradient of a 3x3 / 64 -> 64 'valid' convolution on the tensor cores (training step, csrc/train.cu;
// /root/reference/train.py:92-112 leaves it to TensorFlow's autograd):
//
//     dW[ky][kx][ci][co] = sum over output positions p = (b, y, x) of  x[b][y + ky][x + kx][ci] * dz[p][co]
//
// i.e. nine GEMMs [64 ci] x [64 co] whose K dimension is the position.  tcgen05 wants both operands K-major in shared
// memory ([K / 8][row][8 elements], the layout of the correlation kernels) while the tensors are position-major fp32, so
// "builder" warps make the operands: a lane owns one channel, gathers a run of consecutive pixels of it (coalesced
// 128-byte loads per warp), splits them into bf16 hi / lo and writes one 16-byte piece (eight positions) per plane and
// tap - the transposition happens in registers and the shared-memory stores are conflict free.  Positions are counted
// over rows padded to a multiple of eight pixels (dz = 0 on the padding), so that a piece never leaves its image row:
// ten pixels of an input row then serve the three kx taps of eight positions (30 loads per channel instead of 72).
// The nine taps are nine groups of 64 rows of one A tile (two taps = one M = 128 MMA; a tenth group of zero rows pads
// the last MMA), dz is the B tile (N = 64):
//
//     per 16 positions: 5 row groups x 3 passes (A_lo * B_hi, A_hi * B_lo, A_hi * B_hi) = 15 MMAs into 5 x 64 TMEM columns
//
// A CTA takes a contiguous range of positions (the grid splits K), accumulates it in tensor memory in fp32 and writes its
// [9][64][64] partial result; a second kernel sums the partials in double, like the chunk sums of the fp32 kernel.
// The builders bound the kernel (16 warps, the loads of two tasks in flight per warp): 75 us for the right tower's
// largest layer (185 000 positions) against 464 us for the fp32 shared-memory GEMMs; the first version (8 warps, every
// tap gathered on its own, one task in flight) took 313 us.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {

namespace wg {

constexpr int kTaps = 9, kC = 64;
constexpr int kRows = 640;                         // 9 taps x 64 ci + 64 zero rows
constexpr int kStagePos = 32;                      // positions per stage (two K = 16 steps)
constexpr int kK8 = kStagePos / 8;
constexpr int kAPlane = kK8 * kRows * 16;          // one plane of A: [k8][row][8 x bf16] = 40 KB
constexpr int kBPlane = kK8 * kC * 16;             // one plane of B: 4 KB
constexpr int kStageBytes = 2 * kAPlane + 2 * kBPlane;      // hi + lo: 88 KB
constexpr int kStages = 2;
constexpr int kBars = kStages * kStageBytes;
constexpr int kSmemTotal = kBars + 64;
static_assert(kSmemTotal <= 232448, "shared memory budget of one CTA");
static_assert(kK8 == 4, "the builders' task numbering assumes four groups of eight positions per stage");
constexpr int kBuildWarps = 16;
constexpr int kThreads = 64 + 32 * kBuildWarps;    // warp 0: MMA issuer, warp 1: tensor-memory allocation, then builders
constexpr uint32_t kTmemCols = 512;
constexpr int kGroups = 5;                         // M = 128 row groups of the A tile

struct Params {
    const float *x, *dz;                           // [n][h][w][64], [n][ho][wo][64]
    int n, h, w, ho, wo;
    int gpr;                                       // groups of eight positions per output row = ceil(wo / 8)
    long long groups;                              // n * ho * gpr
    long long per_cta;                             // groups per CTA (a multiple of kK8)
    float *partial;                                // [grid][9][64][64]
};

__global__ __launch_bounds__(kThreads, 1) void wgrad_tc_kernel(Params p)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t s_base = tc::smem_u32(smem);
    const uint32_t bar_full = s_base + kBars;                 // [2] builders -> issuer
    const uint32_t bar_empty = bar_full + 16;                 // [2] MMAs have read the stage
    const uint32_t bar_done = bar_full + 32;                  // all MMAs of the CTA complete
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + kBars + 48);

    if (tid == 0) {
        for (int i = 0; i < kStages; ++i) {
            tc::mbar_init(bar_full + 8 * i, kBuildWarps);
            tc::mbar_init(bar_empty + 8 * i, 1);
        }
        tc::mbar_init(bar_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc::smem_u32(tmem_slot)),
                     "r"(kTmemCols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    // the zero rows of both stages and planes (rows 576..639 of every k8 group) are written once
    for (int i = tid; i < kStages * 2 * kK8 * 64; i += kThreads) {
        const int r = i % 64, k8 = (i / 64) % kK8, pl = (i / (64 * kK8)) % 2, st = i / (64 * kK8 * 2);
        *reinterpret_cast<uint4 *>(smem + st * kStageBytes + pl * kAPlane + (k8 * kRows + 576 + r) * 16) = make_uint4(0, 0, 0, 0);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const long long g0 = (long long)blockIdx.x * p.per_cta;
    const long long g1 = g0 + p.per_cta < p.groups ? g0 + p.per_cta : p.groups;
    const int n_stages = g1 > g0 ? (int)((g1 - g0 + kK8 - 1) / kK8) : 0;

    if (warp == 0) {
        if (lane == 0) {
            // ---- MMA issuer ----
            constexpr uint32_t idesc = tc::make_idesc(128, kC);
            const uint32_t d_hi = tc::desc_hi(128);
            for (int s = 0; s < n_stages; ++s) {
                const int st = s % kStages;
                tc::mbar_wait(bar_full + 8 * st, (s / kStages) & 1);
                tc::tc_fence_after();
                const uint32_t a_base = s_base + st * kStageBytes, b_base = a_base + 2 * kAPlane;
#pragma unroll
                for (int k16 = 0; k16 < kStagePos / 16; ++k16) {
                    const uint32_t b_hi = tc::desc_lo(b_base + k16 * 2 * kC * 16, kC * 16);
                    const uint32_t b_lo = tc::desc_lo(b_base + kBPlane + k16 * 2 * kC * 16, kC * 16);
#pragma unroll
                    for (int g = 0; g < kGroups; ++g) {
                        const uint32_t a_hi = tc::desc_lo(a_base + (k16 * 2 * kRows + 128 * g) * 16, kRows * 16);
                        const uint32_t a_lo = tc::desc_lo(a_base + kAPlane + (k16 * 2 * kRows + 128 * g) * 16, kRows * 16);
                        const uint32_t acc = tmem_base + g * kC;
                        tc::umma_bf16_parts(acc, a_lo, d_hi, b_hi, d_hi, idesc, (s > 0 || k16 > 0) ? 1u : 0u);
                        tc::umma_bf16_parts(acc, a_hi, d_hi, b_lo, d_hi, idesc, 1u);
                        tc::umma_bf16_parts(acc, a_hi, d_hi, b_hi, d_hi, idesc, 1u);
                    }
                }
                tc::tc_commit(bar_empty + 8 * st);
            }
            tc::tc_commit(bar_done);
        }
    } else if (warp >= 2) {
        // ---- builders ----
        const int bw = warp - 2;
        // Task t of a stage: t < 24: (k8 group, ky, channel half) of x - ten pixels, three kx tiles; t >= 24: (k8 group,
        // channel half) of dz.  Every warp takes tasks bw and bw + 16 and issues the loads of both before it converts.
        auto load_task = [&](int t, long long gs, float (&v)[10]) {
            const bool is_b = t >= 24;
            const int k8 = is_b ? (t - 24) >> 1 : t / 6, ky = is_b ? 0 : (t % 6) >> 1;
            const int ch = 32 * (t & 1) + lane;
            const long long gi = gs + k8;
#pragma unroll
            for (int j = 0; j < 10; ++j) v[j] = 0.f;
            if (gi >= g1) return;
            const long long row = gi / p.gpr;                    // b * ho + y
            const int x0 = 8 * (int)(gi - row * p.gpr);
            if (is_b) {
                const float *src = p.dz + ((row * p.wo + x0) * kC + ch);
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (x0 + j < p.wo) v[j] = __ldg(src + j * kC);
            } else {
                const long long b = row / p.ho;
                const int y = (int)(row - b * p.ho);
                const float *src = p.x + (((b * p.h + y + ky) * p.w + x0) * kC + ch);
#pragma unroll
                for (int j = 0; j < 10; ++j)
                    if (x0 + j < p.w) v[j] = __ldg(src + j * kC);
            }
        };
        auto store_task = [&](int t, int st, const float (&v)[10]) {
            const bool is_b = t >= 24;
            const int k8 = is_b ? (t - 24) >> 1 : t / 6, ky = is_b ? 0 : (t % 6) >> 1;
            const int ch = 32 * (t & 1) + lane;
            uint8_t *a_hi = smem + st * kStageBytes, *a_lo = a_hi + kAPlane;
            uint8_t *b_hi = a_hi + 2 * kAPlane, *b_lo = b_hi + kBPlane;
            uint32_t hv[10], lv[10];                              // bf16 bit patterns of the hi and lo parts
#pragma unroll
            for (int j = 0; j < 10; ++j) {
                const __nv_bfloat16 h = __float2bfloat16_rn(v[j]);
                hv[j] = __bfloat16_as_ushort(h);
                lv[j] = __bfloat16_as_ushort(__float2bfloat16_rn(v[j] - __bfloat162float(h)));
            }
            if (is_b) {
                const int idx = (k8 * kC + ch) * 16;
                *reinterpret_cast<uint4 *>(b_hi + idx) = make_uint4(hv[0] | (hv[1] << 16), hv[2] | (hv[3] << 16),
                                                                    hv[4] | (hv[5] << 16), hv[6] | (hv[7] << 16));
                *reinterpret_cast<uint4 *>(b_lo + idx) = make_uint4(lv[0] | (lv[1] << 16), lv[2] | (lv[3] << 16),
                                                                    lv[4] | (lv[5] << 16), lv[6] | (lv[7] << 16));
            } else {
#pragma unroll
                for (int kx = 0; kx < 3; ++kx) {
                    const int idx = (k8 * kRows + (3 * ky + kx) * kC + ch) * 16;
                    *reinterpret_cast<uint4 *>(a_hi + idx) =
                        make_uint4(hv[kx] | (hv[kx + 1] << 16), hv[kx + 2] | (hv[kx + 3] << 16),
                                   hv[kx + 4] | (hv[kx + 5] << 16), hv[kx + 6] | (hv[kx + 7] << 16));
                    *reinterpret_cast<uint4 *>(a_lo + idx) =
                        make_uint4(lv[kx] | (lv[kx + 1] << 16), lv[kx + 2] | (lv[kx + 3] << 16),
                                   lv[kx + 4] | (lv[kx + 5] << 16), lv[kx + 6] | (lv[kx + 7] << 16));
                }
            }
        };
        for (int s = 0; s < n_stages; ++s) {
            const int st = s % kStages;
            const long long gs = g0 + (long long)s * kK8;
            float va[10], vb[10];
            load_task(bw, gs, va);
            load_task(bw + 16, gs, vb);
            if (s >= kStages) tc::mbar_wait(bar_empty + 8 * st, ((s / kStages) - 1) & 1);
            store_task(bw, st, va);
            store_task(bw + 16, st, vb);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> tensor core reads
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(bar_full + 8 * st);
        }
        // ---- epilogue (builder warps 2-5 = tensor-memory lane quarters 2, 3, 0, 1): the CTA's partial sums ----
        if (bw < 4) {
            const int q = warp & 3;
            if (n_stages > 0) {
                tc::mbar_wait(bar_done, 0);
                tc::tc_fence_after();
            }
            const int r = 32 * q + lane;                       // row of the M = 128 group: (tap parity, ci)
            float *out = p.partial + (size_t)blockIdx.x * kTaps * kC * kC;
#pragma unroll 1
            for (int g = 0; g < kGroups; ++g) {
                const int tap = 2 * g + r / kC, ci = r % kC;
#pragma unroll
                for (int c0 = 0; c0 < kC; c0 += 32) {
                    uint32_t v[32];
                    if (n_stages > 0) tc::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + g * kC + c0, v);
                    else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) v[i] = 0;
                    }
                    if (tap < kTaps) {
                        uint4 *dst = reinterpret_cast<uint4 *>(out + ((size_t)tap * kC + ci) * kC + c0);
#pragma unroll
                        for (int i = 0; i < 8; ++i) dst[i] = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                    }
                }
            }
        }
    }

    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc::tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
    }
}

// red_w[i] += sum over the CTAs' partial results (double)
__global__ void wgrad_reduce_kernel(const float *__restrict__ partial, int parts, int total, double *__restrict__ red_w)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    double s = 0.0;
#pragma unroll 8
    for (int c = 0; c < parts; ++c) s += (double)__ldg(partial + (size_t)c * total + i);
    red_w[i] += s;
}

}  // namespace wg

size_t train_wgrad_tc_scratch_bytes(const stereo_ctx *ctx) { return (size_t)ctx->sm_count * wg::kTaps * wg::kC * wg::kC * sizeof(float); }

// dW of a 3x3 / 64 -> 64 layer into red_w (double, [3][3][64][64], accumulated); `scratch`: train_wgrad_tc_scratch_bytes()
int launch_train_wgrad_tc(stereo_ctx *ctx, const float *x, const float *dz, int n, int h, int w, float *scratch,
                          double *red_w)
{
    wg::Params p;
    p.x = x; p.dz = dz;
    p.n = n; p.h = h; p.w = w; p.ho = h - 2; p.wo = w - 2;
    p.gpr = (p.wo + 7) / 8;
    p.groups = (long long)n * p.ho * p.gpr;
    long long per = (p.groups + ctx->sm_count - 1) / ctx->sm_count;
    per = (per + wg::kK8 - 1) / wg::kK8 * wg::kK8;
    p.per_cta = per;
    const int grid = (int)((p.groups + per - 1) / per);
    p.partial = scratch;
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(wg::wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, wg::kSmemTotal));
    wg::wgrad_tc_kernel<<<grid, wg::kThreads, wg::kSmemTotal, ctx->stream>>>(p);
    const int total = wg::kTaps * wg::kC * wg::kC;
    wg::wgrad_reduce_kernel<<<(total + 255) / 256, 256, 0, ctx->stream>>>(scratch, grid, total, red_w);
    ctx->launches += 2;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
