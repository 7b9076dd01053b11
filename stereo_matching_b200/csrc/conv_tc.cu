// Siamese trunk on the 5th-generation tensor cores: 3x3 / 64->64 'valid' convolution as an
// implicit GEMM with tcgen05.mma, accumulators in TMEM, operands staged in shared memory by
// TMA bulk copies (/root/reference/models/model.py:33-39, layers 2-4 of base_model_ws_9).
//
// Precision.  Single-pass bf16 operands miss the 1e-4 cost-volume gate on the ckpt-1000
// weights, so activations and weights are kept as a bf16 (hi, lo) pair
// with x ~= hi + lo and each product is formed as lo*hi + hi*lo + hi*hi in the fp32 TMEM
// accumulator (STEREO_PREC_BF16X3).  STEREO_PREC_BF16 issues only the hi*hi pass.
//
// Data layout.  Activations live in HBM "chunk-planar": [image][y][c/8][x][c%8] bf16, one
// plane for hi and one for lo.  For a fixed (y, c/8) a run of pixels is contiguous, 16 bytes
// per pixel - exactly one row of a no-swizzle K-major UMMA core matrix.  A CTA stages an
// input row as [plane][c/8][130 px][16 B]; the A operand of tap (ky, kx) is then the same
// tile with the start address advanced by kx*16 bytes (ky picks the ring slot), SBO = 128 B,
// LBO = 130*16 B.  No im2col copy is ever made and every input row is fetched once per CTA.
//
// Work split.  A CTA owns a strip of 128 output pixels x all 64 output channels and walks
// down `rows_per_cta` output rows: M = 128 pixels, N = 64 channels, K = 9 taps x 64 channels
// x 3 operand-split passes = 108 MMAs per output row.  Shared memory (227 KB) holds the hi
// weights of all taps (72 KB), a 4-slot ring of input rows (4 x 32.5 KB, hi + lo) and a
// 2-stage ring through which the lo weights stream tap by tap (they are needed by one pass in
// three, so streaming them costs 72 KB of L2 traffic per row instead of 72 KB of residency).
//
// Warp roles (256 threads): warp 0 = TMA producer for input rows, warps 1 and 2 = MMA issuers
// (a single thread cannot issue 108 MMAs per row fast enough; the two issuers split the K loop
// and accumulate into separate TMEM blocks that the epilogue adds), warp 3 = TMA producer for
// the lo-weight ring, warps 4-7 = epilogue (TMEM -> registers -> bias / ReLU -> (hi, lo)
// split or fp32 -> HBM).  Warp 2 also allocates TMEM.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {

namespace tc {

constexpr int kStripW = 128;               // output pixels per CTA row (UMMA M)
constexpr int kPW = kStripW + 2;           // input pixels per staged row
constexpr int kN = 64;                     // output channels (UMMA N)
constexpr int kRunBytes = kPW * 16;        // one (plane, channel-chunk) run of a staged row
constexpr int kPlaneBytes = 8 * kRunBytes; // 8 channel chunks
constexpr int kSlotBytes = 2 * kPlaneBytes;
constexpr int kSlots = 4;
constexpr int kWTapBytes = 8 * kN * 16;    // one tap, one plane: [kc][n][8] bf16 = 8 KB
constexpr int kWHiBytes = 9 * kWTapBytes;  // 73,728
constexpr int kWLoStages = 2;
constexpr int kSmemWLo = kWHiBytes;
constexpr int kSmemRows = kSmemWLo + kWLoStages * kWTapBytes;
constexpr int kSmemBias = kSmemRows + kSlots * kSlotBytes;
constexpr int kSmemBars = kSmemBias + kN * 4;
constexpr int kSmemTotal = kSmemBars + 24 * 8 + 16;
constexpr int kIssuers = 2;
constexpr uint32_t kTmemCols = 256;        // 2 buffers x kIssuers partial accumulators x 64 columns

struct Params {
    const __nv_bfloat16 *in_hi, *in_lo;   // chunk-planar [img][h_in][8][wp_in][8]
    int h_in, w_in, wp_in;
    const __nv_bfloat16 *wimg;            // [plane][tap][kc][n][8] for this layer
    const float *bias;                    // [64]
    __nv_bfloat16 *out_hi, *out_lo;       // chunk-planar [img][h_out][8][wp_out][8] (or null)
    int wp_out;
    float *out_f32;                       // NHWC [img][h_out][w_out][64] (or null)
    int relu, planes, rows_per_cta;
};

__global__ __launch_bounds__(256, 1) void conv3x3_tc_kernel(Params p)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int h_out = p.h_in - 2, w_out = p.w_in - 2;
    const int x0 = blockIdx.x * kStripW;
    const int y0 = blockIdx.y * p.rows_per_cta;
    const int rows_out = min(p.rows_per_cta, h_out - y0);
    const int img = blockIdx.z;
    const int rows_in = rows_out + 2;

    const uint32_t s_whi = smem_u32(smem);
    const uint32_t s_wlo = s_whi + kSmemWLo;
    const uint32_t s_rows = s_whi + kSmemRows;
    float *s_bias = reinterpret_cast<float *>(smem + kSmemBias);
    const uint32_t s_bar = s_whi + kSmemBars;
    const uint32_t bar_full = s_bar;               // [4] input row landed
    const uint32_t bar_empty = s_bar + 32;         // [4] input row consumed by both issuers
    const uint32_t bar_w = s_bar + 64;             // [1] hi weights landed
    const uint32_t bar_acc_full = s_bar + 72;      // [2] both partial accumulators complete
    const uint32_t bar_acc_empty = s_bar + 88;     // [2] epilogue has drained the buffer
    const uint32_t bar_wlo_full = s_bar + 104;     // [2] lo-weight tap landed
    const uint32_t bar_wlo_empty = s_bar + 120;    // [2] lo-weight tap consumed
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + kSmemBars + 24 * 8);

    if (tid == 0) {
        for (int i = 0; i < kSlots; ++i) {
            mbar_init(bar_full + 8 * i, 1);
            mbar_init(bar_empty + 8 * i, kIssuers);
        }
        mbar_init(bar_w, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar_acc_full + 8 * i, kIssuers);
            mbar_init(bar_acc_empty + 8 * i, 4);
            mbar_init(bar_wlo_full + 8 * i, 1);
            mbar_init(bar_wlo_empty + 8 * i, 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < kN) s_bias[tid] = __ldg(p.bias + tid);
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(kTmemCols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint8_t *wsrc = reinterpret_cast<const uint8_t *>(p.wimg);

    if (warp == 0) {
        if (lane == 0) {
            // ---- TMA producer: hi weights once, then the input rows through the ring ----
            mbar_expect_tx(bar_w, kWHiBytes);
            for (int tap = 0; tap < 9; ++tap)
                bulk_g2s(s_whi + tap * kWTapBytes, wsrc + (size_t)tap * kWTapBytes, kWTapBytes, bar_w);
            const size_t row_elems = (size_t)8 * p.wp_in * 8;     // bf16 elements per image row
            for (int r = 0; r < rows_in; ++r) {
                const int slot = r & (kSlots - 1);
                mbar_wait(bar_empty + 8 * slot, ((r >> 2) & 1) ^ 1);
                mbar_expect_tx(bar_full + 8 * slot, (uint32_t)p.planes * kPlaneBytes);
                const size_t row_off = ((size_t)img * p.h_in + (y0 + r)) * row_elems + (size_t)x0 * 8;
                for (int pl = 0; pl < p.planes; ++pl) {
                    const __nv_bfloat16 *src = (pl == 0 ? p.in_hi : p.in_lo) + row_off;
                    const uint32_t dst = s_rows + slot * kSlotBytes + pl * kPlaneBytes;
                    for (int cc = 0; cc < 8; ++cc)
                        bulk_g2s(dst + cc * kRunBytes, src + (size_t)cc * p.wp_in * 8, kRunBytes, bar_full + 8 * slot);
                }
            }
        }
    } else if (warp == 3) {
        if (lane == 0 && p.planes == 2) {
            // ---- TMA producer for the lo weights: one tap per stage, 9 taps per output row ----
            int k = 0;
            for (int j = 0; j < rows_out; ++j)
                for (int tap = 0; tap < 9; ++tap, ++k) {
                    const int st = k & 1;
                    mbar_wait(bar_wlo_empty + 8 * st, ((k >> 1) & 1) ^ 1);
                    mbar_expect_tx(bar_wlo_full + 8 * st, kWTapBytes);
                    bulk_g2s(s_wlo + st * kWTapBytes, wsrc + (size_t)kWHiBytes + (size_t)tap * kWTapBytes, kWTapBytes,
                             bar_wlo_full + 8 * st);
                }
        }
    } else if (warp == 1 || warp == 2) {
        if (lane == 0) {
            // ---- MMA issuers ----
            // issuer 0: A_hi * W_lo for all taps (through the lo-weight ring), A_hi * W_hi taps 0-3
            // issuer 1: A_lo * W_hi for all taps,                             A_hi * W_hi taps 4-8
            const int issuer = warp - 1;
            constexpr uint32_t idesc = make_idesc(kStripW, kN);
            const uint32_t d_hi = desc_hi(128);
            int k = 0;   // lo-weight stage counter (issuer 0)
            mbar_wait(bar_w, 0);
            for (int j = 0; j < rows_out; ++j) {
                const int b = j & 1;
                mbar_wait(bar_acc_empty + 8 * b, ((j >> 1) & 1) ^ 1);
                if (j == 0) {
                    mbar_wait(bar_full + 0, 0);
                    mbar_wait(bar_full + 8, 0);
                }
                mbar_wait(bar_full + 8 * ((j + 2) & 3), ((j + 2) >> 2) & 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(b * kIssuers + issuer) * kN;
                uint32_t acc = 0;
                // one unit = one tap of one operand combination = 4 MMAs (K = 64 channels)
                auto unit = [&](int tap, int apl, uint32_t w_addr) {
                    const int ky = tap / 3, kx = tap - ky * 3;
                    const uint32_t a_lo = desc_lo(s_rows + ((j + ky) & 3) * kSlotBytes + apl * kPlaneBytes + kx * 16, kRunBytes);
                    const uint32_t b_lo = desc_lo(w_addr, kN * 16);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        // K steps add compile-time constants (multiples of 16 bytes, sums below 2^18)
                        umma_bf16_parts(tmem_d, a_lo + ((2 * ks * kRunBytes) >> 4), d_hi,
                                        b_lo + ((2 * ks * (kN * 16)) >> 4), d_hi, idesc, acc);
                        acc = 1;
                    }
                };
                if (p.planes == 2) {
                    if (issuer == 0) {
#pragma unroll 1
                        for (int tap = 0; tap < 9; ++tap, ++k) {
                            const int st = k & 1;
                            mbar_wait(bar_wlo_full + 8 * st, (k >> 1) & 1);
                            tc_fence_after();
                            unit(tap, 0, s_wlo + st * kWTapBytes);
                            tc_commit(bar_wlo_empty + 8 * st);
                        }
                    } else {
#pragma unroll 1
                        for (int tap = 0; tap < 9; ++tap) unit(tap, 1, s_whi + tap * kWTapBytes);
                    }
                }
                const int t0 = issuer == 0 ? 0 : 4, t1 = issuer == 0 ? 4 : 9;
#pragma unroll 1
                for (int tap = t0; tap < t1; ++tap) unit(tap, 0, s_whi + tap * kWTapBytes);
                tc_commit(bar_empty + 8 * (j & 3));      // this issuer no longer needs input row j
                tc_commit(bar_acc_full + 8 * b);         // this issuer's partial accumulator is complete
            }
            // rows j+1, j+2 of the last step are never released through the ring; nothing waits on them
        }
    } else if (warp >= 4) {
        // ---- epilogue: TMEM -> registers -> bias / ReLU -> (hi, lo) or fp32 -> HBM ----
        const int q = warp & 3;                      // TMEM lane quarter this warp may touch
        const int px = x0 + q * 32 + lane;
        for (int j = 0; j < rows_out; ++j) {
            const int b = j & 1;
            mbar_wait(bar_acc_full + 8 * b, (j >> 1) & 1);
            tc_fence_after();
            const int y = y0 + j;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(b * kIssuers) * kN;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                uint32_t v[32], v2[32];
                tmem_ld_32x32(taddr + half * 32, v);
                tmem_ld_32x32(taddr + kN + half * 32, v2);
                if (half == 1) {
                    // everything has been read: the issuers may overwrite this buffer
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_acc_empty + 8 * b);
                }
                if (px < w_out) {
                    float f[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        f[i] = (__uint_as_float(v[i]) + __uint_as_float(v2[i])) + s_bias[half * 32 + i];
                        if (p.relu) f[i] = fmaxf(f[i], 0.f);
                    }
                    if (p.out_f32) {
                        float4 *dst = reinterpret_cast<float4 *>(
                            p.out_f32 + (((size_t)img * h_out + y) * w_out + px) * 64 + half * 32);
#pragma unroll
                        for (int i = 0; i < 8; ++i) dst[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
                    } else {
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            uint32_t hi[4], lo[4];
#pragma unroll
                            for (int e = 0; e < 4; ++e) split_bf16x2(f[8 * c + 2 * e], f[8 * c + 2 * e + 1], hi[e], lo[e]);
                            const size_t off = ((((size_t)img * h_out + y) * 8 + (half * 4 + c)) * p.wp_out + px) * 8;
                            *reinterpret_cast<uint4 *>(p.out_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                            if (p.planes == 2)
                                *reinterpret_cast<uint4 *>(p.out_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                        }
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
    }
}

// fp32 packed weights (one 3x3 64->64 layer: [tap][cin][cout]) -> UMMA B operand images
// [plane][tap][kc][n][8] bf16 with plane 0 = hi, plane 1 = lo
__global__ void weight_image_kernel(const float *__restrict__ w, __nv_bfloat16 *__restrict__ img)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;      // over [tap][kc][n][k8]
    if (i >= 9 * 8 * 64 * 8) return;
    int k8 = i & 7, n = (i >> 3) & 63, kc = (i >> 9) & 7, tap = i >> 12;
    float v = w[((size_t)tap * 64 + kc * 8 + k8) * 64 + n];
    __nv_bfloat16 hi = __float2bfloat16_rn(v);
    __nv_bfloat16 lo = __float2bfloat16_rn(v - __bfloat162float(hi));
    img[i] = hi;
    img[i + 9 * 8 * 64 * 8] = lo;
}

}  // namespace tc

size_t tc_weight_image_elems() { return (size_t)2 * 9 * 8 * 64 * 8; }   // per layer

int launch_weight_images(stereo_ctx *ctx, const float *layer_w, void *img)
{
    tc::weight_image_kernel<<<(9 * 8 * 64 * 8 + 255) / 256, 256, 0, ctx->stream>>>(
        layer_w, static_cast<__nv_bfloat16 *>(img));
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int tc_pitch(int w) { return (w + tc::kStripW - 1) / tc::kStripW * tc::kStripW + 8; }

int launch_conv_tc(stereo_ctx *ctx, const void *in_hi, const void *in_lo, int n, int h_in, int w_in, int wp_in,
                   const void *wimg, const float *bias, void *out_hi, void *out_lo, int wp_out, float *out_f32,
                   int relu, int planes)
{
    STEREO_REQUIRE(h_in >= 3 && w_in >= 3, "input %dx%d smaller than the 3x3 kernel", h_in, w_in);
    const int h_out = h_in - 2, w_out = w_in - 2;
    const int strips = (w_out + tc::kStripW - 1) / tc::kStripW;
    STEREO_REQUIRE(wp_in >= strips * tc::kStripW + 2, "activation pitch %d too small for %d strips", wp_in, strips);
    tc::Params p;
    p.in_hi = static_cast<const __nv_bfloat16 *>(in_hi);
    p.in_lo = static_cast<const __nv_bfloat16 *>(in_lo);
    p.h_in = h_in; p.w_in = w_in; p.wp_in = wp_in;
    p.wimg = static_cast<const __nv_bfloat16 *>(wimg);
    p.bias = bias;
    p.out_hi = static_cast<__nv_bfloat16 *>(out_hi);
    p.out_lo = static_cast<__nv_bfloat16 *>(out_lo);
    p.wp_out = wp_out;
    p.out_f32 = out_f32;
    p.relu = relu; p.planes = planes;
    // one CTA per SM: enough row segments for a few waves, few enough to amortise the resident
    // weights and the two halo rows
    int column_ctas = strips * n;
    int segs = (ctx->sm_count * 2) / column_ctas;     // at most two full waves
    if (segs < 1) segs = 1;
    int rows = (h_out + segs - 1) / segs;
    if (rows < 8) rows = 8;
    p.rows_per_cta = rows;
    segs = (h_out + rows - 1) / rows;
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(tc::conv3x3_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           tc::kSmemTotal));
    ProfScope prof(ctx, STEREO_K_CONV);
    tc::conv3x3_tc_kernel<<<dim3(strips, segs, n), 256, tc::kSmemTotal, ctx->stream>>>(p);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
