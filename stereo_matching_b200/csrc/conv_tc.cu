// Siamese trunk on the 5th-generation tensor cores: 3x3 / 64->64 'valid' convolution as an
// implicit GEMM with tcgen05.mma, accumulators in TMEM, operands staged in shared memory by
// TMA bulk copies (/root/reference/models/model.py:33-39, layers 2-4 of base_model_ws_9).
//
// Precision.  Single-pass bf16 operands miss the 1e-4 cost-volume gate on the ckpt-1000
// weights, so activations and weights are kept as a bf16 (hi, lo) pair
// with x ~= hi + lo and each product is formed as hi*hi + hi*lo + lo*hi in the fp32 TMEM
// accumulator (STEREO_PREC_BF16X3).  STEREO_PREC_BF16 issues only the hi*hi pass.
//
// Data layout.  Activations live in HBM "chunk-planar": [image][y][c/8][x][c%8] bf16, one
// plane for hi and one for lo.  For a fixed (y, c/8) a run of pixels is contiguous, 16 bytes
// per pixel - exactly one row of a no-swizzle K-major UMMA core matrix.  A CTA stages an
// input row as [plane][c/8][130 px][16 B]; the A operand of tap (ky, kx) is then the same
// tile with the start address advanced by kx*16 bytes (ky picks the ring slot), SBO = 128 B,
// LBO = 130*16 B.  No im2col copy is ever made and every input row is fetched once per CTA.
//
// Work split.  A CTA owns a strip of 128 output pixels x 32 output channels and walks down
// `rows_per_cta` output rows: M = 128 pixels, N = 32 channels, K = 9 taps x 64 channels.
// The 32-channel half of the weights (9 taps x hi/lo x 4 KB = 72 KB) stays resident in shared
// memory, input rows stream through a 4-slot ring, and two 32-column TMEM accumulators
// alternate between the MMA issuer and the epilogue warps.
//
// Warp roles (256 threads): warp 0 = TMA producer, warp 1 = MMA issuer, warp 2 = TMEM
// allocator, warps 4-7 = epilogue (TMEM -> registers -> bias/ReLU -> hi/lo split -> HBM).
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {

namespace tc {

constexpr int kStripW = 128;               // output pixels per CTA row (UMMA M)
constexpr int kPW = kStripW + 2;           // input pixels per staged row
constexpr int kNHalf = 32;                 // output channels per CTA (UMMA N)
constexpr int kRunBytes = kPW * 16;        // one (plane, channel-chunk) run of a staged row
constexpr int kPlaneBytes = 8 * kRunBytes; // 8 channel chunks
constexpr int kSlotBytes = 2 * kPlaneBytes;
constexpr int kSlots = 4;
constexpr int kWTapBytes = 8 * kNHalf * 16;            // one tap, one plane: [kc][n][8] bf16
constexpr int kWBytes = 9 * 2 * kWTapBytes;            // 73,728
constexpr int kSmemBars = kWBytes + kSlots * kSlotBytes;
constexpr int kSmemTotal = kSmemBars + 16 * 8 + 16;
constexpr int kIssuers = 2;
constexpr uint32_t kTmemCols = 128;             // 2 buffers x kIssuers partial accumulators x 32 columns

struct Params {
    const __nv_bfloat16 *in_hi, *in_lo;   // chunk-planar [img][h_in][8][wp_in][8]
    int h_in, w_in, wp_in;
    const __nv_bfloat16 *wimg;            // [nh][tap][plane][kc][n][8] for this layer
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
    const int img = blockIdx.z >> 1, nh = blockIdx.z & 1;
    const int rows_in = rows_out + 2;

    const uint32_t s_w = smem_u32(smem);
    const uint32_t s_rows = s_w + kWBytes;
    const uint32_t s_bar = s_w + kSmemBars;
    const uint32_t bar_full = s_bar;              // [4]
    const uint32_t bar_empty = s_bar + 32;        // [4]
    const uint32_t bar_w = s_bar + 64;            // [1]
    const uint32_t bar_acc_full = s_bar + 72;     // [2]
    const uint32_t bar_acc_empty = s_bar + 88;    // [2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + kSmemBars + 16 * 8);

    if (tid == 0) {
        for (int i = 0; i < kSlots; ++i) {
            mbar_init(bar_full + 8 * i, 1);
            mbar_init(bar_empty + 8 * i, kIssuers);
        }
        mbar_init(bar_w, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar_acc_full + 8 * i, kIssuers);
            mbar_init(bar_acc_empty + 8 * i, 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(kTmemCols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ---- TMA producer: weights once, then the input rows through the ring ----
            const uint8_t *wsrc = reinterpret_cast<const uint8_t *>(p.wimg) + (size_t)nh * kWBytes;
            mbar_expect_tx(bar_w, 9u * p.planes * kWTapBytes);
            for (int tap = 0; tap < 9; ++tap)
                bulk_g2s(s_w + tap * 2 * kWTapBytes, wsrc + (size_t)tap * 2 * kWTapBytes, p.planes * kWTapBytes, bar_w);
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
    } else if (warp == 1 || warp == 2) {
        if (lane == 0) {
            // ---- MMA issuers: the K loop (operand-split combination x tap) is dealt out to
            // kIssuers threads, each accumulating into its own TMEM block; the epilogue adds the
            // partial sums.  One thread alone cannot issue 108 small MMAs per row fast enough.
            const int issuer = warp - 1;
            constexpr uint32_t idesc = make_idesc(kStripW, kNHalf);
            const uint32_t a_hi = desc_hi(128), b_hi = desc_hi(128);
            const int n_units = (p.planes == 2 ? 3 : 1) * 9;
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
                const uint32_t tmem_d = tmem_base + (uint32_t)(b * kIssuers + issuer) * kNHalf;
                uint32_t acc = 0;
#pragma unroll 1
                for (int u = issuer; u < n_units; u += kIssuers) {
                    // (activation plane, weight plane): lo*hi, hi*lo, then hi*hi
                    const int combo = u / 9, tap = u - combo * 9;
                    const int ky = tap / 3, kx = tap - ky * 3;
                    const int apl = (p.planes == 2 && combo == 0) ? 1 : 0;
                    const int wpl = (p.planes == 2 && combo == 1) ? 1 : 0;
                    // descriptor lo words; the K steps add compile-time constants (all offsets are
                    // multiples of 16 bytes and stay below 2^18, so the address field never carries)
                    const uint32_t a_lo = desc_lo(s_rows + ((j + ky) & 3) * kSlotBytes + apl * kPlaneBytes + kx * 16, kRunBytes);
                    const uint32_t b_lo = desc_lo(s_w + tap * 2 * kWTapBytes + wpl * kWTapBytes, kNHalf * 16);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        umma_bf16_parts(tmem_d, a_lo + ((2 * ks * kRunBytes) >> 4), a_hi,
                                        b_lo + ((2 * ks * (kNHalf * 16)) >> 4), b_hi, idesc, acc);
                        acc = 1;
                    }
                }
                tc_commit(bar_empty + 8 * (j & 3));      // this issuer no longer needs input row j
                tc_commit(bar_acc_full + 8 * b);         // this issuer's partial accumulator is complete
            }
            // rows j+1, j+2 of the last step are never released through the ring; nothing waits on them
        }
    } else if (warp >= 4) {
        // ---- epilogue: TMEM -> registers -> bias / ReLU -> (hi, lo) or fp32 -> HBM ----
        const int q = warp & 3;                      // TMEM lane quarter this warp may touch
        const int px = x0 + q * 32 + lane;
        float bias[kNHalf];
#pragma unroll
        for (int i = 0; i < kNHalf; ++i) bias[i] = __ldg(p.bias + nh * kNHalf + i);
        for (int j = 0; j < rows_out; ++j) {
            const int b = j & 1;
            mbar_wait(bar_acc_full + 8 * b, (j >> 1) & 1);
            tc_fence_after();
            uint32_t v[kNHalf], v2[kNHalf];
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(b * kIssuers) * kNHalf;
            tmem_ld_32x32(taddr, v);
            tmem_ld_32x32(taddr + kNHalf, v2);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_acc_empty + 8 * b);    // the accumulator may be overwritten

            const int y = y0 + j;
            if (px < w_out) {
                float f[kNHalf];
#pragma unroll
                for (int i = 0; i < kNHalf; ++i) {
                    f[i] = (__uint_as_float(v[i]) + __uint_as_float(v2[i])) + bias[i];
                    if (p.relu) f[i] = fmaxf(f[i], 0.f);
                }
                if (p.out_f32) {
                    float4 *dst = reinterpret_cast<float4 *>(p.out_f32 + (((size_t)img * h_out + y) * w_out + px) * 64 +
                                                             nh * kNHalf);
#pragma unroll
                    for (int i = 0; i < kNHalf / 4; ++i) dst[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
                } else {
#pragma unroll
                    for (int c = 0; c < kNHalf / 8; ++c) {
                        uint32_t hi[4], lo[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            float a0 = f[8 * c + 2 * e], a1 = f[8 * c + 2 * e + 1];
                            __nv_bfloat16 h0 = __float2bfloat16_rn(a0), h1 = __float2bfloat16_rn(a1);
                            __nv_bfloat16 l0 = __float2bfloat16_rn(a0 - __bfloat162float(h0));
                            __nv_bfloat16 l1 = __float2bfloat16_rn(a1 - __bfloat162float(h1));
                            hi[e] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
                            lo[e] = (uint32_t)__bfloat16_as_ushort(l0) | ((uint32_t)__bfloat16_as_ushort(l1) << 16);
                        }
                        const size_t off = ((((size_t)img * h_out + y) * 8 + (nh * 4 + c)) * p.wp_out + px) * 8;
                        *reinterpret_cast<uint4 *>(p.out_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                        if (p.planes == 2)
                            *reinterpret_cast<uint4 *>(p.out_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
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

// fp32 packed weights (one 3x3 64->64 layer: [tap][cin][cout] + bias) -> UMMA B operand images
// [nh][tap][plane][kc][n][8] bf16 with plane 0 = hi, plane 1 = lo
__global__ void weight_image_kernel(const float *__restrict__ w, __nv_bfloat16 *__restrict__ img)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;      // over [nh][tap][kc][n][k8]
    if (i >= 2 * 9 * 8 * 32 * 8) return;
    int k8 = i & 7, n = (i >> 3) & 31, kc = (i >> 8) & 7, tap = (i >> 11) % 9, nh = (i >> 11) / 9;
    float v = w[((size_t)tap * 64 + kc * 8 + k8) * 64 + nh * 32 + n];
    __nv_bfloat16 hi = __float2bfloat16_rn(v);
    __nv_bfloat16 lo = __float2bfloat16_rn(v - __bfloat162float(hi));
    size_t base = (((size_t)nh * 9 + tap) * 2) * (8 * 32 * 8) + ((size_t)kc * 32 + n) * 8 + k8;
    img[base] = hi;
    img[base + 8 * 32 * 8] = lo;
}

}  // namespace tc

size_t tc_weight_image_elems() { return (size_t)2 * 9 * 2 * 8 * 32 * 8; }   // per layer

int launch_weight_images(stereo_ctx *ctx, const float *layer_w, void *img)
{
    tc::weight_image_kernel<<<(2 * 9 * 8 * 32 * 8 + 255) / 256, 256, 0, ctx->stream>>>(
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
    // enough row segments for a few waves of 1 CTA/SM, few enough to amortise weights + halo rows
    int column_ctas = strips * 2 * n;
    int segs = (ctx->sm_count * 3 + column_ctas - 1) / column_ctas;
    int rows = (h_out + segs - 1) / segs;
    if (rows < 8) rows = 8;
    p.rows_per_cta = rows;
    segs = (h_out + rows - 1) / rows;
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(tc::conv3x3_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           tc::kSmemTotal));
    ProfScope prof(ctx, STEREO_K_CONV);
    tc::conv3x3_tc_kernel<<<dim3(strips, segs, 2 * n), 256, tc::kSmemTotal, ctx->stream>>>(p);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
