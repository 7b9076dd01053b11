// 2-CTA (cta_group::2) variant of the tensor-core 3x3 / 64->64 convolution of conv_tc.cu.
//
// Why.  Measured on B200, tcgen05.mma with both operands in shared memory is fed at roughly
// 64-70 B/clk per SM; an M=128 x N=64 x K=16 MMA reads 4 KB of activations and 2 KB of weights
// for 32 cycles of math, so the single-CTA kernel is operand-bandwidth bound (43 % tensor-pipe
// activity), and the lo half of the weights does not even fit next to the hi half.  A CTA pair
// shares the weights: each CTA keeps only its 32-channel half of W_hi and W_lo (72 KB, resident,
// no streaming ring) and the pair issues M=256 MMAs whose B operand is assembled from both halves
// by the hardware.
//
// Layout of the pair.  The two CTAs of a cluster own two adjacent 128-pixel strips of the same
// rows.  Each CTA stages its own input rows (ring of 4) and runs its own epilogue on its own
// 128 TMEM lanes; only the leader CTA (cluster rank 0) issues MMAs.  Synchronisation across the
// pair: CTA 1 forwards "row landed" to the leader with remote mbarrier arrives, the leader's
// tcgen05.commit is multicast to the barriers of both CTAs, and CTA 1's epilogue releases the
// accumulators on the leader's barrier.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {

namespace tc2 {

using namespace tc;

constexpr int kStripW = 128;
constexpr int kPW = kStripW + 2;
constexpr int kN = 64;                      // output channels of the pair's MMA
constexpr int kNH = 32;                     // weight columns kept by one CTA
constexpr int kRunBytes = kPW * 16;
constexpr int kPlaneBytes = 8 * kRunBytes;
constexpr int kSlotBytes = 2 * kPlaneBytes;
constexpr int kSlots = 4;
constexpr int kWTapBytes = 8 * kNH * 16;    // one tap, one plane, one N-half: 4 KB
constexpr int kWPlaneBytes = 9 * kWTapBytes;
constexpr int kWBytes = 2 * kWPlaneBytes;   // hi + lo: 72 KB
constexpr int kSmemRows = kWBytes;
constexpr int kSmemBias = kSmemRows + kSlots * kSlotBytes;
constexpr int kSmemBars = kSmemBias + kN * 4;
constexpr int kSmemTotal = kSmemBars + 32 * 8 + 16;
constexpr int kIssuers = 2;
constexpr uint32_t kTmemCols = 256;

struct Params {
    const __nv_bfloat16 *in_hi, *in_lo;
    int h_in, w_in, wp_in;
    const __nv_bfloat16 *wimg;            // [nh][plane][tap][kc][n32][8]
    const float *bias;
    __nv_bfloat16 *out_hi, *out_lo;
    int wp_out;
    float *out_f32;
    int relu, planes, rows_per_cta;
};

__global__ __cluster_dims__(2, 1, 1) __launch_bounds__(256, 1) void conv3x3_tc2_kernel(Params p)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = cluster_rank();          // 0 = leader (issues the MMAs)
    const int h_out = p.h_in - 2, w_out = p.w_in - 2;
    const int x0 = blockIdx.x * kStripW;           // blockIdx.x = 2 * pair + rank
    const int y0 = blockIdx.y * p.rows_per_cta;
    const int rows_out = min(p.rows_per_cta, h_out - y0);
    const int img = blockIdx.z;
    const int rows_in = rows_out + 2;

    const uint32_t s_w = smem_u32(smem);
    const uint32_t s_rows = s_w + kSmemRows;
    float *s_bias = reinterpret_cast<float *>(smem + kSmemBias);
    const uint32_t s_bar = s_w + kSmemBars;
    const uint32_t bar_full = s_bar;               // [4] own input row landed
    const uint32_t bar_peer_full = s_bar + 32;     // [4] (leader) CTA 1's input row landed
    const uint32_t bar_empty = s_bar + 64;         // [4] input row consumed (multicast commit)
    const uint32_t bar_w = s_bar + 96;             // own weights landed
    const uint32_t bar_peer_w = s_bar + 104;       // (leader) CTA 1's weights landed
    const uint32_t bar_acc_full = s_bar + 112;     // [2] accumulators complete (multicast commit)
    const uint32_t bar_acc_empty = s_bar + 128;    // [2] (leader) both epilogues have drained
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + kSmemBars + 32 * 8);

    if (tid == 0) {
        for (int i = 0; i < kSlots; ++i) {
            mbar_init(bar_full + 8 * i, 1);
            mbar_init(bar_peer_full + 8 * i, 1);
            mbar_init(bar_empty + 8 * i, kIssuers);
        }
        mbar_init(bar_w, 1);
        mbar_init(bar_peer_w, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar_acc_full + 8 * i, kIssuers);
            mbar_init(bar_acc_empty + 8 * i, 8);          // 4 epilogue warps of each CTA
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < kN) s_bias[tid] = __ldg(p.bias + tid);
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(kTmemCols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                            // the peer's barriers exist before anyone arrives remotely
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ---- TMA producer: this CTA's half of the weights, then its input rows ----
            const uint8_t *wsrc = reinterpret_cast<const uint8_t *>(p.wimg) + (size_t)rank * kWBytes;
            mbar_expect_tx(bar_w, (uint32_t)p.planes * kWPlaneBytes);
            for (int pl = 0; pl < p.planes; ++pl)
                for (int tap = 0; tap < 9; ++tap)
                    bulk_g2s(s_w + pl * kWPlaneBytes + tap * kWTapBytes,
                             wsrc + (size_t)pl * kWPlaneBytes + (size_t)tap * kWTapBytes, kWTapBytes, bar_w);
            const size_t row_elems = (size_t)8 * p.wp_in * 8;
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
        if (lane == 0 && rank == 1) {
            // ---- CTA 1: tell the leader when this CTA's weights and rows have landed ----
            mbar_wait(bar_w, 0);
            mbar_arrive_remote(bar_peer_w, 0);
            for (int r = 0; r < rows_in; ++r) {
                const int slot = r & (kSlots - 1);
                mbar_wait(bar_full + 8 * slot, (r >> 2) & 1);
                mbar_arrive_remote(bar_peer_full + 8 * slot, 0);
            }
        }
    } else if (warp == 1 || warp == 2) {
        if (lane == 0 && rank == 0) {
            // ---- MMA issuers (leader CTA only); the K loop is split between two threads ----
            // issuer 0: A_hi * W_lo for all taps, A_hi * W_hi taps 0-3
            // issuer 1: A_lo * W_hi for all taps, A_hi * W_hi taps 4-8
            const int issuer = warp - 1;
            constexpr uint32_t idesc = make_idesc(2 * kStripW, kN);
            const uint32_t d_hi = desc_hi(128);
            mbar_wait(bar_w, 0);
            mbar_wait(bar_peer_w, 0);
            for (int j = 0; j < rows_out; ++j) {
                const int b = j & 1;
                mbar_wait(bar_acc_empty + 8 * b, ((j >> 1) & 1) ^ 1);
                if (j == 0) {
                    for (int s = 0; s < 2; ++s) {
                        mbar_wait(bar_full + 8 * s, 0);
                        mbar_wait(bar_peer_full + 8 * s, 0);
                    }
                }
                mbar_wait(bar_full + 8 * ((j + 2) & 3), ((j + 2) >> 2) & 1);
                mbar_wait(bar_peer_full + 8 * ((j + 2) & 3), ((j + 2) >> 2) & 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(b * kIssuers + issuer) * kN;
                uint32_t acc = 0;
                auto unit = [&](int tap, int apl, int wpl) {
                    const int ky = tap / 3, kx = tap - ky * 3;
                    const uint32_t a_lo = desc_lo(s_rows + ((j + ky) & 3) * kSlotBytes + apl * kPlaneBytes + kx * 16, kRunBytes);
                    const uint32_t b_lo = desc_lo(s_w + wpl * kWPlaneBytes + tap * kWTapBytes, kNH * 16);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        umma_pair(tmem_d, a_lo + ((2 * ks * kRunBytes) >> 4), d_hi,
                                  b_lo + ((2 * ks * (kNH * 16)) >> 4), d_hi, idesc, acc);
                        acc = 1;
                    }
                };
                if (p.planes == 2) {
#pragma unroll 1
                    for (int tap = 0; tap < 9; ++tap) unit(tap, issuer == 0 ? 0 : 1, issuer == 0 ? 1 : 0);
                }
                const int t0 = issuer == 0 ? 0 : 4, t1 = issuer == 0 ? 4 : 9;
#pragma unroll 1
                for (int tap = t0; tap < t1; ++tap) unit(tap, 0, 0);
                tc_commit_pair(bar_empty + 8 * (j & 3));
                tc_commit_pair(bar_acc_full + 8 * b);
            }
        }
    } else if (warp >= 4) {
        // ---- epilogue (both CTAs, each on its own 128 TMEM lanes) ----
        const int q = warp & 3;
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
                    // everything has been read: tell the leader's issuers (local or remote arrive)
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        if (rank == 0) mbar_arrive(bar_acc_empty + 8 * b);
                        else mbar_arrive_remote(bar_acc_empty + 8 * b, 0);
                    }
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
    cluster_sync_all();      // neither CTA frees tensor memory while the pair's MMAs may still touch it
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
    }
}

// fp32 packed weights ([tap][cin][cout]) -> per-CTA B operand images [nh][plane][tap][kc][n32][8]
__global__ void weight_image2_kernel(const float *__restrict__ w, __nv_bfloat16 *__restrict__ img)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;      // over [nh][tap][kc][n32][k8]
    if (i >= 2 * 9 * 8 * 32 * 8) return;
    int k8 = i & 7, n = (i >> 3) & 31, kc = (i >> 8) & 7, tap = (i >> 11) % 9, nh = (i >> 11) / 9;
    float v = w[((size_t)tap * 64 + kc * 8 + k8) * 64 + nh * 32 + n];
    __nv_bfloat16 hi = __float2bfloat16_rn(v);
    __nv_bfloat16 lo = __float2bfloat16_rn(v - __bfloat162float(hi));
    size_t base = (size_t)nh * (2 * 9 * 2048) + (size_t)tap * 2048 + ((size_t)kc * 32 + n) * 8 + k8;
    img[base] = hi;
    img[base + 9 * 2048] = lo;
}

}  // namespace tc2

int launch_weight_images2(stereo_ctx *ctx, const float *layer_w, void *img)
{
    tc2::weight_image2_kernel<<<(2 * 9 * 8 * 32 * 8 + 255) / 256, 256, 0, ctx->stream>>>(
        layer_w, static_cast<__nv_bfloat16 *>(img));
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int tc2_pitch(int w) { return (w + 255) / 256 * 256 + 8; }

int launch_conv_tc2(stereo_ctx *ctx, const void *in_hi, const void *in_lo, int n, int h_in, int w_in, int wp_in,
                    const void *wimg, const float *bias, void *out_hi, void *out_lo, int wp_out, float *out_f32,
                    int relu, int planes)
{
    STEREO_REQUIRE(h_in >= 3 && w_in >= 3, "input %dx%d smaller than the 3x3 kernel", h_in, w_in);
    const int h_out = h_in - 2, w_out = w_in - 2;
    const int strips = (w_out + 255) / 256 * 2;          // an even number of 128-pixel strips
    STEREO_REQUIRE(wp_in >= strips * 128 + 2, "activation pitch %d too small for %d strips", wp_in, strips);
    tc2::Params p;
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
    int column_ctas = strips * n;
    int segs = (ctx->sm_count * 2) / column_ctas;
    if (segs < 1) segs = 1;
    int rows = (h_out + segs - 1) / segs;
    if (rows < 8) rows = 8;
    p.rows_per_cta = rows;
    segs = (h_out + rows - 1) / rows;
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(tc2::conv3x3_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           tc2::kSmemTotal));
    ProfScope prof(ctx, STEREO_K_CONV);
    tc2::conv3x3_tc2_kernel<<<dim3(strips, segs, n), 256, tc2::kSmemTotal, ctx->stream>>>(p);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
