// Row-streaming CTA-pair version of the tensor-core 3x3 / 64->64 convolution (trunk layers 2-4,
// /root/reference/models/model.py:35-47).
//
// conv_tc.cu / conv_tc2.cu compute one output row from three input rows, so every input row is read
// from shared memory three times (once per ky) by MMAs that are only 64 columns wide, and the kernel
// runs at the operand-fetch limit of tcgen05 instead of its math limit.  Here every input row is
// multiplied ONCE, by the weights of all three ky concatenated along N:
//
//     D[256 pixels][3 x 64] += A[row i][kx shift] * [ W(ky_a) | W(ky_b) | W(ky_c) ]          (N = 192)
//
// The three 64-column blocks of D are the accumulators of three consecutive OUTPUT rows
// (i - ky for the three ky).  Output row j lives in block j mod 3, so the ky order inside the
// concatenation rotates with i mod 3; the weights are stored as the five-block sequence
// ky = 0,2,1,0,2 and the B descriptor just starts at block 0, 2 or 1.  After input row i the block of
// output row i-2 is complete: the epilogue warps read it, write zeros back (the next row's MMAs
// accumulate into it again) and release it.
//
// To keep the tensor pipe busy while a block is drained, each CTA pair works on TWO independent
// row segments ("streams") with separate accumulators (2 x 192 tensor-memory columns) and
// alternates between them row by row.
//
// CTA pair: M = 256 = two adjacent 128-pixel strips; CTA c keeps output channels 32c..32c+31 of the
// weights (hi and lo planes, 120 KB) resident and supplies them as its half of B.  Per MMA and CTA
// that is 4 KB of A and 3 KB of B for 96 cycles of math.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {

namespace tc3 {

using namespace tc;

constexpr int kStripW = 128;
constexpr int kPW = kStripW + 2;
constexpr int kRunBytes = kPW * 16;            // one channel chunk of one row: 130 pixels x 8 bf16
constexpr int kPlaneBytes = 8 * kRunBytes;
constexpr int kSlotBytes = 2 * kPlaneBytes;    // hi + lo
constexpr int kSlots = 3;
constexpr int kWRows = 160;                    // 5 blocks (ky = 0,2,1,0,2) x 32 output channels
constexpr int kWKcBytes = kWRows * 16;         // one 8-channel K chunk
constexpr int kWUnitBytes = 8 * kWKcBytes;     // one (plane, kx): 20 KB
constexpr int kWBytes = 6 * kWUnitBytes;       // 2 planes x 3 kx: 120 KB per CTA
constexpr int kSmemRows = kWBytes;
constexpr int kSmemBias = kSmemRows + kSlots * kSlotBytes;
constexpr int kSmemBars = kSmemBias + 64 * 4;
constexpr int kSmemTotal = kSmemBars + 32 * 8 + 16;
constexpr int kAccCols = 192;                  // accumulator columns of one stream
constexpr uint32_t kTmemCols = 512;
constexpr int kThreads = 384;                  // 4 service warps + 8 epilogue warps

struct Params {
    const __nv_bfloat16 *in_hi, *in_lo;
    int h_in, w_in, wp_in;
    const __nv_bfloat16 *wimg;            // [half][plane][kx][kc][row160][8]
    const float *bias;
    __nv_bfloat16 *out_hi, *out_lo;
    int wp_out;
    float *out_f32;
    int relu, planes;
    int rows_per_seg, segs, items;        // work items = (image, row segment); a pair takes two
};

struct Stream {
    int img, y0, rows_out, rows_in;
};

__device__ __forceinline__ Stream make_stream(const Params &p, int item)
{
    Stream s{0, 0, 0, 0};
    if (item < p.items) {
        s.img = item / p.segs;
        s.y0 = (item - s.img * p.segs) * p.rows_per_seg;
        s.rows_out = min(p.rows_per_seg, p.h_in - 2 - s.y0);
        s.rows_in = s.rows_out + 2;
    }
    return s;
}

__global__ __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1) void conv3x3_tc3_kernel(Params p)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = cluster_rank();          // 0 = leader (issues the MMAs)
    const int h_out = p.h_in - 2, w_out = p.w_in - 2;
    const int x0 = blockIdx.x * kStripW;           // blockIdx.x = 2 * pair + rank
    const Stream st0 = make_stream(p, 2 * blockIdx.y), st1 = make_stream(p, 2 * blockIdx.y + 1);
    const int steps = 2 * max(st0.rows_in, st1.rows_in);   // step q: stream q & 1, input row q >> 1

    const uint32_t s_w = smem_u32(smem);
    const uint32_t s_rows = s_w + kSmemRows;
    float *s_bias = reinterpret_cast<float *>(smem + kSmemBias);
    const uint32_t s_bar = s_w + kSmemBars;
    const uint32_t bar_full = s_bar;               // [3] own input row landed
    const uint32_t bar_peer_full = s_bar + 32;     // [3] (leader) CTA 1's input row landed
    const uint32_t bar_empty = s_bar + 64;         // [3] input row consumed (multicast commit)
    const uint32_t bar_w = s_bar + 96;             // own weights landed
    const uint32_t bar_peer_w = s_bar + 104;       // (leader) CTA 1's weights landed
    const uint32_t bar_acc_done = s_bar + 112;     // [2] per stream: MMAs of an input row complete (multicast)
    const uint32_t bar_blk_free = s_bar + 128;     // [2] per stream (leader): finished block drained and zeroed
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + kSmemBars + 32 * 8);

    if (tid == 0) {
        for (int i = 0; i < kSlots; ++i) {
            mbar_init(bar_full + 8 * i, 1);
            mbar_init(bar_peer_full + 8 * i, 1);
            mbar_init(bar_empty + 8 * i, 1);
        }
        mbar_init(bar_w, 1);
        mbar_init(bar_peer_w, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar_acc_done + 8 * i, 1);
            mbar_init(bar_blk_free + 8 * i, 16);          // 8 epilogue warps of each CTA
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 64) s_bias[tid] = __ldg(p.bias + tid);
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
            // ---- TMA producer: this CTA's half of the weights, then the input rows of both streams ----
            const uint8_t *wsrc = reinterpret_cast<const uint8_t *>(p.wimg) + (size_t)rank * kWBytes;
            mbar_expect_tx(bar_w, (uint32_t)p.planes * 3 * kWUnitBytes);
            for (int u = 0; u < p.planes * 3; ++u)
                bulk_g2s(s_w + u * kWUnitBytes, wsrc + (size_t)u * kWUnitBytes, kWUnitBytes, bar_w);
            const size_t row_elems = (size_t)8 * p.wp_in * 8;
            uint32_t cnt = 0;
            for (int q = 0; q < steps; ++q) {
                const Stream &st = (q & 1) ? st1 : st0;
                const int i = q >> 1;
                if (i >= st.rows_in) continue;
                const uint32_t slot = cnt % kSlots, ph = (cnt / kSlots) & 1;
                ++cnt;
                mbar_wait(bar_empty + 8 * slot, ph ^ 1);
                mbar_expect_tx(bar_full + 8 * slot, (uint32_t)p.planes * kPlaneBytes);
                const size_t row_off = ((size_t)st.img * p.h_in + (st.y0 + i)) * row_elems + (size_t)x0 * 8;
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
            uint32_t cnt = 0;
            for (int q = 0; q < steps; ++q) {
                const Stream &st = (q & 1) ? st1 : st0;
                if ((q >> 1) >= st.rows_in) continue;
                const uint32_t slot = cnt % kSlots, ph = (cnt / kSlots) & 1;
                ++cnt;
                mbar_wait(bar_full + 8 * slot, ph);
                mbar_arrive_remote(bar_peer_full + 8 * slot, 0);
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            // ---- MMA issuer (leader CTA only): 36 MMAs of M=256 x N=192 x K=16 per input row ----
            constexpr uint32_t idesc = make_idesc(2 * kStripW, kAccCols);
            const uint32_t d_hi = desc_hi(128);
            mbar_wait(bar_w, 0);
            mbar_wait(bar_peer_w, 0);
            uint32_t cnt = 0;
            for (int q = 0; q < steps; ++q) {
                const int s = q & 1, i = q >> 1;
                const Stream &st = s ? st1 : st0;
                if (i >= st.rows_in) continue;
                const uint32_t slot = cnt % kSlots, ph = (cnt / kSlots) & 1;
                ++cnt;
                if (i > 0) mbar_wait(bar_blk_free + 8 * s, (i - 1) & 1);   // block of output row i-3 is clean again
                mbar_wait(bar_full + 8 * slot, ph);
                mbar_wait(bar_peer_full + 8 * slot, ph);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)s * kAccCols;
                // block b of D must receive ky = (i - b) mod 3: start the B window at block 0, 2 or 1
                const int rho = i % 3;
                const uint32_t win = (rho == 0 ? 0u : (rho == 1 ? 2u : 1u)) * 32 * 16;
                uint32_t acc = i > 0;
                const int combos = p.planes == 2 ? 3 : 1;
#pragma unroll 1
                for (int c = 0; c < combos; ++c) {
                    const int apl = c == 2, wpl = c == 1;       // hi*hi, hi*lo, lo*hi
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) {
                        const uint32_t a_lo = desc_lo(s_rows + slot * kSlotBytes + apl * kPlaneBytes + kx * 16, kRunBytes);
                        const uint32_t b_lo = desc_lo(s_w + (wpl * 3 + kx) * kWUnitBytes + win, kWKcBytes);
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            umma_pair(tmem_d, a_lo + ((2 * ks * kRunBytes) >> 4), d_hi,
                                      b_lo + ((2 * ks * kWKcBytes) >> 4), d_hi, idesc, acc);
                            acc = 1;
                        }
                    }
                }
                tc_commit_pair(bar_empty + 8 * slot);
                tc_commit_pair(bar_acc_done + 8 * s);
            }
        }
    } else if (warp >= 4) {
        // ---- epilogue (both CTAs): warp = (lane quarter, channel half) ----
        const int qd = warp & 3, hf = (warp - 4) >> 2;
        const int px = x0 + qd * 32 + lane;
        for (int q = 0; q < steps; ++q) {
            const int s = q & 1, i = q >> 1;
            const Stream &st = s ? st1 : st0;
            if (i >= st.rows_in) continue;
            mbar_wait(bar_acc_done + 8 * s, i & 1);
            tc_fence_after();
            const int j = i - 2;                              // the output row that is complete now
            const int b = (i + 1) % 3;                        // = j mod 3
            const uint32_t taddr =
                tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)(s * kAccCols + hf * 96 + b * 32);
            uint32_t v[32];
            if (j >= 0) tmem_ld_32x32(taddr, v);
            tmem_zero_32x32(taddr);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (rank == 0) mbar_arrive(bar_blk_free + 8 * s);
                else mbar_arrive_remote(bar_blk_free + 8 * s, 0);
            }
            if (j < 0 || px >= w_out) continue;
            const int y = st.y0 + j;
            float f[32];
#pragma unroll
            for (int k = 0; k < 32; ++k) {
                f[k] = __uint_as_float(v[k]) + s_bias[hf * 32 + k];
                if (p.relu) f[k] = fmaxf(f[k], 0.f);
            }
            if (p.out_f32) {
                float4 *dst = reinterpret_cast<float4 *>(p.out_f32 + (((size_t)st.img * h_out + y) * w_out + px) * 64 +
                                                         hf * 32);
#pragma unroll
                for (int k = 0; k < 8; ++k) dst[k] = make_float4(f[4 * k], f[4 * k + 1], f[4 * k + 2], f[4 * k + 3]);
            } else {
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint32_t hi[4], lo[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) split_bf16x2(f[8 * c + 2 * e], f[8 * c + 2 * e + 1], hi[e], lo[e]);
                    const size_t off = ((((size_t)st.img * h_out + y) * 8 + (hf * 4 + c)) * p.wp_out + px) * 8;
                    *reinterpret_cast<uint4 *>(p.out_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                    if (p.planes == 2)
                        *reinterpret_cast<uint4 *>(p.out_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
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

// fp32 packed weights ([tap][cin][cout]) -> per-CTA B operand images [half][plane][kx][kc][row160][8],
// row = 32 * block + n with block -> ky = 0,2,1,0,2 and cout = 32 * half + n
__global__ void weight_image3_kernel(const float *__restrict__ w, __nv_bfloat16 *__restrict__ img)
{
    const int total = 2 * 3 * 8 * kWRows * 8;               // one plane, both halves
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int k8 = i & 7;
    int t = i >> 3;
    const int row = t % kWRows; t /= kWRows;
    const int kc = t & 7; t >>= 3;
    const int kx = t % 3, half = t / 3;
    const int blk = row >> 5, n = row & 31;
    const int ky = blk == 0 || blk == 3 ? 0 : (blk == 2 ? 1 : 2);
    const float v = w[((size_t)(ky * 3 + kx) * 64 + kc * 8 + k8) * 64 + half * 32 + n];
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    const __nv_bfloat16 lo = __float2bfloat16_rn(v - __bfloat162float(hi));
    const size_t plane_elems = (size_t)3 * 8 * kWRows * 8;
    const size_t base = (size_t)half * 2 * plane_elems + (((size_t)kx * 8 + kc) * kWRows + row) * 8 + k8;
    img[base] = hi;
    img[base + plane_elems] = lo;
}

}  // namespace tc3

// row pitch of the chunk-planar activations: a CTA pair covers 256 pixels
int tc3_pitch(int w) { return (w + 255) / 256 * 256 + 8; }

size_t tc3_weight_image_elems() { return (size_t)2 * tc3::kWBytes / 2; }   // per layer, bf16 elements

int launch_weight_images3(stereo_ctx *ctx, const float *layer_w, void *img)
{
    const int total = 2 * 3 * 8 * tc3::kWRows * 8;
    tc3::weight_image3_kernel<<<(total + 255) / 256, 256, 0, ctx->stream>>>(layer_w, static_cast<__nv_bfloat16 *>(img));
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_conv_tc3(stereo_ctx *ctx, const void *in_hi, const void *in_lo, int n, int h_in, int w_in, int wp_in,
                    const void *wimg, const float *bias, void *out_hi, void *out_lo, int wp_out, float *out_f32,
                    int relu, int planes)
{
    STEREO_REQUIRE(h_in >= 3 && w_in >= 3, "input %dx%d smaller than the 3x3 kernel", h_in, w_in);
    const int h_out = h_in - 2, w_out = w_in - 2;
    const int strips = (w_out + 255) / 256 * 2;          // an even number of 128-pixel strips
    STEREO_REQUIRE(wp_in >= strips * 128 + 2, "activation pitch %d too small for %d strips", wp_in, strips);
    tc3::Params p;
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
    // one wave: every CTA pair takes two row segments; size them so that all pairs are resident at once
    const int pair_slots = ctx->sm_count / 2;
    const int pair_cols = strips / 2;
    int items = 2 * (pair_slots / pair_cols);            // (image, segment) work items in one wave
    if (items < 2) items = 2;
    int segs = items / n;
    if (segs < 1) segs = 1;
    int rows = (h_out + segs - 1) / segs;
    if (rows < 6) rows = 6;
    if (rows > h_out) rows = h_out;
    segs = (h_out + rows - 1) / rows;
    p.rows_per_seg = rows;
    p.segs = segs;
    p.items = segs * n;
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(tc3::conv3x3_tc3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           tc3::kSmemTotal));
    ProfScope prof(ctx, STEREO_K_CONV);
    tc3::conv3x3_tc3_kernel<<<dim3(strips, (p.items + 1) / 2, 1), tc3::kThreads, tc3::kSmemTotal, ctx->stream>>>(p);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
