// Generic 'valid' KS x KS convolution on the tensor cores for the layers of base_model
// (/root/reference/models/model.py:5-28: nine 5x5 layers, 32 or 64 channels) - SURVEY.md section 8(a) row T3.
//
// Same data layout and arithmetic as conv_tc.cu (chunk-planar bf16 (hi, lo) activations, implicit GEMM with
// M = 128 pixels of one output row, accumulation hi*hi + lo*hi + hi*lo in fp32 tensor memory), but nothing is
// resident: a 5x5 / 64 -> 64 layer has 400 KB of (hi, lo) weights, so ALL weights stream through a three-stage
// ring, one (tap, plane) unit of Cin x Cout x 2 B at a time, in the order the single issuer consumes them, once
// per output row (25 taps x 2 planes x 8 KB = 400 KB of L2 traffic for 300 MMAs).  Shared memory holds KS + 1
// input rows (hi + lo) and the ring.
//
// Warp roles: warp 0 = TMA producer of the input rows, warp 3 = TMA producer of the weight ring, warp 1 = MMA
// issuer (whole warp, one elected lane issues), warp 2 allocates tensor memory, warps 4-7 = epilogue.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {

namespace tcg {

using namespace tc;

constexpr int kStripW = 128;
constexpr int kWStages = 3;

struct Params {
    const __nv_bfloat16 *in_hi, *in_lo;   // chunk-planar [img][h_in][CIN/8][wp_in][8]
    int h_in, w_in, wp_in;
    const __nv_bfloat16 *wimg;            // [tap][plane][kc][n][8] for this layer
    const float *bias;                    // [COUT]
    __nv_bfloat16 *out_hi, *out_lo;       // chunk-planar [img][h_out][COUT/8][wp_out][8] (or null)
    int wp_out;
    float *out_f32;                       // NHWC [img][h_out][w_out][COUT] (or null)
    int relu, planes, rows_per_cta;
};

template <int KS, int CIN, int COUT>
struct Cfg {
    static constexpr int kPW = kStripW + KS - 1;          // input pixels per staged row
    static constexpr int kChunks = CIN / 8;
    static constexpr int kRunBytes = kPW * 16;
    static constexpr int kPlaneBytes = kChunks * kRunBytes;
    static constexpr int kSlotBytes = 2 * kPlaneBytes;
    static constexpr int kSlots = KS + 1;
    static constexpr int kUnitBytes = kChunks * COUT * 16; // one tap, one plane: [kc][n][8] bf16
    static constexpr int kSmemW = 0;
    static constexpr int kSmemRows = kSmemW + kWStages * kUnitBytes;
    static constexpr int kSmemBias = kSmemRows + kSlots * kSlotBytes;
    static constexpr int kSmemBars = kSmemBias + COUT * 4;
    static constexpr int kSmemTotal = kSmemBars + 32 * 8 + 16;
    static constexpr uint32_t kTmemCols = 2 * COUT < 32 ? 32 : 2 * COUT;     // two accumulators; 64 or 128: powers of two
    static_assert(kSmemTotal <= 232448, "shared memory budget of one CTA");
    static_assert(CIN % 16 == 0 && COUT % 16 == 0 && COUT <= 64, "channel counts of base_model");
};

template <int KS, int CIN, int COUT>
__global__ __launch_bounds__(256, 1) void conv_tcg_kernel(Params p)
{
    using C = Cfg<KS, CIN, COUT>;
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int h_out = p.h_in - (KS - 1), w_out = p.w_in - (KS - 1);
    const int x0 = blockIdx.x * kStripW;
    const int y0 = blockIdx.y * p.rows_per_cta;
    const int rows_out = min(p.rows_per_cta, h_out - y0);
    const int img = blockIdx.z;
    const int rows_in = rows_out + KS - 1;
    const int units_per_row = KS * KS * p.planes;          // weight units the issuer consumes per output row

    const uint32_t s_w = smem_u32(smem);
    const uint32_t s_rows = s_w + C::kSmemRows;
    float *s_bias = reinterpret_cast<float *>(smem + C::kSmemBias);
    const uint32_t s_bar = s_w + C::kSmemBars;
    const uint32_t bar_full = s_bar;                // [kSlots <= 8] input row landed
    const uint32_t bar_empty = s_bar + 64;          // [kSlots] input row consumed
    const uint32_t bar_w_full = s_bar + 128;        // [3] weight unit landed
    const uint32_t bar_w_empty = s_bar + 152;       // [3] weight unit consumed
    const uint32_t bar_acc_full = s_bar + 176;      // [2]
    const uint32_t bar_acc_empty = s_bar + 192;     // [2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + C::kSmemBars + 32 * 8);

    if (tid == 0) {
        for (int i = 0; i < C::kSlots; ++i) {
            mbar_init(bar_full + 8 * i, 1);
            mbar_init(bar_empty + 8 * i, 1);
        }
        for (int i = 0; i < kWStages; ++i) {
            mbar_init(bar_w_full + 8 * i, 1);
            mbar_init(bar_w_empty + 8 * i, 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar_acc_full + 8 * i, 1);
            mbar_init(bar_acc_empty + 8 * i, 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < COUT) s_bias[tid] = __ldg(p.bias + tid);
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(C::kTmemCols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ---- TMA producer: input rows through a ring of KS + 1 slots ----
            const size_t row_elems = (size_t)C::kChunks * p.wp_in * 8;     // bf16 elements per image row
            for (int r = 0; r < rows_in; ++r) {
                const int slot = r % C::kSlots;
                mbar_wait(bar_empty + 8 * slot, ((r / C::kSlots) & 1) ^ 1);
                mbar_expect_tx(bar_full + 8 * slot, (uint32_t)p.planes * C::kPlaneBytes);
                const size_t row_off = ((size_t)img * p.h_in + (y0 + r)) * row_elems + (size_t)x0 * 8;
                for (int pl = 0; pl < p.planes; ++pl) {
                    const __nv_bfloat16 *src = (pl == 0 ? p.in_hi : p.in_lo) + row_off;
                    const uint32_t dst = s_rows + slot * C::kSlotBytes + pl * C::kPlaneBytes;
                    for (int cc = 0; cc < C::kChunks; ++cc)
                        bulk_g2s(dst + cc * C::kRunBytes, src + (size_t)cc * p.wp_in * 8, C::kRunBytes, bar_full + 8 * slot);
                }
            }
        }
    } else if (warp == 3) {
        if (lane == 0) {
            // ---- TMA producer: weight units in the issuer's order: (tap 0, hi), (tap 0, lo), (tap 1, hi), ... ----
            const uint8_t *wsrc = reinterpret_cast<const uint8_t *>(p.wimg);
            uint32_t u = 0;
            for (int j = 0; j < rows_out; ++j)
                for (int k = 0; k < units_per_row; ++k, ++u) {
                    const int tap = k / p.planes, pl = k % p.planes;
                    const uint32_t st = u % kWStages;
                    mbar_wait(bar_w_empty + 8 * st, ((u / kWStages) & 1) ^ 1);
                    mbar_expect_tx(bar_w_full + 8 * st, C::kUnitBytes);
                    bulk_g2s(s_w + st * C::kUnitBytes, wsrc + ((size_t)tap * 2 + pl) * C::kUnitBytes, C::kUnitBytes,
                             bar_w_full + 8 * st);
                }
        }
    } else if (warp == 1) {
        // ---- MMA issuer: the whole warp runs the loop, one elected lane issues ----
        constexpr uint32_t idesc = make_idesc(kStripW, COUT);
        const uint32_t d_hi = desc_hi(128);
        uint32_t u = 0;                                   // running weight-unit counter
        for (int j = 0; j < rows_out; ++j) {
            const int b = j & 1;
            mbar_wait(bar_acc_empty + 8 * b, ((j >> 1) & 1) ^ 1);
            if (j == 0) {
                for (int r = 0; r < KS - 1; ++r) mbar_wait(bar_full + 8 * r, 0);
            }
            {
                const int r = j + KS - 1;                 // the newest input row this output row needs
                mbar_wait(bar_full + 8 * (r % C::kSlots), (r / C::kSlots) & 1);
            }
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + (uint32_t)b * COUT;
            uint32_t acc = 0;
#pragma unroll 1
            for (int tap = 0; tap < KS * KS; ++tap) {
                const int ky = tap / KS, kx = tap - ky * KS;
                const uint32_t a_base = s_rows + ((j + ky) % C::kSlots) * C::kSlotBytes + kx * 16;
                // unit (tap, hi): A_hi * W_hi, then A_lo * W_hi; unit (tap, lo): A_hi * W_lo
                for (int pl = 0; pl < p.planes; ++pl, ++u) {
                    const uint32_t st = u % kWStages;
                    mbar_wait(bar_w_full + 8 * st, (u / kWStages) & 1);
                    tc_fence_after();
                    const uint32_t b_lo = desc_lo(s_w + st * C::kUnitBytes, COUT * 16);
                    const int n_a = (pl == 0 && p.planes == 2) ? 2 : 1;      // activation planes multiplied with this unit
                    for (int apl = 0; apl < n_a; ++apl) {
                        const uint32_t a_lo = desc_lo(a_base + apl * C::kPlaneBytes, C::kRunBytes);
#pragma unroll
                        for (int ks = 0; ks < CIN / 16; ++ks) {
                            if (elect_one())
                                umma_bf16_parts(tmem_d, a_lo + ((2 * ks * C::kRunBytes) >> 4), d_hi,
                                                b_lo + ((2 * ks * (COUT * 16)) >> 4), d_hi, idesc, acc);
                            acc = 1;
                        }
                    }
                    if (elect_one()) tc_commit(bar_w_empty + 8 * st);
                }
            }
            if (elect_one()) {
                tc_commit(bar_empty + 8 * (j % C::kSlots));     // input row j has had its last use
                tc_commit(bar_acc_full + 8 * b);
            }
        }
    } else if (warp >= 4) {
        // ---- epilogue: TMEM -> registers -> bias / ReLU -> (hi, lo) or fp32 -> HBM ----
        const int q = warp & 3;
        const int px = x0 + q * 32 + lane;
        for (int j = 0; j < rows_out; ++j) {
            const int b = j & 1;
            mbar_wait(bar_acc_full + 8 * b, (j >> 1) & 1);
            tc_fence_after();
            const int y = y0 + j;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)b * COUT;
#pragma unroll
            for (int half = 0; half < COUT / 32; ++half) {
                uint32_t v[32];
                tmem_ld_32x32(taddr + half * 32, v);
                if (half == COUT / 32 - 1) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_acc_empty + 8 * b);
                }
                if (px < w_out) {
                    float f[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        f[i] = __uint_as_float(v[i]) + s_bias[half * 32 + i];
                        if (p.relu) f[i] = fmaxf(f[i], 0.f);
                    }
                    if (p.out_f32) {
                        float4 *dst = reinterpret_cast<float4 *>(
                            p.out_f32 + (((size_t)img * h_out + y) * w_out + px) * COUT + half * 32);
#pragma unroll
                        for (int i = 0; i < 8; ++i) dst[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
                    } else {
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            uint32_t hi[4], lo[4];
#pragma unroll
                            for (int e = 0; e < 4; ++e) split_bf16x2(f[8 * c + 2 * e], f[8 * c + 2 * e + 1], hi[e], lo[e]);
                            const size_t off =
                                ((((size_t)img * h_out + y) * (COUT / 8) + (half * 4 + c)) * p.wp_out + px) * 8;
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
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(C::kTmemCols));
    }
}

// fp32 packed weights ([tap][cin][cout]) -> streaming units [tap][plane][kc][n][8] bf16 (plane 0 = hi, 1 = lo)
__global__ void weight_image_g_kernel(const float *__restrict__ w, int taps, int cin, int cout,
                                      __nv_bfloat16 *__restrict__ img)
{
    const int total = taps * cin * cout;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;      // over [tap][kc][n][k8]
    if (i >= total) return;
    const int k8 = i & 7;
    int r = i >> 3;
    const int n = r % cout;
    r /= cout;
    const int kc = r % (cin / 8), tap = r / (cin / 8);
    const float v = w[((size_t)tap * cin + kc * 8 + k8) * cout + n];
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    const __nv_bfloat16 lo = __float2bfloat16_rn(v - __bfloat162float(hi));
    const size_t unit = (size_t)cin * cout;                   // elements of one (tap, plane)
    const size_t off = ((size_t)kc * cout + n) * 8 + k8;
    img[((size_t)tap * 2) * unit + off] = hi;
    img[((size_t)tap * 2 + 1) * unit + off] = lo;
}

// fp32 NHWC [n][h][w][c] -> chunk-planar bf16 (hi, lo) [n][h][c/8][wp][8] (no margins)
__global__ void nhwc_to_planar_g_kernel(const float *__restrict__ in, size_t rows, int w, int c, int wp,
                                        __nv_bfloat16 *__restrict__ hi, __nv_bfloat16 *__restrict__ lo)
{
    const int chunks = c / 8;
    const size_t total = rows * chunks * w;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int x = (int)(i % w);
        const size_t rest = i / w;
        const int cc = (int)(rest % chunks);
        const size_t row = rest / chunks;
        const float4 *src = reinterpret_cast<const float4 *>(in + (row * w + x) * c + cc * 8);
        const float4 a = __ldg(src), b = __ldg(src + 1);
        uint32_t h4[4], l4[4];
        split_bf16x2(a.x, a.y, h4[0], l4[0]);
        split_bf16x2(a.z, a.w, h4[1], l4[1]);
        split_bf16x2(b.x, b.y, h4[2], l4[2]);
        split_bf16x2(b.z, b.w, h4[3], l4[3]);
        const size_t off = ((row * chunks + cc) * wp + x) * 8;
        *reinterpret_cast<uint4 *>(hi + off) = make_uint4(h4[0], h4[1], h4[2], h4[3]);
        if (lo) *reinterpret_cast<uint4 *>(lo + off) = make_uint4(l4[0], l4[1], l4[2], l4[3]);
    }
}

}  // namespace tcg

size_t tcg_weight_image_elems(int ks, int cin, int cout) { return (size_t)2 * ks * ks * cin * cout; }

// row pitch (pixels) of the chunk-planar activations these kernels read: whole 128-pixel strips plus the halo
int tcg_pitch(int w) { return (w + tcg::kStripW - 1) / tcg::kStripW * tcg::kStripW + 8; }

int launch_weight_image_g(stereo_ctx *ctx, const float *layer_w, int ks, int cin, int cout, void *img)
{
    const int total = ks * ks * cin * cout;
    tcg::weight_image_g_kernel<<<(total + 255) / 256, 256, 0, ctx->stream>>>(layer_w, ks * ks, cin, cout,
                                                                            static_cast<__nv_bfloat16 *>(img));
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_nhwc_to_planar_g(stereo_ctx *ctx, const float *in, int n, int h, int w, int c, int wp, void *hi, void *lo)
{
    ProfScope prof(ctx, STEREO_K_PREP);
    tcg::nhwc_to_planar_g_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(
        in, (size_t)n * h, w, c, wp, static_cast<__nv_bfloat16 *>(hi), static_cast<__nv_bfloat16 *>(lo));
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

template <int KS, int CIN, int COUT>
static int launch_tcg(stereo_ctx *ctx, tcg::Params &p, int n)
{
    using C = tcg::Cfg<KS, CIN, COUT>;
    const int h_out = p.h_in - (KS - 1), w_out = p.w_in - (KS - 1);
    const int strips = (w_out + tcg::kStripW - 1) / tcg::kStripW;
    STEREO_REQUIRE(p.wp_in >= strips * tcg::kStripW + KS - 1, "activation pitch %d too small for %d strips", p.wp_in, strips);
    // one CTA per SM (the row ring fills shared memory): about two waves of row segments
    int segs = (ctx->sm_count * 2) / (strips * n);
    if (segs < 1) segs = 1;
    int rows = (h_out + segs - 1) / segs;
    if (rows < 8) rows = 8;
    p.rows_per_cta = rows;
    segs = (h_out + rows - 1) / rows;
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(tcg::conv_tcg_kernel<KS, CIN, COUT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           C::kSmemTotal));
    ProfScope prof(ctx, STEREO_K_CONV);
    tcg::conv_tcg_kernel<KS, CIN, COUT><<<dim3(strips, segs, n), 256, C::kSmemTotal, ctx->stream>>>(p);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_conv_tcg(stereo_ctx *ctx, int ks, int cin, int cout, const void *in_hi, const void *in_lo, int n, int h_in,
                    int w_in, int wp_in, const void *wimg, const float *bias, void *out_hi, void *out_lo, int wp_out,
                    float *out_f32, int relu, int planes)
{
    STEREO_REQUIRE(h_in >= ks && w_in >= ks, "input %dx%d smaller than the %dx%d kernel", h_in, w_in, ks, ks);
    tcg::Params p;
    p.in_hi = static_cast<const __nv_bfloat16 *>(in_hi);
    p.in_lo = static_cast<const __nv_bfloat16 *>(in_lo);
    p.h_in = h_in; p.w_in = w_in; p.wp_in = wp_in;
    p.wimg = static_cast<const __nv_bfloat16 *>(wimg);
    p.bias = bias;
    p.out_hi = static_cast<__nv_bfloat16 *>(out_hi);
    p.out_lo = static_cast<__nv_bfloat16 *>(out_lo);
    p.wp_out = wp_out;
    p.out_f32 = out_f32;
    p.relu = relu; p.planes = planes; p.rows_per_cta = 0;
    if (ks == 5 && cin == 32 && cout == 32) return launch_tcg<5, 32, 32>(ctx, p, n);
    if (ks == 5 && cin == 32 && cout == 64) return launch_tcg<5, 32, 64>(ctx, p, n);
    if (ks == 5 && cin == 64 && cout == 64) return launch_tcg<5, 64, 64>(ctx, p, n);
    set_error("no tensor-core convolution for ks=%d cin=%d cout=%d", ks, cin, cout);
    return STEREO_E_INVALID;
}

}  // namespace stereo
