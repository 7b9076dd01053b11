// 'valid' KS x KS convolution on the tensor cores, every input row multiplied ONCE - the default for the layers of
// base_model (/root/reference/models/model.py:5-28), SURVEY.md section 8(a) row T3.
//
// conv_tcg.cu issues, per output row and tap, an M128 x N(Cout) x K16 MMA: 4 KB of A for 16-32 clocks of math, so
// it runs at the shared-memory operand-fetch rate (tensor pipe 19-35 % active, ncu).  Here the roles are turned
// round as in conv_tc3.cu: an input row r (shifted by kx) is the A operand of ONE product with the weights of all
// KS vertical taps side by side, B = [W(ky=KS-1,kx) | ... | W(ky=0,kx)] (N = KS * Cout = 160 or 320), whose column
// blocks are the accumulators of the KS output rows r-KS+1 .. r.  Output row j lives in tensor-memory block
// j mod NB (NB = 512 / Cout blocks of Cout columns), so the blocks an input row feeds are consecutive, ascending
// with j, and one or two MMAs (N <= 256, split where the ring of blocks wraps) cover them.  All MMAs accumulate:
// the epilogue drains a finished block, writes zeros back (tcgen05.st) and releases it.
//
// Weights stream from L2 through a three-stage ring, one (kx, plane) unit [kc][KS*Cout][8] (40 KB for 64 -> 64) at
// a time, ten units per input row = 400 KB per 128 output pixels; sharing the units between the CTAs of the left and
// the right image through a cluster halved that L2 traffic and changed nothing measurable (2.67 against 2.69 ms per
// pair), so every CTA streams its own.  Input rows (chunk-planar bf16 hi / lo planes) pass through a ring of three
// slots and are read for exactly one row of MMAs.  Arithmetic and layouts are those of conv_tcg.cu (hi*hi + lo*hi +
// hi*lo into fp32), so are the results up to the order of the fp32 accumulation.
//
// Warp roles: warp 0 = TMA producer of the input rows, warp 3 = TMA producer of the weight ring, warp 1 = MMA
// issuer (whole warp, one elected lane issues), warp 2 allocates tensor memory, warps 4-7 = epilogue.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {

namespace tcr {

using namespace tc;

constexpr int kStripW = 128;
constexpr int kWStages = 3;
constexpr int kSlots = 3;

struct Params {
    const __nv_bfloat16 *in_hi, *in_lo;   // chunk-planar [img][h_in][CIN/8][wp_in][8]
    int h_in, w_in, wp_in;
    const __nv_bfloat16 *wimg;            // [kx][plane][kc][n = (KS-1-ky)*COUT + cout][8] for this layer
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
    static constexpr int kNAll = KS * COUT;                // columns of one weight unit
    static constexpr int kUnitBytes = kChunks * kNAll * 16; // one kx, one plane: [kc][n][8] bf16
    static constexpr int kNB = 512 / COUT;                 // accumulator blocks (output rows in flight)
    static constexpr int kMaxRun = 256 / COUT;             // blocks one MMA may span (N <= 256)
    // all 2 * KS units of a layer stay in shared memory when they fit beside the row slots (32 -> 32: 100 KB);
    // otherwise they stream through kWStages stages
    static constexpr bool kResident = 2 * KS * kUnitBytes <= 112 * 1024;
    static constexpr int kSmemRows = (kResident ? 2 * KS : kWStages) * kUnitBytes;
    static constexpr int kSmemBias = kSmemRows + kSlots * kSlotBytes;
    static constexpr int kSmemBars = kSmemBias + COUT * 4;
    static constexpr int kSmemTotal = kSmemBars + 48 * 8 + 16;
    static_assert(kSmemTotal <= 232448, "shared memory budget of one CTA");
    static_assert(CIN % 16 == 0 && COUT % 32 == 0 && COUT <= 64 && kNB > KS, "channel counts of base_model");
};

template <int KS, int CIN, int COUT>
__global__ __launch_bounds__(256, 1) void conv_tcr_kernel(Params p)
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

    const uint32_t s_w = smem_u32(smem);
    const uint32_t s_rows = s_w + C::kSmemRows;
    float *s_bias = reinterpret_cast<float *>(smem + C::kSmemBias);
    const uint32_t s_bar = s_w + C::kSmemBars;
    const uint32_t bar_full = s_bar;                // [kSlots] input row landed
    const uint32_t bar_empty = s_bar + 32;          // [kSlots] input row consumed
    const uint32_t bar_w_full = s_bar + 64;         // [kWStages] weight unit landed
    const uint32_t bar_w_empty = s_bar + 96;        // [kWStages] weight unit consumed
    const uint32_t bar_acc_full = s_bar + 128;      // [kNB <= 16] output row complete
    const uint32_t bar_acc_empty = s_bar + 256;     // [kNB] block drained and zeroed
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + C::kSmemBars + 48 * 8);

    if (tid == 0) {
        for (int i = 0; i < kSlots; ++i) {
            mbar_init(bar_full + 8 * i, 1);
            mbar_init(bar_empty + 8 * i, 1);
        }
        for (int i = 0; i < kWStages; ++i) {
            mbar_init(bar_w_full + 8 * i, 1);
            mbar_init(bar_w_empty + 8 * i, 1);
        }
        for (int i = 0; i < C::kNB; ++i) {
            mbar_init(bar_acc_full + 8 * i, 1);
            mbar_init(bar_acc_empty + 8 * i, 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < COUT) s_bias[tid] = __ldg(p.bias + tid);
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ---- TMA producer: input rows, each needed for one row of MMAs ----
            const size_t row_elems = (size_t)C::kChunks * p.wp_in * 8;     // bf16 elements per image row
            for (int r = 0; r < rows_in; ++r) {
                const int slot = r % kSlots;
                mbar_wait(bar_empty + 8 * slot, ((r / kSlots) & 1) ^ 1);
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
            // ---- TMA producer: weight units in the issuer's order: per input row (kx 0, hi), (kx 0, lo), (kx 1, hi), ... ----
            const uint8_t *wsrc = reinterpret_cast<const uint8_t *>(p.wimg);
            uint32_t u = 0;
            if (C::kResident) {
                // the whole weight image once: unit (kx, plane) at its own offset, one barrier
                mbar_expect_tx(bar_w_full, 2 * KS * C::kUnitBytes);
                for (int i = 0; i < 2 * KS; ++i)
                    bulk_g2s(s_w + i * C::kUnitBytes, wsrc + (size_t)i * C::kUnitBytes, C::kUnitBytes, bar_w_full);
            } else
            for (int r = 0; r < rows_in; ++r)
                for (int kx = 0; kx < KS; ++kx)
                    for (int pl = 0; pl < p.planes; ++pl, ++u) {
                        const uint32_t st = u % kWStages;
                        mbar_wait(bar_w_empty + 8 * st, ((u / kWStages) & 1) ^ 1);
                        mbar_expect_tx(bar_w_full + 8 * st, C::kUnitBytes);
                        bulk_g2s(s_w + st * C::kUnitBytes, wsrc + ((size_t)kx * 2 + pl) * C::kUnitBytes, C::kUnitBytes,
                                 bar_w_full + 8 * st);
                    }
        }
    } else if (warp == 1) {
        // ---- MMA issuer: the whole warp runs the loop, one elected lane issues ----
        const uint32_t d_hi = desc_hi(128);
        uint32_t u = 0;                                   // running weight-unit counter
        if (C::kResident) mbar_wait(bar_w_full, 0);
        for (int r = 0; r < rows_in; ++r) {
            const int slot = r % kSlots;
            // output rows this input row feeds: j = r - ky for the vertical taps that stay inside the band
            const int j_lo = max(0, r - (KS - 1)), j_hi = min(rows_out - 1, r);
            // one or two MMAs per product: consecutive blocks up to the wrap of the block ring, N <= 256.  Never more
            // than two runs: a wrap leaves two segments of at most kMaxRun blocks, no wrap one segment of at most KS.
            const int seg0 = min(C::kNB - j_lo % C::kNB, j_hi - j_lo + 1);       // blocks up to the wrap / the end
            int len0 = seg0, j1 = j_lo + seg0, len1 = j_hi - j1 + 1;            // second run: after the wrap ...
            if (seg0 > C::kMaxRun) {                                            // ... or the second half of a long segment
                len0 = (seg0 + 1) / 2;
                j1 = j_lo + len0;
                len1 = seg0 - len0;
            }
            // per run: accumulator address, first weight column (tap ky = r - j sits in column block KS-1-ky), shape
            const uint32_t tmem_d0 = tmem_base + (uint32_t)(j_lo % C::kNB) * COUT;
            const uint32_t tmem_d1 = tmem_base + (uint32_t)(j1 % C::kNB) * COUT;
            const uint32_t boff0 = (uint32_t)(KS - 1 - r + j_lo) * COUT;          // in 16-byte rows of the unit
            const uint32_t boff1 = (uint32_t)(KS - 1 - r + j1) * COUT;
            const uint32_t idesc0 = make_idesc(kStripW, len0 * COUT), idesc1 = make_idesc(kStripW, max(len1, 1) * COUT);
            const bool two_runs = len1 > 0;
            // the block of the output row that starts with this input row must have been drained and zeroed
            if (r < rows_out) mbar_wait(bar_acc_empty + 8 * (r % C::kNB), (r / C::kNB) & 1);
            mbar_wait(bar_full + 8 * slot, (r / kSlots) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int kx = 0; kx < KS; ++kx) {
                const uint32_t a_base = s_rows + slot * C::kSlotBytes + kx * 16;
                // unit (kx, hi): A_hi * W_hi, then A_lo * W_hi; unit (kx, lo): A_hi * W_lo
                for (int pl = 0; pl < p.planes; ++pl, ++u) {
                    const uint32_t st = C::kResident ? (uint32_t)(kx * 2 + pl) : u % kWStages;
                    if (!C::kResident) {
                        mbar_wait(bar_w_full + 8 * st, (u / kWStages) & 1);
                        tc_fence_after();
                    }
                    const uint32_t unit = desc_lo(s_w + st * C::kUnitBytes, C::kNAll * 16);
                    const int n_a = (pl == 0 && p.planes == 2) ? 2 : 1;      // activation planes multiplied with this unit
                    for (int apl = 0; apl < n_a; ++apl) {
                        const uint32_t a_lo = desc_lo(a_base + apl * C::kPlaneBytes, C::kRunBytes);
#pragma unroll
                        for (int ks = 0; ks < CIN / 16; ++ks) {
                            const uint32_t a_k = a_lo + ((2 * ks * C::kRunBytes) >> 4);
                            const uint32_t b_k = unit + ((2 * ks * (C::kNAll * 16)) >> 4);
                            if (elect_one()) umma_bf16_parts(tmem_d0, a_k, d_hi, b_k + boff0, d_hi, idesc0, 1u);
                            if (two_runs && elect_one()) umma_bf16_parts(tmem_d1, a_k, d_hi, b_k + boff1, d_hi, idesc1, 1u);
                        }
                    }
                    if (!C::kResident && elect_one()) tc_commit(bar_w_empty + 8 * st);
                }
            }
            if (elect_one()) {
                tc_commit(bar_empty + 8 * slot);                             // the input row has had its only use
                if (r >= KS - 1) tc_commit(bar_acc_full + 8 * ((r - (KS - 1)) % C::kNB));
            }
        }
    } else if (warp >= 4) {
        // ---- epilogue: TMEM -> registers -> zero the block -> bias / ReLU -> (hi, lo) or fp32 -> HBM ----
        const int q = warp & 3;
        const int px = x0 + q * 32 + lane;
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        // all blocks start as zeros (every MMA accumulates)
#pragma unroll 1
        for (int c = 0; c < 512; c += 32) tmem_zero_32x32(lane_base + c);
        tc_fence_before();
        __syncwarp();
        if (lane == 0)
            for (int b = 0; b < C::kNB; ++b) mbar_arrive(bar_acc_empty + 8 * b);
        for (int j = 0; j < rows_out; ++j) {
            const int b = j % C::kNB;
            mbar_wait(bar_acc_full + 8 * b, (j / C::kNB) & 1);
            tc_fence_after();
            const int y = y0 + j;
            const uint32_t taddr = lane_base + (uint32_t)b * COUT;
            uint32_t v[COUT / 32][32];
#pragma unroll
            for (int half = 0; half < COUT / 32; ++half) {
                tmem_ld_32x32(taddr + half * 32, v[half]);
                tmem_zero_32x32(taddr + half * 32);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_acc_empty + 8 * b);
            if (px >= w_out) continue;
#pragma unroll
            for (int half = 0; half < COUT / 32; ++half) {
                float f[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    f[i] = __uint_as_float(v[half][i]) + s_bias[half * 32 + i];
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

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
    }
}

// fp32 packed weights ([tap = ky*KS + kx][cin][cout]) -> streaming units [kx][plane][kc][n][8] bf16 with
// n = (KS-1-ky)*cout + co (plane 0 = hi, 1 = lo)
__global__ void weight_image_r_kernel(const float *__restrict__ w, int ks, int cin, int cout,
                                      __nv_bfloat16 *__restrict__ img)
{
    const int total = ks * ks * cin * cout;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;      // over [kx][kc][ky][co][k8]
    if (i >= total) return;
    const int k8 = i & 7;
    int r = i >> 3;
    const int co = r % cout;
    r /= cout;
    const int ky = r % ks;
    r /= ks;
    const int kc = r % (cin / 8), kx = r / (cin / 8);
    const float v = w[((size_t)(ky * ks + kx) * cin + kc * 8 + k8) * cout + co];
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    const __nv_bfloat16 lo = __float2bfloat16_rn(v - __bfloat162float(hi));
    const size_t unit = (size_t)cin * ks * cout;              // elements of one (kx, plane)
    const size_t off = ((size_t)kc * ks * cout + (size_t)(ks - 1 - ky) * cout + co) * 8 + k8;
    img[((size_t)kx * 2) * unit + off] = hi;
    img[((size_t)kx * 2 + 1) * unit + off] = lo;
}

}  // namespace tcr

int launch_weight_image_r(stereo_ctx *ctx, const float *layer_w, int ks, int cin, int cout, void *img)
{
    const int total = ks * ks * cin * cout;
    tcr::weight_image_r_kernel<<<(total + 255) / 256, 256, 0, ctx->stream>>>(layer_w, ks, cin, cout,
                                                                            static_cast<__nv_bfloat16 *>(img));
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

template <int KS, int CIN, int COUT>
static int launch_tcr(stereo_ctx *ctx, tcr::Params &p, int n)
{
    using C = tcr::Cfg<KS, CIN, COUT>;
    const int h_out = p.h_in - (KS - 1), w_out = p.w_in - (KS - 1);
    const int strips = (w_out + tcr::kStripW - 1) / tcr::kStripW;
    STEREO_REQUIRE(p.wp_in >= strips * tcr::kStripW + KS - 1, "activation pitch %d too small for %d strips", p.wp_in, strips);
    // one CTA per SM (tensor memory and shared memory are both taken whole): about two waves of row bands
    int segs = (ctx->sm_count * 2) / (strips * n);
    if (segs < 1) segs = 1;
    int rows = (h_out + segs - 1) / segs;
    if (rows < 8) rows = 8;
    p.rows_per_cta = rows;
    segs = (h_out + rows - 1) / rows;
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(tcr::conv_tcr_kernel<KS, CIN, COUT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           C::kSmemTotal));
    ProfScope prof(ctx, STEREO_K_CONV);
    tcr::conv_tcr_kernel<KS, CIN, COUT><<<dim3(strips, segs, n), 256, C::kSmemTotal, ctx->stream>>>(p);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// same contract as launch_conv_tcg; `wimg` is the image launch_weight_image_r built
int launch_conv_tcr(stereo_ctx *ctx, int ks, int cin, int cout, const void *in_hi, const void *in_lo, int n, int h_in,
                    int w_in, int wp_in, const void *wimg, const float *bias, void *out_hi, void *out_lo, int wp_out,
                    float *out_f32, int relu, int planes)
{
    STEREO_REQUIRE(h_in >= ks && w_in >= ks, "input %dx%d smaller than the %dx%d kernel", h_in, w_in, ks, ks);
    tcr::Params p;
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
    if (ks == 5 && cin == 32 && cout == 32) return launch_tcr<5, 32, 32>(ctx, p, n);
    if (ks == 5 && cin == 32 && cout == 64) return launch_tcr<5, 32, 64>(ctx, p, n);
    if (ks == 5 && cin == 64 && cout == 64) return launch_tcr<5, 64, 64>(ctx, p, n);
    set_error("no tensor-core convolution for ks=%d cin=%d cout=%d", ks, cin, cout);
    return STEREO_E_INVALID;
}

}  // namespace stereo
