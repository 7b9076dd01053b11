// First trunk layer (3x3, 3 -> 64, ReLU; /root/reference/models/model.py:32) on the tensor cores,
// straight from the uint8 image.
//
// conv1_direct.cu spends 1728 FFMA per output pixel on the CUDA cores (8-9 ms of a 54 ms step).
// K is only 27, but as an implicit GEMM the layer is nearly free on tcgen05: for one output row the
// nine values (3 input rows x 3 colours) under a pixel column form one 16-element K slab
// (9 values + 7 zeros), and the three kx taps are three MMAs whose A operand starts one pixel
// (16 bytes) further along - the same trick as the 64->64 layers.  Per 128-pixel row segment that is
// 9 small MMAs (3 kx x the bf16 hi/lo split); what remains is building the operand and the epilogue,
// i.e. the kernel is bound by its 256 bytes of output per pixel.
//
// Roles inside a CTA (128-pixel strip x a run of output rows):
//   builder warps : read the image bytes, normalise through the per-image table (inference.py:109-110;
//                   zero outside the image = the padding of inference.py:125-133), split into bf16
//                   hi/lo and write the [k/8][pixel][8] operand of the next row into shared memory
//   issuer        : 9 MMAs (M=128, N=64, K=16) per row into a ring of 4 accumulators
//   epilogue warps: tcgen05.ld, bias, ReLU, bf16 hi/lo split, chunk-planar stores (the layout the
//                   tensor-core layers read)
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {

namespace c1 {

using namespace tc;

constexpr int kPad = 4;                          // (patch - 1) / 2 of the 9x9 receptive field
constexpr int kStripW = 128;
constexpr int kPW = kStripW + 2;                 // operand columns incl. the 2-pixel halo of the kx taps
constexpr int kChunkBytes = kPW * 16;            // one 8-element K chunk of all columns: 2080 B
constexpr int kAPlane = 2 * kChunkBytes;         // K = 16
constexpr int kASlot = 2 * kAPlane;              // hi + lo
constexpr int kASlots = 3;
constexpr int kWTap = 2 * 64 * 16;               // one (plane, kx): [2 chunks][64 n][8] = 2 KB
constexpr int kSmemW = 0;                        // [plane][kx] = 12 KB
constexpr int kSmemA = kSmemW + 6 * kWTap;
constexpr int kSmemLut = kSmemA + kASlots * kASlot;
constexpr int kSmemBias = kSmemLut + 256 * 4;
constexpr int kSmemBars = kSmemBias + 64 * 4;
constexpr int kSmemTotal = kSmemBars + 128;
constexpr int kAccBufs = 4;
constexpr uint32_t kTmemCols = 256;
constexpr int kEpiWarps = 8, kBuildWarps = 5;
constexpr int kThreads = 128 + 32 * (kEpiWarps + kBuildWarps);

__global__ __launch_bounds__(kThreads, 2) void conv1_tc_kernel(const uint8_t *__restrict__ img0,
                                                              const uint8_t *__restrict__ img1,
                                                              const ImageStats *__restrict__ stats, int h, int w,
                                                              const __nv_bfloat16 *__restrict__ wimg,
                                                              const float *__restrict__ bias,
                                                              __nv_bfloat16 *__restrict__ out_hi,
                                                              __nv_bfloat16 *__restrict__ out_lo, int wp_out,
                                                              int rows_per_cta)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int im = blockIdx.z;
    const uint8_t *img = im == 0 ? img0 : img1;
    const int h1 = h + 2 * kPad - 2, w1 = w + 2 * kPad - 2;          // layer-1 output size
    const int x0 = blockIdx.x * kStripW;
    const int y0 = blockIdx.y * rows_per_cta;
    const int rows = min(rows_per_cta, h1 - y0);

    const uint32_t s_base = smem_u32(smem);
    const uint32_t s_w = s_base + kSmemW, s_a = s_base + kSmemA;
    float *s_lut = reinterpret_cast<float *>(smem + kSmemLut);
    float *s_bias = reinterpret_cast<float *>(smem + kSmemBias);
    const uint32_t bar_a_full = s_base + kSmemBars;          // [3] operand of a row built
    const uint32_t bar_a_empty = bar_a_full + 24;            // [3] operand consumed (commit)
    const uint32_t bar_acc_full = bar_a_full + 48;           // [4]
    const uint32_t bar_acc_empty = bar_a_full + 80;          // [4]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + kSmemBars + 112);

    // ---- setup: barriers, table, bias, weight operand images ([plane][kx][k/8][n][8], k = 3 ky + c) ----
    if (tid == 0) {
        for (int i = 0; i < kASlots; ++i) {
            mbar_init(bar_a_full + 8 * i, kBuildWarps);
            mbar_init(bar_a_empty + 8 * i, 1);
        }
        for (int i = 0; i < kAccBufs; ++i) {
            mbar_init(bar_acc_full + 8 * i, 1);
            mbar_init(bar_acc_empty + 8 * i, kEpiWarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < 256; i += kThreads) s_lut[i] = stats[im].lut[i];
    if (tid < 64) s_bias[tid] = __ldg(bias + tid);
    for (int i = tid; i < 6 * kWTap / 16; i += kThreads)
        reinterpret_cast<uint4 *>(smem + kSmemW)[i] = __ldg(reinterpret_cast<const uint4 *>(wimg) + i);
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(kTmemCols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // weight images -> visible to the tensor core
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 1) {
        // ---- MMA issuer (whole warp runs the loop, one elected lane issues) ----
        constexpr uint32_t idesc = make_idesc(kStripW, 64);
        const uint32_t a_hi_word = desc_hi(128), b_hi_word = desc_hi(128);
        for (int r = 0; r < rows; ++r) {
            const uint32_t as = r % kASlots, ab = r % kAccBufs;
            mbar_wait(bar_a_full + 8 * as, (r / kASlots) & 1);
            mbar_wait(bar_acc_empty + 8 * ab, ((r / kAccBufs) & 1) ^ 1);
            tc_fence_after();
            uint32_t acc = 0;
#pragma unroll
            for (int combo = 0; combo < 3; ++combo) {
                const int apl = combo == 2, wpl = combo == 1;            // hi*hi, hi*lo, lo*hi
                if (combo > 0 && out_lo == nullptr) break;               // single-pass bf16 mode
#pragma unroll
                for (int kx = 0; kx < 3; ++kx) {
                    if (elect_one())
                        umma_bf16_parts(tmem_base + ab * 64,
                                        desc_lo(s_a + as * kASlot + apl * kAPlane + kx * 16, kChunkBytes), a_hi_word,
                                        desc_lo(s_w + (wpl * 3 + kx) * kWTap, 64 * 16), b_hi_word, idesc, acc);
                    acc = 1;
                }
            }
            if (elect_one()) {
                tc_commit(bar_a_empty + 8 * as);
                tc_commit(bar_acc_full + 8 * ab);
            }
        }
    } else if (warp >= 4 && warp < 4 + kEpiWarps) {
        // ---- epilogue: warp = (lane quarter, channel half) ----
        const int q = warp & 3, hf = (warp - 4) >> 2;
        const int px = x0 + q * 32 + lane;
        for (int r = 0; r < rows; ++r) {
            const uint32_t ab = r % kAccBufs;
            mbar_wait(bar_acc_full + 8 * ab, (r / kAccBufs) & 1);
            tc_fence_after();
            uint32_t v[32];
            tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + ab * 64 + hf * 32, v);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_acc_empty + 8 * ab);
            if (px >= w1) continue;
            const int y = y0 + r;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t hi[4], lo[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int k = 8 * c + 2 * e;
                    split_bf16x2(fmaxf(__uint_as_float(v[k]) + s_bias[hf * 32 + k], 0.f),
                                 fmaxf(__uint_as_float(v[k + 1]) + s_bias[hf * 32 + k + 1], 0.f), hi[e], lo[e]);
                }
                const size_t off = ((((size_t)im * h1 + y) * 8 + (hf * 4 + c)) * wp_out + px) * 8;
                *reinterpret_cast<uint4 *>(out_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                if (out_lo) *reinterpret_cast<uint4 *>(out_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
        }
    } else if (warp >= 4 + kEpiWarps) {
        // ---- operand builders: thread i owns operand column i (padded-image column x0 + i) ----
        const int col = tid - 32 * (4 + kEpiWarps);                  // 0..159, columns >= 130 idle
        const bool active = col < kPW;
        const int xi = x0 + col - kPad;                              // image column
        const bool x_in = active && xi >= 0 && xi < w;
        // raw bytes of one padded-image row at this column (0xffffffff = outside the image), fetched one
        // row ahead of their use so that the load latency overlaps the previous row's work
        auto fetch = [&](int yp) -> uint32_t {
            const int yi = yp - kPad;
            if (!(x_in && yi >= 0 && yi < h)) return 0xffffffffu;
            const uint8_t *px = img + ((size_t)yi * w + xi) * 3;
            return (uint32_t)px[0] | ((uint32_t)px[1] << 8) | ((uint32_t)px[2] << 16);
        };
        // normalised (hi, lo) bf16 of the three colours, packed as hi0|hi1, hi2|0, lo0|lo1, lo2|0
        auto convert = [&](uint32_t raw, uint32_t (&o)[4]) {
            float f0 = 0.f, f1 = 0.f, f2 = 0.f;
            if (raw != 0xffffffffu) {
                f0 = s_lut[raw & 255];
                f1 = s_lut[(raw >> 8) & 255];
                f2 = s_lut[(raw >> 16) & 255];
            }
            split_bf16x2(f0, f1, o[0], o[2]);
            split_bf16x2(f2, 0.f, o[1], o[3]);
        };
        uint32_t r0[4], r1[4], r2[4];
        convert(fetch(y0), r0);
        convert(fetch(y0 + 1), r1);
        uint32_t raw_next = fetch(y0 + 2);
        for (int r = 0; r < rows; ++r) {
            const uint32_t raw = raw_next;
            raw_next = fetch(y0 + r + 3);
            convert(raw, r2);
            const uint32_t as = r % kASlots;
            mbar_wait(bar_a_empty + 8 * as, ((r / kASlots) & 1) ^ 1);
            if (active) {
                // K order: r0c0 r0c1 r0c2 r1c0 r1c1 r1c2 r2c0 r2c1 | r2c2 0 0 0 0 0 0 0
                uint8_t *slot = smem + kSmemA + as * kASlot + col * 16;
#pragma unroll
                for (int pl = 0; pl < 2; ++pl) {
                    const uint32_t a0 = r0[2 * pl], a1 = r0[2 * pl + 1];     // c0|c1, c2|0
                    const uint32_t b0 = r1[2 * pl], b1 = r1[2 * pl + 1];
                    const uint32_t c0 = r2[2 * pl], c1 = r2[2 * pl + 1];
                    uint4 k0;
                    k0.x = a0;                                               // r0c0 r0c1
                    k0.y = (a1 & 0xffffu) | (b0 << 16);                      // r0c2 r1c0
                    k0.z = (b0 >> 16) | (b1 << 16);                          // r1c1 r1c2
                    k0.w = c0;                                               // r2c0 r2c1
                    *reinterpret_cast<uint4 *>(slot + pl * kAPlane) = k0;
                    *reinterpret_cast<uint4 *>(slot + pl * kAPlane + kChunkBytes) = make_uint4(c1 & 0xffffu, 0, 0, 0);
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> tensor core reads
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_a_full + 8 * as);
#pragma unroll
            for (int i = 0; i < 4; ++i) { r0[i] = r1[i]; r1[i] = r2[i]; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
    }
}

// [plane][kx][k/8][n][8] bf16 operand images of the 27x64 weights, k = 3 ky + c (zero for k >= 9)
__global__ void conv1_weight_image_kernel(const float *__restrict__ w, __nv_bfloat16 *__restrict__ img)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;             // over [kx][chunk][n][k8]
    if (i >= 3 * 2 * 64 * 8) return;
    const int k8 = i & 7, n = (i >> 3) & 63, ch = (i >> 9) & 1, kx = i >> 10;
    const int k = ch * 8 + k8;
    float v = 0.f;
    if (k < 9) v = w[(((k / 3) * 3 + kx) * 3 + (k % 3)) * 64 + n];
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    img[i] = hi;
    img[3 * 2 * 64 * 8 + i] = __float2bfloat16_rn(v - __bfloat162float(hi));
}

}  // namespace c1

size_t conv1_tc_weight_image_elems() { return (size_t)2 * 3 * 2 * 64 * 8; }

int launch_conv1_weight_image(stereo_ctx *ctx, const float *layer_w, void *img)
{
    c1::conv1_weight_image_kernel<<<(3 * 2 * 64 * 8 + 255) / 256, 256, 0, ctx->stream>>>(
        layer_w, static_cast<__nv_bfloat16 *>(img));
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_conv1_tc(stereo_ctx *ctx, const uint8_t *img0, const uint8_t *img1, int h, int w, const void *wimg,
                    const float *bias, void *out_hi, void *out_lo, int wp_out, int stats_off)
{
    const int h1 = h + 2 * c1::kPad - 2, w1 = w + 2 * c1::kPad - 2;
    const int strips = (w1 + c1::kStripW - 1) / c1::kStripW;
    // about two CTAs per SM: the row segments of the two images x strips should fill the machine twice
    int segs = (ctx->sm_count * 2) / (strips * 2);
    if (segs < 1) segs = 1;
    int rows = (h1 + segs - 1) / segs;
    if (rows < 8) rows = 8;
    dim3 grid(strips, (h1 + rows - 1) / rows, 2);
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(c1::conv1_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           c1::kSmemTotal));
    ProfScope prof(ctx, STEREO_K_CONV_FIRST);
    c1::conv1_tc_kernel<<<grid, c1::kThreads, c1::kSmemTotal, ctx->stream>>>(
        img0, img1, ctx->stats.as<ImageStats>() + stats_off, h, w, static_cast<const __nv_bfloat16 *>(wimg), bias,
        static_cast<__nv_bfloat16 *>(out_hi),
        static_cast<__nv_bfloat16 *>(out_lo), wp_out, rows);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
