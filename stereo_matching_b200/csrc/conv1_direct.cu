// First trunk layer (3x3, 3 -> 64, ReLU; /root/reference/models/model.py:32) computed straight
// from the uint8 image: normalisation (inference.py:109-110) is a 256-entry table lookup, the zero
// padding of inference.py:125-133 is an index test, and the result is written as the chunk-planar
// bf16 (hi, lo) pair the tensor-core layers read.  The padded fp32 image never exists.
//
// K is only 27, so this layer stays on the CUDA cores in fp32: one thread per output pixel keeps
// all 64 accumulators in registers, walks down ROWS output rows re-using two of the three input
// rows from registers, and reads the weights as shared-memory broadcasts.  Lanes hold consecutive
// pixels, so every output store is a 512-byte contiguous run of one (row, channel-chunk) plane.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {

constexpr int kC1Pad = 4;       // (patch - 1) / 2 of the 9x9 receptive field

// The 27x64 weights and the bias travel as a kernel parameter: they sit in the constant bank, so
// every FFMA takes its weight operand straight from there (uniform across the warp) and the inner
// loop has no load instructions at all.
__global__ __launch_bounds__(128, 3) void conv1_direct_kernel(const uint8_t *__restrict__ img0,
                                                           const uint8_t *__restrict__ img1,
                                                           const ImageStats *__restrict__ stats, int h, int w,
                                                           const __grid_constant__ Conv1Weights cw,
                                                           __nv_bfloat16 *__restrict__ out_hi,
                                                           __nv_bfloat16 *__restrict__ out_lo, int wp_out, int rows_per_cta)
{
    __shared__ float s_lut[256];
    const int tid = threadIdx.x;
    const int im = blockIdx.z;
    const uint8_t *img = im == 0 ? img0 : img1;
    for (int i = tid; i < 256; i += 128) s_lut[i] = stats[im].lut[i];
    __syncthreads();

    const int h1 = h + 2 * kC1Pad - 2, w1o = w + 2 * kC1Pad - 2;      // layer-1 output size
    const int px = blockIdx.x * 128 + tid;
    const int y_begin = blockIdx.y * rows_per_cta;
    const int y_end = min(y_begin + rows_per_cta, h1);
    if (px >= w1o) return;

    // input row of the padded image -> 9 normalised values (3 columns x 3 channels) for this pixel
    auto load_row = [&](int yp, float (&dst)[9]) {
        const int yo = yp - kC1Pad;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
            const int xo = px + kx - kC1Pad;
            const bool in = yo >= 0 && yo < h && xo >= 0 && xo < w;
            const uint8_t *p = img + ((size_t)(in ? yo : 0) * w + (in ? xo : 0)) * 3;
#pragma unroll
            for (int c = 0; c < 3; ++c) dst[kx * 3 + c] = in ? s_lut[p[c]] : 0.f;
        }
    };
    float r0[9], r1[9], r2[9];
    load_row(y_begin, r0);
    load_row(y_begin + 1, r1);
    for (int y = y_begin; y < y_end; ++y) {
        load_row(y + 2, r2);
        float acc[64];
#pragma unroll
        for (int o = 0; o < 64; ++o) acc[o] = cw.b[o];
#pragma unroll
        for (int k = 0; k < 27; ++k) {
            const float a = k < 9 ? r0[k] : (k < 18 ? r1[k - 9] : r2[k - 18]);
#pragma unroll
            for (int o = 0; o < 64; ++o) acc[o] = fmaf(a, cw.w[k * 64 + o], acc[o]);
        }
#pragma unroll
        for (int cc = 0; cc < 8; ++cc) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int e = 0; e < 4; ++e)
                tc::split_bf16x2(fmaxf(acc[8 * cc + 2 * e], 0.f), fmaxf(acc[8 * cc + 2 * e + 1], 0.f), hi[e], lo[e]);
            const size_t off = ((((size_t)im * h1 + y) * 8 + cc) * wp_out + px) * 8;
            *reinterpret_cast<uint4 *>(out_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            if (out_lo) *reinterpret_cast<uint4 *>(out_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
#pragma unroll
        for (int i = 0; i < 9; ++i) { r0[i] = r1[i]; r1[i] = r2[i]; }
    }
}

int launch_conv1_direct(stereo_ctx *ctx, const uint8_t *img0, const uint8_t *img1, int h, int w,
                        const Conv1Weights &cw, void *out_hi, void *out_lo, int wp_out, int stats_off)
{
    const int h1 = h + 2 * kC1Pad - 2, w1o = w + 2 * kC1Pad - 2;
    // 3 CTAs fit on an SM (168 registers): size the row segments so that the grid is one full wave
    const int strips = (w1o + 127) / 128;
    int segs = (ctx->sm_count * 3) / (strips * 2);
    if (segs < 1) segs = 1;
    int rows = (h1 + segs - 1) / segs;
    if (rows < 4) rows = 4;
    dim3 grid(strips, (h1 + rows - 1) / rows, 2);
    ProfScope prof(ctx, STEREO_K_CONV_FIRST);
    conv1_direct_kernel<<<grid, 128, 0, ctx->stream>>>(img0, img1, ctx->stats.as<ImageStats>() + stats_off, h, w, cw,
                                                       static_cast<__nv_bfloat16 *>(out_hi),
                                                       static_cast<__nv_bfloat16 *>(out_lo), wp_out, rows);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
