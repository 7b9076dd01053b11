// fp32 direct convolution on CUDA cores: the STEREO_PREC_FP32 trunk
// (/root/reference/models/model.py:5-41: Conv2D 'valid' + bias [+ ReLU]; the
// inference-mode BatchNormalization is folded into the weights on the host).
//
// One CTA (256 threads) produces an 8 x 32 pixel tile for all COUT channels.
// The input channels are walked in chunks of CK; each chunk's input patch is
// staged in shared memory channel-planar ([c][row][col]) so that a thread reads
// the 8+KS-1 consecutive pixels of one patch row with vector loads and reuses
// them for the KS horizontal taps.  Each thread owns 8 pixels x COUT/8 channels.
#include <cuda_bf16.h>

#include "common.cuh"

namespace stereo {

constexpr int kTH = 8;
constexpr int kTW = 32;

template <int KS, int CK>
struct ConvSmem {
    static constexpr int PH = kTH + KS - 1;
    static constexpr int PW = kTW + KS - 1;
    static constexpr int PWP = (PW + 3) / 4 * 4;
    // plane stride is a multiple of 4 floats (16-byte vector loads) and = 4 mod 32 banks
    static constexpr int PLANE_RAW = PH * PWP;
    static constexpr int PLANE = PLANE_RAW + ((4 - PLANE_RAW % 32) + 32) % 32;
};

template <int KS, int CIN, int COUT, int CK>
__global__ __launch_bounds__(256, 2) void conv_ffma_kernel(const float *__restrict__ in,
                                                           const float *__restrict__ wgt,
                                                           const float *__restrict__ bias,
                                                           float *__restrict__ out, int h_in, int w_in,
                                                           int relu, __nv_bfloat16 *__restrict__ out_hi,
                                                           __nv_bfloat16 *__restrict__ out_lo, int wp_out)
{
    using S = ConvSmem<KS, CK>;
    constexpr int CPT = COUT / 8;          // output channels per thread (4 or 8)
    constexpr int NA = 8 + KS - 1;         // patch-row pixels a thread needs
    constexpr int NA4 = (NA + 3) / 4 * 4;
    extern __shared__ __align__(16) float smem[];
    float *patch = smem;                                   // [CK][PLANE]
    float *wsm = smem + CK * S::PLANE;                     // [KS*KS][CK][COUT]

    const int tid = threadIdx.x;
    const int cg = tid & 7;
    const int pg = tid >> 3;
    const int r = pg >> 2;
    const int col0 = (pg & 3) * 8;
    const int h_out = h_in - KS + 1, w_out = w_in - KS + 1;
    const int x0 = blockIdx.x * kTW, y0 = blockIdx.y * kTH;
    const float *img = in + (size_t)blockIdx.z * h_in * w_in * CIN;

    float acc[8][CPT];
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int o = 0; o < CPT; ++o) acc[j][o] = 0.f;

    for (int c0 = 0; c0 < CIN; c0 += CK) {
        // ---- stage the input patch chunk, transposed to channel-planar ----
        if constexpr (CK % 4 == 0) {
            constexpr int Q = CK / 4;
            for (int it = tid; it < S::PH * S::PW * Q; it += 256) {
                int q = it % Q, pix = it / Q;
                int px = pix % S::PW, py = pix / S::PW;
                int gy = y0 + py, gx = x0 + px;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (gy < h_in && gx < w_in)
                    v = __ldg(reinterpret_cast<const float4 *>(img + ((size_t)gy * w_in + gx) * CIN + c0 + q * 4));
                float *dst = patch + (q * 4) * S::PLANE + py * S::PWP + px;
                dst[0] = v.x;
                dst[S::PLANE] = v.y;
                dst[2 * S::PLANE] = v.z;
                dst[3 * S::PLANE] = v.w;
            }
        } else {
            for (int it = tid; it < S::PH * S::PW * CK; it += 256) {
                int c = it % CK, pix = it / CK;
                int px = pix % S::PW, py = pix / S::PW;
                int gy = y0 + py, gx = x0 + px;
                float v = 0.f;
                if (gy < h_in && gx < w_in) v = __ldg(img + ((size_t)gy * w_in + gx) * CIN + c0 + c);
                patch[c * S::PLANE + py * S::PWP + px] = v;
            }
        }
        // ---- stage the weight chunk [tap][c][cout] ----
        if constexpr ((CK * COUT) % 4 == 0) {
            constexpr int RUN4 = CK * COUT / 4;
            for (int it = tid; it < KS * KS * RUN4; it += 256) {
                int tap = it / RUN4, e = it % RUN4;
                float4 v = __ldg(reinterpret_cast<const float4 *>(wgt + ((size_t)tap * CIN + c0) * COUT) + e);
                reinterpret_cast<float4 *>(wsm + tap * CK * COUT)[e] = v;
            }
        } else {
            for (int it = tid; it < KS * KS * CK * COUT; it += 256) {
                int tap = it / (CK * COUT), e = it % (CK * COUT);
                wsm[it] = __ldg(wgt + ((size_t)tap * CIN + c0) * COUT + e);
            }
        }
        __syncthreads();

#pragma unroll 1
        for (int c = 0; c < CK; ++c) {
#pragma unroll
            for (int ky = 0; ky < KS; ++ky) {
                float a[NA4];
                const float *prow = patch + c * S::PLANE + (r + ky) * S::PWP + col0;
#pragma unroll
                for (int v = 0; v < NA4 / 4; ++v) {
                    float4 t = *reinterpret_cast<const float4 *>(prow + 4 * v);
                    a[4 * v] = t.x; a[4 * v + 1] = t.y; a[4 * v + 2] = t.z; a[4 * v + 3] = t.w;
                }
#pragma unroll
                for (int kx = 0; kx < KS; ++kx) {
                    const float *wp = wsm + ((ky * KS + kx) * CK + c) * COUT + cg * 4;
                    float w[CPT];
                    float4 w0 = *reinterpret_cast<const float4 *>(wp);
                    w[0] = w0.x; w[1] = w0.y; w[2] = w0.z; w[3] = w0.w;
                    if constexpr (CPT == 8) {
                        float4 w1 = *reinterpret_cast<const float4 *>(wp + 32);
                        w[4] = w1.x; w[5] = w1.y; w[6] = w1.z; w[7] = w1.w;
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j)
#pragma unroll
                        for (int o = 0; o < CPT; ++o) acc[j][o] = fmaf(a[kx + j], w[o], acc[j][o]);
                }
            }
        }
        __syncthreads();
    }

    // ---- epilogue: bias, optional ReLU, NHWC store ----
    const int gy = y0 + r;
    if (gy >= h_out) return;
    float b[CPT];
#pragma unroll
    for (int o = 0; o < 4; ++o) b[o] = __ldg(bias + cg * 4 + o);
    if constexpr (CPT == 8) {
#pragma unroll
        for (int o = 0; o < 4; ++o) b[4 + o] = __ldg(bias + 32 + cg * 4 + o);
    }
    float *orow = out + ((size_t)blockIdx.z * h_out + gy) * w_out * COUT;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        int gx = x0 + col0 + j;
        if (gx >= w_out) break;
        float v[CPT];
#pragma unroll
        for (int o = 0; o < CPT; ++o) {
            v[o] = acc[j][o] + b[o];
            if (relu) v[o] = fmaxf(v[o], 0.f);
        }
        if (out_hi != nullptr) {
            // chunk-planar bf16 (hi, lo) for the tensor-core layers: [img][y][c/8][x][c%8]
            if constexpr (COUT == 64) {
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    uint32_t hi[2], lo[2];
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        float a0 = v[4 * g + 2 * e], a1 = v[4 * g + 2 * e + 1];
                        __nv_bfloat16 h0 = __float2bfloat16_rn(a0), h1 = __float2bfloat16_rn(a1);
                        __nv_bfloat16 l0 = __float2bfloat16_rn(a0 - __bfloat162float(h0));
                        __nv_bfloat16 l1 = __float2bfloat16_rn(a1 - __bfloat162float(h1));
                        hi[e] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
                        lo[e] = (uint32_t)__bfloat16_as_ushort(l0) | ((uint32_t)__bfloat16_as_ushort(l1) << 16);
                    }
                    int cc = 4 * g + (cg >> 1);
                    size_t off = ((((size_t)blockIdx.z * h_out + gy) * 8 + cc) * wp_out + gx) * 8 + (cg & 1) * 4;
                    *reinterpret_cast<uint2 *>(out_hi + off) = make_uint2(hi[0], hi[1]);
                    if (out_lo != nullptr) *reinterpret_cast<uint2 *>(out_lo + off) = make_uint2(lo[0], lo[1]);
                }
            } else if constexpr (COUT == 32) {
                // four chunks of eight channels; this thread holds channels 4 cg .. 4 cg + 3
                uint32_t hi[2], lo[2];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    float a0 = v[2 * e], a1 = v[2 * e + 1];
                    __nv_bfloat16 h0 = __float2bfloat16_rn(a0), h1 = __float2bfloat16_rn(a1);
                    __nv_bfloat16 l0 = __float2bfloat16_rn(a0 - __bfloat162float(h0));
                    __nv_bfloat16 l1 = __float2bfloat16_rn(a1 - __bfloat162float(h1));
                    hi[e] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
                    lo[e] = (uint32_t)__bfloat16_as_ushort(l0) | ((uint32_t)__bfloat16_as_ushort(l1) << 16);
                }
                size_t off = ((((size_t)blockIdx.z * h_out + gy) * 4 + (cg >> 1)) * wp_out + gx) * 8 + (cg & 1) * 4;
                *reinterpret_cast<uint2 *>(out_hi + off) = make_uint2(hi[0], hi[1]);
                if (out_lo != nullptr) *reinterpret_cast<uint2 *>(out_lo + off) = make_uint2(lo[0], lo[1]);
            }
            continue;
        }
        float *p = orow + (size_t)gx * COUT + cg * 4;
        *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]);
        if constexpr (CPT == 8) *reinterpret_cast<float4 *>(p + 32) = make_float4(v[4], v[5], v[6], v[7]);
    }
}

template <int KS, int CIN, int COUT, int CK>
static int launch_one(stereo_ctx *ctx, const LayerDesc &L, const float *weights, const float *in, int n,
                      int h_in, int w_in, float *out, void *out_hi = nullptr, void *out_lo = nullptr,
                      int wp_out = 0)
{
    using S = ConvSmem<KS, CK>;
    constexpr size_t smem = sizeof(float) * (CK * S::PLANE + KS * KS * CK * COUT);
    static_assert(S::PLANE % 4 == 0, "plane stride must keep 16-byte alignment");
    auto kern = conv_ffma_kernel<KS, CIN, COUT, CK>;
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int h_out = h_in - KS + 1, w_out = w_in - KS + 1;
    dim3 grid((w_out + kTW - 1) / kTW, (h_out + kTH - 1) / kTH, n);
    ProfScope prof(ctx, CIN == 3 ? STEREO_K_CONV_FIRST : STEREO_K_CONV);
    kern<<<grid, 256, smem, ctx->stream>>>(in, weights + L.w_off, weights + L.b_off, out, h_in, w_in, L.relu,
                                           static_cast<__nv_bfloat16 *>(out_hi), static_cast<__nv_bfloat16 *>(out_lo),
                                           wp_out);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_conv_ffma(stereo_ctx *ctx, const LayerDesc &L, const float *weights, const float *in, int n,
                     int h_in, int w_in, float *out)
{
    STEREO_REQUIRE(h_in >= L.ks && w_in >= L.ks, "input %dx%d smaller than the %dx%d kernel", h_in, w_in, L.ks, L.ks);
    if (L.ks == 3 && L.cin == 3 && L.cout == 64) return launch_one<3, 3, 64, 3>(ctx, L, weights, in, n, h_in, w_in, out);
    if (L.ks == 3 && L.cin == 64 && L.cout == 64) return launch_one<3, 64, 64, 16>(ctx, L, weights, in, n, h_in, w_in, out);
    if (L.ks == 5 && L.cin == 3 && L.cout == 32) return launch_one<5, 3, 32, 3>(ctx, L, weights, in, n, h_in, w_in, out);
    if (L.ks == 5 && L.cin == 32 && L.cout == 32) return launch_one<5, 32, 32, 8>(ctx, L, weights, in, n, h_in, w_in, out);
    if (L.ks == 5 && L.cin == 32 && L.cout == 64) return launch_one<5, 32, 64, 8>(ctx, L, weights, in, n, h_in, w_in, out);
    if (L.ks == 5 && L.cin == 64 && L.cout == 64) return launch_one<5, 64, 64, 8>(ctx, L, weights, in, n, h_in, w_in, out);
    set_error("no fp32 convolution kernel for ks=%d cin=%d cout=%d", L.ks, L.cin, L.cout);
    return STEREO_E_INVALID;
}

// First layer of base_model_ws_9 / base_model with the output written as chunk-planar bf16 (hi, lo)
int launch_conv_first_split(stereo_ctx *ctx, const LayerDesc &L, const float *weights, const float *in, int n,
                            int h_in, int w_in, void *out_hi, void *out_lo, int wp_out)
{
    if (L.ks == 5 && L.cin == 3 && L.cout == 32)        // first layer of base_model
        return launch_one<5, 3, 32, 3>(ctx, L, weights, in, n, h_in, w_in, nullptr, out_hi, out_lo, wp_out);
    STEREO_REQUIRE(L.ks == 3 && L.cin == 3 && L.cout == 64, "split output is only built for the first layers (3x3 3->64, 5x5 3->32)");
    return launch_one<3, 3, 64, 3>(ctx, L, weights, in, n, h_in, w_in, nullptr, out_hi, out_lo, wp_out);
}

}  // namespace stereo
