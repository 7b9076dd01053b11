// Image normalisation + zero padding (/root/reference/inference.py:109-133).
//
// The reference computes (x - x.mean()) / x.std() over all pixels and channels
// of a uint8 image in float64 and stores the result in a float32 canvas.  A
// uint8 image has only 256 distinct values, so the kernel sums the image
// exactly in integers, builds a 256-entry float LUT in float64 and the
// normalise+pad pass is a table lookup.
#include "common.cuh"

namespace stereo {

__global__ void image_sums_kernel(const uint8_t *__restrict__ rgb, size_t per_image, ImageStats *stats,
                                  int stat_stride = 1)
{
    const uint8_t *img = rgb + (size_t)blockIdx.y * per_image;
    unsigned long long s = 0, ss = 0;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    // 16-byte vector loads over the aligned body, scalar head/tail
    size_t head = (16 - ((size_t)img & 15)) & 15;
    if (head > per_image) head = per_image;
    size_t body = (per_image - head) / 16;
    const uint4 *v = reinterpret_cast<const uint4 *>(img + head);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < body; i += stride) {
        uint4 q = __ldg(v + i);
        unsigned w[4] = {q.x, q.y, q.z, q.w};
        unsigned ls = 0, lss = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                unsigned x = (w[k] >> (8 * b)) & 0xffu;
                ls += x;
                lss += x * x;
            }
        }
        s += ls;
        ss += lss;
    }
    size_t tail0 = head + body * 16;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < head + (per_image - tail0); i += stride) {
        size_t idx = i < head ? i : tail0 + (i - head);
        unsigned x = img[idx];
        s += x;
        ss += x * x;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_down_sync(0xffffffffu, s, o);
        ss += __shfl_down_sync(0xffffffffu, ss, o);
    }
    __shared__ unsigned long long sh[2][32];
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { sh[0][warp] = s; sh[1][warp] = ss; }
    __syncthreads();
    if (warp == 0) {
        int nw = blockDim.x >> 5;
        s = lane < nw ? sh[0][lane] : 0;
        ss = lane < nw ? sh[1][lane] : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            s += __shfl_down_sync(0xffffffffu, s, o);
            ss += __shfl_down_sync(0xffffffffu, ss, o);
        }
        if (lane == 0) {
            atomicAdd(&stats[blockIdx.y * stat_stride].sum, s);
            atomicAdd(&stats[blockIdx.y * stat_stride].sumsq, ss);
        }
    }
}

__global__ void build_lut_kernel(ImageStats *stats, double count)
{
    ImageStats &st = stats[blockIdx.x];
    // population variance = (N*SS - S^2) / N^2, numerator exact in 128-bit integers
    unsigned __int128 nss = (unsigned __int128)(unsigned long long)count * st.sumsq;
    unsigned __int128 s2 = (unsigned __int128)st.sum * st.sum;
    unsigned __int128 num = nss - s2;
    double numd = (double)(unsigned long long)(num >> 64) * 18446744073709551616.0 +
                  (double)(unsigned long long)num;
    double mean = (double)st.sum / count;
    double sd = sqrt(numd / (count * count));
    int v = threadIdx.x;
    st.lut[v] = (float)(((double)v - mean) / sd);
}

__global__ void normalize_pad_kernel(const uint8_t *__restrict__ rgb, const ImageStats *__restrict__ stats,
                                     int h, int w, int pad, float *__restrict__ out)
{
    __shared__ float lut[256];
    int img = blockIdx.y;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) lut[i] = stats[img].lut[i];
    __syncthreads();
    int hp = h + 2 * pad, wp = w + 2 * pad;
    size_t total = (size_t)hp * wp * 3;
    const uint8_t *src = rgb + (size_t)img * h * w * 3;
    float *dst = out + (size_t)img * total;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int c = (int)(i % 3);
        size_t p = i / 3;
        int x = (int)(p % wp) - pad;
        int y = (int)(p / wp) - pad;
        float v = 0.f;
        if (x >= 0 && x < w && y >= 0 && y < h) v = lut[src[((size_t)y * w + x) * 3 + c]];
        dst[i] = v;
    }
}

// sums + LUT of n pairs whose left and right images are stored in two separate arrays, into
// ctx->stats[2i] (left of pair i) and ctx->stats[2i + 1] (right): three launches for the whole batch
int launch_pair_stats(stereo_ctx *ctx, const uint8_t *left, const uint8_t *right, int n, int h, int w)
{
    STEREO_TRY(ctx->stats.reserve(sizeof(ImageStats) * 2 * (size_t)n));
    ImageStats *stats = ctx->stats.as<ImageStats>();
    ProfScope prof(ctx, STEREO_K_PREP);
    STEREO_CUDA_CHECK(cudaMemsetAsync(stats, 0, sizeof(ImageStats) * 2 * (size_t)n, ctx->stream));
    size_t per_image = (size_t)h * w * 3;
    int bx = (int)((per_image / 16 + 255) / 256);
    const int cap = n >= 8 ? 32 : ctx->sm_count * 2;
    if (bx > cap) bx = cap;
    if (bx < 1) bx = 1;
    image_sums_kernel<<<dim3(bx, n), 256, 0, ctx->stream>>>(left, per_image, stats, 2);
    image_sums_kernel<<<dim3(bx, n), 256, 0, ctx->stream>>>(right, per_image, stats + 1, 2);
    build_lut_kernel<<<2 * n, 256, 0, ctx->stream>>>(stats, (double)per_image);
    ctx->launches += 3;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_normalize_pad(stereo_ctx *ctx, const uint8_t *rgb, int n, int h, int w, int patch, float *out)
{
    STEREO_REQUIRE(patch >= 1 && (patch & 1), "patch size must be odd and positive, got %d", patch);
    STEREO_TRY(ctx->stats.reserve(sizeof(ImageStats) * (size_t)n));
    ImageStats *stats = ctx->stats.as<ImageStats>();
    ProfScope prof(ctx, STEREO_K_PREP);
    STEREO_CUDA_CHECK(cudaMemsetAsync(stats, 0, sizeof(ImageStats) * (size_t)n, ctx->stream));
    size_t per_image = (size_t)h * w * 3;
    int bx = (int)((per_image / 16 + 255) / 256);
    if (bx > ctx->sm_count * 4) bx = ctx->sm_count * 4;
    if (bx < 1) bx = 1;
    image_sums_kernel<<<dim3(bx, n), 256, 0, ctx->stream>>>(rgb, per_image, stats);
    build_lut_kernel<<<n, 256, 0, ctx->stream>>>(stats, (double)per_image);
    int pad = (patch - 1) / 2;
    size_t total = (size_t)(h + 2 * pad) * (w + 2 * pad) * 3;
    int gx = (int)((total + 255) / 256);
    if (gx > ctx->sm_count * 8) gx = ctx->sm_count * 8;
    normalize_pad_kernel<<<dim3(gx, n), 256, 0, ctx->stream>>>(rgb, stats, h, w, pad, out);
    ctx->launches += 3;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// normalise + pad one image with statistics that are already in ctx->stats[stats_index] (launch_pair_stats)
int launch_pad_with_stats(stereo_ctx *ctx, const uint8_t *rgb, int h, int w, int patch, int stats_index, float *out)
{
    STEREO_REQUIRE(patch >= 1 && (patch & 1), "patch size must be odd and positive, got %d", patch);
    ProfScope prof(ctx, STEREO_K_PREP);
    const int pad = (patch - 1) / 2;
    const size_t total = (size_t)(h + 2 * pad) * (w + 2 * pad) * 3;
    int gx = (int)((total + 255) / 256);
    if (gx > ctx->sm_count * 8) gx = ctx->sm_count * 8;
    normalize_pad_kernel<<<dim3(gx, 1), 256, 0, ctx->stream>>>(rgb, ctx->stats.as<ImageStats>() + stats_index, h, w, pad, out);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
