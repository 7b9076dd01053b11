// The post-processing chain of /root/reference/post.lua + smooth/nyu.lua, i.e. the
// mc-cnn "adcensus" kernels the reference loads from an external library
// (smooth/init.lua:6), rebuilt for sm_100a.
//
// Layout decision: the reference shuttles every volume between [H][W][D] (files,
// SGM) and [D][H][W] (CBCA, arg-min) with three permute+clone copies per volume
// (post.lua:119-120, nyu.lua:52,64).  Here every kernel works on [H][W][D] with
// D fastest - warps run along D, so CBCA gathers, SGM scanline steps and the
// arg-min all read 128-byte lines - and no transpose is ever made.
//
// Arithmetic follows the restated CUDA original operation for operation (same
// fp32 operations in the same order) so that results match oracle/post_oracle.c
// bit for bit; the only place where nvcc's FMA contraction is part of the
// original's arithmetic (mean2d) uses an explicit fmaf, everything else is
// compiled without contraction.
#include <algorithm>

#include "common.cuh"
#include "exact_div.cuh"
#include "median25.cuh"
#include "sgm_step.cuh"

namespace stereo {

// ------------------------------------------------------------------------------------------
// P0  guide image: byte image -> float, subtract mean, divide by the unbiased std
//     (post.lua:91-94).  256 possible values -> exact histogram, float64 moments, LUT.
// ------------------------------------------------------------------------------------------
struct GuideStats {
    unsigned long long hist[256];
    float lut[256];
};

__global__ void guide_hist_kernel(const uint8_t *__restrict__ img, size_t n, GuideStats *st)
{
    __shared__ unsigned int sh[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        atomicAdd(&sh[img[i]], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i < 256; i += blockDim.x)
        if (sh[i]) atomicAdd(&st->hist[i], (unsigned long long)sh[i]);
}

__global__ void guide_lut_kernel(GuideStats *st, double n)
{
    __shared__ double red[256];
    int v = threadIdx.x;
    double cnt = (double)st->hist[v];
    red[v] = cnt * (double)v;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (v < s) red[v] += red[v + s];
        __syncthreads();
    }
    float mean = (float)(red[0] / n);
    __syncthreads();
    float shifted = __fadd_rn((float)v, -mean);          // x:add(-mean) in fp32
    red[v] = cnt * (double)shifted;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (v < s) red[v] += red[v + s];
        __syncthreads();
    }
    double m2 = red[0] / n;                               // mean of the shifted image (about 0)
    __syncthreads();
    red[v] = cnt * ((double)shifted - m2) * ((double)shifted - m2);
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (v < s) red[v] += red[v + s];
        __syncthreads();
    }
    float sd = (float)sqrt(red[0] / (n - 1.0));           // torch :std() is unbiased
    st->lut[v] = __fdiv_rn(shifted, sd);
}

__global__ void guide_apply_kernel(const uint8_t *__restrict__ img, size_t n, const GuideStats *__restrict__ st,
                                   float *__restrict__ out)
{
    __shared__ float lut[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) lut[i] = st->lut[i];
    __syncthreads();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = lut[img[i]];
}

int launch_guide(stereo_ctx *ctx, const uint8_t *img, int h, int w, float *out)
{
    STEREO_TRY(ctx->stats.reserve(sizeof(GuideStats)));
    GuideStats *st = ctx->stats.as<GuideStats>();
    size_t n = (size_t)h * w;
    ProfScope prof(ctx, STEREO_K_POST_MAPS);
    STEREO_CUDA_CHECK(cudaMemsetAsync(st, 0, sizeof(GuideStats), ctx->stream));
    int blocks = (int)((n + 255) / 256);
    if (blocks > ctx->sm_count * 4) blocks = ctx->sm_count * 4;
    guide_hist_kernel<<<blocks, 256, 0, ctx->stream>>>(img, n, st);
    guide_lut_kernel<<<1, 256, 0, ctx->stream>>>(st, (double)n);
    guide_apply_kernel<<<blocks, 256, 0, ctx->stream>>>(img, n, st, out);
    ctx->launches += 3;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// ------------------------------------------------------------------------------------------
// P1  adcensus.cross (nyu.lua:33-34): four arms per pixel, exclusive end coordinate.
//     arms[a][y][x], a = 0 left, 1 right, 2 up, 3 down, stored as int16.
// ------------------------------------------------------------------------------------------
__global__ void cross_kernel(const float *__restrict__ img, int h, int w, int L1, float tau1,
                             short *__restrict__ arms)
{
    size_t id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t hw = (size_t)h * w;
    if (id >= 4 * hw) return;
    int arm = (int)(id / hw);
    int rem = (int)(id % hw);
    int x = rem % w, y = rem / w;
    int dx = 0, dy = 0;
    if (arm == 0) dx = -1;
    else if (arm == 1) dx = 1;
    else if (arm == 2) dy = -1;
    else dy = 1;
    float centre = img[rem];
    int xx = x + dx, yy = y + dy;
    for (;; xx += dx, yy += dy) {
        if (xx < 0 || xx >= w || yy < 0 || yy >= h) break;
        int dist = max(abs(xx - x), abs(yy - y));
        if (dist == 1) continue;
        if (fabsf(centre - img[yy * w + xx]) >= tau1) break;
        if (dist >= L1) break;
    }
    arms[id] = (short)(dy == 0 ? xx : yy);
}

int launch_cross(stereo_ctx *ctx, const float *img, int h, int w, int L1, float tau1, short *arms)
{
    size_t total = 4 * (size_t)h * w;
    ProfScope prof(ctx, STEREO_K_POST_MAPS);
    cross_kernel<<<(unsigned)((total + 255) / 256), 256, 0, ctx->stream>>>(img, h, w, L1, tau1, arms);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// ------------------------------------------------------------------------------------------
// P2  adcensus.cbca (nyu.lua:37,39) on [H][W][D].  One warp per (pixel, 32 disparities):
//     the pixel's own arms are warp-uniform, the matched pixel's arms vary per lane; the
//     loops run over the own-arm window and each lane predicates on its intersection.
//     Summation order (rows ascending, columns ascending) is that of the original.
// ------------------------------------------------------------------------------------------
__global__ __launch_bounds__(256) void cbca_kernel(const short *__restrict__ x0c, const short *__restrict__ x1c,
                                                   const float *__restrict__ vol, float *__restrict__ out,
                                                   int h, int w, int D, int direction)
{
    const int lane = threadIdx.x & 31;
    const int chunks = (D + 31) >> 5;
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    size_t hw = (size_t)h * w;
    if (warp >= hw * chunks) return;
    int chunk = (int)(warp % chunks);
    size_t pix = warp / chunks;
    int x = (int)(pix % w), y = (int)(pix / w);
    int d = chunk * 32 + lane;
    if (d >= D) return;
    size_t id = pix * D + d;
    int xm = x + d * direction;
    if (xm < 0 || xm >= w) {
        out[id] = vol[id];
        return;
    }
    const short *a0 = x0c, *a1 = x1c;
    int up0 = a0[2 * hw + pix], down0 = a0[3 * hw + pix];
    int yy_s = max(up0, (int)a1[2 * hw + (size_t)y * w + xm]);
    int yy_t = min(down0, (int)a1[3 * hw + (size_t)y * w + xm]);
    float sum = 0.f;
    int cnt = 0;
    const int shift = d * direction;
    for (int yy = yy_s + 1; yy < yy_t; ++yy) {
        size_t rp = (size_t)yy * w;
        int xx_s = max((int)a0[rp + x], (int)a1[rp + xm] - shift);
        int xx_t = min((int)a0[hw + rp + x], (int)a1[hw + rp + xm] - shift);
        const float *vrow = vol + rp * D + d;
        for (int xx = xx_s + 1; xx < xx_t; ++xx) {
            sum = __fadd_rn(sum, vrow[(size_t)xx * D]);
            ++cnt;
        }
    }
    out[id] = __fdiv_rn(sum, (float)cnt);
}

int launch_cbca(stereo_ctx *ctx, const short *x0c, const short *x1c, const float *vol, float *out, int h, int w,
                int D, int direction)
{
    size_t warps = (size_t)h * w * ((D + 31) / 32);
    size_t blocks = (warps + 7) / 8;
    ProfScope prof(ctx, STEREO_K_CBCA);
    cbca_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(x0c, x1c, vol, out, h, w, D, direction);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// Separable form of the same aggregation.  The support region of (y, x, d) is a union of row
// segments: rows yy between the vertical arm ends, and on row yy the segment fixed by the arms
// at (yy, x) and (yy, x + d*dir) - which depends on (yy, x, d) only.  So
//   hsum[yy][x][d] = sum over that segment,  hcnt = its length            (horizontal pass)
//   out[y][x][d]   = sum_yy hsum[yy][x][d] / sum_yy hcnt[yy][x][d]        (vertical pass)
// is the same sum with <= 9 + 9 instead of <= 81 loads per voxel.  Mathematically identical to
// cbca_kernel; the fp32 additions associate per row first, so results differ from the
// original's strictly sequential order in the last bits (tests bound it at 2e-6 relative).
// block = one pixel (blockIdx.x = x, blockIdx.y = y), thread = disparity: no index decoding
__global__ __launch_bounds__(512) void cbca_h_kernel(const short *__restrict__ x0c, const short *__restrict__ x1c,
                                                     const float *__restrict__ vol, float *__restrict__ hsum,
                                                     uint8_t *__restrict__ hcnt, int h, int w, int D, int direction)
{
    const int d = threadIdx.x;
    if (d >= D) return;
    const int x = blockIdx.x, y = blockIdx.y;
    const int hw = h * w;
    const int rowpix = y * w, pix = rowpix + x;
    const int id = pix * D + d;
    const int shift = d * direction;
    const int xm = x + shift;
    if (xm < 0 || xm >= w) {           // copied through by the vertical pass
        hsum[id] = vol[id];
        hcnt[id] = 0;
        return;
    }
    const int xx_s = max((int)x0c[pix], (int)x1c[rowpix + xm] - shift);
    const int xx_t = min((int)x0c[hw + pix], (int)x1c[hw + rowpix + xm] - shift);
    const float *vp = vol + (rowpix + xx_s + 1) * D + d;
    float sum = 0.f;
    for (int n = xx_t - xx_s - 1; n > 0; --n, vp += D) sum = __fadd_rn(sum, *vp);
    hsum[id] = sum;
    hcnt[id] = (uint8_t)(xx_t - xx_s - 1);
}

__global__ __launch_bounds__(512) void cbca_v_kernel(const short *__restrict__ x0c, const short *__restrict__ x1c,
                                                     const float *__restrict__ hsum, const uint8_t *__restrict__ hcnt,
                                                     float *__restrict__ out, int h, int w, int D, int direction)
{
    const int d = threadIdx.x;
    if (d >= D) return;
    const int x = blockIdx.x, y = blockIdx.y;
    const int hw = h * w;
    const int pix = y * w + x;
    const int id = pix * D + d;
    const int xm = x + d * direction;
    if (xm < 0 || xm >= w) {
        out[id] = hsum[id];
        return;
    }
    const int yy_s = max((int)x0c[2 * hw + pix], (int)x1c[2 * hw + y * w + xm]);
    const int yy_t = min((int)x0c[3 * hw + pix], (int)x1c[3 * hw + y * w + xm]);
    const int pitch = w * D;
    int off = ((yy_s + 1) * w + x) * D + d;
    float sum = 0.f;
    int cnt = 0;
    for (int n = yy_t - yy_s - 1; n > 0; --n, off += pitch) {
        sum = __fadd_rn(sum, hsum[off]);
        cnt += hcnt[off];
    }
    out[id] = __fdiv_rn(sum, (float)cnt);
}

// Register-blocked versions for arm lengths up to L1 = 5 (windows of at most 9): a thread owns
// one disparity of FOUR neighbouring pixels, loads the 12 values their windows cover once and
// forms the four sums from registers (same ascending order of additions as above).
constexpr int kCbR = 4;                  // window half width = L1 - 1
constexpr int kCbB = 4;                  // pixels per thread
constexpr int kCbN = kCbB + 2 * kCbR;    // values loaded

__global__ __launch_bounds__(512) void cbca_h4_kernel(const short *__restrict__ x0c, const short *__restrict__ x1c,
                                                      const float *__restrict__ vol, float *__restrict__ hsum,
                                                      uint8_t *__restrict__ hcnt, int h, int w, int D, int direction)
{
    const int d = threadIdx.x;
    if (d >= D) return;
    const int xb = blockIdx.x * kCbB, y = blockIdx.y;
    const int hw = h * w, rowpix = y * w;
    const int shift = d * direction;
    float v[kCbN];
#pragma unroll
    for (int j = 0; j < kCbN; ++j) {
        const int col = xb - kCbR + j;
        v[j] = (col >= 0 && col < w) ? vol[(rowpix + col) * D + d] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < kCbB; ++i) {
        const int x = xb + i;
        if (x >= w) break;
        const int pix = rowpix + x, id = pix * D + d;
        const int xm = x + shift;
        if (xm < 0 || xm >= w) {
            hsum[id] = v[kCbR + i];
            hcnt[id] = 0;
            continue;
        }
        const int xx_s = max((int)x0c[pix], (int)x1c[rowpix + xm] - shift);
        const int xx_t = min((int)x0c[hw + pix], (int)x1c[hw + rowpix + xm] - shift);
        float sum = 0.f;
#pragma unroll
        for (int j = i; j <= i + 2 * kCbR; ++j) {       // columns x-R .. x+R
            const int col = xb - kCbR + j;
            if (col > xx_s && col < xx_t) sum = __fadd_rn(sum, v[j]);
        }
        hsum[id] = sum;
        hcnt[id] = (uint8_t)(xx_t - xx_s - 1);
    }
}

__global__ __launch_bounds__(512) void cbca_v4_kernel(const short *__restrict__ x0c, const short *__restrict__ x1c,
                                                      const float *__restrict__ hsum, const uint8_t *__restrict__ hcnt,
                                                      float *__restrict__ out, int h, int w, int D, int direction)
{
    const int d = threadIdx.x;
    if (d >= D) return;
    const int x = blockIdx.x, yb = blockIdx.y * kCbB;
    const int hw = h * w;
    const int xm = x + d * direction;
    if (xm < 0 || xm >= w) {
#pragma unroll
        for (int i = 0; i < kCbB; ++i) {
            const int y = yb + i;
            if (y < h) out[(y * w + x) * D + d] = hsum[(y * w + x) * D + d];
        }
        return;
    }
    float v[kCbN];
    int c[kCbN];
#pragma unroll
    for (int j = 0; j < kCbN; ++j) {
        const int row = yb - kCbR + j;
        const bool in = row >= 0 && row < h;
        const int off = (row * w + x) * D + d;
        v[j] = in ? hsum[off] : 0.f;
        c[j] = in ? (int)hcnt[off] : 0;
    }
#pragma unroll
    for (int i = 0; i < kCbB; ++i) {
        const int y = yb + i;
        if (y >= h) break;
        const int pix = y * w + x;
        const int yy_s = max((int)x0c[2 * hw + pix], (int)x1c[2 * hw + y * w + xm]);
        const int yy_t = min((int)x0c[3 * hw + pix], (int)x1c[3 * hw + y * w + xm]);
        float sum = 0.f;
        int cnt = 0;
#pragma unroll
        for (int j = i; j <= i + 2 * kCbR; ++j) {       // rows y-R .. y+R
            const int row = yb - kCbR + j;
            if (row > yy_s && row < yy_t) {
                sum = __fadd_rn(sum, v[j]);
                cnt += c[j];
            }
        }
        out[pix * D + d] = __fdiv_rn(sum, (float)cnt);
    }
}

int launch_cbca_separable(stereo_ctx *ctx, const short *x0c, const short *x1c, const float *vol, float *hsum,
                          uint8_t *hcnt, float *out, int h, int w, int D, int direction, int L1)
{
    STEREO_REQUIRE((size_t)h * w * D < (size_t)1 << 31, "volume too large for 32-bit indexing");
    STEREO_REQUIRE(D <= 512, "separable CBCA runs one thread per disparity: D <= 512, got %d", D);
    const int nt = (D + 31) / 32 * 32;
    ProfScope prof(ctx, STEREO_K_CBCA);
    if (L1 - 1 <= kCbR) {
        cbca_h4_kernel<<<dim3((w + kCbB - 1) / kCbB, h), nt, 0, ctx->stream>>>(x0c, x1c, vol, hsum, hcnt, h, w, D, direction);
        cbca_v4_kernel<<<dim3(w, (h + kCbB - 1) / kCbB), nt, 0, ctx->stream>>>(x0c, x1c, hsum, hcnt, out, h, w, D, direction);
    } else {
        cbca_h_kernel<<<dim3(w, h), nt, 0, ctx->stream>>>(x0c, x1c, vol, hsum, hcnt, h, w, D, direction);
        cbca_v_kernel<<<dim3(w, h), nt, 0, ctx->stream>>>(x0c, x1c, hsum, hcnt, out, h, w, D, direction);
    }
    ctx->launches += 2;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// ------------------------------------------------------------------------------------------
// post.lua:146-147  vol / vol:std()  - unbiased std over the whole volume (zero band
// included).  Two-stage float64 reduction through a scratch array of per-block partials;
// the partials are combined in a fixed order, so the result is deterministic.
// result[0] = std (float), result[1] = mean (float)
// ------------------------------------------------------------------------------------------
__global__ void volume_moments_kernel(const float *__restrict__ vol, size_t n, double *__restrict__ partial)
{
    double s = 0.0, ss = 0.0;
    const size_t stride = (size_t)gridDim.x * blockDim.x, tid0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (((uintptr_t)vol & 15) == 0) {
        // 16-byte loads, two of them in flight per thread, four independent double accumulations
        const float4 *v4 = reinterpret_cast<const float4 *>(vol);
        const size_t n4 = n / 4;
        double s1 = 0.0, ss1 = 0.0, s2 = 0.0, ss2 = 0.0, s3 = 0.0, ss3 = 0.0;
        size_t i = tid0;
        for (; i + stride < n4; i += 2 * stride) {
            const float4 a = __ldg(v4 + i), b = __ldg(v4 + i + stride);
            const double a0 = a.x, a1 = a.y, a2 = a.z, a3 = a.w, b0 = b.x, b1 = b.y, b2 = b.z, b3 = b.w;
            s += a0; ss += a0 * a0; s1 += a1; ss1 += a1 * a1; s2 += a2; ss2 += a2 * a2; s3 += a3; ss3 += a3 * a3;
            s += b0; ss += b0 * b0; s1 += b1; ss1 += b1 * b1; s2 += b2; ss2 += b2 * b2; s3 += b3; ss3 += b3 * b3;
        }
        if (i < n4) {
            const float4 a = __ldg(v4 + i);
            const double a0 = a.x, a1 = a.y, a2 = a.z, a3 = a.w;
            s += a0; ss += a0 * a0; s1 += a1; ss1 += a1 * a1; s2 += a2; ss2 += a2 * a2; s3 += a3; ss3 += a3 * a3;
        }
        for (size_t k = 4 * n4 + tid0; k < n; k += stride) {          // the last n % 4 elements
            const double v = (double)vol[k];
            s += v;
            ss += v * v;
        }
        s = (s + s1) + (s2 + s3);
        ss = (ss + ss1) + (ss2 + ss3);
    } else {
        for (size_t i = tid0; i < n; i += stride) {
            double v = (double)vol[i];
            s += v;
            ss += v * v;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_down_sync(0xffffffffu, s, o);
        ss += __shfl_down_sync(0xffffffffu, ss, o);
    }
    __shared__ double sh[2][32];
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { sh[0][warp] = s; sh[1][warp] = ss; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { a += sh[0][i]; b += sh[1][i]; }
        partial[2 * blockIdx.x] = a;
        partial[2 * blockIdx.x + 1] = b;
    }
}

__global__ void volume_std_final_kernel(const double *__restrict__ partial, int nblocks, double n,
                                        float *__restrict__ result)
{
    // one warp: lane l adds partials l, l + 32, ... in order, then a fixed shuffle tree - deterministic, and
    // ~40 dependent loads instead of ~2 400
    if (blockIdx.x != 0 || threadIdx.x >= 32) return;
    double s = 0.0, ss = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 32) { s += partial[2 * i]; ss += partial[2 * i + 1]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_down_sync(0xffffffffu, s, o);
        ss += __shfl_down_sync(0xffffffffu, ss, o);
    }
    if (threadIdx.x != 0) return;
    double mean = s / n;
    double var = (ss - s * mean) / (n - 1.0);
    result[0] = (float)sqrt(var > 0.0 ? var : 0.0);
    result[1] = (float)mean;
}

constexpr int kStdBlocks = 1184;   // 148 SMs x 8

int launch_volume_std_with(stereo_ctx *ctx, const float *vol, size_t n, double *partial, float *result)
{
    ProfScope prof(ctx, STEREO_K_VOLSTD);
    volume_moments_kernel<<<kStdBlocks, 256, 0, ctx->stream>>>(vol, n, partial);
    volume_std_final_kernel<<<1, 32, 0, ctx->stream>>>(partial, kStdBlocks, (double)n, result);
    ctx->launches += 2;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// post.lua:125 + :146 from one read of the volume: the arg max map of every pixel and the unbiased std (aligned
// volumes with D % 4 == 0, D <= 256; the caller falls back to the two stand-alone kernels otherwise)
int launch_volume_std_wta(stereo_ctx *ctx, const float *vol, size_t n_pixels, int D, float *result, uint8_t *wta, int slot)
{
    const size_t area = 2 * (size_t)std::max(kStdBlocks, wta_moments_max_blocks(ctx));
    STEREO_TRY(ctx->scratch.reserve(sizeof(double) * 2 * area));
    double *partial = ctx->scratch.as<double>() + (slot ? area : 0);
    int nblocks = 0;
    {
        ProfScope prof(ctx, STEREO_K_WTA);
        STEREO_TRY(launch_wta_moments(ctx, vol, n_pixels, D, 0, wta, partial, &nblocks));
    }
    ProfScope prof(ctx, STEREO_K_VOLSTD);
    volume_std_final_kernel<<<1, 32, 0, ctx->stream>>>(partial, nblocks, (double)n_pixels * D, result);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// `slot` (0 or 1) selects one of two partial-sum areas so that two streams can reduce at once
int launch_volume_std(stereo_ctx *ctx, const float *vol, size_t n, float *result, int slot)
{
    const size_t area = 2 * (size_t)std::max(kStdBlocks, wta_moments_max_blocks(ctx));
    STEREO_TRY(ctx->scratch.reserve(sizeof(double) * 2 * area));
    return launch_volume_std_with(ctx, vol, n, ctx->scratch.as<double>() + (slot ? area : 0), result);
}

// ------------------------------------------------------------------------------------------
// stereo_volume_std_static: the same reduction for callers that must not allocate.
//
// Who that is: a perception loop that carves up device memory once at start-up and forbids
// any later cudaMalloc (it synchronises the whole device and may fail once memory is
// fragmented), or a host process whose own pool already owns all of HBM, so that an
// allocation made behind its back by this library would simply fail.  For them the per-block
// partial sums live in a fixed array that is part of the library image; every GPU that loads
// the library carries these 64 KB until the process exits.
// ------------------------------------------------------------------------------------------
constexpr int kStaticStdMaxBlocks = 4096;
constexpr size_t kStaticStdElemsPerBlock = 65536;     // so at most 268,435,456 elements (1 GiB of fp32)
static __device__ __align__(256) double g_static_std_scratch[2 * kStaticStdMaxBlocks];

// In host code `g_static_std_scratch` is only a handle for the symbol, not a device address;
// the address is looked up once per device and kept here.
constexpr int kMaxDevices = 64;
static double *g_static_std_addr[kMaxDevices] = {};

}  // namespace stereo

extern "C" int stereo_volume_std_static(const float *vol, size_t n, float *result, void *cuda_stream)
{
    using namespace stereo;
    STEREO_REQUIRE(vol != nullptr && result != nullptr, "null pointer");
    STEREO_REQUIRE(n >= 2, "the unbiased std needs at least 2 elements, got %zu", n);
    const size_t max_n = (size_t)kStaticStdMaxBlocks * kStaticStdElemsPerBlock;
    if (n > max_n) {
        set_error("stereo_volume_std_static handles at most %zu elements (its scratch array is fixed at %d partial "
                  "sums); got %zu - use stereo_volume_std",
                  max_n, kStaticStdMaxBlocks, n);
        return STEREO_E_LIMIT;
    }
    int dev = -1;
    STEREO_CUDA_CHECK(cudaGetDevice(&dev));
    STEREO_REQUIRE(dev >= 0 && dev < kMaxDevices, "device index %d is beyond the %d this entry point tracks", dev,
                   kMaxDevices);
    if (g_static_std_addr[dev] == nullptr) {
        // first call on this device (with it current): resolve the symbol to a device pointer
        double *addr = nullptr;
        STEREO_CUDA_CHECK(cudaGetSymbolAddress((void **)&addr, g_static_std_scratch));
        STEREO_REQUIRE(addr != nullptr, "cudaGetSymbolAddress returned a null address");
        g_static_std_addr[dev] = addr;
    }
    double *scratch = g_static_std_addr[dev];
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    const int nblocks = (int)((n + kStaticStdElemsPerBlock - 1) / kStaticStdElemsPerBlock);
    // Nothing has to be cleared: block b writes slots 2b and 2b+1 before the final kernel reads
    // exactly the first 2*nblocks slots, so whatever the previous call left behind is never read.
    volume_moments_kernel<<<nblocks, 256, 0, stream>>>(vol, n, scratch);
    volume_std_final_kernel<<<1, 32, 0, stream>>>(scratch, nblocks, (double)n, result);
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

namespace stereo {

// elementwise helpers of post.lua:146-153 and nyu.lua:61,64
__global__ void scale_kernel(const float *__restrict__ in, size_t n, const float *__restrict__ stdp, int mode,
                             float *__restrict__ out)
{
    // mode 0: x / std   mode 1: -(x / std)   mode 2: x / 4
    // One divisor for the whole volume: div_exact (exact_div.cuh) gives the bits of IEEE division at five FMA-class
    // instructions per voxel; with __fdiv_rn (reciprocal approximation, range check, slow-path call per element) this
    // pass ran at 3.8 TB/s instead of at the speed of a copy.
    const ConstDivisor sd = make_divisor(mode == 2 ? 4.f : stdp[0]);
    const size_t stride = (size_t)gridDim.x * blockDim.x, tid0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const float sign = mode == 1 ? -1.f : 1.f;
    if ((((uintptr_t)in | (uintptr_t)out) & 15) == 0) {
        const float4 *in4 = reinterpret_cast<const float4 *>(in);
        float4 *out4 = reinterpret_cast<float4 *>(out);
        const size_t n4 = n / 4;
        for (size_t i = tid0; i < n4; i += stride) {
            float4 v = __ldg(in4 + i);
            v.x = __fmul_rn(div_exact(v.x, sd), sign);
            v.y = __fmul_rn(div_exact(v.y, sd), sign);
            v.z = __fmul_rn(div_exact(v.z, sd), sign);
            v.w = __fmul_rn(div_exact(v.w, sd), sign);
            out4[i] = v;
        }
        for (size_t i = 4 * n4 + tid0; i < n; i += stride) out[i] = __fmul_rn(div_exact(in[i], sd), sign);
    } else {
        for (size_t i = tid0; i < n; i += stride) out[i] = __fmul_rn(div_exact(in[i], sd), sign);
    }
}

// The same pass for a volume whose [x][d] rows are not 16-byte aligned (D % 4 != 0, the reference's default 201):
// the output gets rows of Dp = D rounded up to 4 floats, the Dp - D slots behind every row filled with kSgmBig
// ("no such disparity": never the minimum of a scanline step, and it stays that way through every addition), so
// that the aligned SGM kernels can run on it as a volume of Dp disparities.
__global__ void scale_pad_kernel(const float *__restrict__ in, size_t n_pixels, int D, int Dp,
                                 const float *__restrict__ stdp, int mode, float *__restrict__ out)
{
    // a warp per pixel: lane q writes float4 q, q + 32, ... of the padded row (16-byte stores), reading four
    // consecutive 4-byte words of the unpadded one
    const ConstDivisor sd = make_divisor(stdp[0]);
    const float sign = mode == 1 ? -1.f : 1.f;
    const int lane = threadIdx.x & 31, q4 = Dp >> 2;
    const size_t warps = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t px = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; px < n_pixels; px += warps) {
        const float *src = in + px * D;
        float4 *dst = reinterpret_cast<float4 *>(out + px * Dp);
        for (int q = lane; q < q4; q += 32) {
            float v[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int d = 4 * q + e;
                v[e] = d < D ? __fmul_rn(div_exact(__ldg(src + d), sd), sign) : kSgmBig;
            }
            dst[q] = make_float4(v[0], v[1], v[2], v[3]);
        }
    }
}

int launch_scale_pad(stereo_ctx *ctx, const float *in, size_t n_pixels, int D, int Dp, const float *stdp, int mode,
                     float *out)
{
    ProfScope prof(ctx, STEREO_K_VOLSTD);
    scale_pad_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(in, n_pixels, D, Dp, stdp, mode, out);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_scale(stereo_ctx *ctx, const float *in, size_t n, const float *stdp, int mode, float *out)
{
    ProfScope prof(ctx, STEREO_K_VOLSTD);
    scale_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(in, n, stdp, mode, out);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// ------------------------------------------------------------------------------------------
// P3  adcensus.sgm2 (nyu.lua:59-60): semi-global matching with intensity-adaptive penalties.
//     The four paths of the original (-> <- down up) run in sgm_ring.cu.  This file holds the
//     DIAGONAL paths of the 8-path variant, which has no counterpart in the reference (its chain
//     is 4-path, nyu.lua:59-64; BASELINE.json configs[3] names an 8-path chain): one WARP owns a
//     diagonal line and walks it, each lane keeps KPL consecutive disparities of the previous
//     step's path costs in registers, d-1 / d+1 across lanes are two shuffles, the minimum over D
//     a shuffle reduction, and the costs and running sums of the next PF steps are in flight
//     while a step is computed.  Every pass adds its path costs to the running sum in place
//     (same thread reads and writes a voxel); the last one scales by 1/8 and can write the
//     arg min map.  Step arithmetic: sgm_step() of sgm_step.cuh, the original's, with no alpha1
//     scaling (the original applies it on the two vertical paths only).
//     Path order of the sum: -> <- down up, then 4: down-right, 5: down-left, 6: up-right, 7: up-left.
// ------------------------------------------------------------------------------------------
template <int SD, int KPL, int PF, bool VEC>
__global__ __launch_bounds__(32) void sgm_diag_kernel(const float *__restrict__ x0, const float *__restrict__ x1,
                                                      const float *__restrict__ cvol, float *acc, int h, int w, int D,
                                                      SgmParams p, int direction, float final_scale, int last,
                                                      uint8_t *wta)
{
    static_assert(SD >= 4 && SD < 8, "diagonal paths only");
    constexpr int dx = (SD == 4 || SD == 6) ? 1 : -1;
    constexpr int dy = SD < 6 ? 1 : -1;
    const int lane = threadIdx.x & 31;
    // line -> first pixel: the first row in path order, then the rest of the first column
    const int line = blockIdx.x;
    const int y_first = dy > 0 ? 0 : h - 1, x_first = dx > 0 ? 0 : w - 1;
    int x = line < w ? line : x_first;
    int y = line < w ? y_first : y_first + dy * (line - w + 1);
    const int steps = min(dx > 0 ? w - x : x + 1, dy > 0 ? h - y : y + 1);
    const int d0 = lane * KPL;
    const SgmPenalties<SD> pen(p);

    const ptrdiff_t vstride = ((ptrdiff_t)dy * w + dx) * D;     // floats between consecutive steps
    const size_t base = ((size_t)y * w + x) * D + d0;
    const float *cptr = cvol + base;
    float *aptr = acc + base;

    bool valid[KPL];
#pragma unroll
    for (int k = 0; k < KPL; ++k) valid[k] = d0 + k < D;

    auto load_row = [&](const float *src, float (&dst)[KPL]) {
        if (VEC) {
#pragma unroll
            for (int j = 0; j < KPL / 4; ++j) {
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (valid[4 * j]) v = *(reinterpret_cast<const float4 *>(src) + j);
                dst[4 * j] = v.x; dst[4 * j + 1] = v.y; dst[4 * j + 2] = v.z; dst[4 * j + 3] = v.w;
            }
        } else {
#pragma unroll
            for (int k = 0; k < KPL; ++k) dst[k] = valid[k] ? src[k] : 0.f;
        }
    };

    float cbuf[PF][KPL], abuf[PF][KPL];
#pragma unroll
    for (int i = 0; i < PF; ++i) {
        if (i < steps) {
            load_row(cptr + i * vstride, cbuf[i]);
            load_row(aptr + i * vstride, abuf[i]);
        }
    }

    float prev[KPL], i1p[KPL];
#pragma unroll
    for (int k = 0; k < KPL; ++k) { prev[k] = kSgmBig; i1p[k] = 0.f; }
    const bool has_below = lane > 0, has_above = d0 + KPL < D;

    for (int g = 0; g < steps; g += PF) {
#pragma unroll
        for (int i = 0; i < PF; ++i) {
            const int step = g + i;
            if (step >= steps) break;
            // guide values at the matched pixels of this step
            float i1c[KPL];
            const float *g1 = x1 + (size_t)y * w;
            const int xx0 = x + d0 * direction;
#pragma unroll
            for (int k = 0; k < KPL; ++k) {
                const int xx = xx0 + k * direction;
                i1c[k] = (unsigned)xx < (unsigned)w ? __ldg(g1 + xx) : 0.f;
            }
            float c[KPL];
#pragma unroll
            for (int k = 0; k < KPL; ++k) c[k] = cbuf[i][k];
            float val[KPL];
            if (step == 0) {
#pragma unroll
                for (int k = 0; k < KPL; ++k) val[k] = valid[k] ? c[k] : kSgmBig;
            } else {
                const float D1 = fabsf(__ldg(x0 + (size_t)y * w + x) - __ldg(x0 + (size_t)(y - dy) * w + (x - dx)));
                sgm_step<SD, KPL>(pen, D1, c, prev, i1c, i1p, valid, xx0, direction, w, has_below, has_above, val);
            }
            // accumulate into the running sum
            float res[KPL];
            float best = kSgmBig;
            int best_k = 0;
#pragma unroll
            for (int k = 0; k < KPL; ++k) {
                const float sum = __fadd_rn(abuf[i][k], val[k]);
                res[k] = last ? __fmul_rn(sum, final_scale) : sum;
                if (valid[k] && res[k] < best) { best = res[k]; best_k = k; }
            }
            if (VEC) {
#pragma unroll
                for (int j = 0; j < KPL / 4; ++j)
                    if (valid[4 * j])
                        reinterpret_cast<float4 *>(aptr)[j] = make_float4(res[4 * j], res[4 * j + 1], res[4 * j + 2], res[4 * j + 3]);
            } else {
#pragma unroll
                for (int k = 0; k < KPL; ++k)
                    if (valid[k]) aptr[k] = res[k];
            }
            if (last && wta != nullptr) {
                int key = __float_as_int(__fadd_rn(best, 0.f));
                key ^= (key >> 31) & 0x7fffffff;
                const int mn = __reduce_min_sync(0xffffffffu, key);
                const unsigned winners = __ballot_sync(0xffffffffu, key == mn);
                if (lane == __ffs(winners) - 1) wta[(size_t)y * w + x] = (uint8_t)(d0 + best_k);
            }
#pragma unroll
            for (int k = 0; k < KPL; ++k) { prev[k] = val[k]; i1p[k] = i1c[k]; }
            // refill this ring slot with the step PF ahead
            if (step + PF < steps) {
                load_row(cptr + (ptrdiff_t)PF * vstride, cbuf[i]);
                load_row(aptr + (ptrdiff_t)PF * vstride, abuf[i]);
            }
            cptr += vstride;
            aptr += vstride;
            x += dx;
            y += dy;
        }
    }
}

template <int KPL, int PF, bool VEC>
static int launch_sgm_diagonals(stereo_ctx *ctx, const float *x0, const float *x1, const float *cvol, float *acc, int h,
                                int w, int D, const SgmParams &p, int direction, const SgmExtras &ex)
{
    const int lines = w + h - 1;
    sgm_diag_kernel<4, KPL, PF, VEC><<<lines, 32, 0, ctx->stream>>>(x0, x1, cvol, acc, h, w, D, p, direction, 1.f, 0, nullptr);
    sgm_diag_kernel<5, KPL, PF, VEC><<<lines, 32, 0, ctx->stream>>>(x0, x1, cvol, acc, h, w, D, p, direction, 1.f, 0, nullptr);
    sgm_diag_kernel<6, KPL, PF, VEC><<<lines, 32, 0, ctx->stream>>>(x0, x1, cvol, acc, h, w, D, p, direction, 1.f, 0, nullptr);
    sgm_diag_kernel<7, KPL, PF, VEC><<<lines, 32, 0, ctx->stream>>>(x0, x1, cvol, acc, h, w, D, p, direction,
                                                                    ex.final_scale, 1, ex.wta);
    ctx->launches += 4;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// cvol: pre-scaled costs; out: result; scratch: one more volume (the right->left path costs)
int launch_sgm(stereo_ctx *ctx, const float *x0, const float *x1, const float *cvol, float *out, float *scratch,
               int h, int w, int D, const SgmParams &p, int direction, const SgmExtras &ex)
{
    if (D > 512) {
        set_error("SGM supports disp_range <= 512 (16 disparities per lane), got %d", D);
        return STEREO_E_LIMIT;
    }
    STEREO_REQUIRE(ex.paths == 4 || ex.paths == 8, "sgm_paths must be 4 (the reference) or 8, got %d", ex.paths);
    STEREO_REQUIRE(ex.wta == nullptr || D <= 256, "the arg min map is bytes: D=%d > 256", D);
    ProfScope prof(ctx, STEREO_K_SGM);
    if (ex.paths == 4) return launch_sgm_ring(ctx, x0, x1, cvol, out, scratch, h, w, D, p, direction, ex);
    // eight paths: the four of the original leave their unscaled sum in `out`, the diagonals add to it in place
    SgmExtras four = ex;
    four.final_scale = 1.f;
    four.wta = nullptr;
    STEREO_TRY(launch_sgm_ring(ctx, x0, x1, cvol, out, scratch, h, w, D, p, direction, four));
    SgmExtras diag = ex;
    diag.final_scale = 0.125f;                       // x / 8 == x * 0.125 exactly
    const bool vec = D % 4 == 0 && ((uintptr_t)cvol % 16 == 0) && ((uintptr_t)out % 16 == 0);
    if (D <= 128) return vec ? launch_sgm_diagonals<4, 4, true>(ctx, x0, x1, cvol, out, h, w, D, p, direction, diag)
                             : launch_sgm_diagonals<4, 4, false>(ctx, x0, x1, cvol, out, h, w, D, p, direction, diag);
    if (D <= 256) return vec ? launch_sgm_diagonals<8, 4, true>(ctx, x0, x1, cvol, out, h, w, D, p, direction, diag)
                             : launch_sgm_diagonals<8, 4, false>(ctx, x0, x1, cvol, out, h, w, D, p, direction, diag);
    return vec ? launch_sgm_diagonals<16, 2, true>(ctx, x0, x1, cvol, out, h, w, D, p, direction, diag)
               : launch_sgm_diagonals<16, 2, false>(ctx, x0, x1, cvol, out, h, w, D, p, direction, diag);
}

// ------------------------------------------------------------------------------------------
// post.lua:98-115,188-204  optional box aggregation: nn.SpatialAveragePooling(cost_w, cost_h, 1, 1, pad)
// with setCountExcludePad() on every disparity plane.  One thread per voxel (d fastest, so the <= kw*kh loads
// of a warp are contiguous runs), same summation order as the original (row by row), sum / in-image taps.
// ------------------------------------------------------------------------------------------
__global__ void box_agg_kernel(const float *__restrict__ in, int h, int w, int D, int kw, int kh,
                               float *__restrict__ out)
{
    const size_t total = (size_t)h * w * D;
    const int pw = (kw - 1) / 2, ph = (kh - 1) / 2;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int d = (int)(i % D);
        const size_t p = i / D;
        const int x = (int)(p % w), y = (int)(p / w);
        const int y0 = max(y - ph, 0), x0 = max(x - pw, 0);
        const int y1 = min(min(y - ph + kh, h + ph), h), x1 = min(min(x - pw + kw, w + pw), w);
        float s = 0.f;
        for (int yy = y0; yy < y1; ++yy)
            for (int xx = x0; xx < x1; ++xx) s = __fadd_rn(s, in[((size_t)yy * w + xx) * D + d]);
        out[i] = __fdiv_rn(s, (float)((y1 - y0) * (x1 - x0)));
    }
}

int launch_box_agg(stereo_ctx *ctx, const float *in, int h, int w, int D, int kw, int kh, float *out)
{
    ProfScope prof(ctx, STEREO_K_CBCA);
    box_agg_kernel<<<ctx->sm_count * 16, 256, 0, ctx->stream>>>(in, h, w, D, kw, kh, out);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// ------------------------------------------------------------------------------------------
// P5-P9  per-pixel map kernels (nyu.lua:91-115)
// ------------------------------------------------------------------------------------------
__global__ void outlier_kernel(const float *__restrict__ d0, const float *__restrict__ d1, int total, int w,
                               int disp_max, float *__restrict__ outlier)
{
    int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= total) return;
    int x = id % w;
    int d0i = (int)d0[id];
    float res;
    if (x - d0i < 0) {
        res = 1.f;
    } else if (fabsf(d0[id] - d1[id - d0i]) < 1.1f) {
        res = 0.f;
    } else {
        res = 1.f;
        for (int d = 0; d < disp_max; ++d)
            if (x - d >= 0 && fabsf((float)d - d1[id - d]) < 1.1f) { res = 2.f; break; }
    }
    outlier[id] = res;
}

__global__ void interpolate_occlusion_kernel(const float *__restrict__ d0, const float *__restrict__ outlier,
                                             int total, int w, float *__restrict__ out)
{
    int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= total) return;
    if (outlier[id] != 1.f) { out[id] = d0[id]; return; }
    int x = id % w;
    int dx = 0;
    while (x + dx >= 0 && outlier[id + dx] != 0.f) --dx;
    if (x + dx < 0) {
        dx = 0;
        while (x + dx < w && outlier[id + dx] != 0.f) ++dx;
    }
    out[id] = x + dx < w ? d0[id + dx] : d0[id];
}

__device__ __forceinline__ void insertion_sort(float *v, int n)
{
    for (int i = 1; i < n; ++i) {
        float t = v[i];
        int j = i - 1;
        while (j >= 0 && v[j] > t) { v[j + 1] = v[j]; --j; }
        v[j + 1] = t;
    }
}

__global__ void interpolate_mismatch_kernel(const float *__restrict__ d0, const float *__restrict__ outlier,
                                            int h, int w, float *__restrict__ out)
{
    const float dirs[32] = {0.f, 1.f, -0.5f, 1.f, -1.f, 1.f, -1.f, 0.5f, -1.f, 0.f, -1.f, -0.5f, -1.f, -1.f, -0.5f, -1.f,
                            0.f, -1.f, 0.5f, -1.f, 1.f, -1.f, 1.f, -0.5f, 1.f, 0.f, 1.f, 0.5f, 1.f, 1.f, 0.5f, 1.f};
    int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= h * w) return;
    if (outlier[id] != 2.f) { out[id] = d0[id]; return; }
    float vals[16];
    int n = 0;
    int x = id % w, y = id / w;
    for (int k = 0; k < 16; ++k) {
        float ddx = dirs[2 * k], ddy = dirs[2 * k + 1];
        float xx = (float)x, yy = (float)y;
        int xi = (int)roundf(xx), yi = (int)roundf(yy);
        while (0 <= yi && yi < h && 0 <= xi && xi < w && outlier[yi * w + xi] == 2.f) {
            xx = __fadd_rn(xx, ddx);
            yy = __fadd_rn(yy, ddy);
            xi = (int)roundf(xx);
            yi = (int)roundf(yy);
        }
        if (0 <= yi && yi < h && 0 <= xi && xi < w) vals[n++] = d0[yi * w + xi];
    }
    if (n == 0) { out[id] = d0[id]; return; }
    insertion_sort(vals, n);
    out[id] = vals[n / 2];
}

__global__ void subpixel_kernel(const float *__restrict__ d0, const float *__restrict__ vol, int total,
                                int disp_max, float *__restrict__ out)
{
    int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= total) return;
    int d = (int)d0[id];
    float res = (float)d;
    if (1 <= d && d < disp_max - 1) {
        const float *c = vol + (size_t)id * disp_max + d;      // [H][W][D] layout
        float cn = c[-1], cz = c[0], cp = c[1];
        float denom = __fmul_rn(2.f, __fadd_rn(__fadd_rn(cp, cn), -__fmul_rn(2.f, cz)));
        if (denom > 1e-5f) {
            float t = __fdiv_rn(__fadd_rn(cp, -cn), denom);
            t = t < -1.f ? -1.f : (t > 1.f ? 1.f : t);
            res = __fadd_rn((float)d, -t);
        }
    }
    out[id] = res;
}

__global__ void median2d_kernel(const float *__restrict__ img, int h, int w, int radius, float *__restrict__ out)
{
    int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= h * w) return;
    int x = id % w, y = id / w;
    if (radius == 2 && x >= 2 && x < w - 2 && y >= 2 && y < h - 2) {
        // the 5 x 5 window of nyu.lua:107 inside the image: median of 25 by a compare-exchange network in registers
        // (median25.cuh) instead of an insertion sort in local memory - the same value, whatever the order
        float v[25];
#pragma unroll
        for (int dx = 0; dx < 5; ++dx)
#pragma unroll
            for (int dy = 0; dy < 5; ++dy) v[dx * 5 + dy] = __ldg(img + (y - 2 + dy) * w + (x - 2 + dx));
        out[id] = median25(v);
        return;
    }
    float xs[11 * 11];
    int n = 0;
    for (int xx = x - radius; xx <= x + radius; ++xx)
        for (int yy = y - radius; yy <= y + radius; ++yy)
            if (0 <= xx && xx < w && 0 <= yy && yy < h) xs[n++] = img[yy * w + xx];
    insertion_sort(xs, n);
    out[id] = xs[n / 2];
}

// mean2d: a 16 x 32 pixel tile + halo staged in shared memory; taps walk columns then rows like the original so
// that the fp32 sums associate identically.  A thread owns kBlurPix vertically adjacent pixels: the value one row
// further down serves tap oy of one pixel and tap oy + 1 of the pixel above it, and the kernel weight of a tap is
// the same for all of them, so a tap costs 2 / kBlurPix shared-memory loads instead of 2 (the round-2 ncu table
// had this kernel at 91 % of the L1 / shared-memory throughput: 0.24 ms for 1 369 taps per pixel).
constexpr int kBlurTileW = 16, kBlurTileH = 32, kBlurPix = 4;

__global__ __launch_bounds__(kBlurTileW * kBlurTileH / kBlurPix) void mean2d_kernel(const float *__restrict__ img,
                                                                                     const float *__restrict__ kernel, int h,
                                                                                     int w, int radius, float alpha2,
                                                                                     float *__restrict__ out)
{
    extern __shared__ float sm[];
    const int side_w = kBlurTileW + 2 * radius, side_h = kBlurTileH + 2 * radius, pitch = side_w + 1;
    const int ksz = 2 * radius + 1;
    float *tile = sm;                  // [side_h][pitch], NaN outside the image
    float *ker = sm + side_h * pitch;
    const int x0 = blockIdx.x * kBlurTileW - radius, y0 = blockIdx.y * kBlurTileH - radius;
    const int nthr = blockDim.x * blockDim.y, tid = threadIdx.y * blockDim.x + threadIdx.x;
    for (int i = tid; i < side_w * side_h; i += nthr) {
        int ty = i / side_w, tx = i % side_w;
        int gx = x0 + tx, gy = y0 + ty;
        tile[ty * pitch + tx] = (gx >= 0 && gx < w && gy >= 0 && gy < h) ? img[gy * w + gx] : __int_as_float(0x7fc00000);
    }
    for (int i = tid; i < ksz * ksz; i += nthr) ker[i] = kernel[i];
    __syncthreads();
    const int x = blockIdx.x * kBlurTileW + threadIdx.x, yb = threadIdx.y * kBlurPix;      // first pixel row inside the tile
    if (x >= w) return;
    float centre[kBlurPix], sum[kBlurPix], cnt[kBlurPix];
#pragma unroll
    for (int q = 0; q < kBlurPix; ++q) {
        centre[q] = tile[(yb + q + radius) * pitch + threadIdx.x + radius];
        sum[q] = 0.f;
        cnt[q] = 0.f;
    }
    int i = 0;
    for (int ox = 0; ox < ksz; ++ox) {
        const float *col = tile + yb * pitch + threadIdx.x + ox;
        float win[kBlurPix];                       // win[q] = tap (ox, oy) of pixel q = tile row yb + q + oy
#pragma unroll
        for (int q = 0; q < kBlurPix - 1; ++q) win[q + 1] = col[q * pitch];
        for (int oy = 0; oy < ksz; ++oy, ++i) {
#pragma unroll
            for (int q = 0; q < kBlurPix - 1; ++q) win[q] = win[q + 1];
            win[kBlurPix - 1] = col[(oy + kBlurPix - 1) * pitch];
            const float k = ker[i];
#pragma unroll
            for (int q = 0; q < kBlurPix; ++q) {
                if (fabsf(win[q] - centre[q]) < alpha2) {        // false for the NaN padding
                    sum[q] = fmaf(win[q], k, sum[q]);            // the original is built with -fmad=true
                    cnt[q] = __fadd_rn(cnt[q], k);
                }
            }
        }
    }
#pragma unroll
    for (int q = 0; q < kBlurPix; ++q) {
        const int y = blockIdx.y * kBlurTileH + yb + q;
        if (y < h) out[y * w + x] = __fdiv_rn(sum[q], cnt[q]);
    }
}

__global__ void u8_to_float_kernel(const uint8_t *__restrict__ in, int n, float *__restrict__ out)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (float)in[i];
}

__global__ void float_to_u8_kernel(const float *__restrict__ in, int n, float scale, uint8_t *__restrict__ out)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (uint8_t)(int)__fmul_rn(in[i], scale);     // :byte() truncation
}

static inline unsigned blocks_for(int n) { return (unsigned)((n + 255) / 256); }

int launch_outlier(stereo_ctx *ctx, const float *d0, const float *d1, int h, int w, int disp_max, float *outlier)
{
    ProfScope prof(ctx, STEREO_K_POST_MAPS);
    outlier_kernel<<<blocks_for(h * w), 256, 0, ctx->stream>>>(d0, d1, h * w, w, disp_max, outlier);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_interpolate(stereo_ctx *ctx, const float *d0, const float *outlier, int h, int w, float *tmp, float *out)
{
    ProfScope prof(ctx, STEREO_K_POST_MAPS);
    interpolate_occlusion_kernel<<<blocks_for(h * w), 256, 0, ctx->stream>>>(d0, outlier, h * w, w, tmp);
    interpolate_mismatch_kernel<<<blocks_for(h * w), 256, 0, ctx->stream>>>(tmp, outlier, h, w, out);
    ctx->launches += 2;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_subpixel(stereo_ctx *ctx, const float *d0, const float *vol, int h, int w, int disp_max, float *out)
{
    ProfScope prof(ctx, STEREO_K_POST_MAPS);
    subpixel_kernel<<<blocks_for(h * w), 256, 0, ctx->stream>>>(d0, vol, h * w, disp_max, out);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_median(stereo_ctx *ctx, const float *img, int h, int w, int kernel_size, float *out)
{
    STEREO_REQUIRE(kernel_size >= 1 && kernel_size <= 11, "median kernel size must be in [1, 11], got %d", kernel_size);
    ProfScope prof(ctx, STEREO_K_POST_MAPS);
    median2d_kernel<<<blocks_for(h * w), 256, 0, ctx->stream>>>(img, h, w, kernel_size / 2, out);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_mean2d(stereo_ctx *ctx, const float *img, const float *kernel, int h, int w, int radius, float alpha2,
                  float *out)
{
    const int side_w = kBlurTileW + 2 * radius, side_h = kBlurTileH + 2 * radius, ksz = 2 * radius + 1;
    size_t smem = sizeof(float) * ((size_t)side_h * (side_w + 1) + (size_t)ksz * ksz);
    if (smem > 200 * 1024) {
        set_error("bilateral radius %d needs %zu B of shared memory", radius, smem);
        return STEREO_E_LIMIT;
    }
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(mean2d_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((w + kBlurTileW - 1) / kBlurTileW, (h + kBlurTileH - 1) / kBlurTileH);
    ProfScope prof(ctx, STEREO_K_POST_MAPS);
    mean2d_kernel<<<grid, dim3(kBlurTileW, kBlurTileH / kBlurPix), smem, ctx->stream>>>(img, kernel, h, w, radius, alpha2, out);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_u8_to_float(stereo_ctx *ctx, const uint8_t *in, int n, float *out)
{
    u8_to_float_kernel<<<blocks_for(n), 256, 0, ctx->stream>>>(in, n, out);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_float_to_u8(stereo_ctx *ctx, const float *in, int n, float scale, uint8_t *out)
{
    float_to_u8_kernel<<<blocks_for(n), 256, 0, ctx->stream>>>(in, n, scale, out);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
