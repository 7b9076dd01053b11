// Left/right correlation cost volumes and winner-take-all
// (/root/reference/inference.py:139-183), fp32 on CUDA cores.
//
//   left [y][x][d] = <L[y][x], R[y][x-d]>   (x >= d,    else exactly 0)
//   right[y][x][d] = <L[y][x+d], R[y][x]>   (x+d < w,   else exactly 0)
//
// Both volumes are the same banded Gram matrix of one image row, indexed two
// ways.  One CTA handles 64 consecutive x of one row of one side: it stages the
// 64 "own" feature vectors and the 64+D-1 candidate vectors of the other view
// transposed in shared memory, forms the 64 x (64+D-1) product block with 8x8
// register tiles (one 32-column slab per warp), parks it in shared memory and
// then either streams the band out as [x][d] rows (coalesced, D fastest - the
// .bin layout of inference.py:210-231) or reduces it to the arg max per pixel
// without ever touching HBM (fused correlation + WTA).
#include "common.cuh"

namespace stereo {

constexpr int kCorrTX = 64;   // pixels per CTA
constexpr int kPA = 68;       // pitch of the transposed own-feature tile

__device__ __forceinline__ void stage_transposed(const float *__restrict__ row, int w, int c, int xstart,
                                                 int ncols, float *dst, int pitch, int tid, int nthreads)
{
    // item -> (x_lo = it % 8, quad = (it / 8) % (c/4), x_hi): each pixel's 4 quads are
    // contiguous in global memory, the 8 x of a lane group hit distinct banks.
    int q4 = c >> 2;
    int items = ((ncols + 7) >> 3) * 8 * q4;
    for (int it = tid; it < items; it += nthreads) {
        int xl = it & 7;
        int q = (it >> 3) % q4;
        int xh = (it >> 3) / q4;
        int u = xh * 8 + xl;
        if (u >= ncols) continue;
        int gx = xstart + u;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (gx >= 0 && gx < w) v = __ldg(reinterpret_cast<const float4 *>(row + (size_t)gx * c) + q);
        float *p = dst + (q * 4) * pitch + u;
        p[0] = v.x;
        p[pitch] = v.y;
        p[2 * pitch] = v.z;
        p[3 * pitch] = v.w;
    }
}

template <bool WRITE_VOL, bool WRITE_WTA>
__global__ __launch_bounds__(512, 1) void corr_ffma_kernel(const float *__restrict__ lmap,
                                                            const float *__restrict__ rmap, int h, int w,
                                                            int c, int D, int pb, float *__restrict__ left_vol,
                                                            float *__restrict__ right_vol,
                                                            uint8_t *__restrict__ left_disp,
                                                            uint8_t *__restrict__ right_disp)
{
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, nthreads = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, nwarps = nthreads >> 5;
    const int side = blockIdx.z & 1, img = blockIdx.z >> 1;
    const int y = blockIdx.y, x0 = blockIdx.x * kCorrTX;
    const int nb = kCorrTX + D - 1;
    const size_t row_off = ((size_t)img * h + y) * w * c;
    const float *own = (side == 0 ? lmap : rmap) + row_off;
    const float *oth = (side == 0 ? rmap : lmap) + row_off;
    const int wstart = side == 0 ? x0 - (D - 1) : x0;

    float *at = smem;                // [c][kPA]
    float *bt = smem + c * kPA;      // [c][pb]
    stage_transposed(own, w, c, x0, kCorrTX, at, kPA, tid, nthreads);
    stage_transposed(oth, w, c, wstart, nb, bt, pb, tid, nthreads);
    __syncthreads();

    const int a = lane >> 2, b = lane & 3;
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    {
        const float *ap = at + 4 * a;
        const float *bp = bt + 32 * warp + 4 * b;
#pragma unroll 4
        for (int k = 0; k < c; ++k) {
            float4 a0 = *reinterpret_cast<const float4 *>(ap + k * kPA);
            float4 a1 = *reinterpret_cast<const float4 *>(ap + k * kPA + 32);
            float4 b0 = *reinterpret_cast<const float4 *>(bp + k * pb);
            float4 b1 = *reinterpret_cast<const float4 *>(bp + k * pb + 16);
            float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
    }
    __syncthreads();   // operands are dead: reuse the space for the product block

    float *ss = smem;  // [64][pb]
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        int xi = (i < 4 ? 4 * a + i : 32 + 4 * a + (i - 4));
        float *p = ss + xi * pb + 32 * warp + 4 * b;
        *reinterpret_cast<float4 *>(p) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        *reinterpret_cast<float4 *>(p + 16) = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
    }
    __syncthreads();

    // band extraction: left d = xi + D-1 - u, right d = u - xi
    for (int xi = warp; xi < kCorrTX; xi += nwarps) {
        int x = x0 + xi;
        if (x >= w) break;
        const float *srow = ss + xi * pb;
        size_t pix = ((size_t)img * h + y) * w + x;
        float best = 0.f;
        int best_d = 0x7fffffff;
        for (int d = lane; d < D; d += 32) {
            int u = side == 0 ? xi + D - 1 - d : xi + d;
            float v = srow[u];
            if (WRITE_VOL) {
                float *vol = side == 0 ? left_vol : right_vol;
                vol[pix * D + d] = v;
            }
            if (WRITE_WTA) {
                if (best_d == 0x7fffffff || v > best) { best = v; best_d = d; }
            }
        }
        if (WRITE_WTA) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                float ov = __shfl_xor_sync(0xffffffffu, best, o);
                int od = __shfl_xor_sync(0xffffffffu, best_d, o);
                if (od != 0x7fffffff && (best_d == 0x7fffffff || ov > best || (ov == best && od < best_d))) {
                    best = ov;
                    best_d = od;
                }
            }
            if (lane == 0) (side == 0 ? left_disp : right_disp)[pix] = (uint8_t)(best_d & 255);
        }
    }
}

int launch_corr_ffma(stereo_ctx *ctx, const float *lmap, const float *rmap, int n, int h, int w, int c,
                     int D, float *left_vol, float *right_vol, uint8_t *left_disp, uint8_t *right_disp)
{
    STEREO_REQUIRE(c >= 4 && c % 4 == 0, "feature channels must be a positive multiple of 4, got %d", c);
    int nb = kCorrTX + D - 1;
    int nwarps = (nb + 31) / 32;
    int pb = nwarps * 32 + 4;
    size_t smem = sizeof(float) * (size_t)c * (kPA + pb);
    size_t smem_prod = sizeof(float) * (size_t)kCorrTX * pb;
    if (smem_prod > smem) smem = smem_prod;
    if (nwarps > 16 || smem > 227 * 1024) {
        set_error("disp_range %d with %d channels exceeds the correlation kernel's tile (needs %zu B shared memory)",
                  D, c, smem);
        return STEREO_E_LIMIT;
    }
    bool vol = left_vol != nullptr, wta = left_disp != nullptr;
    STEREO_REQUIRE(vol || wta, "nothing to compute");
    STEREO_REQUIRE((left_vol == nullptr) == (right_vol == nullptr) && (left_disp == nullptr) == (right_disp == nullptr),
                   "left and right outputs must be given together");
    dim3 grid((w + kCorrTX - 1) / kCorrTX, h, 2 * n);
    auto launch = [&](auto kern) -> int {
        STEREO_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ProfScope prof(ctx, STEREO_K_CORR);
        kern<<<grid, nwarps * 32, smem, ctx->stream>>>(lmap, rmap, h, w, c, D, pb, left_vol, right_vol,
                                                      left_disp, right_disp);
        ctx->launches += 1;
        STEREO_CUDA_CHECK(cudaGetLastError());
        return STEREO_OK;
    };
    if (vol && wta) return launch(corr_ffma_kernel<true, true>);
    if (vol) return launch(corr_ffma_kernel<true, false>);
    return launch(corr_ffma_kernel<false, true>);
}

// ---- standalone WTA over a materialised volume (inference.py:169-183; post.lua argmin) ----
// one warp per pixel, coalesced reads along D, shuffle arg-reduction, lowest index wins ties
__global__ void wta_kernel(const float *__restrict__ vol, size_t npix, int D, int take_min,
                           uint8_t *__restrict__ disp)
{
    int lane = threadIdx.x & 31;
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t p = warp; p < npix; p += nwarp) {
        const float *row = vol + p * D;
        float best = 0.f;
        int best_d = 0x7fffffff;
        if (D <= 256) {
            // all of the lane's (up to eight) loads in flight before the first compare
            float v8[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v8[i] = lane + 32 * i < D ? __ldg(row + lane + 32 * i) : 0.f;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int d = lane + 32 * i;
                const float v = take_min ? -v8[i] : v8[i];
                if (d < D && (best_d == 0x7fffffff || v > best)) { best = v; best_d = d; }
            }
        } else {
            for (int d = lane; d < D; d += 32) {
                float v = __ldg(row + d);
                if (take_min) v = -v;
                if (best_d == 0x7fffffff || v > best) { best = v; best_d = d; }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, best, o);
            int od = __shfl_xor_sync(0xffffffffu, best_d, o);
            if (od != 0x7fffffff && (best_d == 0x7fffffff || ov > best || (ov == best && od < best_d))) {
                best = ov;
                best_d = od;
            }
        }
        if (lane == 0) disp[p] = (uint8_t)(best_d & 255);
    }
}

// Vector variant for D % 4 == 0, D <= 256: half a warp per pixel, four 16-byte loads per lane
// in flight before any compare, 4 shuffle rounds.  Same result as wta_kernel.
// MOMENTS: the same pass also accumulates the sum and the sum of squares of every voxel (double; per-block
// partials in block order, combined by volume_std_final_kernel): the volume std of post.lua:146 and the stage image
// of post.lua:125 from ONE read of the volume.
template <bool MOMENTS>
__global__ __launch_bounds__(256) void wta_vec_kernel(const float4 *__restrict__ vol, size_t npix, int D4,
                                                      int take_min, uint8_t *__restrict__ disp,
                                                      double *__restrict__ partial)
{
    double m_s = 0.0, m_ss = 0.0;
    const int sub = threadIdx.x & 15;
    size_t grp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 4;
    const size_t ngrp = ((size_t)gridDim.x * blockDim.x) >> 4;
    const float sign = take_min ? -1.f : 1.f;
    // every group runs the same number of iterations so that the shuffles stay converged
    const size_t iters = (npix + ngrp - 1) / ngrp;
    for (size_t it = 0; it < iters; ++it, grp += ngrp) {
        const bool live = grp < npix;
        const float4 *row = vol + (live ? grp : 0) * D4;
        float4 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int q = sub + 16 * j;
            v[j] = (live && q < D4) ? __ldg(row + q) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (MOMENTS) {
            // rows past the end and columns past D were loaded as zeros: they add nothing
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const double a = v[j].x, b = v[j].y, c = v[j].z, d = v[j].w;
                m_s += (a + b) + (c + d);
                m_ss += (a * a + b * b) + (c * c + d * d);
            }
        }
        float best = 0.f;
        int best_d = 0x7fffffff;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int q = sub + 16 * j;
            if (q < D4) {
                const float e[4] = {v[j].x * sign, v[j].y * sign, v[j].z * sign, v[j].w * sign};
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (best_d == 0x7fffffff || e[k] > best) { best = e[k]; best_d = 4 * q + k; }
            }
        }
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, best, o);
            int od = __shfl_xor_sync(0xffffffffu, best_d, o);
            if (od != 0x7fffffff && (best_d == 0x7fffffff || ov > best || (ov == best && od < best_d))) {
                best = ov;
                best_d = od;
            }
        }
        if (live && sub == 0) disp[grp] = (uint8_t)(best_d & 255);
    }
    if (MOMENTS) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m_s += __shfl_down_sync(0xffffffffu, m_s, o);
            m_ss += __shfl_down_sync(0xffffffffu, m_ss, o);
        }
        __shared__ double sh[2][8];
        if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = m_s; sh[1][threadIdx.x >> 5] = m_ss; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double a = 0.0, b = 0.0;
            for (int i = 0; i < 8; ++i) { a += sh[0][i]; b += sh[1][i]; }
            partial[2 * blockIdx.x] = a;
            partial[2 * blockIdx.x + 1] = b;
        }
    }
}

// arg max / arg min map and the moment partials of a 16-byte aligned volume with D % 4 == 0, D <= 256, in one pass;
// *nblocks = number of (sum, sum of squares) pairs written to `partial` (room for 2 * wta_moments_max_blocks())
int wta_moments_max_blocks(const stereo_ctx *ctx) { return ctx->sm_count * 16; }

int launch_wta_moments(stereo_ctx *ctx, const float *vol, size_t npix, int D, int take_min, uint8_t *disp,
                       double *partial, int *nblocks)
{
    STEREO_REQUIRE(D % 4 == 0 && D <= 256 && (uintptr_t)vol % 16 == 0, "wta + moments pass needs an aligned volume");
    int vblocks = (int)((npix + 15) / 16);
    if (vblocks > wta_moments_max_blocks(ctx)) vblocks = wta_moments_max_blocks(ctx);
    wta_vec_kernel<true><<<vblocks, 256, 0, ctx->stream>>>(reinterpret_cast<const float4 *>(vol), npix, D / 4, take_min,
                                                           disp, partial);
    *nblocks = vblocks;
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

int launch_wta(stereo_ctx *ctx, const float *vol, int n, int h, int w, int D, int take_min, uint8_t *disp)
{
    size_t npix = (size_t)n * h * w;
    if (D % 4 == 0 && D <= 256 && (uintptr_t)vol % 16 == 0) {
        int vblocks = (int)((npix + 15) / 16);
        int vcap = ctx->sm_count * 16;
        if (vblocks > vcap) vblocks = vcap;
        ProfScope prof(ctx, STEREO_K_WTA);
        wta_vec_kernel<false><<<vblocks, 256, 0, ctx->stream>>>(reinterpret_cast<const float4 *>(vol), npix, D / 4,
                                                                take_min, disp, nullptr);
        ctx->launches += 1;
        STEREO_CUDA_CHECK(cudaGetLastError());
        return STEREO_OK;
    }
    int blocks = (int)((npix + 7) / 8);
    int cap = ctx->sm_count * 16;
    if (blocks > cap) blocks = cap;
    ProfScope prof(ctx, STEREO_K_WTA);
    wta_kernel<<<blocks, 256, 0, ctx->stream>>>(vol, npix, D, take_min, disp);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// ---- left-right consistency (adcensus.outlier_detection, smooth/nyu.lua:95) ----
// The reference searches, for every inconsistent left pixel x, all d in [0, min(D, x+1)) for a right
// pixel x-d whose disparity is within 1 of d (-> "mismatch", else "occlusion").  Seen from the right
// image that is a scatter: right pixel x' with disparity v explains the left positions x'+d for
// d in {v-1, v, v+1} and 0 <= d < D.  One CTA per image row marks those positions in shared memory
// (3 stores per pixel instead of a D-step search per pixel), then classifies the row.
constexpr int kLrMaxW = 8192;

__global__ __launch_bounds__(256) void lr_check_kernel(const uint8_t *__restrict__ d0, const uint8_t *__restrict__ d1,
                                                       int w, int D, uint8_t *__restrict__ outlier)
{
    __shared__ uint8_t s_d1[kLrMaxW];
    __shared__ uint8_t s_mark[kLrMaxW];
    const size_t row = (size_t)blockIdx.x * w;
    for (int x = threadIdx.x; x < w; x += blockDim.x) {
        s_d1[x] = d1[row + x];
        s_mark[x] = 0;
    }
    __syncthreads();
    for (int x = threadIdx.x; x < w; x += blockDim.x) {
        const int v = s_d1[x];
#pragma unroll
        for (int d = v - 1; d <= v + 1; ++d)
            if (d >= 0 && d < D && x + d < w) s_mark[x + d] = 1;
    }
    __syncthreads();
    for (int x = threadIdx.x; x < w; x += blockDim.x) {
        const int d0i = d0[row + x];
        uint8_t res;
        if (x - d0i < 0) res = 1;
        else if (abs(d0i - (int)s_d1[x - d0i]) <= 1) res = 0;
        else res = s_mark[x] ? 2 : 1;
        outlier[row + x] = res;
    }
}

int launch_lr_check(stereo_ctx *ctx, const uint8_t *d0, const uint8_t *d1, int n, int h, int w, int D,
                    uint8_t *outlier)
{
    STEREO_REQUIRE(w <= kLrMaxW, "image width %d exceeds the LR-check row buffer (%d)", w, kLrMaxW);
    ProfScope prof(ctx, STEREO_K_LR);
    lr_check_kernel<<<(unsigned)(n * h), 256, 0, ctx->stream>>>(d0, d1, w, D, outlier);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
