// Entry points of the post chain (post.lua + smooth/nyu.lua): parameter defaults, the
// three smooth.nyu.* functions as host-pointer calls, and the whole per-pair pipeline.
#include <algorithm>
#include <cmath>
#include <vector>

#include "common.cuh"

namespace stereo {

static int check_params(const stereo_post_params *p, int h, int w, int D)
{
    STEREO_REQUIRE(p != nullptr, "null post parameters");
    STEREO_REQUIRE(h > 0 && w > 0 && D > 0, "h, w, D must be positive (got %d, %d, %d)", h, w, D);
    STEREO_REQUIRE(p->L1 >= 1 && p->sgm_i >= 1, "L1 and sgm_i must be >= 1");
    STEREO_REQUIRE(p->sgm_paths == 4 || p->sgm_paths == 8, "sgm_paths must be 4 (the reference) or 8, got %d", p->sgm_paths);
    STEREO_REQUIRE(w < 32768 && h < 32768, "image dimensions must fit the int16 arm encoding");
    return STEREO_OK;
}

static SgmParams sgm_params(const stereo_post_params *p)
{
    return SgmParams{p->pi1, p->pi2, p->tau_so, p->alpha1, p->sgm_q1, p->sgm_q2};
}

// smooth/nyu.lua:77-89 gaussian(sigma): double arithmetic, stored as float (:cuda())
static std::vector<float> gaussian_kernel(double sigma, int *radius)
{
    int kr = (int)std::ceil(sigma * 3.0);
    int ks = 2 * kr + 1;
    std::vector<float> k((size_t)ks * ks);
    double s = sigma;
    for (int i = 0; i < ks; ++i)
        for (int j = 0; j < ks; ++j) {
            double y = i - kr, x = j - kr;
            k[(size_t)i * ks + j] = (float)std::exp(-(x * x + y * y) / (2.0 * s * s));
        }
    *radius = kr;
    return k;
}

// Carve the small per-pair buffers out of one allocation.
struct PostMaps {
    float *guide_l, *guide_r, *d0, *d1, *t0, *t1, *outlier, *stdp, *gauss;
    short *arms_l, *arms_r;
    uint8_t *u8a, *u8b;
};

static int carve_maps(stereo_ctx *ctx, int h, int w, int gauss_n, PostMaps *m)
{
    size_t hw = (size_t)h * w;
    size_t hw4 = (hw + 3) & ~(size_t)3;
    size_t g4 = ((size_t)gauss_n + 3) & ~(size_t)3;
    size_t bytes = sizeof(float) * (7 * hw4 + 4 + g4) + sizeof(short) * 8 * hw4 + 2 * hw4;
    STEREO_TRY(ctx->post_maps.reserve(bytes));
    float *f = ctx->post_maps.as<float>();
    m->guide_l = f; f += hw4;
    m->guide_r = f; f += hw4;
    m->d0 = f; f += hw4;
    m->d1 = f; f += hw4;
    m->t0 = f; f += hw4;
    m->t1 = f; f += hw4;
    m->outlier = f; f += hw4;
    m->stdp = f; f += 4;
    m->gauss = f; f += g4;
    short *s = reinterpret_cast<short *>(f);
    m->arms_l = s; s += 4 * hw4;
    m->arms_r = s; s += 4 * hw4;
    uint8_t *u = reinterpret_cast<uint8_t *>(s);
    m->u8a = u; u += hw4;
    m->u8b = u;
    return STEREO_OK;
}

// one adcensus.cbca call in either summation order; `tmp` is a spare volume for the separable form
// The SGM ring kernels fetch the 16-byte aligned superset of every run of an unaligned volume (sgm_ring.cu): up to
// 12 bytes past the last voxel, which this slack keeps inside the allocation whatever its size.
static constexpr size_t kVolSlack = 256;

static int cbca_dev(stereo_ctx *ctx, const stereo_post_params *p, const short *x0c, const short *x1c, const float *vol,
                    float *tmp, float *out, int h, int w, int D, int direction, int slot = 0, int in_pitch = 0)
{
    if (in_pitch != 0 && in_pitch != D) {
        STEREO_REQUIRE(!p->cbca_exact && D <= 256 && p->L1 <= 5, "padded input rows are read by the fused CBCA kernel only");
        return launch_cbca_fused(ctx, x0c, x1c, vol, out, h, w, D, direction, p->L1, in_pitch);
    }
    if (p->cbca_exact || D > 512) return launch_cbca(ctx, x0c, x1c, vol, out, h, w, D, direction);
    // one fused pass (cbca_fused.cu); arms longer than 5 take the two-pass separable kernels
    if (D <= 256 && p->L1 <= 5) return launch_cbca_fused(ctx, x0c, x1c, vol, out, h, w, D, direction, p->L1);
    STEREO_TRY(ctx->post_cnt[slot].reserve((size_t)h * w * D));
    return launch_cbca_separable(ctx, x0c, x1c, vol, tmp, ctx->post_cnt[slot].as<uint8_t>(), out, h, w, D, direction,
                                 p->L1);
}

// Runs the launches issued during its lifetime on another stream.
struct StreamSwap {
    stereo_ctx *ctx;
    cudaStream_t saved;
    StreamSwap(stereo_ctx *c, cudaStream_t s) : ctx(c), saved(c->stream) { c->stream = s; }
    ~StreamSwap() { ctx->stream = saved; }
};

// nyu.lua:91-115 on device maps
static int nyu_post_dev(stereo_ctx *ctx, PostMaps &m, const float *left_vol, int h, int w, int D,
                        const stereo_post_params *p, float **final_out)
{
    STEREO_TRY(launch_outlier(ctx, m.d0, m.d1, h, w, D, m.outlier));
    float *cur = m.d0, *other = m.t0;
    if (p->occlusion == 1) {
        STEREO_TRY(launch_interpolate(ctx, cur, m.outlier, h, w, m.t1, other));
        std::swap(cur, other);
    }
    if (p->subpixel == 1) {
        STEREO_TRY(launch_subpixel(ctx, cur, left_vol, h, w, D, other));
        std::swap(cur, other);
    }
    if (p->median == 1) {
        STEREO_TRY(launch_median(ctx, cur, h, w, 5, other));
        std::swap(cur, other);
    }
    if (p->bilateral == 1) {
        int radius = 0;
        std::vector<float> k = gaussian_kernel(p->blur_sigma, &radius);
        STEREO_CUDA_CHECK(cudaMemcpyAsync(m.gauss, k.data(), sizeof(float) * k.size(), cudaMemcpyHostToDevice,
                                          ctx->stream));
        // the host vector must outlive the copy
        STEREO_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
        STEREO_TRY(launch_mean2d(ctx, cur, m.gauss, h, w, radius, p->blur_t, other));
        std::swap(cur, other);
    }
    *final_out = cur;
    return STEREO_OK;
}

static int gauss_count(const stereo_post_params *p)
{
    int kr = (int)std::ceil(p->blur_sigma * 3.0);
    return (2 * kr + 1) * (2 * kr + 1);
}

static int post_dev(stereo_ctx *ctx, const float *left_cost, const float *right_cost, const uint8_t *left_disp,
                    const uint8_t *right_disp, int h, int w, int D, const stereo_post_params *p,
                    uint8_t *const stages[8], float *final_disp)
{
    STEREO_TRY(check_params(p, h, w, D));
    STEREO_REQUIRE(left_cost && right_cost && left_disp && right_disp && stages, "null pointer");
    STEREO_REQUIRE(D <= 256, "the post chain keeps disparity maps as bytes like post.lua (:byte()); D=%d > 256", D);
    const size_t hw = (size_t)h * w, nvox = hw * D;
    PostMaps m;
    STEREO_TRY(carve_maps(ctx, h, w, gauss_count(p), &m));
    // STEREO_POST_SERIAL=1 (measurement knob), or kernel-class profiling switched on: both sides on one stream, so
    // that event times are of kernels running alone (as stereo_pairs does with its lanes)
    static const bool serial_env = [] { const char *e = getenv("STEREO_POST_SERIAL"); return e && atoi(e) == 1; }();
    const bool serial = serial_env || ctx->prof.on;
    const bool fused_cbca = !p->cbca_exact && D <= 256 && p->L1 <= 5;
    // everything is reserved up front: a reallocation (cudaFree) in the middle of the two-stream section would
    // serialise it.  Four working volumes per side: result, SGM scratch, and what they read.
    // A range that is not a multiple of 4 (the reference's default 201) makes every [x][d] row of a volume start off
    // the 16-byte grid, which costs the SGM kernels a third of their speed (aligned supersets, 4-byte accesses).  When
    // SGM sits between two fused CBCA calls, the scale pass in front of it writes rows padded to Dp = 4 * ceil(D / 4)
    // floats with "no such disparity" in the padding, SGM runs as a range of Dp (the padded slots never win a minimum
    // and never turn into NaN: kSgmBig, then +inf in the sums), and the CBCA behind it reads rows of Dp.  Same bits
    // for every real disparity.  STEREO_POST_PAD=0 switches it off (cross-check).
    static const bool pad_ok = [] { const char *e = getenv("STEREO_POST_PAD"); return !(e && atoi(e) == 0); }();
    const int Dp = (D + 3) & ~3;
    const bool padded = pad_ok && Dp != D && p->sgm == 1 && p->nyu_cost_agg_2 > 0 && fused_cbca;
    const size_t nvox_alloc = hw * (size_t)(padded ? Dp : D);
    for (int i = 0; i < 8; ++i) STEREO_TRY(ctx->post_vol[i].reserve(sizeof(float) * nvox_alloc + kVolSlack));
    if (!fused_cbca && !p->cbca_exact)
        for (int i = 0; i < 2; ++i) STEREO_TRY(ctx->post_cnt[i].reserve(nvox));
    STEREO_TRY(ctx->scratch.reserve(sizeof(double) * 4 * (size_t)std::max(1184, wta_moments_max_blocks(ctx))));
    if (!ctx->aux_stream) STEREO_CUDA_CHECK(cudaStreamCreateWithFlags(&ctx->aux_stream, cudaStreamNonBlocking));
    if (!ctx->ev_fork) STEREO_CUDA_CHECK(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
    if (!ctx->ev_join) STEREO_CUDA_CHECK(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
    STEREO_TRY(launch_guide(ctx, left_disp, h, w, m.guide_l));
    STEREO_TRY(launch_guide(ctx, right_disp, h, w, m.guide_r));
    STEREO_TRY(launch_cross(ctx, m.guide_l, h, w, p->L1, p->tau1, m.arms_l));
    STEREO_TRY(launch_cross(ctx, m.guide_r, h, w, p->L1, p->tau1, m.arms_r));
    const SgmParams sp = sgm_params(p);
    const bool vec_ok = D % 4 == 0;              // working volumes are 256-byte aligned allocations

    // The left and the right volume are independent until smooth.nyu.post: the left chain is issued on the
    // context's stream, the right chain on a second stream.  Every stage reads `cur` and writes a free buffer;
    // the buffer it read goes back to the free list unless it is the caller's (read-only) input.
    // Passes over a volume that only exist to look at it are merged where that costs nothing: the stage image
    // and the std of the first CBCA's output come from ONE read (post.lua:125 + :146), the stage image of the SGM
    // result is written by the last SGM pass (post.lua:159).  Fusing more into the producers was measured and
    // dropped - see DESIGN.md section 4 and profiles/r02_post_chain_ab.md.
    STEREO_CUDA_CHECK(cudaEventRecord(ctx->ev_fork, ctx->stream));
    if (!serial) STEREO_CUDA_CHECK(cudaStreamWaitEvent(ctx->aux_stream, ctx->ev_fork, 0));
    const float *left_final = nullptr;
    for (int side = 0; side < 2; ++side) {
        StreamSwap swap(ctx, side == 0 || serial ? ctx->stream : ctx->aux_stream);
        std::vector<float *> pool;
        for (int i = 0; i < 4; ++i) pool.push_back(ctx->post_vol[4 * side + i].as<float>());
        const int direction = side == 0 ? -1 : 1;
        const float *input = side == 0 ? left_cost : right_cost;
        const float *cur = input;
        uint8_t *final_map = side == 0 ? m.u8a : m.u8b;
        bool have_final_map = false;                 // final_map holds the arg min of `cur`
        auto advance = [&](float *dst) {
            if (cur != input) pool.push_back(const_cast<float *>(cur));
            cur = dst;
            have_final_map = false;
        };
        auto take = [&]() {
            float *dst = pool.back();
            pool.pop_back();
            return dst;
        };
        // nyu_cost_agg_1 (post.lua:116-133); both volumes are called with (x0c, x1c), nyu.lua:37,39
        for (int i = 0; i < p->nyu_cost_agg_1; ++i) {
            float *dst = take();
            float *tmp = take();
            STEREO_TRY(cbca_dev(ctx, p, m.arms_l, m.arms_r, cur, tmp, dst, h, w, D, direction, side));
            pool.push_back(tmp);
            advance(dst);
        }
        // post.lua:125 stage image (lu:max(2)) and post.lua:146-147 lu:std(): one pass when both are wanted
        float *stdp = m.stdp + side * 2;
        const bool aligned_cur = vec_ok && ((uintptr_t)cur % 16 == 0);
        if (p->nyu_cost_agg_1 > 0 && stages[side] && aligned_cur) {
            STEREO_TRY(launch_volume_std_wta(ctx, cur, hw, D, stdp, stages[side], side));
        } else {
            if (p->nyu_cost_agg_1 > 0 && stages[side]) STEREO_TRY(launch_wta(ctx, cur, 1, h, w, D, 0, stages[side]));
            STEREO_TRY(launch_volume_std(ctx, cur, nvox, stdp, side));
        }
        {   // post.lua:146,152: lu = (lu / lu:std()) * -1   (without SGM: lu / lu:std())
            float *dst = take();
            if (padded) STEREO_TRY(launch_scale_pad(ctx, cur, hw, D, Dp, stdp, 1, dst));
            else STEREO_TRY(launch_scale(ctx, cur, nvox, stdp, p->sgm == 1 ? 1 : 0, dst));
            advance(dst);
        }
        const int Dsgm = padded ? Dp : D;        // the range the SGM kernels see
        if (p->sgm == 1) {
            // post.lua:152-155 + nyu.lua:47-75
            for (int it = 0; it < p->sgm_i; ++it) {
                const bool last_iter = it == p->sgm_i - 1;
                float *dst = take();
                float *tmp = take();
                SgmExtras ex;
                ex.paths = p->sgm_paths == 8 ? 8 : 4;
                ex.final_scale = ex.paths == 8 ? 0.125f : 0.25f;
                // post.lua:159 lu:min(2): written by the last pass when somebody wants it
                const bool want_map = last_iter && (stages[2 + side] || p->nyu_cost_agg_2 <= 0);
                ex.wta = want_map ? final_map : nullptr;
                STEREO_TRY(launch_sgm(ctx, m.guide_l, m.guide_r, cur, dst, tmp, h, w, Dsgm, sp, direction, ex));
                pool.push_back(tmp);
                advance(dst);
                have_final_map = want_map;
            }
            if (stages[2 + side])
                STEREO_CUDA_CHECK(cudaMemcpyAsync(stages[2 + side], final_map, hw, cudaMemcpyDeviceToDevice, ctx->stream));
        }
        // nyu_cost_agg_2 (post.lua:171-186); the first call reads the SGM result (rows of Dsgm floats)
        for (int i = 0; i < p->nyu_cost_agg_2; ++i) {
            float *dst = take();
            float *tmp = take();
            STEREO_TRY(cbca_dev(ctx, p, m.arms_l, m.arms_r, cur, tmp, dst, h, w, D, direction, side, i == 0 ? Dsgm : D));
            pool.push_back(tmp);
            advance(dst);
        }
        // post.lua:226-229: disp = argmin - 1 (0-based here); the stage image of cross_agg 2 is the same map
        if (!have_final_map) STEREO_TRY(launch_wta(ctx, cur, 1, h, w, D, 1, final_map));
        if (p->nyu_cost_agg_2 > 0 && stages[4 + side])
            STEREO_CUDA_CHECK(cudaMemcpyAsync(stages[4 + side], final_map, hw, cudaMemcpyDeviceToDevice, ctx->stream));
        if (side == 0) left_final = cur;
        else if (!serial) STEREO_CUDA_CHECK(cudaEventRecord(ctx->ev_join, ctx->stream));      // still the second stream here
    }
    if (!serial) STEREO_CUDA_CHECK(cudaStreamWaitEvent(ctx->stream, ctx->ev_join, 0));
    if (p->post == 1) {
        STEREO_TRY(launch_u8_to_float(ctx, m.u8a, (int)hw, m.d0));
        STEREO_TRY(launch_u8_to_float(ctx, m.u8b, (int)hw, m.d1));
        float *fin = nullptr;
        STEREO_TRY(nyu_post_dev(ctx, m, left_final, h, w, D, p, &fin));
        if (stages[6]) STEREO_TRY(launch_float_to_u8(ctx, fin, (int)hw, 1.f, stages[6]));
        if (stages[7]) STEREO_TRY(launch_float_to_u8(ctx, m.outlier, (int)hw, 127.f, stages[7]));
        if (final_disp)
            STEREO_CUDA_CHECK(cudaMemcpyAsync(final_disp, fin, sizeof(float) * hw, cudaMemcpyDeviceToDevice, ctx->stream));
    }
    return STEREO_OK;
}

static int to_dev(stereo_ctx *ctx, void *dst, const void *src, size_t bytes)
{
    STEREO_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return STEREO_OK;
}

static int to_host(stereo_ctx *ctx, void *dst, const void *src, size_t bytes)
{
    STEREO_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return STEREO_OK;
}

}  // namespace stereo

using namespace stereo;

extern "C" {

int stereo_post_default_params(stereo_post_params *p)
{
    STEREO_REQUIRE(p != nullptr, "null pointer");
    p->L1 = 5;
    p->tau1 = 0.03f;
    p->sgm_i = 1;
    p->pi1 = 2.3f;
    p->pi2 = 24.25f;
    p->tau_so = 0.08f;
    p->alpha1 = 1.75f;
    p->sgm_q1 = 3.f;
    p->sgm_q2 = 2.f;
    p->occlusion = 1;
    p->subpixel = 1;
    p->median = 1;
    p->bilateral = 1;
    p->blur_sigma = 5.99;
    p->blur_t = 5.f;
    p->nyu_cost_agg_1 = 1;
    p->sgm = 1;
    p->nyu_cost_agg_2 = 1;
    p->post = 1;
    p->cbca_exact = 0;
    p->sgm_paths = 4;
    return STEREO_OK;
}

int stereo_guide_image(stereo_ctx *ctx, const uint8_t *img, int h, int w, float *out)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_REQUIRE(img && out && h > 0 && w > 0, "bad argument");
    size_t hw = (size_t)h * w;
    STEREO_TRY(ctx->stage_in[0].reserve(hw));
    STEREO_TRY(ctx->stage_vol[0].reserve(sizeof(float) * hw));
    STEREO_TRY(to_dev(ctx, ctx->stage_in[0].ptr, img, hw));
    STEREO_TRY(launch_guide(ctx, ctx->stage_in[0].as<uint8_t>(), h, w, ctx->stage_vol[0].as<float>()));
    STEREO_TRY(to_host(ctx, out, ctx->stage_vol[0].ptr, sizeof(float) * hw));
    return finish(ctx);
}

int stereo_volume_std(stereo_ctx *ctx, const float *vol, size_t n, float *std_out)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_REQUIRE(vol && std_out && n > 1, "bad argument");
    STEREO_TRY(ctx->stage_vol[0].reserve(sizeof(float) * n));
    STEREO_TRY(ctx->stage_out[0].reserve(16));
    STEREO_TRY(to_dev(ctx, ctx->stage_vol[0].ptr, vol, sizeof(float) * n));
    STEREO_TRY(launch_volume_std(ctx, ctx->stage_vol[0].as<float>(), n, ctx->stage_out[0].as<float>()));
    STEREO_TRY(to_host(ctx, std_out, ctx->stage_out[0].ptr, sizeof(float)));
    return finish(ctx);
}

int stereo_box_agg_dev(stereo_ctx *ctx, const float *vol, int h, int w, int D, int kw, int kh, int iters, float *out,
                       float *scratch)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_REQUIRE(vol && out && h > 0 && w > 0 && D > 0, "bad argument");
    STEREO_REQUIRE(kw >= 1 && kh >= 1 && (kw & 1) && (kh & 1) && iters >= 1, "window %dx%d must be odd and positive, iters >= 1",
                   kw, kh);
    STEREO_REQUIRE(iters == 1 || scratch, "more than one pass needs a scratch volume");
    // passes ping-pong between `out` and `scratch` so that the last one lands in `out`
    const float *src = vol;
    for (int i = 0; i < iters; ++i) {
        float *dst = ((iters - 1 - i) & 1) ? scratch : out;
        STEREO_TRY(launch_box_agg(ctx, src, h, w, D, kw, kh, dst));
        src = dst;
    }
    return STEREO_OK;
}

int stereo_box_agg(stereo_ctx *ctx, const float *vol, int h, int w, int D, int kw, int kh, int iters, float *out)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_REQUIRE(vol && out && h > 0 && w > 0 && D > 0, "bad argument");
    const size_t bytes = sizeof(float) * (size_t)h * w * D;
    STEREO_TRY(ctx->stage_vol[0].reserve(bytes));
    STEREO_TRY(ctx->stage_vol[1].reserve(bytes));
    STEREO_TRY(ctx->post_vol[0].reserve(bytes));
    STEREO_TRY(to_dev(ctx, ctx->stage_vol[0].ptr, vol, bytes));
    STEREO_TRY(stereo_box_agg_dev(ctx, ctx->stage_vol[0].as<float>(), h, w, D, kw, kh, iters, ctx->stage_vol[1].as<float>(),
                                  ctx->post_vol[0].as<float>()));
    STEREO_TRY(to_host(ctx, out, ctx->stage_vol[1].ptr, bytes));
    return finish(ctx);
}

int stereo_nyu_cross_agg(stereo_ctx *ctx, const float *l_img, const float *r_img, float *left_vol, float *right_vol,
                         int h, int w, int D, int iters, const stereo_post_params *p)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_params(p, h, w, D));
    STEREO_REQUIRE(l_img && r_img && left_vol && right_vol && iters >= 0, "bad argument");
    const size_t hw = (size_t)h * w, nvox = hw * D;
    PostMaps m;
    STEREO_TRY(carve_maps(ctx, h, w, 1, &m));
    for (int i = 0; i < 3; ++i) STEREO_TRY(ctx->post_vol[i].reserve(sizeof(float) * nvox + kVolSlack));
    STEREO_TRY(to_dev(ctx, m.guide_l, l_img, sizeof(float) * hw));
    STEREO_TRY(to_dev(ctx, m.guide_r, r_img, sizeof(float) * hw));
    STEREO_TRY(launch_cross(ctx, m.guide_l, h, w, p->L1, p->tau1, m.arms_l));
    STEREO_TRY(launch_cross(ctx, m.guide_r, h, w, p->L1, p->tau1, m.arms_r));
    for (int side = 0; side < 2; ++side) {
        float *host = side == 0 ? left_vol : right_vol;
        float *a = ctx->post_vol[0].as<float>(), *b = ctx->post_vol[1].as<float>();
        STEREO_TRY(to_dev(ctx, a, host, sizeof(float) * nvox));
        for (int i = 0; i < iters; ++i) {
            // both calls pass (x0c, x1c), nyu.lua:37,39
            STEREO_TRY(cbca_dev(ctx, p, m.arms_l, m.arms_r, a, ctx->post_vol[2].as<float>(), b, h, w, D,
                                side == 0 ? -1 : 1));
            std::swap(a, b);
        }
        STEREO_TRY(to_host(ctx, host, a, sizeof(float) * nvox));
        STEREO_TRY(finish(ctx));
    }
    return STEREO_OK;
}

int stereo_nyu_sgm(stereo_ctx *ctx, const float *l_img, const float *r_img, float *vol, int h, int w, int D,
                   int direction, const stereo_post_params *p)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_params(p, h, w, D));
    STEREO_REQUIRE(l_img && r_img && vol && (direction == -1 || direction == 1), "bad argument");
    const size_t hw = (size_t)h * w, nvox = hw * D;
    PostMaps m;
    STEREO_TRY(carve_maps(ctx, h, w, 1, &m));
    for (int i = 0; i < 3; ++i) STEREO_TRY(ctx->post_vol[i].reserve(sizeof(float) * nvox + kVolSlack));
    STEREO_TRY(to_dev(ctx, m.guide_l, l_img, sizeof(float) * hw));
    STEREO_TRY(to_dev(ctx, m.guide_r, r_img, sizeof(float) * hw));
    float *a = ctx->post_vol[0].as<float>(), *b = ctx->post_vol[1].as<float>();
    STEREO_TRY(to_dev(ctx, a, vol, sizeof(float) * nvox));
    const SgmParams sp = sgm_params(p);
    SgmExtras ex;
    ex.paths = p->sgm_paths == 8 ? 8 : 4;
    ex.final_scale = ex.paths == 8 ? 0.125f : 0.25f;
    for (int it = 0; it < p->sgm_i; ++it) {
        STEREO_TRY(launch_sgm(ctx, m.guide_l, m.guide_r, a, b, ctx->post_vol[2].as<float>(), h, w, D, sp, direction, ex));
        std::swap(a, b);
    }
    STEREO_TRY(to_host(ctx, vol, a, sizeof(float) * nvox));
    return finish(ctx);
}

int stereo_nyu_post(stereo_ctx *ctx, const float *disp_left, const float *disp_right, const float *left_vol, int h,
                    int w, int D, const stereo_post_params *p, float *final_disp, float *outlier)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_params(p, h, w, D));
    STEREO_REQUIRE(disp_left && disp_right && left_vol && final_disp && outlier, "null pointer");
    const size_t hw = (size_t)h * w, nvox = hw * D;
    PostMaps m;
    STEREO_TRY(carve_maps(ctx, h, w, gauss_count(p), &m));
    STEREO_TRY(ctx->post_vol[0].reserve(sizeof(float) * nvox));
    STEREO_TRY(to_dev(ctx, m.d0, disp_left, sizeof(float) * hw));
    STEREO_TRY(to_dev(ctx, m.d1, disp_right, sizeof(float) * hw));
    STEREO_TRY(to_dev(ctx, ctx->post_vol[0].ptr, left_vol, sizeof(float) * nvox));
    float *fin = nullptr;
    STEREO_TRY(nyu_post_dev(ctx, m, ctx->post_vol[0].as<float>(), h, w, D, p, &fin));
    STEREO_TRY(to_host(ctx, final_disp, fin, sizeof(float) * hw));
    STEREO_TRY(to_host(ctx, outlier, m.outlier, sizeof(float) * hw));
    return finish(ctx);
}

int stereo_post_dev(stereo_ctx *ctx, const float *left_cost, const float *right_cost, const uint8_t *left_disp,
                    const uint8_t *right_disp, int h, int w, int D, const stereo_post_params *p,
                    uint8_t *const stages[8], float *final_disp)
{
    STEREO_TRY(check_ctx(ctx));
    return post_dev(ctx, left_cost, right_cost, left_disp, right_disp, h, w, D, p, stages, final_disp);
}

// inference.py and post.lua back to back on the device: the two cost volumes are materialised into library
// buffers, post-processed and never leave the GPU, so the reference's 850 MB of .bin files per pair
// (inference.py:210-231 -> post.lua:89-90) become an optional export.
// Device pointers.  nyu_post [n][h][w] uint8 (post.lua:238 `:byte()`), outlier and left_wta may be null.
int stereo_pairs_post_dev(stereo_ctx *ctx, int net, const uint8_t *left_rgb, const uint8_t *right_rgb, int n, int h,
                          int w, int D, int precision, const stereo_post_params *p, uint8_t *nyu_post,
                          uint8_t *outlier, uint8_t *left_wta)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_params(p, h, w, D));
    STEREO_REQUIRE(left_rgb && right_rgb && nyu_post && n > 0, "bad argument");
    const size_t hw = (size_t)h * w, nvox = hw * D;
    STEREO_TRY(ctx->stage_vol[0].reserve(sizeof(float) * nvox));
    STEREO_TRY(ctx->stage_vol[1].reserve(sizeof(float) * nvox));
    STEREO_TRY(ctx->pp_maps.reserve(10 * hw));
    uint8_t *maps = ctx->pp_maps.as<uint8_t>();          // [0] left WTA, [1] right WTA, [2..9] the eight stage images
    for (int i = 0; i < n; ++i) {
        STEREO_TRY(stereo_pairs_dev(ctx, net, left_rgb + (size_t)i * hw * 3, right_rgb + (size_t)i * hw * 3, 1, h, w, D,
                                    precision, maps, maps + hw, nullptr, ctx->stage_vol[0].as<float>(),
                                    ctx->stage_vol[1].as<float>()));
        // only the final maps are asked for: the six intermediate stage images (one WTA over a volume each) are skipped
        uint8_t *stages[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
        stages[6] = nyu_post + (size_t)i * hw;                                   // "nyu_post"
        if (outlier) stages[7] = outlier + (size_t)i * hw;                      // "outlier" (label * 127)
        STEREO_TRY(post_dev(ctx, ctx->stage_vol[0].as<float>(), ctx->stage_vol[1].as<float>(), maps, maps + hw, h, w, D, p,
                            stages, nullptr));
        if (left_wta)
            STEREO_CUDA_CHECK(cudaMemcpyAsync(left_wta + (size_t)i * hw, maps, hw, cudaMemcpyDeviceToDevice, ctx->stream));
    }
    return STEREO_OK;
}

// The same with host pointers.
int stereo_pairs_post(stereo_ctx *ctx, int net, const uint8_t *left_rgb, const uint8_t *right_rgb, int n, int h, int w,
                      int D, int precision, const stereo_post_params *p, uint8_t *nyu_post, uint8_t *outlier,
                      uint8_t *left_wta)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_REQUIRE(left_rgb && right_rgb && nyu_post && n > 0 && h > 0 && w > 0, "bad argument");
    const size_t hw = (size_t)h * w, img_b = (size_t)n * hw * 3, map_b = (size_t)n * hw;
    STEREO_TRY(ctx->stage_in[0].reserve(img_b));
    STEREO_TRY(ctx->stage_in[1].reserve(img_b));
    for (int i = 0; i < 3; ++i) STEREO_TRY(ctx->stage_out[i].reserve(map_b));
    STEREO_TRY(to_dev(ctx, ctx->stage_in[0].ptr, left_rgb, img_b));
    STEREO_TRY(to_dev(ctx, ctx->stage_in[1].ptr, right_rgb, img_b));
    STEREO_TRY(stereo_pairs_post_dev(ctx, net, ctx->stage_in[0].as<uint8_t>(), ctx->stage_in[1].as<uint8_t>(), n, h, w, D,
                                     precision, p, ctx->stage_out[0].as<uint8_t>(),
                                     outlier ? ctx->stage_out[1].as<uint8_t>() : nullptr,
                                     left_wta ? ctx->stage_out[2].as<uint8_t>() : nullptr));
    STEREO_TRY(to_host(ctx, nyu_post, ctx->stage_out[0].ptr, map_b));
    if (outlier) STEREO_TRY(to_host(ctx, outlier, ctx->stage_out[1].ptr, map_b));
    if (left_wta) STEREO_TRY(to_host(ctx, left_wta, ctx->stage_out[2].ptr, map_b));
    return finish(ctx);
}

int stereo_post(stereo_ctx *ctx, const float *left_cost, const float *right_cost, const uint8_t *left_disp,
                const uint8_t *right_disp, int h, int w, int D, const stereo_post_params *p,
                uint8_t *const stages[8], float *final_disp)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_params(p, h, w, D));
    STEREO_REQUIRE(left_cost && right_cost && left_disp && right_disp && stages, "null pointer");
    const size_t hw = (size_t)h * w, nvox = hw * D;
    STEREO_TRY(ctx->stage_vol[0].reserve(sizeof(float) * nvox));
    STEREO_TRY(ctx->stage_vol[1].reserve(sizeof(float) * nvox));
    STEREO_TRY(ctx->stage_in[0].reserve(hw));
    STEREO_TRY(ctx->stage_in[1].reserve(hw));
    STEREO_TRY(ctx->stage_out[0].reserve(8 * hw));
    STEREO_TRY(ctx->stage_out[1].reserve(sizeof(float) * hw));
    STEREO_TRY(to_dev(ctx, ctx->stage_vol[0].ptr, left_cost, sizeof(float) * nvox));
    STEREO_TRY(to_dev(ctx, ctx->stage_vol[1].ptr, right_cost, sizeof(float) * nvox));
    STEREO_TRY(to_dev(ctx, ctx->stage_in[0].ptr, left_disp, hw));
    STEREO_TRY(to_dev(ctx, ctx->stage_in[1].ptr, right_disp, hw));
    uint8_t *dstages[8];
    for (int i = 0; i < 8; ++i) dstages[i] = stages[i] ? ctx->stage_out[0].as<uint8_t>() + i * hw : nullptr;
    STEREO_TRY(post_dev(ctx, ctx->stage_vol[0].as<float>(), ctx->stage_vol[1].as<float>(),
                        ctx->stage_in[0].as<uint8_t>(), ctx->stage_in[1].as<uint8_t>(), h, w, D, p, dstages,
                        final_disp ? ctx->stage_out[1].as<float>() : nullptr));
    for (int i = 0; i < 8; ++i)
        if (stages[i]) STEREO_TRY(to_host(ctx, stages[i], dstages[i], hw));
    if (final_disp) STEREO_TRY(to_host(ctx, final_disp, ctx->stage_out[1].ptr, sizeof(float) * hw));
    return finish(ctx);
}

}  // extern "C"
