// C ABI of libstereo_b200.so: context management, argument checking, host<->device
// staging for the host-pointer entry points, and the per-pair pipeline.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>

#include "common.cuh"

namespace stereo {

static thread_local char g_error[512] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

int DevBuf::reserve(size_t want)
{
    if (want <= bytes) return STEREO_OK;
    if (ptr) {
        cudaFree(ptr);
        ptr = nullptr;
        bytes = 0;
    }
    size_t rounded = (want + 255) & ~(size_t)255;
    cudaError_t err = cudaMalloc(&ptr, rounded);
    if (err != cudaSuccess) {
        ptr = nullptr;
        (void)cudaGetLastError();
        set_error("cudaMalloc(%zu bytes) failed: %s", rounded, cudaGetErrorString(err));
        return STEREO_E_NOMEM;
    }
    bytes = rounded;
    return STEREO_OK;
}

void DevBuf::release()
{
    if (ptr) cudaFree(ptr);
    ptr = nullptr;
    bytes = 0;
}

NetDesc net_desc(int net)
{
    NetDesc d;
    auto add = [&](int ks, int cin, int cout, int relu) {
        LayerDesc &L = d.layer[d.n_layers++];
        L.ks = ks; L.cin = cin; L.cout = cout; L.relu = relu;
        L.w_off = d.n_floats;
        d.n_floats += (size_t)ks * ks * cin * cout;
        L.b_off = d.n_floats;
        d.n_floats += cout;
        d.shrink += ks - 1;
    };
    if (net == STEREO_NET_WIN9) {
        // models/model.py:30-41
        add(3, 3, 64, 1); add(3, 64, 64, 1); add(3, 64, 64, 1); add(3, 64, 64, 0);
    } else if (net == STEREO_NET_WIN37) {
        // models/model.py:5-28 (the dead conv8/bn8 pair at :21-22 is not in the graph)
        add(5, 3, 32, 1); add(5, 32, 32, 1); add(5, 32, 32, 1); add(5, 32, 64, 1);
        add(5, 64, 64, 1); add(5, 64, 64, 1); add(5, 64, 64, 1); add(5, 64, 64, 1); add(5, 64, 64, 0);
    }
    return d;
}

int check_ctx(stereo_ctx *ctx)
{
    if (!ctx) {
        set_error("null context");
        return STEREO_E_INVALID;
    }
    STEREO_CUDA_CHECK(cudaSetDevice(ctx->device));
    return STEREO_OK;
}

static int check_dims(int n, int h, int w)
{
    STEREO_REQUIRE(n > 0 && h > 0 && w > 0, "n, h, w must be positive (got %d, %d, %d)", n, h, w);
    return STEREO_OK;
}

// Planar (hi, lo) feature output of the tensor-core trunk, consumed by the tensor-core correlation.
struct PlanarOut {
    uint16_t *hi = nullptr, *lo = nullptr;   // already offset by the left margin
    int wf = 0;
};

// `u8_left`/`u8_right` (tensor-core path, n == 2 only): the original uint8 images of one pair; layer 1
// is then computed straight from them (their ImageStats must be in ctx->stats) and `in` is not read.
static int trunk_dev(stereo_ctx *ctx, int net, const float *in, int n, int h, int w, float *out, int precision,
                     const PlanarOut *planar = nullptr, const uint8_t *u8_left = nullptr,
                     const uint8_t *u8_right = nullptr, int stats_off = 0, int lane = 0)
{
    STEREO_REQUIRE(net == STEREO_NET_WIN9 || net == STEREO_NET_WIN37, "unknown net %d", net);
    if (!ctx->has_weights[net]) {
        set_error("weights for net %d have not been set (call stereo_set_weights first)", net);
        return STEREO_E_STATE;
    }
    STEREO_REQUIRE(precision == STEREO_PREC_FP32 || precision == STEREO_PREC_BF16X3 || precision == STEREO_PREC_BF16,
                   "unknown precision mode %d", precision);
    NetDesc nd = net_desc(net);
    STEREO_REQUIRE(h > nd.shrink && w > nd.shrink, "input %dx%d is too small for a receptive field of %d", h, w,
                   nd.shrink + 1);
    const float *wts = ctx->weights[net].as<float>();
    if (precision != STEREO_PREC_FP32) {
        if (net == STEREO_NET_WIN37) {
            // base_model: layer 1 (K = 75) on the CUDA cores in fp32, converted to chunk-planar bf16 (hi, lo);
            // layers 2-9 on tcgen05 with all weights streamed (conv_tcg.cu)
            STEREO_REQUIRE(in != nullptr, "base_model takes the normalised, padded fp32 image");
            const int planes = precision == STEREO_PREC_BF16X3 ? 2 : 1;
            const int h1 = h - 4, w1 = w - 4;
            const int wp = tcg_pitch(w1);
            const size_t plane_elems = (size_t)n * h1 * 8 * wp * 8;            // up to 64 channels
            DevBuf *act = lane ? ctx->act_b + 2 * (lane - 1) : ctx->act;
            STEREO_TRY(act[0].reserve(2 * plane_elems * 2));
            STEREO_TRY(act[1].reserve(2 * plane_elems * 2));
            uint16_t *buf_hi[2] = {act[0].as<uint16_t>(), act[1].as<uint16_t>()};
            uint16_t *buf_lo[2] = {buf_hi[0] + plane_elems, buf_hi[1] + plane_elems};
            // layer 1 writes the chunk-planar (hi, lo) planes itself
            STEREO_TRY(launch_conv_first_split(ctx, nd.layer[0], wts, in, n, h, w, buf_hi[0], planes == 2 ? buf_lo[0] : nullptr, wp));
            // STEREO_CONV5_KERNEL=1: conv_tcg.cu (one MMA per tap); default: conv_tcr.cu (each input row once)
            static const bool per_tap = [] { const char *e = getenv("STEREO_CONV5_KERNEL"); return e && atoi(e) == 1; }();
            size_t img_total = 0;
            for (int l = 1; l < nd.n_layers; ++l) img_total += tcg_weight_image_elems(nd.layer[l].ks, nd.layer[l].cin, nd.layer[l].cout);
            const uint16_t *wimg = ctx->wsplit[net].as<uint16_t>() + (per_tap ? 0 : img_total);
            int ch = h1, cw = w1, cur = 0;
            for (int l = 1; l < nd.n_layers; ++l) {
                const LayerDesc &L = nd.layer[l];
                const bool last = l + 1 == nd.n_layers;
                void *o_hi = buf_hi[cur ^ 1], *o_lo = buf_lo[cur ^ 1];
                int o_wp = wp;
                float *o_f32 = nullptr;
                if (last) {
                    if (planar) { o_hi = planar->hi; o_lo = planar->lo; o_wp = planar->wf; }
                    else { o_hi = nullptr; o_lo = nullptr; o_f32 = out; }
                }
                STEREO_TRY((per_tap ? launch_conv_tcg : launch_conv_tcr)(ctx, L.ks, L.cin, L.cout, buf_hi[cur], buf_lo[cur], n, ch,
                                                                         cw, wp, wimg, wts + L.b_off, o_hi, o_lo, o_wp,
                                                                         o_f32, L.relu, planes));
                wimg += tcg_weight_image_elems(L.ks, L.cin, L.cout);
                ch -= L.ks - 1;
                cw -= L.ks - 1;
                cur ^= 1;
            }
            return STEREO_OK;
        }
        // base_model_ws_9: layers 2-4 on tcgen05 (conv_tc3.cu: CTA pair, every input row multiplied once, N = 192)
        const int planes = precision == STEREO_PREC_BF16X3 ? 2 : 1;
        const int wp = tc3_pitch(w - 2);
        const size_t plane_elems = (size_t)n * (h - 2) * 8 * wp * 8;
        DevBuf *act = lane ? ctx->act_b + 2 * (lane - 1) : ctx->act;
        STEREO_TRY(act[0].reserve(2 * plane_elems * 2));
        STEREO_TRY(act[1].reserve(2 * plane_elems * 2));
        uint16_t *a_hi = act[0].as<uint16_t>(), *a_lo = a_hi + plane_elems;
        uint16_t *b_hi = act[1].as<uint16_t>(), *b_lo = b_hi + plane_elems;
        const size_t wl = tc3_weight_image_elems();
        const uint16_t *wimg = ctx->wsplit[net].as<uint16_t>();
        auto launch_conv_tc = launch_conv_tc3;
        if (u8_left)
            // layer 1 straight from the uint8 images on the tensor cores (conv1_tc.cu)
            STEREO_TRY(launch_conv1_tc(ctx, u8_left, u8_right, h - nd.shrink, w - nd.shrink,
                                       ctx->wsplit[net].as<uint16_t>() + 3 * tc3_weight_image_elems(),
                                       wts + nd.layer[0].b_off, a_hi, planes == 2 ? a_lo : nullptr, wp, stats_off));
        else
            STEREO_TRY(launch_conv_first_split(ctx, nd.layer[0], wts, in, n, h, w, a_hi, planes == 2 ? a_lo : nullptr, wp));
        STEREO_TRY(launch_conv_tc(ctx, a_hi, a_lo, n, h - 2, w - 2, wp, wimg, wts + nd.layer[1].b_off, b_hi, b_lo, wp,
                                  nullptr, 1, planes));
        STEREO_TRY(launch_conv_tc(ctx, b_hi, b_lo, n, h - 4, w - 4, wp, wimg + wl, wts + nd.layer[2].b_off, a_hi, a_lo,
                                  wp, nullptr, 1, planes));
        if (planar)
            STEREO_TRY(launch_conv_tc(ctx, a_hi, a_lo, n, h - 6, w - 6, wp, wimg + 2 * wl, wts + nd.layer[3].b_off,
                                      planar->hi, planar->lo, planar->wf, nullptr, 0, planes));
        else
            STEREO_TRY(launch_conv_tc(ctx, a_hi, a_lo, n, h - 6, w - 6, wp, wimg + 2 * wl, wts + nd.layer[3].b_off,
                                      nullptr, nullptr, wp, out, 0, planes));
        return STEREO_OK;
    }
    size_t max_act = 0;
    {
        int ch = h, cw = w;
        for (int l = 0; l + 1 < nd.n_layers; ++l) {
            ch -= nd.layer[l].ks - 1;
            cw -= nd.layer[l].ks - 1;
            size_t b = sizeof(float) * (size_t)n * ch * cw * nd.layer[l].cout;
            if (b > max_act) max_act = b;
        }
    }
    STEREO_TRY(ctx->act[0].reserve(max_act));
    STEREO_TRY(ctx->act[1].reserve(max_act));
    const float *cur = in;
    int ch = h, cw = w;
    for (int l = 0; l < nd.n_layers; ++l) {
        float *dst = (l + 1 == nd.n_layers) ? out : ctx->act[l & 1].as<float>();
        STEREO_TRY(launch_conv_ffma(ctx, nd.layer[l], wts, cur, n, ch, cw, dst));
        ch -= nd.layer[l].ks - 1;
        cw -= nd.layer[l].ks - 1;
        cur = dst;
    }
    return STEREO_OK;
}

// Chunk-planar bf16 feature buffers with zero margins: [plane][side][n][h][8][wf][8].
struct PlanarFeat {
    uint16_t *l_hi, *l_lo, *r_hi, *r_lo;
    int wf, ml;
};

static int ensure_planar(stereo_ctx *ctx, int n, int h, int w, int D, PlanarFeat *f, int lane = 0)
{
    DevBuf &buf = lane ? ctx->featp_b[lane - 1] : ctx->featp;
    int *cached = lane ? ctx->featp_b_geom[lane - 1] : ctx->featp_geom;
    corr_tc_geometry(w, D, &f->wf, &f->ml);
    const size_t side_elems = (size_t)n * h * 8 * f->wf * 8;
    const size_t bytes = 4 * side_elems * sizeof(uint16_t);
    const bool grown = bytes > buf.bytes;
    STEREO_TRY(buf.reserve(bytes));
    const int geom[5] = {n, h, w, f->wf, f->ml};
    bool same = !grown;
    for (int i = 0; i < 5; ++i) same = same && cached[i] == geom[i];
    if (!same) {
        // margins must read as exact zeros (the zero-filled slots of inference.py:140-141)
        STEREO_CUDA_CHECK(cudaMemsetAsync(buf.ptr, 0, buf.bytes, ctx->stream));
        for (int i = 0; i < 5; ++i) cached[i] = geom[i];
    }
    uint16_t *base = buf.as<uint16_t>();
    f->l_hi = base;
    f->r_hi = base + side_elems;
    f->l_lo = base + 2 * side_elems;
    f->r_lo = base + 3 * side_elems;
    return STEREO_OK;
}

// Correlation from caller-provided fp32 NHWC feature maps.
static int corr_dev(stereo_ctx *ctx, const float *lmap, const float *rmap, int n, int h, int w, int c, int D,
                    int precision, float *left_vol, float *right_vol, uint8_t *left_disp, uint8_t *right_disp)
{
    if (precision == STEREO_PREC_FP32 || c != 64 || D > corr_tc_max_disp())
        return launch_corr_ffma(ctx, lmap, rmap, n, h, w, c, D, left_vol, right_vol, left_disp, right_disp);
    PlanarFeat f;
    STEREO_TRY(ensure_planar(ctx, n, h, w, D, &f));
    const int planes = precision == STEREO_PREC_BF16X3 ? 2 : 1;
    STEREO_TRY(launch_nhwc_to_planar(ctx, lmap, n, h, w, f.wf, f.ml, f.l_hi, planes == 2 ? f.l_lo : nullptr));
    STEREO_TRY(launch_nhwc_to_planar(ctx, rmap, n, h, w, f.wf, f.ml, f.r_hi, planes == 2 ? f.r_lo : nullptr));
    return launch_corr_tc(ctx, f.l_hi, f.l_lo, f.r_hi, f.r_lo, n, h, w, f.wf, f.ml, D, planes, left_vol, right_vol,
                          left_disp, right_disp);
}

static int pairs_dev(stereo_ctx *ctx, int net, const uint8_t *left_rgb, const uint8_t *right_rgb, int n, int h,
                     int w, int D, int precision, uint8_t *left_disp, uint8_t *right_disp, uint8_t *lr_outlier,
                     float *left_vol, float *right_vol)
{
    STEREO_REQUIRE(net == STEREO_NET_WIN9 || net == STEREO_NET_WIN37, "unknown net %d", net);
    STEREO_REQUIRE(D > 0, "disp_range must be positive, got %d", D);
    // inference.py:202-207 writes the maps as 8-bit PNGs; the 3-pixel error of a wrapped map would be meaningless
    STEREO_REQUIRE(D <= 256, "disparity maps are bytes like the reference's PNGs; disp_range %d > 256", D);
    STEREO_REQUIRE(left_rgb && right_rgb && left_disp && right_disp, "null image or disparity pointer");
    NetDesc nd = net_desc(net);
    const int patch = nd.shrink + 1;
    const int hp = h + nd.shrink, wp = w + nd.shrink;
    const size_t pad_img = (size_t)hp * wp * 3;
    const size_t feat_img = (size_t)h * w * 64;
    STEREO_TRY(ctx->padded.reserve(sizeof(float) * 2 * pad_img * 4));      // one padded pair per lane
    STEREO_TRY(ctx->feat.reserve(sizeof(float) * 2 * feat_img));
    float *padded = ctx->padded.as<float>();
    float *feat = ctx->feat.as<float>();
    const size_t img_bytes = (size_t)h * w * 3;
    const size_t map = (size_t)h * w;
    const bool tc_corr = precision != STEREO_PREC_FP32 && D <= corr_tc_max_disp();
    const bool from_u8 = tc_corr && net == STEREO_NET_WIN9;      // layer 1 reads the uint8 images itself
    PlanarFeat pf[4];
    // Pairs alternate between up to four streams ("lanes") with their own activation and feature buffers: the
    // first CTAs of one pair's next kernel fill the SMs that the tail of another pair's kernel leaves idle, and
    // launch gaps disappear.  Kernel-class profiling wants every kernel alone on the GPU: one lane while it is on.
    static const int max_lanes = [] { const char *e = getenv("STEREO_PAIR_LANES"); int v = e ? atoi(e) : 2; return v < 1 ? 1 : (v > 4 ? 4 : v); }();
    int lanes = (tc_corr && !ctx->prof.on) ? max_lanes : 1;
    if (lanes > n) lanes = n;
    // Fused path: the features of `group` consecutive pairs stay in the lane's planar buffer and ONE correlation
    // launch serves them - a single pair is 375 rows for 148 persistent CTAs (the third round runs half empty).
    // The trunk writes left and right of a pair as neighbouring images: [hi | lo][pair][side][h][8][wf][8].
    static const int max_group = [] { const char *e = getenv("STEREO_CORR_GROUP"); int v = e ? atoi(e) : 4; return v < 1 ? 1 : (v > 16 ? 16 : v); }();
    int group = 1;
    if (tc_corr && !left_vol && !right_vol && corr_both_enabled()) {
        group = n / lanes < max_group ? n / lanes : max_group;
        if (group < 1) group = 1;
    }
    cudaStream_t lane_stream[4] = {ctx->stream, ctx->stream, ctx->stream, ctx->stream};
    if (tc_corr) {
        // tensor-core path: only the image statistics (all pairs at once); layer 1 reads the uint8 images itself
        STEREO_TRY(launch_pair_stats(ctx, left_rgb, right_rgb, n, h, w));
        STEREO_TRY(ensure_planar(ctx, group, h, w, D, &pf[0], 0));
        if (lanes > 1) {
            if (!ctx->ev_fork) STEREO_CUDA_CHECK(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
            if (!ctx->ev_join) STEREO_CUDA_CHECK(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
            STEREO_CUDA_CHECK(cudaEventRecord(ctx->ev_fork, ctx->stream));
            for (int l = 1; l < lanes; ++l) {
                if (!ctx->lane_streams[l - 1]) {
                    STEREO_CUDA_CHECK(cudaStreamCreateWithFlags(&ctx->lane_streams[l - 1], cudaStreamNonBlocking));
                    STEREO_CUDA_CHECK(cudaEventCreateWithFlags(&ctx->lane_join[l - 1], cudaEventDisableTiming));
                }
                lane_stream[l] = ctx->lane_streams[l - 1];
                STEREO_CUDA_CHECK(cudaStreamWaitEvent(lane_stream[l], ctx->ev_fork, 0));
            }
            for (int l = 1; l < lanes; ++l) {
                cudaStream_t saved = ctx->stream;
                ctx->stream = lane_stream[l];                   // the first-use memset of the margins goes to the lane
                int rc = ensure_planar(ctx, group, h, w, D, &pf[l], l);
                ctx->stream = saved;
                STEREO_TRY(rc);
            }
        }
    }
    struct LaneStream {                      // launches go to ctx->stream: point it at the lane's stream
        stereo_ctx *ctx;
        cudaStream_t saved;
        LaneStream(stereo_ctx *c, cudaStream_t s) : ctx(c), saved(c->stream) { c->stream = s; }
        ~LaneStream() { ctx->stream = saved; }
    };
    for (int i = 0; i < n; ++i) {
        const int lane = (i / group) % lanes, slot = i % group;
        LaneStream on_lane(ctx, lane_stream[lane]);
        // left and right of one pair go through the trunk as a batch of two
        float *padded_lane = padded + (size_t)lane * 2 * pad_img;
        if (tc_corr && !from_u8) {
            // base_model: the statistics of all pairs are there already, the lane has its own padded images
            STEREO_TRY(launch_pad_with_stats(ctx, left_rgb + i * img_bytes, h, w, patch, 2 * i, padded_lane));
            STEREO_TRY(launch_pad_with_stats(ctx, right_rgb + i * img_bytes, h, w, patch, 2 * i + 1, padded_lane + pad_img));
        } else if (!from_u8) {
            STEREO_TRY(launch_normalize_pad(ctx, left_rgb + i * img_bytes, 1, h, w, patch, padded_lane));
            STEREO_TRY(launch_normalize_pad(ctx, right_rgb + i * img_bytes, 1, h, w, patch, padded_lane + pad_img));
        }
        float *lv = left_vol ? left_vol + i * map * D : nullptr;
        float *rv = right_vol ? right_vol + i * map * D : nullptr;
        if (tc_corr) {
            // the trunk's last layer writes the chunk-planar bf16 features the correlation reads
            // (image 0 = left, image 1 = right are adjacent sides of one plane)
            const PlanarFeat &f = pf[lane];
            const size_t img_elems = (size_t)h * 8 * f.wf * 8;
            const int planes = precision == STEREO_PREC_BF16X3 ? 2 : 1;
            PlanarOut po;
            po.wf = f.wf;
            if (group == 1) {
                po.hi = f.l_hi + (size_t)f.ml * 8;
                po.lo = f.l_lo + (size_t)f.ml * 8;
            } else {
                // hi plane = the buffer's two hi regions (2 * group images), lo plane = its two lo regions
                po.hi = f.l_hi + 2 * slot * img_elems + (size_t)f.ml * 8;
                po.lo = f.l_lo + 2 * slot * img_elems + (size_t)f.ml * 8;
            }
            STEREO_TRY(trunk_dev(ctx, net, padded_lane, 2, hp, wp, nullptr, precision, &po,
                                 from_u8 ? left_rgb + i * img_bytes : nullptr, from_u8 ? right_rgb + i * img_bytes : nullptr,
                                 2 * i, lane));
            if (group == 1) {
                STEREO_TRY(launch_corr_tc(ctx, f.l_hi, f.l_lo, f.r_hi, f.r_lo, 1, h, w, f.wf, f.ml, D, planes, lv, rv,
                                          left_disp + i * map, right_disp + i * map));
            } else if (slot == group - 1 || i == n - 1) {
                const int first = i - slot;
                STEREO_TRY(launch_corr_both(ctx, f.l_hi, f.l_lo, f.l_hi + img_elems, f.l_lo + img_elems, slot + 1, h, w, f.wf,
                                            f.ml, D, planes, left_disp + first * map, right_disp + first * map,
                                            2 * img_elems));
            }
        } else {
            STEREO_TRY(trunk_dev(ctx, net, padded_lane, 2, hp, wp, feat, precision));
            STEREO_TRY(corr_dev(ctx, feat, feat + feat_img, 1, h, w, 64, D, precision, lv, rv, left_disp + i * map,
                                right_disp + i * map));
        }
    }
    for (int l = 1; l < lanes; ++l) {
        STEREO_CUDA_CHECK(cudaEventRecord(ctx->lane_join[l - 1], lane_stream[l]));
        STEREO_CUDA_CHECK(cudaStreamWaitEvent(ctx->stream, ctx->lane_join[l - 1], 0));
    }
    if (lr_outlier) STEREO_TRY(launch_lr_check(ctx, left_disp, right_disp, n, h, w, D, lr_outlier));
    return STEREO_OK;
}

// ---- host staging helpers ----
static int h2d(stereo_ctx *ctx, DevBuf &buf, const void *src, size_t bytes)
{
    STEREO_TRY(buf.reserve(bytes));
    STEREO_CUDA_CHECK(cudaMemcpyAsync(buf.ptr, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return STEREO_OK;
}

static int d2h(stereo_ctx *ctx, void *dst, const void *src, size_t bytes)
{
    STEREO_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return STEREO_OK;
}

int finish(stereo_ctx *ctx)
{
    STEREO_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    return STEREO_OK;
}

}  // namespace stereo

using namespace stereo;

extern "C" {

const char *stereo_last_error(void) { return g_error; }

int stereo_version(void) { return 100; }

int stereo_create(int device, stereo_ctx **out)
{
    STEREO_REQUIRE(out != nullptr, "null output pointer");
    *out = nullptr;
    int count = 0;
    STEREO_CUDA_CHECK(cudaGetDeviceCount(&count));
    STEREO_REQUIRE(device >= 0 && device < count, "device %d out of range (%d visible)", device, count);
    STEREO_CUDA_CHECK(cudaSetDevice(device));
    cudaDeviceProp prop;
    STEREO_CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        set_error("device %d is sm_%d%d; this library contains sm_100a code only", device, prop.major, prop.minor);
        return STEREO_E_INVALID;
    }
    stereo_ctx *ctx = new (std::nothrow) stereo_ctx();
    if (!ctx) {
        set_error("out of host memory");
        return STEREO_E_NOMEM;
    }
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    cudaError_t err = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking);
    if (err != cudaSuccess) {
        set_error("cudaStreamCreate failed: %s", cudaGetErrorString(err));
        delete ctx;
        return STEREO_E_CUDA;
    }
    ctx->stream = ctx->own_stream;
    *out = ctx;
    return STEREO_OK;
}

int stereo_destroy(stereo_ctx *ctx)
{
    if (!ctx) return STEREO_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    DevBuf *bufs[] = {&ctx->weights[0], &ctx->weights[1], &ctx->wsplit[0], &ctx->wsplit[1], &ctx->stats,
                      &ctx->padded, &ctx->act[0], &ctx->act[1], &ctx->feat, &ctx->stage_in[0], &ctx->stage_in[1],
                      &ctx->stage_out[0], &ctx->stage_out[1], &ctx->stage_out[2], &ctx->stage_vol[0],
                      &ctx->stage_vol[1], &ctx->scratch, &ctx->post_vol[0], &ctx->post_vol[1], &ctx->post_vol[2], &ctx->post_vol[3], &ctx->post_vol[4], &ctx->post_vol[5], &ctx->post_vol[6], &ctx->post_vol[7],
                      &ctx->post_cnt[0], &ctx->post_cnt[1], &ctx->sgm_aux[0], &ctx->sgm_aux[1], &ctx->cbca_rec[0], &ctx->cbca_rec[1],
                      &ctx->post_maps, &ctx->pp_maps, &ctx->featp, &ctx->featp_b[0], &ctx->featp_b[1], &ctx->featp_b[2], &ctx->act_b[0],
                      &ctx->act_b[1], &ctx->act_b[2], &ctx->act_b[3], &ctx->act_b[4], &ctx->act_b[5]};
    for (DevBuf *b : bufs) b->release();
    if (ctx->prof.start) {
        for (int i = 0; i < Profiler::kMaxPairs; ++i) {
            cudaEventDestroy(ctx->prof.start[i]);
            cudaEventDestroy(ctx->prof.stop[i]);
        }
        delete[] ctx->prof.start;
        delete[] ctx->prof.stop;
        delete[] ctx->prof.cls;
    }
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    if (ctx->aux_stream) cudaStreamDestroy(ctx->aux_stream);
    for (int i = 0; i < 16; ++i) {
        if (ctx->ev_in[i]) cudaEventDestroy(ctx->ev_in[i]);
        if (ctx->ev_out[i]) cudaEventDestroy(ctx->ev_out[i]);
    }
    for (int i = 0; i < 3; ++i) {
        if (ctx->lane_join[i]) cudaEventDestroy(ctx->lane_join[i]);
        if (ctx->lane_streams[i]) cudaStreamDestroy(ctx->lane_streams[i]);
    }
    if (ctx->copy_in) cudaStreamDestroy(ctx->copy_in);
    if (ctx->copy_out) cudaStreamDestroy(ctx->copy_out);
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    delete ctx;
    return STEREO_OK;
}

int stereo_set_stream(stereo_ctx *ctx, void *cuda_stream)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    ctx->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
    return STEREO_OK;
}

int stereo_sync(stereo_ctx *ctx)
{
    STEREO_TRY(check_ctx(ctx));
    return finish(ctx);
}

int64_t stereo_launch_count(const stereo_ctx *ctx) { return ctx ? ctx->launches : 0; }

int stereo_profile_enable(stereo_ctx *ctx, int on)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    Profiler &p = ctx->prof;
    if (on && !p.start) {
        p.start = new (std::nothrow) cudaEvent_t[Profiler::kMaxPairs];
        p.stop = new (std::nothrow) cudaEvent_t[Profiler::kMaxPairs];
        p.cls = new (std::nothrow) int[Profiler::kMaxPairs];
        if (!p.start || !p.stop || !p.cls) {
            set_error("out of host memory");
            return STEREO_E_NOMEM;
        }
        for (int i = 0; i < Profiler::kMaxPairs; ++i) {
            STEREO_CUDA_CHECK(cudaEventCreate(&p.start[i]));
            STEREO_CUDA_CHECK(cudaEventCreate(&p.stop[i]));
        }
    }
    p.on = on != 0;
    p.used = 0;
    for (int k = 0; k < STEREO_K_COUNT; ++k) p.launches[k] = 0;
    return STEREO_OK;
}

int stereo_profile_read(stereo_ctx *ctx, int kernel_class, double *total_ms, int64_t *launches)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_REQUIRE(kernel_class >= 0 && kernel_class < STEREO_K_COUNT && total_ms && launches, "bad argument");
    STEREO_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    Profiler &p = ctx->prof;
    double sum = 0.0;
    int64_t timed = 0;
    for (int i = 0; i < p.used; ++i) {
        if (p.cls[i] != kernel_class) continue;
        float ms = 0.f;
        STEREO_CUDA_CHECK(cudaEventElapsedTime(&ms, p.start[i], p.stop[i]));
        sum += ms;
        ++timed;
    }
    // launches beyond the event pool are extrapolated from the timed ones
    int64_t all = p.launches[kernel_class];
    *total_ms = timed > 0 ? sum * ((double)all / (double)timed) : 0.0;
    *launches = all;
    return STEREO_OK;
}

const char *stereo_profile_name(int kernel_class)
{
    static const char *names[STEREO_K_COUNT] = {"prep", "conv_first", "conv", "corr", "wta", "lr_check",
                                                "cbca", "volume_std", "sgm", "post_maps"};
    return kernel_class >= 0 && kernel_class < STEREO_K_COUNT ? names[kernel_class] : "?";
}

int stereo_set_weights(stereo_ctx *ctx, int net, const float *packed, size_t n_floats)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_REQUIRE(net == STEREO_NET_WIN9 || net == STEREO_NET_WIN37, "unknown net %d", net);
    STEREO_REQUIRE(packed != nullptr, "null weight pointer");
    NetDesc nd = net_desc(net);
    STEREO_REQUIRE(n_floats == nd.n_floats, "net %d expects %zu packed floats, got %zu", net, nd.n_floats, n_floats);
    STEREO_TRY(ctx->weights[net].reserve(sizeof(float) * n_floats));
    STEREO_CUDA_CHECK(cudaMemcpyAsync(ctx->weights[net].ptr, packed, sizeof(float) * n_floats,
                                      cudaMemcpyHostToDevice, ctx->stream));
    if (net == STEREO_NET_WIN9) {
        // UMMA B-operand images (bf16 hi/lo) of the three 64->64 layers (conv_tc3.cu), then layer 1's (conv1_tc.cu)
        const size_t wl3 = tc3_weight_image_elems();
        STEREO_TRY(ctx->wsplit[net].reserve((3 * wl3 + conv1_tc_weight_image_elems()) * 2));
        STEREO_TRY(launch_conv1_weight_image(ctx, ctx->weights[net].as<float>() + nd.layer[0].w_off,
                                             ctx->wsplit[net].as<uint16_t>() + 3 * wl3));
        for (int l = 1; l < 4; ++l)
            STEREO_TRY(launch_weight_images3(ctx, ctx->weights[net].as<float>() + nd.layer[l].w_off,
                                             ctx->wsplit[net].as<uint16_t>() + (l - 1) * wl3));
    }
    if (net == STEREO_NET_WIN37) {
        // streaming units of conv_tcg.cu for layers 2-9, one after the other
        size_t total = 0;
        for (int l = 1; l < nd.n_layers; ++l) total += tcg_weight_image_elems(nd.layer[l].ks, nd.layer[l].cin, nd.layer[l].cout);
        // two images of the same size: per-tap units (conv_tcg.cu), then per-kx units (conv_tcr.cu)
        STEREO_TRY(ctx->wsplit[net].reserve(2 * total * 2));
        size_t off = 0;
        for (int l = 1; l < nd.n_layers; ++l) {
            const LayerDesc &L = nd.layer[l];
            STEREO_TRY(launch_weight_image_g(ctx, ctx->weights[net].as<float>() + L.w_off, L.ks, L.cin, L.cout,
                                             ctx->wsplit[net].as<uint16_t>() + off));
            STEREO_TRY(launch_weight_image_r(ctx, ctx->weights[net].as<float>() + L.w_off, L.ks, L.cin, L.cout,
                                             ctx->wsplit[net].as<uint16_t>() + total + off));
            off += tcg_weight_image_elems(L.ks, L.cin, L.cout);
        }
    }
    STEREO_TRY(finish(ctx));
    ctx->has_weights[net] = true;
    return STEREO_OK;
}

int stereo_normalize_pad_dev(stereo_ctx *ctx, const uint8_t *rgb, int n, int h, int w, int patch, float *out)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_dims(n, h, w));
    STEREO_REQUIRE(rgb && out, "null pointer");
    return launch_normalize_pad(ctx, rgb, n, h, w, patch, out);
}

int stereo_normalize_pad(stereo_ctx *ctx, const uint8_t *rgb, int n, int h, int w, int patch, float *out)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_dims(n, h, w));
    STEREO_REQUIRE(rgb && out, "null pointer");
    STEREO_REQUIRE(patch >= 1 && (patch & 1), "patch size must be odd and positive, got %d", patch);
    size_t in_b = (size_t)n * h * w * 3;
    size_t out_b = sizeof(float) * (size_t)n * (h + patch - 1) * (w + patch - 1) * 3;
    STEREO_TRY(h2d(ctx, ctx->stage_in[0], rgb, in_b));
    STEREO_TRY(ctx->stage_vol[0].reserve(out_b));
    STEREO_TRY(launch_normalize_pad(ctx, ctx->stage_in[0].as<uint8_t>(), n, h, w, patch, ctx->stage_vol[0].as<float>()));
    STEREO_TRY(d2h(ctx, out, ctx->stage_vol[0].ptr, out_b));
    return finish(ctx);
}

int stereo_trunk_dev(stereo_ctx *ctx, int net, const float *nhwc_in, int n, int h, int w, float *nhwc_out,
                     int precision)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_dims(n, h, w));
    STEREO_REQUIRE(nhwc_in && nhwc_out, "null pointer");
    return trunk_dev(ctx, net, nhwc_in, n, h, w, nhwc_out, precision);
}

int stereo_trunk(stereo_ctx *ctx, int net, const float *nhwc_in, int n, int h, int w, float *nhwc_out,
                 int precision)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_dims(n, h, w));
    STEREO_REQUIRE(nhwc_in && nhwc_out, "null pointer");
    STEREO_REQUIRE(net == STEREO_NET_WIN9 || net == STEREO_NET_WIN37, "unknown net %d", net);
    NetDesc nd = net_desc(net);
    STEREO_REQUIRE(h > nd.shrink && w > nd.shrink, "input %dx%d is too small for a receptive field of %d", h, w,
                   nd.shrink + 1);
    size_t in_b = sizeof(float) * (size_t)n * h * w * 3;
    size_t out_b = sizeof(float) * (size_t)n * (h - nd.shrink) * (w - nd.shrink) * 64;
    STEREO_TRY(h2d(ctx, ctx->stage_in[0], nhwc_in, in_b));
    STEREO_TRY(ctx->stage_vol[0].reserve(out_b));
    STEREO_TRY(trunk_dev(ctx, net, ctx->stage_in[0].as<float>(), n, h, w, ctx->stage_vol[0].as<float>(), precision));
    STEREO_TRY(d2h(ctx, nhwc_out, ctx->stage_vol[0].ptr, out_b));
    return finish(ctx);
}

static int corr_args(int n, int h, int w, int c, int D, int precision)
{
    STEREO_TRY(check_dims(n, h, w));
    STEREO_REQUIRE(c > 0 && D > 0, "channels and disp_range must be positive (got %d, %d)", c, D);
    STEREO_REQUIRE(precision == STEREO_PREC_FP32 || precision == STEREO_PREC_BF16X3 || precision == STEREO_PREC_BF16,
                   "unknown precision mode %d", precision);
    return STEREO_OK;
}

int stereo_cost_volume_dev(stereo_ctx *ctx, const float *lmap, const float *rmap, int n, int h, int w, int c,
                           int D, float *left_vol, float *right_vol, int precision)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(corr_args(n, h, w, c, D, precision));
    STEREO_REQUIRE(lmap && rmap && left_vol && right_vol, "null pointer");
    return corr_dev(ctx, lmap, rmap, n, h, w, c, D, precision, left_vol, right_vol, nullptr, nullptr);
}

int stereo_cost_volume(stereo_ctx *ctx, const float *lmap, const float *rmap, int n, int h, int w, int c, int D,
                       float *left_vol, float *right_vol, int precision)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(corr_args(n, h, w, c, D, precision));
    STEREO_REQUIRE(lmap && rmap && left_vol && right_vol, "null pointer");
    size_t map_b = sizeof(float) * (size_t)n * h * w * c;
    size_t vol_b = sizeof(float) * (size_t)n * h * w * D;
    STEREO_TRY(h2d(ctx, ctx->stage_in[0], lmap, map_b));
    STEREO_TRY(h2d(ctx, ctx->stage_in[1], rmap, map_b));
    STEREO_TRY(ctx->stage_vol[0].reserve(vol_b));
    STEREO_TRY(ctx->stage_vol[1].reserve(vol_b));
    STEREO_TRY(corr_dev(ctx, ctx->stage_in[0].as<float>(), ctx->stage_in[1].as<float>(), n, h, w, c, D, precision,
                        ctx->stage_vol[0].as<float>(), ctx->stage_vol[1].as<float>(), nullptr, nullptr));
    STEREO_TRY(d2h(ctx, left_vol, ctx->stage_vol[0].ptr, vol_b));
    STEREO_TRY(d2h(ctx, right_vol, ctx->stage_vol[1].ptr, vol_b));
    return finish(ctx);
}

int stereo_wta_dev(stereo_ctx *ctx, const float *vol, int n, int h, int w, int D, int take_min, uint8_t *disp)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_dims(n, h, w));
    STEREO_REQUIRE(D > 0 && vol && disp, "bad argument");
    STEREO_REQUIRE(D <= 256, "disparity maps are bytes; disp_range %d > 256", D);
    return launch_wta(ctx, vol, n, h, w, D, take_min, disp);
}

int stereo_wta(stereo_ctx *ctx, const float *vol, int n, int h, int w, int D, int take_min, uint8_t *disp)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_dims(n, h, w));
    STEREO_REQUIRE(D > 0 && vol && disp, "bad argument");
    STEREO_REQUIRE(D <= 256, "disparity maps are bytes; disp_range %d > 256", D);
    size_t vol_b = sizeof(float) * (size_t)n * h * w * D;
    size_t map_b = (size_t)n * h * w;
    STEREO_TRY(h2d(ctx, ctx->stage_vol[0], vol, vol_b));
    STEREO_TRY(ctx->stage_out[0].reserve(map_b));
    STEREO_TRY(launch_wta(ctx, ctx->stage_vol[0].as<float>(), n, h, w, D, take_min, ctx->stage_out[0].as<uint8_t>()));
    STEREO_TRY(d2h(ctx, disp, ctx->stage_out[0].ptr, map_b));
    return finish(ctx);
}

int stereo_corr_wta_dev(stereo_ctx *ctx, const float *lmap, const float *rmap, int n, int h, int w, int c, int D,
                        uint8_t *left_disp, uint8_t *right_disp, int precision)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(corr_args(n, h, w, c, D, precision));
    STEREO_REQUIRE(lmap && rmap && left_disp && right_disp, "null pointer");
    STEREO_REQUIRE(D <= 256, "disparity maps are bytes; disp_range %d > 256", D);
    return corr_dev(ctx, lmap, rmap, n, h, w, c, D, precision, nullptr, nullptr, left_disp, right_disp);
}

int stereo_corr_wta(stereo_ctx *ctx, const float *lmap, const float *rmap, int n, int h, int w, int c, int D,
                    uint8_t *left_disp, uint8_t *right_disp, int precision)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(corr_args(n, h, w, c, D, precision));
    STEREO_REQUIRE(lmap && rmap && left_disp && right_disp, "null pointer");
    STEREO_REQUIRE(D <= 256, "disparity maps are bytes; disp_range %d > 256", D);
    size_t map_b = sizeof(float) * (size_t)n * h * w * c;
    size_t out_b = (size_t)n * h * w;
    STEREO_TRY(h2d(ctx, ctx->stage_in[0], lmap, map_b));
    STEREO_TRY(h2d(ctx, ctx->stage_in[1], rmap, map_b));
    STEREO_TRY(ctx->stage_out[0].reserve(out_b));
    STEREO_TRY(ctx->stage_out[1].reserve(out_b));
    STEREO_TRY(corr_dev(ctx, ctx->stage_in[0].as<float>(), ctx->stage_in[1].as<float>(), n, h, w, c, D, precision,
                        nullptr, nullptr, ctx->stage_out[0].as<uint8_t>(), ctx->stage_out[1].as<uint8_t>()));
    STEREO_TRY(d2h(ctx, left_disp, ctx->stage_out[0].ptr, out_b));
    STEREO_TRY(d2h(ctx, right_disp, ctx->stage_out[1].ptr, out_b));
    return finish(ctx);
}

int stereo_lr_check_dev(stereo_ctx *ctx, const uint8_t *d0, const uint8_t *d1, int n, int h, int w, int D,
                        uint8_t *outlier)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_dims(n, h, w));
    STEREO_REQUIRE(D > 0 && d0 && d1 && outlier, "bad argument");
    return launch_lr_check(ctx, d0, d1, n, h, w, D, outlier);
}

int stereo_lr_check(stereo_ctx *ctx, const uint8_t *d0, const uint8_t *d1, int n, int h, int w, int D,
                    uint8_t *outlier)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_dims(n, h, w));
    STEREO_REQUIRE(D > 0 && d0 && d1 && outlier, "bad argument");
    size_t b = (size_t)n * h * w;
    STEREO_TRY(h2d(ctx, ctx->stage_in[0], d0, b));
    STEREO_TRY(h2d(ctx, ctx->stage_in[1], d1, b));
    STEREO_TRY(ctx->stage_out[0].reserve(b));
    STEREO_TRY(launch_lr_check(ctx, ctx->stage_in[0].as<uint8_t>(), ctx->stage_in[1].as<uint8_t>(), n, h, w, D,
                               ctx->stage_out[0].as<uint8_t>()));
    STEREO_TRY(d2h(ctx, outlier, ctx->stage_out[0].ptr, b));
    return finish(ctx);
}

int stereo_pairs_dev(stereo_ctx *ctx, int net, const uint8_t *left_rgb, const uint8_t *right_rgb, int n, int h,
                     int w, int D, int precision, uint8_t *left_disp, uint8_t *right_disp, uint8_t *lr_outlier,
                     float *left_vol, float *right_vol)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_dims(n, h, w));
    return pairs_dev(ctx, net, left_rgb, right_rgb, n, h, w, D, precision, left_disp, right_disp, lr_outlier,
                     left_vol, right_vol);
}

int stereo_pairs(stereo_ctx *ctx, int net, const uint8_t *left_rgb, const uint8_t *right_rgb, int n, int h, int w,
                 int D, int precision, uint8_t *left_disp, uint8_t *right_disp, uint8_t *lr_outlier,
                 float *left_vol, float *right_vol)
{
    STEREO_TRY(check_ctx(ctx));
    STEREO_TRY(check_dims(n, h, w));
    STEREO_REQUIRE(left_rgb && right_rgb && left_disp && right_disp, "null image or disparity pointer");
    STEREO_REQUIRE((left_vol == nullptr) == (right_vol == nullptr), "left and right volumes must be given together");
    size_t img_b = (size_t)n * h * w * 3, map_b = (size_t)n * h * w;
    size_t vol_b = sizeof(float) * (size_t)n * h * w * D;
    if (!left_vol && n > 1) {
        // Maps only: cut the batch into up to 16 chunks and run upload, compute and download of neighbouring
        // chunks at the same time (copy engines on their own streams, ordered by events).  The staging
        // buffers hold the whole batch, so chunks never alias.
        STEREO_TRY(ctx->stage_in[0].reserve(img_b));
        STEREO_TRY(ctx->stage_in[1].reserve(img_b));
        for (int i = 0; i < 3; ++i) STEREO_TRY(ctx->stage_out[i].reserve(map_b));
        if (!ctx->copy_in) {
            STEREO_CUDA_CHECK(cudaStreamCreateWithFlags(&ctx->copy_in, cudaStreamNonBlocking));
            STEREO_CUDA_CHECK(cudaStreamCreateWithFlags(&ctx->copy_out, cudaStreamNonBlocking));
            for (int i = 0; i < 16; ++i) {
                STEREO_CUDA_CHECK(cudaEventCreateWithFlags(&ctx->ev_in[i], cudaEventDisableTiming));
                STEREO_CUDA_CHECK(cudaEventCreateWithFlags(&ctx->ev_out[i], cudaEventDisableTiming));
            }
        }
        const int chunk = (n + 15) / 16;
        const size_t img1 = (size_t)h * w * 3, map1 = (size_t)h * w;
        uint8_t *d_l = ctx->stage_in[0].as<uint8_t>(), *d_r = ctx->stage_in[1].as<uint8_t>();
        uint8_t *d_ld = ctx->stage_out[0].as<uint8_t>(), *d_rd = ctx->stage_out[1].as<uint8_t>();
        uint8_t *d_lr = ctx->stage_out[2].as<uint8_t>();
        // uploads start once everything queued on the compute stream so far (which may still read the
        // staging buffers) is done
        STEREO_CUDA_CHECK(cudaEventRecord(ctx->ev_out[0], ctx->stream));
        STEREO_CUDA_CHECK(cudaStreamWaitEvent(ctx->copy_in, ctx->ev_out[0], 0));
        int k = 0;
        for (int i0 = 0; i0 < n; i0 += chunk, ++k) {
            const int m = n - i0 < chunk ? n - i0 : chunk;
            STEREO_CUDA_CHECK(cudaMemcpyAsync(d_l + i0 * img1, left_rgb + i0 * img1, m * img1, cudaMemcpyHostToDevice, ctx->copy_in));
            STEREO_CUDA_CHECK(cudaMemcpyAsync(d_r + i0 * img1, right_rgb + i0 * img1, m * img1, cudaMemcpyHostToDevice, ctx->copy_in));
            STEREO_CUDA_CHECK(cudaEventRecord(ctx->ev_in[k], ctx->copy_in));
        }
        k = 0;
        for (int i0 = 0; i0 < n; i0 += chunk, ++k) {
            const int m = n - i0 < chunk ? n - i0 : chunk;
            STEREO_CUDA_CHECK(cudaStreamWaitEvent(ctx->stream, ctx->ev_in[k], 0));
            STEREO_TRY(pairs_dev(ctx, net, d_l + i0 * img1, d_r + i0 * img1, m, h, w, D, precision, d_ld + i0 * map1,
                                 d_rd + i0 * map1, lr_outlier ? d_lr + i0 * map1 : nullptr, nullptr, nullptr));
            STEREO_CUDA_CHECK(cudaEventRecord(ctx->ev_out[k], ctx->stream));
            STEREO_CUDA_CHECK(cudaStreamWaitEvent(ctx->copy_out, ctx->ev_out[k], 0));
            STEREO_CUDA_CHECK(cudaMemcpyAsync(left_disp + i0 * map1, d_ld + i0 * map1, m * map1, cudaMemcpyDeviceToHost, ctx->copy_out));
            STEREO_CUDA_CHECK(cudaMemcpyAsync(right_disp + i0 * map1, d_rd + i0 * map1, m * map1, cudaMemcpyDeviceToHost, ctx->copy_out));
            if (lr_outlier)
                STEREO_CUDA_CHECK(cudaMemcpyAsync(lr_outlier + i0 * map1, d_lr + i0 * map1, m * map1, cudaMemcpyDeviceToHost, ctx->copy_out));
        }
        STEREO_CUDA_CHECK(cudaStreamSynchronize(ctx->copy_out));
        return finish(ctx);
    }
    STEREO_TRY(h2d(ctx, ctx->stage_in[0], left_rgb, img_b));
    STEREO_TRY(h2d(ctx, ctx->stage_in[1], right_rgb, img_b));
    for (int i = 0; i < 3; ++i) STEREO_TRY(ctx->stage_out[i].reserve(map_b));
    if (left_vol) {
        STEREO_TRY(ctx->stage_vol[0].reserve(vol_b));
        STEREO_TRY(ctx->stage_vol[1].reserve(vol_b));
    }
    STEREO_TRY(pairs_dev(ctx, net, ctx->stage_in[0].as<uint8_t>(), ctx->stage_in[1].as<uint8_t>(), n, h, w, D,
                         precision, ctx->stage_out[0].as<uint8_t>(), ctx->stage_out[1].as<uint8_t>(),
                         lr_outlier ? ctx->stage_out[2].as<uint8_t>() : nullptr,
                         left_vol ? ctx->stage_vol[0].as<float>() : nullptr,
                         left_vol ? ctx->stage_vol[1].as<float>() : nullptr));
    STEREO_TRY(d2h(ctx, left_disp, ctx->stage_out[0].ptr, map_b));
    STEREO_TRY(d2h(ctx, right_disp, ctx->stage_out[1].ptr, map_b));
    if (lr_outlier) STEREO_TRY(d2h(ctx, lr_outlier, ctx->stage_out[2].ptr, map_b));
    if (left_vol) {
        STEREO_TRY(d2h(ctx, left_vol, ctx->stage_vol[0].ptr, vol_b));
        STEREO_TRY(d2h(ctx, right_vol, ctx->stage_vol[1].ptr, vol_b));
    }
    return finish(ctx);
}

}  // extern "C"
