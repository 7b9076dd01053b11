// Shared declarations of libstereo_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/stereo_b200.h"

namespace stereo {

void set_error(const char *fmt, ...);

#define STEREO_CUDA_CHECK(expr)                                                              \
    do {                                                                                     \
        cudaError_t err__ = (expr);                                                          \
        if (err__ != cudaSuccess) {                                                          \
            stereo::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(err__),     \
                              __FILE__, __LINE__);                                           \
            return err__ == cudaErrorMemoryAllocation ? STEREO_E_NOMEM : STEREO_E_CUDA;      \
        }                                                                                    \
    } while (0)

#define STEREO_TRY(expr)              \
    do {                              \
        int rc__ = (expr);            \
        if (rc__ != STEREO_OK) return rc__; \
    } while (0)

#define STEREO_REQUIRE(cond, ...)          \
    do {                                   \
        if (!(cond)) {                     \
            stereo::set_error(__VA_ARGS__); \
            return STEREO_E_INVALID;       \
        }                                  \
    } while (0)

// Device buffer that only ever grows; owned by a context.
struct DevBuf {
    void *ptr = nullptr;
    size_t bytes = 0;
    int reserve(size_t want);
    void release();
    template <typename T> T *as() const { return static_cast<T *>(ptr); }
};

constexpr int kMaxLayers = 9;

struct LayerDesc {
    int ks, cin, cout, relu;
    size_t w_off, b_off;  // float offsets into the packed weight buffer
};

struct NetDesc {
    int n_layers = 0;
    int shrink = 0;  // total border lost by the valid convolutions
    LayerDesc layer[kMaxLayers];
    size_t n_floats = 0;
};

NetDesc net_desc(int net);

// exact integer moments of a uint8 image and the 256-entry normalisation table built from them
struct ImageStats {
    unsigned long long sum;
    unsigned long long sumsq;
    float lut[256];
};

// CUDA-event timing of kernel classes (stereo_profile_*).
struct Profiler {
    static constexpr int kMaxPairs = 8192;
    bool on = false;
    int used = 0;
    cudaEvent_t *start = nullptr, *stop = nullptr;
    int *cls = nullptr;
    int64_t launches[STEREO_K_COUNT] = {};
};

}  // namespace stereo

struct stereo_ctx {
    int device = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    int sm_count = 148;
    int64_t launches = 0;
    stereo::Profiler prof;
    bool has_weights[2] = {false, false};
    stereo::DevBuf weights[2];   // packed fp32, layout of stereo_set_weights
    stereo::DevBuf wsplit[2];    // tensor-core operand images of the weights (bf16 hi/lo)
    stereo::DevBuf stats;        // per-image u64 sums + 256-entry LUTs
    stereo::DevBuf padded;       // normalised, padded images
    stereo::DevBuf act[2];       // ping-pong activations
    stereo::DevBuf act_b[6];     // the same for lanes 1..3 of stereo_pairs (pairs alternate between streams)
    stereo::DevBuf feat;         // trunk outputs of one pair
    stereo::DevBuf stage_in[2];  // H2D staging of the host-pointer entry points
    stereo::DevBuf stage_out[3];
    stereo::DevBuf stage_vol[2];
    stereo::DevBuf featp;        // chunk-planar bf16 (hi, lo) features with zero margins
    int featp_geom[5] = {0, 0, 0, 0, 0};
    stereo::DevBuf featp_b[3];   // lanes 1..3
    int featp_b_geom[3][5] = {};
    cudaStream_t lane_streams[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t lane_join[3] = {nullptr, nullptr, nullptr};
    stereo::DevBuf scratch;      // per-block partials of the volume reduction
    stereo::DevBuf post_vol[8];  // working volumes of the post chain: 4 per side (the sides run on two streams)
    stereo::DevBuf cbca_rec[2];  // packed relative arms of the fused CBCA kernel, per side
    stereo::DevBuf sgm_aux[2];   // classified guide differences of the SGM ring kernels, per side
    stereo::DevBuf post_cnt[2];  // row-segment lengths of the separable CBCA (uint8 per voxel), per side
    cudaStream_t aux_stream = nullptr;   // second stream of the post chain
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    stereo::DevBuf post_maps;    // guide images, arms, float maps, std scalars
    stereo::DevBuf pp_maps;      // stereo_pairs_post: WTA maps and stage images of the pair in flight
    // host-pointer stereo_pairs: H2D and D2H run on their own streams, chunk by chunk, under the compute
    cudaStream_t copy_in = nullptr, copy_out = nullptr;
    cudaEvent_t ev_in[16] = {}, ev_out[16] = {};
};

namespace stereo {

// Brackets the launches issued during its lifetime with a start/stop event pair.
struct ProfScope {
    stereo_ctx *ctx;
    int slot = -1;
    ProfScope(stereo_ctx *c, int kernel_class) : ctx(c)
    {
        Profiler &p = ctx->prof;
        if (!p.on) return;
        p.launches[kernel_class] += 1;
        if (p.used >= Profiler::kMaxPairs) return;
        slot = p.used++;
        p.cls[slot] = kernel_class;
        cudaEventRecord(p.start[slot], ctx->stream);
    }
    ~ProfScope()
    {
        if (slot >= 0) cudaEventRecord(ctx->prof.stop[slot], ctx->stream);
    }
};

// prep.cu
int launch_normalize_pad(stereo_ctx *ctx, const uint8_t *rgb, int n, int h, int w, int patch, float *out);
int launch_pad_with_stats(stereo_ctx *ctx, const uint8_t *rgb, int h, int w, int patch, int stats_index, float *out);
int launch_pair_stats(stereo_ctx *ctx, const uint8_t *left, const uint8_t *right, int n, int h, int w);
// conv1_tc.cu
size_t conv1_tc_weight_image_elems();
int launch_conv1_weight_image(stereo_ctx *ctx, const float *layer_w, void *img);
int launch_conv1_tc(stereo_ctx *ctx, const uint8_t *img0, const uint8_t *img1, int h, int w, const void *wimg,
                    const float *bias, void *out_hi, void *out_lo, int wp_out, int stats_off);
// conv_ffma.cu
int launch_conv_ffma(stereo_ctx *ctx, const LayerDesc &L, const float *weights, const float *in,
                     int n, int h_in, int w_in, float *out);
int launch_conv_first_split(stereo_ctx *ctx, const LayerDesc &L, const float *weights, const float *in, int n,
                            int h_in, int w_in, void *out_hi, void *out_lo, int wp_out);
// conv_tcg.cu
size_t tcg_weight_image_elems(int ks, int cin, int cout);
int tcg_pitch(int w);
int launch_weight_image_g(stereo_ctx *ctx, const float *layer_w, int ks, int cin, int cout, void *img);
int launch_nhwc_to_planar_g(stereo_ctx *ctx, const float *in, int n, int h, int w, int c, int wp, void *hi, void *lo);
int launch_conv_tcg(stereo_ctx *ctx, int ks, int cin, int cout, const void *in_hi, const void *in_lo, int n, int h_in,
                    int w_in, int wp_in, const void *wimg, const float *bias, void *out_hi, void *out_lo, int wp_out,
                    float *out_f32, int relu, int planes);
// conv_tcr.cu (every input row multiplied once; same contract, its own weight image)
int launch_weight_image_r(stereo_ctx *ctx, const float *layer_w, int ks, int cin, int cout, void *img);
int launch_conv_tcr(stereo_ctx *ctx, int ks, int cin, int cout, const void *in_hi, const void *in_lo, int n, int h_in,
                    int w_in, int wp_in, const void *wimg, const float *bias, void *out_hi, void *out_lo, int wp_out,
                    float *out_f32, int relu, int planes);
int corr_sweep_margin();
// corr_both.cu: both disparity maps from ONE sweep per row (fused WTA only).  img_stride: elements between consecutive
// images of one side (0: the images of a side are contiguous); corr_both_enabled(): launch_corr_tc would pick it.
int launch_corr_both(stereo_ctx *ctx, const void *l_hi, const void *l_lo, const void *r_hi, const void *r_lo, int n,
                     int h, int w, int wf, int ml, int D, int planes, uint8_t *left_disp, uint8_t *right_disp,
                     size_t img_stride = 0);
bool corr_both_enabled();
// train_wgrad_tc.cu: weight gradient of a 3x3 / 64 -> 64 layer on the tensor cores, accumulated into red_w (double)
size_t train_wgrad_tc_scratch_bytes(const stereo_ctx *ctx);
int launch_train_wgrad_tc(stereo_ctx *ctx, const float *x, const float *dz, int n, int h, int w, float *scratch,
                          double *red_w);
int launch_corr_sweep(stereo_ctx *ctx, const void *l_hi, const void *l_lo, const void *r_hi, const void *r_lo, int n,
                      int h, int w, int wf, int ml, int D, int planes, float *left_vol, float *right_vol,
                      uint8_t *left_disp, uint8_t *right_disp);
int tc3_pitch(int w);
size_t tc3_weight_image_elems();
int launch_weight_images3(stereo_ctx *ctx, const float *layer_w, void *img);
int launch_conv_tc3(stereo_ctx *ctx, const void *in_hi, const void *in_lo, int n, int h_in, int w_in, int wp_in,
                    const void *wimg, const float *bias, void *out_hi, void *out_lo, int wp_out, float *out_f32,
                    int relu, int planes);
// corr_tc.cu
int corr_tc_max_disp();
void corr_tc_geometry(int w, int D, int *wf, int *ml);
int launch_nhwc_to_planar(stereo_ctx *ctx, const float *in, int n, int h, int w, int wf, int ml, void *hi, void *lo);
int launch_corr_tc(stereo_ctx *ctx, const void *l_hi, const void *l_lo, const void *r_hi, const void *r_lo, int n,
                   int h, int w, int wf, int ml, int D, int planes, float *left_vol, float *right_vol,
                   uint8_t *left_disp, uint8_t *right_disp);
// corr.cu
int launch_corr_ffma(stereo_ctx *ctx, const float *lmap, const float *rmap, int n, int h, int w, int c,
                     int D, float *left_vol, float *right_vol, uint8_t *left_disp, uint8_t *right_disp);
int launch_wta(stereo_ctx *ctx, const float *vol, int n, int h, int w, int D, int take_min, uint8_t *disp);
int wta_moments_max_blocks(const stereo_ctx *ctx);
int launch_wta_moments(stereo_ctx *ctx, const float *vol, size_t npix, int D, int take_min, uint8_t *disp,
                       double *partial, int *nblocks);
int launch_lr_check(stereo_ctx *ctx, const uint8_t *d0, const uint8_t *d1, int n, int h, int w, int D,
                    uint8_t *outlier);
// post.cu
struct SgmParams {
    float pi1, pi2, tau_so, alpha1, q1, q2;
};
int launch_guide(stereo_ctx *ctx, const uint8_t *img, int h, int w, float *out);
int launch_cross(stereo_ctx *ctx, const float *img, int h, int w, int L1, float tau1, short *arms);
int launch_cbca(stereo_ctx *ctx, const short *x0c, const short *x1c, const float *vol, float *out, int h, int w,
                int D, int direction);
int launch_cbca_separable(stereo_ctx *ctx, const short *x0c, const short *x1c, const float *vol, float *hsum,
                          uint8_t *hcnt, float *out, int h, int w, int D, int direction, int L1);
// cbca_fused.cu
// in_pitch: floats between the [d] rows of two neighbouring pixels of `vol` (0: D; the output always has pitch D)
int launch_cbca_fused(stereo_ctx *ctx, const short *x0c, const short *x1c, const float *vol, float *out, int h, int w,
                      int D, int direction, int L1, int in_pitch = 0);
int launch_volume_std(stereo_ctx *ctx, const float *vol, size_t n, float *result, int slot = 0);
// the same with the arg max over D of every [x][d] row written to `wta` (uint8 [h*w]; D <= 256) in the same pass
int launch_volume_std_wta(stereo_ctx *ctx, const float *vol, size_t n_pixels, int D, float *result, uint8_t *wta, int slot);
int launch_box_agg(stereo_ctx *ctx, const float *in, int h, int w, int D, int kw, int kh, float *out);
int launch_scale(stereo_ctx *ctx, const float *in, size_t n, const float *stdp, int mode, float *out);
int launch_scale_pad(stereo_ctx *ctx, const float *in, size_t n_pixels, int D, int Dp, const float *stdp, int mode,
                     float *out);
// What the SGM kernels fuse in besides the four paths (all optional):
//   final_scale  factor of the last path's sum: 0.25 (nyu.lua:61,64 `/ 4`), 0.125 after eight paths
//   wta          uint8 [h][w]: arg min over D of the final result (the nyu_sgm stage image, post.lua:159); D <= 256
//   paths        4 (the reference) or 8 (adds the diagonals; no counterpart in the reference)
struct SgmExtras {
    float final_scale = 0.25f;
    uint8_t *wta = nullptr;
    int paths = 4;
};
// sgm_ring.cu: the four paths of the original (-> <- down up); with ex.paths == 8 the sum is left unscaled
int launch_sgm_ring(stereo_ctx *ctx, const float *x0, const float *x1, const float *cvol, float *out, float *scratch,
                    int h, int w, int D, const SgmParams &p, int direction, const SgmExtras &ex);
int launch_sgm(stereo_ctx *ctx, const float *x0, const float *x1, const float *cvol, float *out, float *scratch,
               int h, int w, int D, const SgmParams &p, int direction, const SgmExtras &ex = SgmExtras());
int launch_outlier(stereo_ctx *ctx, const float *d0, const float *d1, int h, int w, int disp_max, float *outlier);
int launch_interpolate(stereo_ctx *ctx, const float *d0, const float *outlier, int h, int w, float *tmp, float *out);
int launch_subpixel(stereo_ctx *ctx, const float *d0, const float *vol, int h, int w, int disp_max, float *out);
int launch_median(stereo_ctx *ctx, const float *img, int h, int w, int kernel_size, float *out);
int launch_mean2d(stereo_ctx *ctx, const float *img, const float *kernel, int h, int w, int radius, float alpha2,
                  float *out);
int launch_u8_to_float(stereo_ctx *ctx, const uint8_t *in, int n, float *out);
int launch_float_to_u8(stereo_ctx *ctx, const float *in, int n, float scale, uint8_t *out);
// api.cu helpers shared with post_api.cu
int check_ctx(stereo_ctx *ctx);
int finish(stereo_ctx *ctx);

}  // namespace stereo
