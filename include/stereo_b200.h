/*
 * stereo_b200.h - C ABI of libstereo_b200.so, the sm_100a implementation of the
 * Luo/Schwing/Urtasun stereo hot path of datvuthanh/Stereo-Matching.
 *
 * The reference has no FFI of its own: its hot path is Python calling TensorFlow
 * ops (inference.py) and Lua calling the external mc-cnn CUDA library
 * (post.lua / smooth/nyu.lua).  Each entry point below names the reference
 * lines it replaces.  The Python package `stereo_matching_b200` binds these
 * with ctypes and mirrors the reference's own surface on top (see
 * INTEGRATION.md).
 *
 * Conventions
 *   - every function returns 0 on success and a negative STEREO_E_* code on
 *     failure; stereo_last_error() returns a message for the calling thread.
 *     Nothing throws across the boundary.
 *   - a stereo_ctx is bound to one CUDA device and one stream.  It is not
 *     thread-safe; use one context per GPU / host thread.
 *   - functions without a suffix take HOST pointers, copy in and out and
 *     return after the result is in the caller's buffer.  `_dev` variants take
 *     DEVICE pointers on the context's device, enqueue work on the context's
 *     stream and return without synchronising (call stereo_sync).
 *   - images are NHWC, volumes are [n][H][W][D] float32 with D fastest, exactly
 *     the layout inference.py writes to its .bin files (inference.py:210-231).
 */
#ifndef STEREO_B200_H
#define STEREO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct stereo_ctx stereo_ctx;

enum {
    STEREO_OK = 0,
    STEREO_E_INVALID = -1, /* bad argument (null pointer, non-positive size, unknown enum) */
    STEREO_E_CUDA = -2,    /* a CUDA runtime call or kernel failed */
    STEREO_E_STATE = -3,   /* weights for the requested net have not been set */
    STEREO_E_NOMEM = -4,   /* device allocation failed */
    STEREO_E_LIMIT = -5    /* a size exceeds what a kernel supports */
};

/* net_type of inference.py:46,81-84 */
enum { STEREO_NET_WIN9 = 0 /* base_model_ws_9 */, STEREO_NET_WIN37 = 1 /* base_model */ };

/* arithmetic of the trunk / correlation contractions */
enum {
    STEREO_PREC_FP32 = 0,   /* fp32 FMA on CUDA cores */
    STEREO_PREC_BF16X3 = 1, /* tcgen05 tensor cores, 3-pass bf16 split, fp32 accumulate (parity mode) */
    STEREO_PREC_BF16 = 2    /* tcgen05 tensor cores, single bf16 pass (does not meet the 1e-4 gate) */
};

const char *stereo_last_error(void);
int stereo_version(void);

/* One context per GPU.  Replaces the implicit TF / torch7 device state. */
int stereo_create(int device, stereo_ctx **out);
int stereo_destroy(stereo_ctx *ctx);
/* Run on a caller-owned cudaStream_t (pass NULL to go back to the context's own stream). */
int stereo_set_stream(stereo_ctx *ctx, void *cuda_stream);
int stereo_sync(stereo_ctx *ctx);
/* Number of kernels this context has launched since creation (bench bookkeeping). */
int64_t stereo_launch_count(const stereo_ctx *ctx);

/* Weights of one trunk, replacing `ckpt.restore(manager.latest_checkpoint)`
 * (inference.py:88-91).  `packed` holds, for every conv layer in order, the
 * kernel in HWIO order [k][k][Cin][Cout] followed by the bias [Cout]; the
 * inference-mode BatchNormalization of models/model.py has already been folded
 * into them by the host (stereo_matching_b200/models/model.py).
 * WIN9: 4 layers 3x3, 3->64->64->64->64.  WIN37: 9 layers 5x5,
 * 3->32->32->32->64->64->64->64->64->64. */
int stereo_set_weights(stereo_ctx *ctx, int net, const float *packed, size_t n_floats);

/* inference.py:109-133: per image (x - mean) / std over all pixels and
 * channels (population std, float64), zero padded by (patch-1)/2 on each side.
 * rgb: uint8 [n][h][w][3] -> out: float32 [n][h+patch-1][w+patch-1][3]. */
int stereo_normalize_pad(stereo_ctx *ctx, const uint8_t *rgb, int n, int h, int w, int patch,
                         float *out);
int stereo_normalize_pad_dev(stereo_ctx *ctx, const uint8_t *rgb, int n, int h, int w, int patch,
                             float *out);

/* `model(x, training=False)` of inference.py:136-137 / models/model.py:5-41.
 * in: float32 [n][h][w][3] -> out: float32 [n][h-8][w-8][64] (WIN9) or
 * [n][h-36][w-36][64] (WIN37). */
int stereo_trunk(stereo_ctx *ctx, int net, const float *nhwc_in, int n, int h, int w,
                 float *nhwc_out, int precision);
int stereo_trunk_dev(stereo_ctx *ctx, int net, const float *nhwc_in, int n, int h, int w,
                     float *nhwc_out, int precision);

/* The disparity loop of inference.py:139-150, materialised.
 * lmap, rmap: float32 [n][h][w][c];  left_vol, right_vol: float32 [n][h][w][D]
 *   left [y][x][d] = <L[y][x], R[y][x-d]>  for x >= d,   else exactly 0
 *   right[y][x][d] = <L[y][x+d], R[y][x]>  for x+d < w,  else exactly 0 */
int stereo_cost_volume(stereo_ctx *ctx, const float *lmap, const float *rmap, int n, int h, int w,
                       int c, int D, float *left_vol, float *right_vol, int precision);
int stereo_cost_volume_dev(stereo_ctx *ctx, const float *lmap, const float *rmap, int n, int h,
                           int w, int c, int D, float *left_vol, float *right_vol, int precision);

/* tf.argmax(cost_volume, axis=2) -> uint8 of inference.py:169-183: arg max over
 * all D slots (zero-filled ones included), lowest index wins ties.
 * With take_min != 0 it is the `vol:min(2)` of post.lua:159,178,226.
 * vol: float32 [n][h][w][D] -> disp: uint8 [n][h][w]. */
int stereo_wta(stereo_ctx *ctx, const float *vol, int n, int h, int w, int D, int take_min,
               uint8_t *disp);
int stereo_wta_dev(stereo_ctx *ctx, const float *vol, int n, int h, int w, int D, int take_min,
                   uint8_t *disp);

/* inference.py:139-183 without ever writing the volumes: correlation and WTA
 * fused.  lmap, rmap: float32 [n][h][w][c] -> left_disp, right_disp: uint8 [n][h][w]. */
int stereo_corr_wta(stereo_ctx *ctx, const float *lmap, const float *rmap, int n, int h, int w,
                    int c, int D, uint8_t *left_disp, uint8_t *right_disp, int precision);
int stereo_corr_wta_dev(stereo_ctx *ctx, const float *lmap, const float *rmap, int n, int h, int w,
                        int c, int D, uint8_t *left_disp, uint8_t *right_disp, int precision);

/* Left-right consistency check = adcensus.outlier_detection (smooth/nyu.lua:95):
 * 0 match, 1 occlusion, 2 mismatch.  d0, d1: uint8 [n][h][w] -> outlier uint8 [n][h][w]. */
int stereo_lr_check(stereo_ctx *ctx, const uint8_t *d0, const uint8_t *d1, int n, int h, int w,
                    int D, uint8_t *outlier);
int stereo_lr_check_dev(stereo_ctx *ctx, const uint8_t *d0, const uint8_t *d1, int n, int h, int w,
                        int D, uint8_t *outlier);

/* One iteration of the per-pair loop inference.py:99-183 for n pairs: normalise,
 * pad, both trunk forwards, fused correlation + WTA, LR check.
 * left_rgb, right_rgb: uint8 [n][h][w][3];  outputs uint8 [n][h][w];
 * lr_outlier may be NULL.  If left_vol/right_vol are non-NULL the volumes
 * (float32 [n][h][w][D]) are materialised as well, as inference.py:210-231 writes them. */
int stereo_pairs(stereo_ctx *ctx, int net, const uint8_t *left_rgb, const uint8_t *right_rgb, int n,
                 int h, int w, int D, int precision, uint8_t *left_disp, uint8_t *right_disp,
                 uint8_t *lr_outlier, float *left_vol, float *right_vol);
int stereo_pairs_dev(stereo_ctx *ctx, int net, const uint8_t *left_rgb, const uint8_t *right_rgb,
                     int n, int h, int w, int D, int precision, uint8_t *left_disp,
                     uint8_t *right_disp, uint8_t *lr_outlier, float *left_vol, float *right_vol);

/* ---- post.lua / smooth/nyu.lua ------------------------------------------------------------
 * The reference reaches these kernels through the external mc-cnn library
 * (smooth/init.lua:6).  All volumes here are float32 [h][w][D] (the .bin layout); the
 * reference's [1][D][h][w] views are a host-side transpose away. */
typedef struct stereo_post_params {
    int L1;          /* smooth/nyu.lua:3  cross arm length limit */
    float tau1;      /* :4  cross intensity threshold */
    int sgm_i;       /* :7  SGM sweeps */
    float pi1, pi2, tau_so, alpha1, sgm_q1, sgm_q2; /* :9-14 */
    int occlusion, subpixel, median, bilateral;     /* :17-20 */
    double blur_sigma;                              /* :21  (a Lua double; the Gaussian of nyu.lua:77-89 is built in double) */
    float blur_t;                                   /* :22 */
    int nyu_cost_agg_1, sgm, nyu_cost_agg_2, post;  /* post.lua:37-40 */
    int cbca_exact;  /* 0: separable CBCA (row sums first; same sum, last-bit differences);
                        1: the original's strictly sequential summation order (slower) */
    int sgm_paths;   /* 4: the reference (nyu.lua:59-64: -> <- down up, sum / 4).  8: the four diagonal paths
                        are added (same step, no alpha1 scaling) and the sum is divided by 8 - an extension
                        named in BASELINE.json configs[3] with NO counterpart in the reference; results of
                        an 8-path chain are never parity figures */
} stereo_post_params;
/* the constants of smooth/nyu.lua:3-22 and the option defaults of post.lua:10-49 */
int stereo_post_default_params(stereo_post_params *p);

/* post.lua:91-94: byte image -> (x - mean) / std with torch's unbiased std. u8 [h][w] -> f32 [h][w] */
int stereo_guide_image(stereo_ctx *ctx, const uint8_t *img, int h, int w, float *out);

/* smooth.nyu.cross_agg(l_img, r_img, left_vol, right_vol, iter), smooth/nyu.lua:24-45
 * (adcensus.cross + adcensus.cbca).  Guide images f32 [h][w]; volumes updated in place. */
int stereo_nyu_cross_agg(stereo_ctx *ctx, const float *l_img, const float *r_img, float *left_vol,
                         float *right_vol, int h, int w, int D, int iters, const stereo_post_params *p);

/* smooth.nyu.sgm(l_img, r_img, vol, direction), smooth/nyu.lua:47-75 (adcensus.sgm2, 4 paths,
 * result = accumulated path costs / 4; p->sgm_paths == 8: see stereo_post_params).  direction -1 for
 * the left volume, +1 for the right.  D <= 512.  When D % 4 != 0 the kernels fetch the 16-byte
 * aligned superset of every [x][d] run: up to 12 bytes before the first and after the last voxel of
 * the (library-owned) device copy are read, never used; the host array is copied, not touched. */
int stereo_nyu_sgm(stereo_ctx *ctx, const float *l_img, const float *r_img, float *vol, int h, int w,
                   int D, int direction, const stereo_post_params *p);

/* smooth.nyu.post(disp, left_vol), smooth/nyu.lua:91-115: outlier_detection,
 * interpolate_occlusion, interpolate_mismatch, subpixel_enchancement, median2d(5),
 * mean2d(gaussian(blur_sigma), blur_t).  disp maps f32 [h][w] (0-based disparities);
 * final f32 [h][w]; outlier f32 [h][w] in {0,1,2}. */
int stereo_nyu_post(stereo_ctx *ctx, const float *disp_left, const float *disp_right,
                    const float *left_vol, int h, int w, int D, const stereo_post_params *p,
                    float *final_disp, float *outlier);

/* post.lua:146-147 `vol:std()`: unbiased std over the whole volume. */
int stereo_volume_std(stereo_ctx *ctx, const float *vol, size_t n, float *std_out);

/* No counterpart in the reference.  The unbiased std of `stereo_volume_std` for callers that
 * cannot allocate device memory: a loop that fixes its memory map at start-up and bans later
 * cudaMalloc calls (they stall the device and can fail under fragmentation), or a host whose
 * own pool already holds all of HBM.  This call allocates nothing - no cudaMalloc, no
 * cudaFree, no library allocator: the per-block partial sums go to a 64 KB array built into
 * the library (resident on every GPU that loads it until exit).
 *   vol, result : DEVICE pointers on the current device; result receives 2 floats,
 *                 result[0] = std (unbiased), result[1] = mean.  Work is enqueued on
 *                 `cuda_stream` (a cudaStream_t, NULL = default stream) and not synchronised.
 *   n           : 2 <= n <= 268,435,456 elements; larger inputs are refused with
 *                 STEREO_E_LIMIT (use stereo_volume_std).
 * The first call on each device looks up the array's device address (that device must be
 * current); make that first call before any stream capture begins.  All calls share the one
 * array: two calls whose kernels overlap on the GPU (different streams, no ordering) race on
 * it and give wrong results - serialise them.  Nothing needs clearing between calls. */
int stereo_volume_std_static(const float *vol, size_t n, float *result, void *cuda_stream);

/* The loop body of post.lua:78-243 for one pair: CBCA -> /std -> negate -> SGM -> CBCA ->
 * LR check, occlusion / mismatch fill, sub-pixel, median, bilateral.  Inputs are what
 * inference.py wrote: the two volumes and the two uint8 WTA maps, which post.lua loads as its
 * guide images (post.lua:91-92).  stages[8] receive the uint8 maps post.lua saves, in order:
 * nyu_cost_img, right_nyu_cost_img, nyu_sgm_img, right_nyu_sgm_img, nyu_cost_img_2,
 * right_nyu_cost_img_2, nyu_post, outlier (x127); any entry may be NULL.  final_disp
 * (optional) receives the float map before the :byte() truncation of post.lua:238. */
int stereo_post(stereo_ctx *ctx, const float *left_cost, const float *right_cost,
                const uint8_t *left_disp, const uint8_t *right_disp, int h, int w, int D,
                const stereo_post_params *p, uint8_t *const stages[8], float *final_disp);
int stereo_post_dev(stereo_ctx *ctx, const float *left_cost, const float *right_cost,
                    const uint8_t *left_disp, const uint8_t *right_disp, int h, int w, int D,
                    const stereo_post_params *p, uint8_t *const stages[8], float *final_disp);

/* Per-kernel timing with CUDA events recorded on the context's stream around every
 * launch of a kernel class (no counterpart in the reference, which only prints
 * torch.tic()/toc() wall times, post.lua:100-235).  Enabling resets the counters;
 * reading synchronises the stream. */
enum {
    STEREO_K_PREP = 0,       /* image sums + LUT + normalise/pad */
    STEREO_K_CONV_FIRST = 1, /* first trunk layer (Cin = 3) */
    STEREO_K_CONV = 2,       /* remaining trunk layers */
    STEREO_K_CORR = 3,       /* correlation (+ fused WTA) */
    STEREO_K_WTA = 4,        /* standalone arg max / arg min over D */
    STEREO_K_LR = 5,         /* left-right consistency */
    STEREO_K_CBCA = 6,       /* cross-based cost aggregation */
    STEREO_K_VOLSTD = 7,     /* whole-volume std + scaling */
    STEREO_K_SGM = 8,        /* semi-global matching, all four paths */
    STEREO_K_POST_MAPS = 9,  /* guide image, cross arms and the per-pixel post kernels */
    STEREO_K_COUNT = 10
};
int stereo_profile_enable(stereo_ctx *ctx, int on);
int stereo_profile_read(stereo_ctx *ctx, int kernel_class, double *total_ms, int64_t *launches);
const char *stereo_profile_name(int kernel_class);

/* inference.py (trunk, both cost volumes, WTA) and post.lua (CBCA, SGM, CBCA, LR check, fills, sub-pixel, median,
 * bilateral) back to back for n pairs; the volumes stay on the device (the reference passes them through
 * left_cost/*.bin and right_cost/*.bin, inference.py:210-231 -> post.lua:89-90).  nyu_post [n][h][w] uint8 is
 * post.lua:238's image; outlier (labels * 127, post.lua:241) and left_wta (inference.py:202) may be null. */
int stereo_pairs_post(stereo_ctx *ctx, int net, const uint8_t *left_rgb, const uint8_t *right_rgb, int n, int h, int w,
                      int disp_range, int precision, const stereo_post_params *params, uint8_t *nyu_post,
                      uint8_t *outlier, uint8_t *left_wta);
int stereo_pairs_post_dev(stereo_ctx *ctx, int net, const uint8_t *left_rgb, const uint8_t *right_rgb, int n, int h,
                          int w, int disp_range, int precision, const stereo_post_params *params, uint8_t *nyu_post,
                          uint8_t *outlier, uint8_t *left_wta);

/* post.lua:98-115,188-204 (options cost_agg, cost_agg_2, cost_w, cost_h): `iters` passes of
 * nn.SpatialAveragePooling(kw, kh, 1, 1, (kw-1)/2, (kh-1)/2) with setCountExcludePad() over every disparity plane of
 * a [h][w][D] volume.  The device variant needs a scratch volume when iters > 1. */
int stereo_box_agg(stereo_ctx *ctx, const float *vol, int h, int w, int D, int kw, int kh, int iters, float *out);
int stereo_box_agg_dev(stereo_ctx *ctx, const float *vol, int h, int w, int D, int kw, int kh, int iters, float *out,
                       float *scratch);

/* ---------------------------------------------------------------------------------------------
 * Patch training.  Replaces the body of the training loop of
 * /root/reference/train.py:92-112: model(lpatch, training=True), model(rpatch, training=True),
 * map_inner_product (train.py:11-17), softmax cross-entropy (train.py:106), tape.gradient and
 * optimizer.apply_gradients with tf.keras.optimizers.Adam (train.py:80,109-112).
 * Parameters are one flat fp32 array in the order of Keras' trainable_variables: per layer kernel
 * [k][k][Cin][Cout], bias, gamma, beta; the moving statistics (mean, variance per layer) are a second
 * array.  A trainer belongs to the context it was created on and uses its stream (plus one internal stream on which
 * the left tower runs next to the right one; both are joined before stereo_train_forward_backward returns).
 * ------------------------------------------------------------------------------------------- */
typedef struct stereo_trainer stereo_trainer;
int stereo_train_create(stereo_ctx *ctx, int net, int batch, int disp_range, stereo_trainer **out);
int stereo_train_destroy(stereo_trainer *tr);
size_t stereo_train_param_count(const stereo_trainer *tr);
size_t stereo_train_stat_count(const stereo_trainer *tr);
int stereo_train_patch_size(const stereo_trainer *tr);
/* host arrays; adam_m / adam_v may be null (zeros) */
int stereo_train_set_state(stereo_trainer *tr, const float *params, const float *stats, const float *adam_m,
                           const float *adam_v, long long step);
int stereo_train_get_state(stereo_trainer *tr, float *params, float *stats, float *adam_m, float *adam_v,
                           long long *step);
/* left [B][p][p][3], right [B][p][p+D-1][3], targets [B][D]: host fp32.  loss = batch mean; correct3 = number of
 * patches with |argmax(logits) - argmax(targets)| <= 3 (train.py:117-119); grads_out (host, may be null). */
int stereo_train_forward_backward(stereo_trainer *tr, const float *left, const float *right, const float *targets,
                                  float *loss, int *correct3, float *grads_out);
/* device address of the flat gradient array, for an in-place all-reduce between ranks */
int stereo_train_grads_dev(stereo_trainer *tr, float **grads);
/* one Adam step with the stored gradients (or grads_dev, a device array, if not null) */
int stereo_train_apply(stereo_trainer *tr, const float *grads_dev, float lr);

#ifdef __cplusplus
}
#endif
#endif /* STEREO_B200_H */
