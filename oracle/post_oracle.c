/*
 * TEST INFRASTRUCTURE ONLY - CPU restatement of the post-processing kernels the
 * reference drives through the external mc-cnn `libadcensus.so`
 * (/root/reference/smooth/init.lua:6, call sites smooth/nyu.lua:33-40,59,95-111).
 *
 * The library itself (jzbontar/mc-cnn `adcensus.cu`, commit unpinned, see
 * /root/reference/smooth/README.md:1) is NOT under /root/reference and cannot be
 * fetched, so each function below restates the published algorithm (Zbontar &
 * LeCun, JMLR 2016, section 4; Mei et al. 2011 for the cross construction) as
 * used by the reference's call sites.  Only tests/, __graft_entry__.smoke() and
 * bench.py's CPU-baseline legs may load this file; the product never does.
 *
 * Parity status: outlier_detection is pinned bit-exactly by the 40 PNG triples
 * the reference ships (results/nyu_cost_img_2, right_nyu_cost_img_2 -> outlier);
 * occlusion/mismatch fill + median + bilateral are soft-pinned by
 * results/nyu_post; cross, cbca, sgm2 and subpixel are PARITY UNPINNED (no
 * shipped cost volumes).  See tests/test_post_oracle.py.
 *
 * Layouts follow the reference: images [H][W]; arms [4][H][W]; CBCA volumes
 * [D][H][W] (post.lua:119-120); SGM volumes [H][W][D] (nyu.lua:52).
 * All arithmetic is fp32 like the CUDA original.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

static inline float fmaxf2(float a, float b) { return a > b ? a : b; }
static inline float fminf2(float a, float b) { return a < b ? a : b; }

/* nyu.lua:33-34  adcensus.cross(img, out, L1, tau1)
 * Four arms per pixel (0:left 1:right 2:up 3:down).  Walk away from the pixel;
 * the neighbour at distance 1 is always accepted; stop at the image edge, at
 * the first pixel whose intensity differs by >= tau1, or once distance >= L1.
 * Stores the EXCLUSIVE end coordinate (x for arms 0/1, y for arms 2/3). */
void oracle_cross(const float *img, float *out, int H, int W, int L1, float tau1)
{
#pragma omp parallel for collapse(2)
    for (int arm = 0; arm < 4; ++arm)
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) {
                int dx = 0, dy = 0;
                if (arm == 0) dx = -1; else if (arm == 1) dx = 1;
                else if (arm == 2) dy = -1; else dy = 1;
                int xx = x + dx, yy = y + dy;
                float centre = img[y * W + x];
                for (;; xx += dx, yy += dy) {
                    if (xx < 0 || xx >= W || yy < 0 || yy >= H) break;
                    int ax = abs(xx - x), ay = abs(yy - y);
                    int dist = ax > ay ? ax : ay;
                    if (dist == 1) continue;
                    if (fabsf(centre - img[yy * W + xx]) >= tau1) break;
                    if (dist >= L1) break;
                }
                out[((size_t)arm * H + y) * W + x] = (float)(dy == 0 ? xx : yy);
            }
}

/* nyu.lua:37,39  adcensus.cbca(x0c, x1c, vol, out, direction)
 * Cross-based aggregation of one [D][H][W] volume: average over the region
 * that is the intersection of the pixel's cross (arms x0c) and the cross of
 * the matched pixel x + d*direction in the other view (arms x1c, shifted back).
 * Voxels whose match falls outside the image are copied. */
void oracle_cbca(const float *x0c, const float *x1c, const float *vol, float *out,
                 int D, int H, int W, int direction)
{
#pragma omp parallel for collapse(2) schedule(dynamic, 4)
    for (int d = 0; d < D; ++d)
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) {
                size_t id = ((size_t)d * H + y) * W + x;
                int xm = x + d * direction;
                if (xm < 0 || xm >= W) { out[id] = vol[id]; continue; }
                float sum = 0.f;
                int cnt = 0;
                int yy_s = (int)fmaxf2(x0c[((size_t)2 * H + y) * W + x], x1c[((size_t)2 * H + y) * W + xm]);
                int yy_t = (int)fminf2(x0c[((size_t)3 * H + y) * W + x], x1c[((size_t)3 * H + y) * W + xm]);
                for (int yy = yy_s + 1; yy < yy_t; ++yy) {
                    int xx_s = (int)fmaxf2(x0c[((size_t)0 * H + yy) * W + x],
                                           x1c[((size_t)0 * H + yy) * W + xm] - d * direction);
                    int xx_t = (int)fminf2(x0c[((size_t)1 * H + yy) * W + x],
                                           x1c[((size_t)1 * H + yy) * W + xm] - d * direction);
                    for (int xx = xx_s + 1; xx < xx_t; ++xx) {
                        sum += vol[((size_t)d * H + yy) * W + xx];
                        ++cnt;
                    }
                }
                out[id] = sum / (float)cnt;
            }
}

/* nyu.lua:59-60  adcensus.sgm2(x0, x1, vol, out, tmp, pi1, pi2, tau_so, alpha1, q1, q2, direction)
 * Four scan directions (0: left->right, 1: right->left, 2: top->bottom,
 * 3: bottom->top); `out` accumulates the four path costs (the Lua wrapper
 * divides by 4).  vol/out are [H][W][D].  Penalties adapt to the intensity
 * steps D1 (in x0 along the path) and D2 (in x1 at the matched pixel). */
static void sgm_step(const float *x0, const float *x1, const float *vol, float *out,
                     const float *prev, float *cur, int H, int W, int D,
                     int x, int y, int dx, int dy, int sgm_direction, int direction,
                     float pi1, float pi2, float tau_so, float alpha1, float q1, float q2,
                     int first)
{
    size_t base = ((size_t)y * W + x) * D;
    if (first) {
        for (int d = 0; d < D; ++d) {
            float v = vol[base + d];
            out[base + d] += v;
            cur[d] = v;
        }
        return;
    }
    float pmin = prev[0];
    for (int d = 1; d < D; ++d) if (prev[d] < pmin) pmin = prev[d];
    int ind2 = y * W + x;
    float D1 = fabsf(x0[ind2] - x0[ind2 - dy * W - dx]);
    for (int d = 0; d < D; ++d) {
        float D2;
        int xx = x + d * direction;
        if (xx < 0 || xx >= W || xx - dx < 0 || xx - dx >= W) D2 = 10.f;
        else D2 = fabsf(x1[ind2 + d * direction] - x1[ind2 + d * direction - dy * W - dx]);
        float P1, P2;
        if (D1 < tau_so && D2 < tau_so) { P1 = pi1; P2 = pi2; }
        else if (D1 > tau_so && D2 > tau_so) { P1 = pi1 / (q1 * q2); P2 = pi2 / (q1 * q2); }
        else { P1 = pi1 / q1; P2 = pi2 / q1; }
        float cost = fminf2(prev[d], pmin + P2);
        if (d - 1 >= 0) cost = fminf2(cost, prev[d - 1] + (sgm_direction == 2 ? P1 / alpha1 : P1));
        if (d + 1 < D) cost = fminf2(cost, prev[d + 1] + (sgm_direction == 3 ? P1 / alpha1 : P1));
        float v = vol[base + d] + cost - pmin;
        out[base + d] += v;
        cur[d] = v;
    }
}

static void sgm_diagonals(const float *x0, const float *x1, const float *vol, float *out,
                          int H, int W, int D, float pi1, float pi2, float tau_so, float alpha1,
                          float q1, float q2, int direction);

/* `paths` = 4: the reference (nyu.lua:59-60).  `paths` = 8 is NOT in the reference: the four
 * diagonal paths (4: down-right, 5: down-left, 6: up-right, 7: up-left) are added after the
 * four of the original with the same step (no alpha1 scaling, which the original applies to
 * the two vertical paths only); the caller divides by 8.  BASELINE.json configs[3] names an
 * 8-path chain; it is kept apart from every parity figure. */
void oracle_sgm2_paths(const float *x0, const float *x1, const float *vol, float *out,
                       int H, int W, int D, float pi1, float pi2, float tau_so, float alpha1,
                       float q1, float q2, int direction, int paths);

void oracle_sgm2(const float *x0, const float *x1, const float *vol, float *out,
                 int H, int W, int D, float pi1, float pi2, float tau_so, float alpha1,
                 float q1, float q2, int direction)
{
    oracle_sgm2_paths(x0, x1, vol, out, H, W, D, pi1, pi2, tau_so, alpha1, q1, q2, direction, 4);
}

void oracle_sgm2_paths(const float *x0, const float *x1, const float *vol, float *out,
                       int H, int W, int D, float pi1, float pi2, float tau_so, float alpha1,
                       float q1, float q2, int direction, int paths)
{
    /* horizontal paths: scanlines (rows) are independent */
    for (int sd = 0; sd < 2; ++sd) {
#pragma omp parallel
        {
            float *a = (float *)malloc(sizeof(float) * D), *b = (float *)malloc(sizeof(float) * D);
#pragma omp for schedule(static)
            for (int y = 0; y < H; ++y) {
                float *prev = a, *cur = b;
                for (int step = 0; step < W; ++step) {
                    int x = sd == 0 ? step : W - 1 - step;
                    int dx = sd == 0 ? 1 : -1;
                    sgm_step(x0, x1, vol, out, prev, cur, H, W, D, x, y, dx, 0, sd, direction,
                             pi1, pi2, tau_so, alpha1, q1, q2, step == 0);
                    float *t = prev; prev = cur; cur = t;
                }
            }
            free(a); free(b);
        }
    }
    /* vertical paths: columns are independent */
    for (int sd = 2; sd < 4; ++sd) {
#pragma omp parallel
        {
            float *a = (float *)malloc(sizeof(float) * D), *b = (float *)malloc(sizeof(float) * D);
#pragma omp for schedule(static)
            for (int x = 0; x < W; ++x) {
                float *prev = a, *cur = b;
                for (int step = 0; step < H; ++step) {
                    int y = sd == 2 ? step : H - 1 - step;
                    int dy = sd == 2 ? 1 : -1;
                    sgm_step(x0, x1, vol, out, prev, cur, H, W, D, x, y, 0, dy, sd, direction,
                             pi1, pi2, tau_so, alpha1, q1, q2, step == 0);
                    float *t = prev; prev = cur; cur = t;
                }
            }
            free(a); free(b);
        }
    }
    if (paths == 8) sgm_diagonals(x0, x1, vol, out, H, W, D, pi1, pi2, tau_so, alpha1, q1, q2, direction);
}

/* the four diagonal paths: image rows in path order, the previous row's path costs kept whole */
static void sgm_diagonals(const float *x0, const float *x1, const float *vol, float *out,
                          int H, int W, int D, float pi1, float pi2, float tau_so, float alpha1,
                          float q1, float q2, int direction)
{
    float *rows[2];
    rows[0] = (float *)malloc(sizeof(float) * (size_t)W * D);
    rows[1] = (float *)malloc(sizeof(float) * (size_t)W * D);
    for (int sd = 4; sd < 8; ++sd) {
        int dx = (sd == 4 || sd == 6) ? 1 : -1;
        int dy = sd < 6 ? 1 : -1;
        for (int step = 0; step < H; ++step) {
            int y = dy > 0 ? step : H - 1 - step;
            const float *prow = rows[(step + 1) & 1];
            float *crow = rows[step & 1];
#pragma omp parallel for schedule(static)
            for (int x = 0; x < W; ++x) {
                int first = step == 0 || x - dx < 0 || x - dx >= W;
                sgm_step(x0, x1, vol, out, first ? crow : prow + (size_t)(x - dx) * D, crow + (size_t)x * D,
                         H, W, D, x, y, dx, dy, sd, direction, pi1, pi2, tau_so, alpha1, q1, q2, first);
            }
        }
    }
    free(rows[0]); free(rows[1]);
}

/* nyu.lua:95  adcensus.outlier_detection(d0, d1, outlier, disp_max)
 * 0 = match, 1 = occlusion, 2 = mismatch.  Pinned bit-exactly (K5). */
void oracle_outlier_detection(const float *d0, const float *d1, float *outlier,
                              int H, int W, int disp_max)
{
#pragma omp parallel for
    for (int id = 0; id < H * W; ++id) {
        int x = id % W;
        int d0i = (int)d0[id];
        if (x - d0i < 0) { outlier[id] = 1.f; continue; }
        if (fabsf(d0[id] - d1[id - d0i]) < 1.1f) { outlier[id] = 0.f; continue; }
        outlier[id] = 1.f;
        for (int d = 0; d < disp_max; ++d)
            if (x - d >= 0 && fabsf((float)d - d1[id - d]) < 1.1f) { outlier[id] = 2.f; break; }
    }
}

/* nyu.lua:98  adcensus.interpolate_occlusion(d0, outlier)
 * Occluded pixels take the disparity of the nearest match to the left on the
 * same row; if there is none, the nearest match to the right; else unchanged. */
void oracle_interpolate_occlusion(const float *d0, const float *outlier, float *out, int H, int W)
{
#pragma omp parallel for
    for (int id = 0; id < H * W; ++id) {
        if (outlier[id] != 1.f) { out[id] = d0[id]; continue; }
        int x = id % W;
        int dx = 0;
        while (x + dx >= 0 && outlier[id + dx] != 0.f) --dx;
        if (x + dx < 0) {
            dx = 0;
            while (x + dx < W && outlier[id + dx] != 0.f) ++dx;
        }
        out[id] = x + dx < W ? d0[id + dx] : d0[id];
    }
}

static void sort_floats(float *v, int n)
{
    for (int i = 1; i < n; ++i) {
        float t = v[i];
        int j = i - 1;
        while (j >= 0 && v[j] > t) { v[j + 1] = v[j]; --j; }
        v[j + 1] = t;
    }
}

/* nyu.lua:99  adcensus.interpolate_mismatch(d0, outlier)
 * Mismatched pixels look along 16 directions for the first pixel that is not a
 * mismatch and take the median (index n/2 of the sorted hits). */
void oracle_interpolate_mismatch(const float *d0, const float *outlier, float *out, int H, int W)
{
    static const float dirs[32] = {
        0.f, 1.f, -0.5f, 1.f, -1.f, 1.f, -1.f, 0.5f, -1.f, 0.f, -1.f, -0.5f, -1.f, -1.f, -0.5f, -1.f,
        0.f, -1.f, 0.5f, -1.f, 1.f, -1.f, 1.f, -0.5f, 1.f, 0.f, 1.f, 0.5f, 1.f, 1.f, 0.5f, 1.f};
#pragma omp parallel for
    for (int id = 0; id < H * W; ++id) {
        if (outlier[id] != 2.f) { out[id] = d0[id]; continue; }
        float vals[16];
        int n = 0;
        int x = id % W, y = id / W;
        for (int k = 0; k < 16; ++k) {
            float dx = dirs[2 * k], dy = dirs[2 * k + 1];
            float xx = (float)x, yy = (float)y;
            int xi = (int)roundf(xx), yi = (int)roundf(yy);
            while (0 <= yi && yi < H && 0 <= xi && xi < W && outlier[yi * W + xi] == 2.f) {
                xx += dx; yy += dy;
                xi = (int)roundf(xx); yi = (int)roundf(yy);
            }
            if (0 <= yi && yi < H && 0 <= xi && xi < W) vals[n++] = d0[yi * W + xi];
        }
        if (n == 0) { out[id] = d0[id]; continue; }
        sort_floats(vals, n);
        out[id] = vals[n / 2];
    }
}

/* nyu.lua:103  adcensus.subpixel_enchancement(d0, left_vol, disp_max)
 * Parabola fit through the costs at d-1, d, d+1 of the [D][H][W] volume. */
void oracle_subpixel(const float *d0, const float *vol, float *out, int H, int W, int disp_max)
{
    size_t hw = (size_t)H * W;
#pragma omp parallel for
    for (int id = 0; id < H * W; ++id) {
        int d = (int)d0[id];
        out[id] = (float)d;
        if (1 <= d && d < disp_max - 1) {
            float cn = vol[(size_t)(d - 1) * hw + id];
            float cz = vol[(size_t)d * hw + id];
            float cp = vol[(size_t)(d + 1) * hw + id];
            float denom = 2.f * (cp + cn - 2.f * cz);
            if (denom > 1e-5f) {
                float t = (cp - cn) / denom;
                t = t < -1.f ? -1.f : (t > 1.f ? 1.f : t);
                out[id] = (float)d - t;
            }
        }
    }
}

/* nyu.lua:107  adcensus.median2d(img, 5): median over the in-image part of the
 * window, element n/2 of the sorted samples. */
void oracle_median2d(const float *img, float *out, int H, int W, int kernel_size)
{
    int r = kernel_size / 2;
#pragma omp parallel for
    for (int id = 0; id < H * W; ++id) {
        int x = id % W, y = id / W;
        float xs[11 * 11];
        int n = 0;
        for (int xx = x - r; xx <= x + r; ++xx)
            for (int yy = y - r; yy <= y + r; ++yy)
                if (0 <= xx && xx < W && 0 <= yy && yy < H) xs[n++] = img[yy * W + xx];
        sort_floats(xs, n);
        out[id] = xs[n / 2];
    }
}

/* nyu.lua:111  adcensus.mean2d(img, gaussian(blur_sigma), blur_t): bilateral-like
 * mean over in-image neighbours whose value is within blur_t of the centre. */
void oracle_mean2d(const float *img, const float *kernel, float *out, int H, int W,
                   int kernel_radius, float alpha2)
{
#pragma omp parallel for schedule(dynamic, 64)
    for (int id = 0; id < H * W; ++id) {
        int x = id % W, y = id / W;
        float sum = 0.f, cnt = 0.f;
        float centre = img[id];
        int i = 0;
        for (int xx = x - kernel_radius; xx <= x + kernel_radius; ++xx)
            for (int yy = y - kernel_radius; yy <= y + kernel_radius; ++yy, ++i)
                if (0 <= xx && xx < W && 0 <= yy && yy < H &&
                    fabsf(img[yy * W + xx] - centre) < alpha2) {
                    /* the CUDA original is built with nvcc's default -fmad=true, which
                     * contracts this multiply-add; fmaf mirrors that */
                    sum = fmaf(img[yy * W + xx], kernel[i], sum);
                    cnt += kernel[i];
                }
        out[id] = sum / cnt;
    }
}

int oracle_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* post.lua:98-115,188-204  nn.SpatialAveragePooling(cost_w, cost_h, 1, 1, pad_w, pad_h) with
 * setCountExcludePad(), applied to every disparity plane (torch7 THCUNN AvePoolForward: window clipped to the
 * image, values summed row by row in fp32, divided by the number of in-image taps).
 * Layout here [H][W][D] (D fastest), like every volume of this oracle. */
void oracle_box_agg(const float *in, float *out, int H, int W, int D, int kw, int kh)
{
    const int pw = (kw - 1) / 2, ph = (kh - 1) / 2;
#pragma omp parallel for
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            int y0 = y - ph, x0 = x - pw, y1 = y0 + kh, x1 = x0 + kw;
            if (y1 > H + ph) y1 = H + ph;
            if (x1 > W + pw) x1 = W + pw;
            if (y0 < 0) y0 = 0;
            if (x0 < 0) x0 = 0;
            if (y1 > H) y1 = H;
            if (x1 > W) x1 = W;
            const float cnt = (float)((y1 - y0) * (x1 - x0));
            for (int d = 0; d < D; ++d) {
                float s = 0.f;
                for (int yy = y0; yy < y1; ++yy)
                    for (int xx = x0; xx < x1; ++xx) s += in[((size_t)yy * W + xx) * D + d];
                out[((size_t)y * W + x) * D + d] = s / cnt;
            }
        }
}
