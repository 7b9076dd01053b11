// Patch training of the Siamese trunk on the device (/root/reference/train.py:92-112, models/model.py,
// data_handler.py:75-85): forward of both towers in training mode (conv 'valid' -> ReLU -> BatchNorm
// with batch statistics), inner product of the left feature with the D right features, softmax
// cross-entropy against the soft 5-tap labels, backward through both towers (shared weights, gradients
// summed), Adam.  SURVEY.md section 8(f) rank 1.
//
// Everything is fp32 on the CUDA cores: the tensors are small (B = 128 patches of 9 x 9 and 9 x 209) and
// the step is latency-, not throughput-bound; channel reductions (BatchNorm statistics, weight gradients)
// accumulate in double.  The gradient buffer is one flat device array in the order of Keras'
// `trainable_variables` (per layer: kernel [k][k][Cin][Cout], bias, gamma, beta), so that data-parallel
// training is ONE all-reduce over it (113 088 floats for base_model_ws_9).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"

namespace stereo {

constexpr float kBnEps = 1e-3f;        // Keras BatchNormalization default (models/model.py)
constexpr float kBnMomentum = 0.99f;   // Keras default
constexpr int kRedW = 256;             // offset of the weight-gradient sums in the double scratch (after the BatchNorm sums)
constexpr int kTrainTcMinWidth = 96;   // rows at least this wide go to the tensor-core convolution (256-pixel tiles)

struct TrainLayer {
    int ks, cin, cout, relu;
    size_t w_off, b_off, g_off, be_off;     // offsets into the flat parameter / gradient arrays
    size_t mean_off, var_off;               // offsets into the moving statistics
};

// activations of one tower that the backward pass needs
struct Tower {
    DevBuf x[kMaxLayers + 1];      // x[l] = input of layer l (x[0] = patches), x[L] = features
    DevBuf a[kMaxLayers];          // conv + bias (+ ReLU) output, before BatchNorm
    DevBuf bn[kMaxLayers];         // per channel: batch mean, inverse std, batch variance  (3 * cout floats)
    int h[kMaxLayers + 1], w[kMaxLayers + 1];
    // scratch of this tower's passes (the towers run on two streams at the same time)
    DevBuf red;                    // double scratch for reductions (>= n_params + 1024)
    DevBuf red_part;               // per-block partial sums of the channel reductions (double)
    DevBuf d_act[2];               // ping-pong gradient wrt activations
    DevBuf d_pad, w_flip;          // backward-data as a forward convolution: padded dz, flipped weights
    DevBuf tc_act, tc_wimg;        // tensor-core route: chunk-planar bf16 (hi, lo) input, UMMA weight image
    DevBuf wg_scratch;             // tensor-core weight gradient: per-CTA partial sums
    DevBuf grads;                  // this tower's share of the gradient (the left tower; the right one adds into tr->grads)
};

}  // namespace stereo

struct stereo_trainer {
    stereo_ctx *ctx = nullptr;
    int net = 0, batch = 0, D = 0, patch = 0, n_layers = 0;
    stereo::TrainLayer layer[stereo::kMaxLayers];
    size_t n_params = 0, n_stats = 0;
    stereo::DevBuf params, grads, adam_m, adam_v, stats;
    stereo::DevBuf logits, dlogits, loss_buf, targets;
    stereo::DevBuf dfeat[2];       // gradient wrt the left / right feature tensors
    stereo::DevBuf zero_bias;
    stereo::Tower tower[2];
    long long step = 0;
    cudaStream_t side_stream = nullptr;         // the left tower runs here while the right one runs on ctx->stream
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    bool warm = false;                          // a step has run: every buffer has its size
    cudaGraphExec_t graph_exec = nullptr;       // the captured step
    long long graph_launches = 0;               // kernel launches inside it
};

namespace stereo {

// ------------------------------------------------------------------------------------------
// forward kernels
// ------------------------------------------------------------------------------------------
// out[p][co] = bias[co] + sum_{ky,kx,ci} in[(y+ky, x+kx)][ci] * w[ky][kx][ci][co]   (+ ReLU)
__global__ void train_conv_fwd_kernel(const float *__restrict__ in, const float *__restrict__ w,
                                      const float *__restrict__ bias, int n, int h, int wd, int ks, int cin, int cout,
                                      int relu, float *__restrict__ out)
{
    const int ho = h - ks + 1, wo = wd - ks + 1;
    const size_t total = (size_t)n * ho * wo * cout;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int co = (int)(i % cout);
        size_t p = i / cout;
        const int x = (int)(p % wo);
        p /= wo;
        const int y = (int)(p % ho), b = (int)(p / ho);
        float acc = bias[co];
        for (int ky = 0; ky < ks; ++ky)
            for (int kx = 0; kx < ks; ++kx) {
                const float *ip = in + (((size_t)b * h + y + ky) * wd + x + kx) * cin;
                const float *wp = w + ((size_t)(ky * ks + kx) * cin) * cout + co;
                for (int ci = 0; ci < cin; ++ci) acc = fmaf(ip[ci], wp[(size_t)ci * cout], acc);
            }
        out[i] = relu ? fmaxf(acc, 0.f) : acc;
    }
}

// batch mean / biased variance -> (mean, invstd, variance)
__global__ void train_bn_finalize_kernel(const double *__restrict__ red, double m, int c, float *__restrict__ bn)
{
    const int ch = threadIdx.x;
    if (ch >= c) return;
    const double mean = red[ch] / m;
    double var = red[c + ch] / m - mean * mean;
    if (var < 0.0) var = 0.0;
    bn[ch] = (float)mean;
    bn[c + ch] = (float)(1.0 / sqrt(var + (double)kBnEps));
    bn[2 * c + ch] = (float)var;
}

// Moving statistics as Keras updates them: the model is called on the left batch, then on the right one
// (train.py:92-95), each call one momentum step.  The towers run at the same time, so the two steps are applied
// here, in that order, once both batches' statistics are there.
__global__ void train_update_stats_kernel(const float *__restrict__ bn_left, const float *__restrict__ bn_right, int c,
                                          float *__restrict__ moving_mean, float *__restrict__ moving_var)
{
    const int ch = threadIdx.x;
    if (ch >= c) return;
    float mm = moving_mean[ch], mv = moving_var[ch];
    mm = mm * kBnMomentum + bn_left[ch] * (1.f - kBnMomentum);
    mv = mv * kBnMomentum + bn_left[2 * c + ch] * (1.f - kBnMomentum);
    mm = mm * kBnMomentum + bn_right[ch] * (1.f - kBnMomentum);
    mv = mv * kBnMomentum + bn_right[2 * c + ch] * (1.f - kBnMomentum);
    moving_mean[ch] = mm;
    moving_var[ch] = mv;
}

// ---- vectorised forms (c % 4 == 0: every layer's output has 32 or 64 channels) ----------------------------------
// A thread owns four consecutive channels (quad q = tid % (c / 4)) of the rows r, r + rows, ... (r = tid / (c / 4)):
// one 16-byte load per row and tensor, double sums in registers, one shared-memory reduction over the block's row
// groups, and the block's sums go to its own row of a partial buffer; train_sum_partials_kernel adds the rows up.
// No atomics: the scalar forms issued two double atomics per THREAD on 128 addresses (2 400 per address and launch,
// 40-80 us for 47 MB), and even one per channel and block still serialised at ~30 ns each (20-35 us per launch
// whatever the tensor size - ncu round 2).
template <int NV>      // doubles per thread
__device__ __forceinline__ void block_reduce_to_partial(double (&v)[NV], int c4, double *__restrict__ part, int comp_stride)
{
    // v[k]: k = 4 * component + channel-in-quad; component j goes to part[j * comp_stride + channel]
    __shared__ double sh[256 * NV];
    const int q = threadIdx.x % c4, r = threadIdx.x / c4, rows = blockDim.x / c4;
#pragma unroll
    for (int k = 0; k < NV; ++k) sh[(k * rows + r) * c4 + q] = v[k];
    __syncthreads();
    if (r == 0) {
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            double s = 0.0;
            for (int rr = 0; rr < rows; ++rr) s += sh[(k * rows + rr) * c4 + q];
            part[(k >> 2) * comp_stride + 4 * q + (k & 3)] = s;
        }
    }
}

// out[i] (+)= sum over the `parts` rows of partial[part][i], i < n: block b sums the eight values 8 b .. 8 b + 7, thread
// (g, j) of it adds the rows g, g + 128, ... (a single block took 8 us per call: it waits out the L2 latency of its ~40
// loads per thread one batch after the other).
__global__ __launch_bounds__(1024) void train_sum_partials_kernel(const double *__restrict__ partial, int parts, int n,
                                                                  double *__restrict__ out, int add)
{
    __shared__ double sh[1024];
    const int j = threadIdx.x & 7, g = threadIdx.x >> 3, i = 8 * blockIdx.x + j;
    double s = 0.0;
    if (i < n) {
#pragma unroll 4
        for (int pr = g; pr < parts; pr += 128) s += __ldg(partial + (size_t)pr * n + i);
    }
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 512; o >= 8; o >>= 1) {
        if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x < 8 && i < n) out[i] = add ? out[i] + sh[threadIdx.x] : sh[threadIdx.x];
}

__global__ __launch_bounds__(256) void train_moments4_kernel(const float4 *__restrict__ a, size_t m, int c4,
                                                             double *__restrict__ red)
{
    const int q = threadIdx.x % c4, r = threadIdx.x / c4, rows = blockDim.x / c4;
    double v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll 2
    for (size_t p = (size_t)blockIdx.x * rows + r; p < m; p += (size_t)gridDim.x * rows) {
        const float4 x = __ldg(a + p * c4 + q);
        const double d0 = x.x, d1 = x.y, d2 = x.z, d3 = x.w;
        v[0] += d0; v[1] += d1; v[2] += d2; v[3] += d3;
        v[4] += d0 * d0; v[5] += d1 * d1; v[6] += d2 * d2; v[7] += d3 * d3;
    }
    block_reduce_to_partial<8>(v, c4, red + (size_t)blockIdx.x * 8 * c4, 4 * c4);
}

__global__ __launch_bounds__(256) void train_bn_apply4_kernel(const float4 *__restrict__ a, size_t total4, int c4,
                                                              const float *__restrict__ bn, const float *__restrict__ gamma,
                                                              const float *__restrict__ beta, float4 *__restrict__ y)
{
    // the grid stride is a multiple of c4 (host), so a thread keeps its channel quad
    const size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int ch = (int)(i0 % c4) * 4, c = 4 * c4;
    float mean[4], scale[4], shift[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { mean[k] = bn[ch + k]; scale[k] = bn[c + ch + k]; shift[k] = beta[ch + k]; }
    const float g0 = gamma[ch], g1 = gamma[ch + 1], g2 = gamma[ch + 2], g3 = gamma[ch + 3];
    for (size_t i = i0; i < total4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 x = __ldg(a + i);
        float4 o;
        o.x = g0 * ((x.x - mean[0]) * scale[0]) + shift[0];
        o.y = g1 * ((x.y - mean[1]) * scale[1]) + shift[1];
        o.z = g2 * ((x.z - mean[2]) * scale[2]) + shift[2];
        o.w = g3 * ((x.w - mean[3]) * scale[3]) + shift[3];
        y[i] = o;
    }
}

// ------------------------------------------------------------------------------------------
// loss: logits[b][d] = <L[b], R[b][d]>, softmax cross-entropy with soft labels (train.py:11-17,106)
// one block per patch pair; writes dlogits = (softmax - t) / B, loss_b, |argmax - argmax(t)| <= 3
// ------------------------------------------------------------------------------------------
__global__ void train_loss_kernel(const float *__restrict__ lf, const float *__restrict__ rf,
                                  const float *__restrict__ targets, int D, int c, int batch,
                                  float *__restrict__ logits, float *__restrict__ dlogits, float *__restrict__ loss_out)
{
    extern __shared__ float sh[];        // [D] logits, then reduction scratch
    float *lg = sh, *scr = sh + ((D + 1) & ~1);      // 8-byte aligned: reused as double scratch
    const int b = blockIdx.x, tid = threadIdx.x, nt = blockDim.x;
    const float *l = lf + (size_t)b * c;
    for (int d = tid; d < D; d += nt) {
        const float *r = rf + ((size_t)b * D + d) * c;
        float acc = 0.f;
        for (int k = 0; k < c; ++k) acc = fmaf(l[k], r[k], acc);
        lg[d] = acc;
        logits[(size_t)b * D + d] = acc;
    }
    __syncthreads();
    // max, argmax (lowest index), label argmax
    float mx = -INFINITY;
    int am = 0x7fffffff;
    float tmx = -INFINITY;
    int tam = 0x7fffffff;
    for (int d = tid; d < D; d += nt) {
        if (lg[d] > mx) { mx = lg[d]; am = d; }
        const float t = targets[(size_t)b * D + d];
        if (t > tmx) { tmx = t; tam = d; }
    }
    float *s_v = scr, *s_t = scr + nt;
    int *s_i = reinterpret_cast<int *>(scr + 2 * nt), *s_ti = reinterpret_cast<int *>(scr + 3 * nt);
    s_v[tid] = mx; s_i[tid] = am; s_t[tid] = tmx; s_ti[tid] = tam;
    __syncthreads();
    for (int o = nt / 2; o > 0; o >>= 1) {
        if (tid < o) {
            if (s_v[tid + o] > s_v[tid] || (s_v[tid + o] == s_v[tid] && s_i[tid + o] < s_i[tid])) { s_v[tid] = s_v[tid + o]; s_i[tid] = s_i[tid + o]; }
            if (s_t[tid + o] > s_t[tid] || (s_t[tid + o] == s_t[tid] && s_ti[tid + o] < s_ti[tid])) { s_t[tid] = s_t[tid + o]; s_ti[tid] = s_ti[tid + o]; }
        }
        __syncthreads();
    }
    mx = s_v[0];
    const int pred = s_i[0], label = s_ti[0];
    __syncthreads();
    // sum of exp, sum of t * (x - max)
    double se = 0.0, st = 0.0, tsum = 0.0;
    for (int d = tid; d < D; d += nt) {
        const float z = lg[d] - mx;
        se += exp((double)z);
        const double t = (double)targets[(size_t)b * D + d];
        st += t * (double)z;
        tsum += t;
    }
    double *s_d = reinterpret_cast<double *>(scr);
    s_d[tid] = se; s_d[nt + tid] = st; s_d[2 * nt + tid] = tsum;
    __syncthreads();
    for (int o = nt / 2; o > 0; o >>= 1) {
        if (tid < o) { s_d[tid] += s_d[tid + o]; s_d[nt + tid] += s_d[nt + tid + o]; s_d[2 * nt + tid] += s_d[2 * nt + tid + o]; }
        __syncthreads();
    }
    se = s_d[0]; st = s_d[nt]; tsum = s_d[2 * nt];
    const double lse = log(se);
    for (int d = tid; d < D; d += nt) {
        const double sm = exp((double)(lg[d] - mx)) / se;
        dlogits[(size_t)b * D + d] = (float)((sm * tsum - (double)targets[(size_t)b * D + d]) / (double)batch);
    }
    if (tid == 0) {
        loss_out[2 * b] = (float)(tsum * lse - st);                // -sum t * log softmax
        loss_out[2 * b + 1] = abs(pred - label) <= 3 ? 1.f : 0.f;  // train.py:117-119
    }
}

// dL[b][k] = sum_d dlogits[b][d] * R[b][d][k];  dR[b][d][k] = dlogits[b][d] * L[b][k]
__global__ void train_loss_bwd_kernel(const float *__restrict__ lf, const float *__restrict__ rf,
                                      const float *__restrict__ dlogits, int D, int c, float *__restrict__ dl,
                                      float *__restrict__ dr)
{
    const int b = blockIdx.x, k = threadIdx.x;
    if (k >= c) return;
    const float lk = lf[(size_t)b * c + k];
    float acc = 0.f;
    for (int d = 0; d < D; ++d) {
        const float g = dlogits[(size_t)b * D + d];
        acc = fmaf(g, rf[((size_t)b * D + d) * c + k], acc);
        dr[((size_t)b * D + d) * c + k] = g * lk;
    }
    dl[(size_t)b * c + k] = acc;
}

// ------------------------------------------------------------------------------------------
// backward kernels
// ------------------------------------------------------------------------------------------
// red[ch] += sum dy, red[c + ch] += sum dy * xhat (vectorised form of train_bn_bwd_reduce_kernel)
__global__ __launch_bounds__(256) void train_bn_bwd_reduce4_kernel(const float4 *__restrict__ dy, const float4 *__restrict__ a,
                                                                   size_t m, int c4, const float *__restrict__ bn,
                                                                   double *__restrict__ red)
{
    const int q = threadIdx.x % c4, r = threadIdx.x / c4, rows = blockDim.x / c4, c = 4 * c4;
    float mean[4], invstd[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { mean[k] = bn[4 * q + k]; invstd[k] = bn[c + 4 * q + k]; }
    double v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll 2
    for (size_t p = (size_t)blockIdx.x * rows + r; p < m; p += (size_t)gridDim.x * rows) {
        const float4 g = __ldg(dy + p * c4 + q), x = __ldg(a + p * c4 + q);
        v[0] += (double)g.x; v[1] += (double)g.y; v[2] += (double)g.z; v[3] += (double)g.w;
        v[4] += (double)g.x * (double)((x.x - mean[0]) * invstd[0]);
        v[5] += (double)g.y * (double)((x.y - mean[1]) * invstd[1]);
        v[6] += (double)g.z * (double)((x.z - mean[2]) * invstd[2]);
        v[7] += (double)g.w * (double)((x.w - mean[3]) * invstd[3]);
    }
    block_reduce_to_partial<8>(v, c4, red + (size_t)blockIdx.x * 2 * c, c);
}

// da = gamma * invstd * (dy - dbeta / M - xhat * dgamma / M), through the ReLU mask; the same pass sums dz per channel
// into `red_bias` (the bias gradient), which the scalar path computed in a launch of its own.
__global__ __launch_bounds__(256) void train_bn_bwd_apply4_kernel(const float4 *__restrict__ dy, const float4 *__restrict__ a,
                                                                  size_t total4, int c4, double m, const float *__restrict__ bn,
                                                                  const float *__restrict__ gamma, const double *__restrict__ red,
                                                                  int relu, float4 *__restrict__ dz, double *__restrict__ red_bias)
{
    const size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;     // grid stride: a multiple of c4
    const int ch = (int)(i0 % c4) * 4, c = 4 * c4;
    float mean[4], invstd[4], gs[4], db[4], dg[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        mean[k] = bn[ch + k];
        invstd[k] = bn[c + ch + k];
        gs[k] = gamma[ch + k] * invstd[k];
        db[k] = (float)(red[ch + k] / m);
        dg[k] = (float)(red[c + ch + k] / m);
    }
    double v[4] = {0, 0, 0, 0};
    for (size_t i = i0; i < total4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 g = __ldg(dy + i), x = __ldg(a + i);
        float4 o;
        o.x = gs[0] * (g.x - db[0] - ((x.x - mean[0]) * invstd[0]) * dg[0]);
        o.y = gs[1] * (g.y - db[1] - ((x.y - mean[1]) * invstd[1]) * dg[1]);
        o.z = gs[2] * (g.z - db[2] - ((x.z - mean[2]) * invstd[2]) * dg[2]);
        o.w = gs[3] * (g.w - db[3] - ((x.w - mean[3]) * invstd[3]) * dg[3]);
        if (relu) {
            if (!(x.x > 0.f)) o.x = 0.f;
            if (!(x.y > 0.f)) o.y = 0.f;
            if (!(x.z > 0.f)) o.z = 0.f;
            if (!(x.w > 0.f)) o.w = 0.f;
        }
        dz[i] = o;
        v[0] += (double)o.x; v[1] += (double)o.y; v[2] += (double)o.z; v[3] += (double)o.w;
    }
    block_reduce_to_partial<4>(v, c4, red_bias + (size_t)blockIdx.x * c, c);
}

// dst[i] += red[i] for two ranges in one launch (BatchNorm: dbeta, dgamma; convolution: dW, db)
__global__ void train_accumulate2_kernel(const double *__restrict__ red_a, int na, float *__restrict__ dst_a,
                                         const double *__restrict__ red_b, int nb, float *__restrict__ dst_b)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < na) dst_a[i] += (float)red_a[i];
    else if (i < na + nb) dst_b[i - na] += (float)red_b[i - na];
}

// dx[b][yy][xx][ci] = sum_{ky,kx,co} dz[b][yy - ky][xx - kx][co] * w[ky][kx][ci][co] is the 'valid' convolution of
// dz, zero-padded by ks - 1 on every side, with the flipped and transposed weights
// w'[ky'][kx'][co][ci] = w[ks-1-ky'][ks-1-kx'][ci][co]: it runs through train_conv.  This kernel makes the zero-padded copy.
__global__ void train_pad4_kernel(const float4 *__restrict__ dz, int n, int ho, int wo, int c4, int pad, float4 *__restrict__ out)
{
    const int hp = ho + 2 * pad, wp = wo + 2 * pad;
    const size_t total = (size_t)n * hp * wp * c4;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int q = (int)(i % c4);
        size_t p = i / c4;
        const int x = (int)(p % wp) - pad;
        p /= wp;
        const int y = (int)(p % hp) - pad, b = (int)(p / hp);
        out[i] = (x >= 0 && x < wo && y >= 0 && y < ho) ? __ldg(dz + (((size_t)b * ho + y) * wo + x) * c4 + q)
                                                       : make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

// dW[t][ci][co] += sum_p x[p + t][ci] * dz[p][co].  A shared-memory GEMM per tap: grid = (taps, position
// chunks), 256 threads, every thread a 4 x 4 block of the (<= 64 x 64) [ci][co] tile; the positions of a chunk
// are accumulated in fp32 (<= kBwdChunk terms; the launch picks the chunk), the chunks in double (atomics).
constexpr int kBwdChunk = 2048;
__global__ __launch_bounds__(256) void train_conv_bwd_w_kernel(const float *__restrict__ x, const float *__restrict__ dz,
                                                               int n, int h, int wd, int ks, int cin, int cout,
                                                               int chunk, double *__restrict__ red_w)
{
    // 32 positions x 64 channels of x (shifted by the tap) and of dz; rows are 256 bytes, so a thread reads its four
    // input and four output channels as one 16-byte word each (the 16 lanes that share ci0 read the same word)
    __shared__ __align__(16) float xs[32][64], ds[32][64];
    const int ho = h - ks + 1, wo = wd - ks + 1;
    const int t = blockIdx.x, ky = t / ks, kx = t % ks;
    const size_t m = (size_t)n * ho * wo;
    const size_t p0 = (size_t)blockIdx.y * chunk, p1 = p0 + chunk < m ? p0 + chunk : m;
    const int tid = threadIdx.x, ci0 = (tid >> 4) * 4, co0 = (tid & 15) * 4;
    // loader role: channel c of positions pb + lp, lp = tid / 64 + 4 i; (b, yo, xo) of the position walk along
    const int c = tid & 63, lp0 = tid >> 6;
    size_t pl = p0 + lp0;
    int xo = (int)(pl % wo);
    size_t q = pl / wo;
    int yo = (int)(q % ho), b = (int)(q / ho);
    float acc[4][4] = {};
    for (size_t pb = p0; pb < p1; pb += 32) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int pp = lp0 + 4 * i;
            float xv = 0.f, dv = 0.f;
            if (pl < p1) {
                if (c < cin) xv = x[(((size_t)b * h + yo + ky) * wd + xo + kx) * cin + c];
                if (c < cout) dv = dz[pl * cout + c];
            }
            xs[pp][c] = xv;
            ds[pp][c] = dv;
            pl += 4;
            xo += 4;
            while (xo >= wo) {
                xo -= wo;
                if (++yo == ho) { yo = 0; ++b; }
            }
        }
        __syncthreads();
#pragma unroll 8
        for (int pp = 0; pp < 32; ++pp) {
            const float4 a4 = *reinterpret_cast<const float4 *>(&xs[pp][ci0]);
            const float4 b4 = *reinterpret_cast<const float4 *>(&ds[pp][co0]);
            const float a[4] = {a4.x, a4.y, a4.z, a4.w}, bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (ci0 + i < cin && co0 + j < cout)
                atomicAdd(red_w + ((size_t)t * cin + ci0 + i) * cout + co0 + j, (double)acc[i][j]);
}

// The same for a first layer (ks * ks * cin <= 32, i.e. 3 input channels): all taps at once.  With r = t * cin + ci
// the gradient is ONE [r][co] product over the positions, dW[r][co] += sum_p patch[p][r] * dz[p][co]; running it as
// ks * ks products of 64 x 64 tiles would spend 21 of 22 multiplications on zero padding.
__global__ __launch_bounds__(256) void train_conv_bwd_w_small_kernel(const float *__restrict__ x, const float *__restrict__ dz,
                                                                     int n, int h, int wd, int ks, int cin, int cout,
                                                                     int chunk, double *__restrict__ red_w)
{
    __shared__ __align__(16) float xs[32][32], ds[32][64];
    const int ho = h - ks + 1, wo = wd - ks + 1, rows = ks * ks * cin;
    const size_t m = (size_t)n * ho * wo;
    const size_t p0 = (size_t)blockIdx.x * chunk, p1 = p0 + chunk < m ? p0 + chunk : m;
    const int tid = threadIdx.x, r0 = (tid >> 4) * 2, co0 = (tid & 15) * 4;
    // loader roles: dz channel c of positions lp0 + 4 i; patch entry pr of positions pq0 + 8 i
    const int c = tid & 63, lp0 = tid >> 6;
    const int pr = tid & 31, pq0 = tid >> 5;
    const int pt = pr / cin, pci = pr - pt * cin, pky = pt / ks, pkx = pt - pky * ks;
    float acc[2][4] = {};
    for (size_t pb = p0; pb < p1; pb += 32) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int pp = lp0 + 4 * i;
            const size_t p = pb + pp;
            ds[pp][c] = (p < p1 && c < cout) ? dz[p * cout + c] : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int pp = pq0 + 8 * i;
            const size_t p = pb + pp;
            float xv = 0.f;
            if (p < p1 && pr < rows) {
                const int xo = (int)(p % wo);
                const size_t q = p / wo;
                const int yo = (int)(q % ho), b = (int)(q / ho);
                xv = x[(((size_t)b * h + yo + pky) * wd + xo + pkx) * cin + pci];
            }
            xs[pp][pr] = xv;
        }
        __syncthreads();
#pragma unroll 8
        for (int pp = 0; pp < 32; ++pp) {
            const float2 a2 = *reinterpret_cast<const float2 *>(&xs[pp][r0]);
            const float4 b4 = *reinterpret_cast<const float4 *>(&ds[pp][co0]);
            const float a[2] = {a2.x, a2.y}, bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (r0 + i < rows && co0 + j < cout) atomicAdd(red_w + (size_t)(r0 + i) * cout + co0 + j, (double)acc[i][j]);
}

__global__ void train_flip_weights_kernel(const float *__restrict__ w, int ks, int cin, int cout, float *__restrict__ wt)
{
    const int total = ks * ks * cin * cout;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int ci = i % cin;
    int r = i / cin;
    const int co = r % cout;
    r /= cout;
    const int kx = r % ks, ky = r / ks;
    wt[i] = w[(((size_t)(ks - 1 - ky) * ks + (ks - 1 - kx)) * cin + ci) * cout + co];
}

__global__ void train_add_kernel(const float *__restrict__ src, size_t n, float *__restrict__ dst)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] += src[i];
}

// Adam as tf.keras.optimizers.Adam applies it (train.py:80,112): lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t)
__global__ void train_adam_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m,
                                  float *__restrict__ v, size_t n, float lr_t, float b1, float b2, float eps)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float gi = g[i];
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] -= lr_t * mi / (sqrtf(vi) + eps);
}

__global__ void train_reduce_loss_kernel(const float *__restrict__ per_pair, int batch, float *__restrict__ out)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double s = 0.0, c = 0.0;
    for (int b = 0; b < batch; ++b) { s += (double)per_pair[2 * b]; c += (double)per_pair[2 * b + 1]; }
    out[0] = (float)(s / batch);
    out[1] = (float)c;
}

static inline unsigned grid_for(size_t total, int block, int cap)
{
    size_t g = (total + block - 1) / block;
    return (unsigned)(g < 1 ? 1 : (g > (size_t)cap ? (size_t)cap : g));
}

// Wide 3x3 / 64 -> 64 convolutions (the right tower forward, and its backward-data as a forward convolution of the
// padded dz with the flipped kernel) on the tensor cores: the inference kernel conv_tc3.cu (tcgen05, CTA pair,
// split-bf16 operands with fp32 accumulation).  The fp32 NHWC input is converted to its chunk-planar (hi, lo) layout,
// the UMMA weight image is rebuilt from the current weights (they change every step), the output is fp32 NHWC again
// because BatchNorm and the weight gradient read it.  Columns behind the image width hold whatever an earlier, wider
// layer left there: they only reach output pixels the kernel does not store.
static int train_conv_tc(stereo_trainer *tr, Tower &tw, int relu, const float *w, const float *bias, const float *in, int n,
                         int h, int wd, float *out)
{
    stereo_ctx *ctx = tr->ctx;
    const int wp = tc3_pitch(wd);
    const size_t plane_elems = (size_t)n * h * 8 * wp * 8;
    const size_t had = tw.tc_act.bytes;
    STEREO_TRY(tw.tc_act.reserve(2 * plane_elems * sizeof(uint16_t)));
    if (tw.tc_act.bytes != had) STEREO_CUDA_CHECK(cudaMemsetAsync(tw.tc_act.ptr, 0, tw.tc_act.bytes, ctx->stream));
    STEREO_TRY(tw.tc_wimg.reserve(tc3_weight_image_elems() * sizeof(uint16_t)));
    uint16_t *hi = tw.tc_act.as<uint16_t>(), *lo = hi + plane_elems;
    STEREO_TRY(launch_nhwc_to_planar(ctx, in, n, h, wd, wp, 0, hi, lo));
    STEREO_TRY(launch_weight_images3(ctx, w, tw.tc_wimg.ptr));
    return launch_conv_tc3(ctx, hi, lo, n, h, wd, wp, tw.tc_wimg.ptr, bias, nullptr, nullptr, wp, out, relu, 2);
}

// 'valid' convolution + bias (+ ReLU): the tensor-core route above for wide 64 -> 64 layers (STEREO_TRAIN_CONV=ffma
// keeps them on the CUDA cores), else the tiled fp32 kernels of the inference path (conv_ffma.cu) where one is
// instantiated for the shape, the plain kernel above otherwise.  `wbase + w_off` = kernel, `wbase + b_off` = bias.
static int train_conv(stereo_trainer *tr, Tower &tw, int ks, int cin, int cout, int relu, const float *wbase, size_t w_off,
                      size_t b_off, const float *in, int n, int h, int w, float *out)
{
    stereo_ctx *ctx = tr->ctx;
    static const bool no_tc = [] { const char *e = getenv("STEREO_TRAIN_CONV"); return e && (e[0] == 'f' || e[0] == 'p'); }();
    if (!no_tc && ks == 3 && cin == 64 && cout == 64 && w >= kTrainTcMinWidth && h >= 3)
        return train_conv_tc(tr, tw, relu, wbase + w_off, wbase + b_off, in, n, h, w, out);
    const bool tiled = (ks == 3 && cout == 64 && (cin == 3 || cin == 64)) ||
                       (ks == 5 && ((cin == 3 && cout == 32) || (cin == 32 && cout == 32) || (cin == 32 && cout == 64) ||
                                    (cin == 64 && cout == 64)));
    // (STEREO_TRAIN_CONV=plain forces the plain kernel: used to tell round-off from a bug)
    static const bool plain = [] { const char *e = getenv("STEREO_TRAIN_CONV"); return e && e[0] == 'p'; }();
    if (tiled && !plain) {
        LayerDesc L;
        L.ks = ks; L.cin = cin; L.cout = cout; L.relu = relu; L.w_off = w_off; L.b_off = b_off;
        return launch_conv_ffma(ctx, L, wbase, in, n, h, w, out);
    }
    const size_t total = (size_t)n * (h - ks + 1) * (w - ks + 1) * cout;
    train_conv_fwd_kernel<<<grid_for(total, 256, ctx->sm_count * 16), 256, 0, ctx->stream>>>(
        in, wbase + w_off, wbase + b_off, n, h, w, ks, cin, cout, relu, out);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// ------------------------------------------------------------------------------------------
// one tower forward (training mode) and backward
// ------------------------------------------------------------------------------------------
static int tower_forward(stereo_trainer *tr, Tower &tw, const float *patches, int h, int w)
{
    stereo_ctx *ctx = tr->ctx;
    const float *P = tr->params.as<float>();
    double *red = tw.red.as<double>();
    const int n = tr->batch;
    tw.h[0] = h; tw.w[0] = w;
    const float *cur = patches;
    for (int l = 0; l < tr->n_layers; ++l) {
        const TrainLayer &L = tr->layer[l];
        const int ho = tw.h[l] - L.ks + 1, wo = tw.w[l] - L.ks + 1;
        tw.h[l + 1] = ho; tw.w[l + 1] = wo;
        const size_t m = (size_t)n * ho * wo, total = m * L.cout;
        STEREO_TRY(tw.a[l].reserve(total * sizeof(float)));
        STEREO_TRY(tw.x[l + 1].reserve(total * sizeof(float)));
        STEREO_TRY(tw.bn[l].reserve(3 * L.cout * sizeof(float)));
        STEREO_TRY(train_conv(tr, tw, L.ks, L.cin, L.cout, L.relu, P, L.w_off, L.b_off, cur, n, tw.h[l], tw.w[l],
                              tw.a[l].as<float>()));
        const int c4 = L.cout / 4, rows4 = 256 / c4;
        double *part = tw.red_part.as<double>();
        const unsigned gm = grid_for(m, rows4, ctx->sm_count * 2);
        train_moments4_kernel<<<gm, 256, 0, ctx->stream>>>(reinterpret_cast<const float4 *>(tw.a[l].as<float>()), m, c4, part);
        train_sum_partials_kernel<<<(2 * L.cout + 7) / 8, 1024, 0, ctx->stream>>>(part, (int)gm, 2 * L.cout, red, 0);
        train_bn_finalize_kernel<<<1, L.cout, 0, ctx->stream>>>(red, (double)m, L.cout, tw.bn[l].as<float>());
        train_bn_apply4_kernel<<<grid_for(total / 4, 256, ctx->sm_count * 8), 256, 0, ctx->stream>>>(
            reinterpret_cast<const float4 *>(tw.a[l].as<float>()), total / 4, c4, tw.bn[l].as<float>(), P + L.g_off,
            P + L.be_off, reinterpret_cast<float4 *>(tw.x[l + 1].as<float>()));
        ctx->launches += 4;
        cur = tw.x[l + 1].as<float>();
    }
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// dfeat: gradient wrt the tower's output (same shape as x[L]); gradients are ADDED to G
static int tower_backward(stereo_trainer *tr, Tower &tw, const float *patches, const float *dfeat, float *G)
{
    stereo_ctx *ctx = tr->ctx;
    const float *P = tr->params.as<float>();
    double *red = tw.red.as<double>();
    const int n = tr->batch;
    const float *dy = dfeat;
    for (int l = tr->n_layers - 1; l >= 0; --l) {
        const TrainLayer &L = tr->layer[l];
        const int ho = tw.h[l + 1], wo = tw.w[l + 1];
        const size_t m = (size_t)n * ho * wo, total = m * L.cout;
        const float *a = tw.a[l].as<float>();
        const float *xin = l == 0 ? patches : tw.x[l].as<float>();
        float *dz = tw.d_act[0].as<float>();                 // gradient wrt conv output (after the ReLU mask)
        // Scratch: red[0, 2c) BatchNorm sums, red[kRedW, kRedW + wn) dW, then db.  BatchNorm: dgamma, dbeta, da
        const size_t wn = (size_t)L.ks * L.ks * L.cin * L.cout;
        double *red_w = red + kRedW, *red_b = red_w + wn;
        STEREO_CUDA_CHECK(cudaMemsetAsync(red, 0, (kRedW + wn + L.cout) * sizeof(double), ctx->stream));
        const int c4 = L.cout / 4, rows4 = 256 / c4;
        double *part = tw.red_part.as<double>();
        const unsigned gr = grid_for(m, rows4, ctx->sm_count * 2);
        train_bn_bwd_reduce4_kernel<<<gr, 256, 0, ctx->stream>>>(
            reinterpret_cast<const float4 *>(dy), reinterpret_cast<const float4 *>(a), m, c4, tw.bn[l].as<float>(), part);
        train_sum_partials_kernel<<<(2 * L.cout + 7) / 8, 1024, 0, ctx->stream>>>(part, (int)gr, 2 * L.cout, red, 0);
        // (parameter order per layer: kernel, bias, gamma, beta; red[0, c) = dbeta, red[c, 2c) = dgamma)
        train_accumulate2_kernel<<<1, 2 * L.cout, 0, ctx->stream>>>(red + L.cout, L.cout, G + L.g_off, red, L.cout, G + L.be_off);
        // da through the ReLU mask = dz; the same pass sums dz per channel (db)
        const unsigned ga = grid_for(total / 4, 256, ctx->sm_count * 4);
        train_bn_bwd_apply4_kernel<<<ga, 256, 0, ctx->stream>>>(
            reinterpret_cast<const float4 *>(dy), reinterpret_cast<const float4 *>(a), total / 4, c4, (double)m,
            tw.bn[l].as<float>(), P + L.g_off, red, L.relu, reinterpret_cast<float4 *>(dz), part);
        train_sum_partials_kernel<<<(L.cout + 7) / 8, 1024, 0, ctx->stream>>>(part, (int)ga, L.cout, red_b, 0);
        // conv: dW (double accumulation over position chunks)
        // positions per block (accumulated in fp32; the blocks in double): at most kBwdChunk, fewer when that would
        // leave SMs idle (the left tower, the first layer's single product): about eight blocks per SM
        const bool small = L.ks * L.ks * L.cin <= 32 && L.cout <= 64;
        const size_t products = small ? 1 : (size_t)L.ks * L.ks;
        size_t chunk = (m * products / ((size_t)ctx->sm_count * 8) + 31) / 32 * 32;
        chunk = chunk < 128 ? 128 : (chunk > (size_t)kBwdChunk ? (size_t)kBwdChunk : chunk);
        const int chunks = (int)((m + chunk - 1) / chunk);
        // 3x3 / 64 -> 64: positions as the K dimension of tcgen05 MMAs (train_wgrad_tc.cu); STEREO_TRAIN_CONV=ffma: fp32
        static const bool no_tc = [] { const char *e = getenv("STEREO_TRAIN_CONV"); return e && (e[0] == 'f' || e[0] == 'p'); }();
        if (!no_tc && L.ks == 3 && L.cin == 64 && L.cout == 64) {
            STEREO_TRY(tw.wg_scratch.reserve(train_wgrad_tc_scratch_bytes(ctx)));
            STEREO_TRY(launch_train_wgrad_tc(ctx, xin, dz, n, tw.h[l], tw.w[l], tw.wg_scratch.as<float>(), red_w));
        } else if (small)
            train_conv_bwd_w_small_kernel<<<chunks, 256, 0, ctx->stream>>>(xin, dz, n, tw.h[l], tw.w[l], L.ks, L.cin, L.cout,
                                                                           (int)chunk, red_w);
        else
            train_conv_bwd_w_kernel<<<dim3(L.ks * L.ks, chunks), 256, 0, ctx->stream>>>(
                xin, dz, n, tw.h[l], tw.w[l], L.ks, L.cin, L.cout, (int)chunk, red_w);
        // kernel and bias are neighbours in the gradient array, like dW and db in the scratch
        train_accumulate2_kernel<<<(unsigned)((wn + L.cout + 255) / 256), 256, 0, ctx->stream>>>(red_w, (int)wn, G + L.w_off,
                                                                                           red_b, L.cout, G + L.b_off);
        ctx->launches += 7;
        if (l > 0) {
            float *dx = tw.d_act[1].as<float>();
            const int pad = L.ks - 1, hp = ho + 2 * pad, wp = wo + 2 * pad;
            const size_t tpad = (size_t)n * hp * wp * L.cout;
            STEREO_TRY(tw.d_pad.reserve(tpad * sizeof(float)));
            STEREO_TRY(tw.w_flip.reserve((wn + 64) * sizeof(float)));     // flipped kernel, then a zero bias
            train_pad4_kernel<<<grid_for(tpad / 4, 256, ctx->sm_count * 8), 256, 0, ctx->stream>>>(
                reinterpret_cast<const float4 *>(dz), n, ho, wo, c4, pad, reinterpret_cast<float4 *>(tw.d_pad.as<float>()));
            train_flip_weights_kernel<<<(unsigned)((wn + 255) / 256), 256, 0, ctx->stream>>>(P + L.w_off, L.ks, L.cin, L.cout,
                                                                                            tw.w_flip.as<float>());
            STEREO_CUDA_CHECK(cudaMemsetAsync(tw.w_flip.as<float>() + wn, 0, 64 * sizeof(float), ctx->stream));
            STEREO_TRY(train_conv(tr, tw, L.ks, L.cout, L.cin, 0, tw.w_flip.as<float>(), 0, wn, tw.d_pad.as<float>(), n, hp, wp, dx));
            ctx->launches += 2;
            dy = dx;          // d_act[1] carries the gradient down; d_act[0] is rewritten as the next dz
        }
    }
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

// Everything of a step between the uploads and the downloads, enqueued on ctx->stream (and the side stream)
static int enqueue_step(stereo_trainer *tr)
{
    stereo_ctx *ctx = tr->ctx;
    cudaStream_t s = ctx->stream;
    const int B = tr->batch, D = tr->D, p = tr->patch, pw = p + D - 1;
    STEREO_CUDA_CHECK(cudaMemsetAsync(tr->grads.ptr, 0, tr->n_params * sizeof(float), s));
    STEREO_CUDA_CHECK(cudaMemsetAsync(tr->tower[0].grads.ptr, 0, tr->n_params * sizeof(float), s));
    // The left tower (9 x 9 patches: small, latency-bound launches) runs on a stream of its own next to the right one;
    // kernel-class profiling wants every kernel alone on the GPU: one stream while it is on.
    cudaStream_t side = ctx->prof.on ? s : tr->side_stream;
    struct OnStream {                        // launches go to ctx->stream
        stereo_ctx *ctx;
        cudaStream_t saved;
        OnStream(stereo_ctx *c, cudaStream_t st) : ctx(c), saved(c->stream) { c->stream = st; }
        ~OnStream() { ctx->stream = saved; }
    };
    auto fork = [&]() -> int {
        if (side == s) return STEREO_OK;
        STEREO_CUDA_CHECK(cudaEventRecord(tr->ev_fork, s));
        STEREO_CUDA_CHECK(cudaStreamWaitEvent(side, tr->ev_fork, 0));
        return STEREO_OK;
    };
    auto join = [&]() -> int {
        if (side == s) return STEREO_OK;
        STEREO_CUDA_CHECK(cudaEventRecord(tr->ev_join, side));
        STEREO_CUDA_CHECK(cudaStreamWaitEvent(s, tr->ev_join, 0));
        return STEREO_OK;
    };
    STEREO_TRY(fork());
    {
        OnStream left(ctx, side);
        STEREO_TRY(tower_forward(tr, tr->tower[0], tr->tower[0].x[0].as<float>(), p, p));
    }
    STEREO_TRY(tower_forward(tr, tr->tower[1], tr->tower[1].x[0].as<float>(), p, pw));
    STEREO_TRY(join());
    for (int l = 0; l < tr->n_layers; ++l) {
        const TrainLayer &Ly = tr->layer[l];
        train_update_stats_kernel<<<1, Ly.cout, 0, s>>>(tr->tower[0].bn[l].as<float>(), tr->tower[1].bn[l].as<float>(), Ly.cout,
                                                        tr->stats.as<float>() + Ly.mean_off, tr->stats.as<float>() + Ly.var_off);
    }
    ctx->launches += tr->n_layers;
    const int L = tr->n_layers, c = tr->layer[L - 1].cout;
    STEREO_REQUIRE(tr->tower[0].h[L] == 1 && tr->tower[0].w[L] == 1 && tr->tower[1].h[L] == 1 && tr->tower[1].w[L] == D,
                   "unexpected feature shape");
    const float *lf = tr->tower[0].x[L].as<float>(), *rf = tr->tower[1].x[L].as<float>();
    const int nt = 256;
    const size_t sh = (((size_t)D + 1) & ~(size_t)1) * sizeof(float) + 6 * nt * sizeof(float);
    train_loss_kernel<<<B, nt, sh, s>>>(lf, rf, tr->targets.as<float>(), D, c, B, tr->logits.as<float>(),
                                        tr->dlogits.as<float>(), tr->loss_buf.as<float>());
    train_reduce_loss_kernel<<<1, 32, 0, s>>>(tr->loss_buf.as<float>(), B, tr->loss_buf.as<float>() + 2 * B);
    DevBuf &dl = tr->dfeat[0], &dr = tr->dfeat[1];
    STEREO_TRY(dl.reserve((size_t)B * c * sizeof(float)));
    STEREO_TRY(dr.reserve((size_t)B * D * c * sizeof(float)));
    train_loss_bwd_kernel<<<B, c, 0, s>>>(lf, rf, tr->dlogits.as<float>(), D, c, dl.as<float>(), dr.as<float>());
    ctx->launches += 3;
    STEREO_CUDA_CHECK(cudaGetLastError());
    STEREO_TRY(fork());
    {
        OnStream left(ctx, side);
        STEREO_TRY(tower_backward(tr, tr->tower[0], tr->tower[0].x[0].as<float>(), dl.as<float>(), tr->tower[0].grads.as<float>()));
    }
    STEREO_TRY(tower_backward(tr, tr->tower[1], tr->tower[1].x[0].as<float>(), dr.as<float>(), tr->grads.as<float>()));
    STEREO_TRY(join());
    // shared weights: the gradient is the sum of the towers' (right + left, the order of the sequential version)
    train_add_kernel<<<(unsigned)((tr->n_params + 255) / 256), 256, 0, s>>>(tr->tower[0].grads.as<float>(), tr->n_params,
                                                                         tr->grads.as<float>());
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo

using namespace stereo;

extern "C" {

int stereo_train_create(stereo_ctx *ctx, int net, int batch, int disp_range, stereo_trainer **out)
{
    STEREO_REQUIRE(ctx != nullptr && out != nullptr, "null pointer");
    STEREO_REQUIRE(net == STEREO_NET_WIN9 || net == STEREO_NET_WIN37, "unknown net %d", net);
    STEREO_REQUIRE(batch >= 2 && disp_range >= 5 && (disp_range & 1), "need batch >= 2 and an odd disp_range >= 5 (got %d, %d)",
                   batch, disp_range);
    STEREO_CUDA_CHECK(cudaSetDevice(ctx->device));
    stereo_trainer *tr = new stereo_trainer();
    tr->ctx = ctx;
    tr->net = net;
    tr->batch = batch;
    tr->D = disp_range;
    const NetDesc nd = net_desc(net);
    tr->n_layers = nd.n_layers;
    tr->patch = nd.shrink + 1;
    size_t po = 0, so = 0;
    for (int l = 0; l < nd.n_layers; ++l) {
        TrainLayer &L = tr->layer[l];
        L.ks = nd.layer[l].ks; L.cin = nd.layer[l].cin; L.cout = nd.layer[l].cout; L.relu = nd.layer[l].relu;
        L.w_off = po; po += (size_t)L.ks * L.ks * L.cin * L.cout;
        L.b_off = po; po += L.cout;
        L.g_off = po; po += L.cout;
        L.be_off = po; po += L.cout;
        L.mean_off = so; so += L.cout;
        L.var_off = so; so += L.cout;
    }
    tr->n_params = po;
    tr->n_stats = so;
    int rc = STEREO_OK;
    auto need = [&](DevBuf &b, size_t bytes) { if (rc == STEREO_OK) rc = b.reserve(bytes); };
    need(tr->params, po * sizeof(float));
    need(tr->grads, po * sizeof(float));
    need(tr->adam_m, po * sizeof(float));
    need(tr->adam_v, po * sizeof(float));
    need(tr->stats, so * sizeof(float));
    const size_t pw = (size_t)tr->patch + disp_range - 1;
    const size_t act = (size_t)batch * (tr->patch - nd.layer[0].ks + 1) * (pw - nd.layer[0].ks + 1) * 64;
    for (int t = 0; t < 2; ++t) {
        Tower &tw = tr->tower[t];
        const size_t a = t == 0 ? (size_t)batch * (tr->patch - nd.layer[0].ks + 1) * (tr->patch - nd.layer[0].ks + 1) * 64 : act;
        need(tw.red, (po + 1024) * sizeof(double));
        need(tw.red_part, (size_t)ctx->sm_count * 8 * 128 * sizeof(double));      // <= 8 blocks per SM x 2 x 64 channels
        need(tw.d_act[0], a * sizeof(float));
        need(tw.d_act[1], a * sizeof(float));
    }
    need(tr->tower[0].grads, po * sizeof(float));
    need(tr->logits, (size_t)batch * disp_range * sizeof(float));
    need(tr->dlogits, (size_t)batch * disp_range * sizeof(float));
    need(tr->targets, (size_t)batch * disp_range * sizeof(float));
    need(tr->loss_buf, ((size_t)2 * batch + 2) * sizeof(float));
    need(tr->zero_bias, 64 * sizeof(float));
    need(tr->tower[0].x[0], (size_t)batch * tr->patch * tr->patch * 3 * sizeof(float));
    need(tr->tower[1].x[0], (size_t)batch * tr->patch * pw * 3 * sizeof(float));
    if (rc != STEREO_OK) { delete tr; return rc; }
    cudaMemsetAsync(tr->adam_m.ptr, 0, po * sizeof(float), ctx->stream);
    cudaMemsetAsync(tr->adam_v.ptr, 0, po * sizeof(float), ctx->stream);
    cudaMemsetAsync(tr->params.ptr, 0, po * sizeof(float), ctx->stream);
    cudaMemsetAsync(tr->stats.ptr, 0, so * sizeof(float), ctx->stream);
    cudaMemsetAsync(tr->zero_bias.ptr, 0, 64 * sizeof(float), ctx->stream);
    STEREO_TRY(finish(ctx));
    STEREO_CUDA_CHECK(cudaStreamCreateWithFlags(&tr->side_stream, cudaStreamNonBlocking));
    STEREO_CUDA_CHECK(cudaEventCreateWithFlags(&tr->ev_fork, cudaEventDisableTiming));
    STEREO_CUDA_CHECK(cudaEventCreateWithFlags(&tr->ev_join, cudaEventDisableTiming));
    *out = tr;
    return STEREO_OK;
}

int stereo_train_destroy(stereo_trainer *tr)
{
    if (!tr) return STEREO_OK;
    cudaSetDevice(tr->ctx->device);
    cudaStreamSynchronize(tr->ctx->stream);
    if (tr->side_stream) { cudaStreamSynchronize(tr->side_stream); cudaStreamDestroy(tr->side_stream); }
    if (tr->graph_exec) cudaGraphExecDestroy(tr->graph_exec);
    if (tr->ev_fork) cudaEventDestroy(tr->ev_fork);
    if (tr->ev_join) cudaEventDestroy(tr->ev_join);
    DevBuf *bufs[] = {&tr->params, &tr->grads, &tr->adam_m, &tr->adam_v, &tr->stats, &tr->logits, &tr->dlogits, &tr->loss_buf,
                      &tr->targets, &tr->dfeat[0], &tr->dfeat[1], &tr->zero_bias};
    for (DevBuf *b : bufs) b->release();
    for (int t = 0; t < 2; ++t) {
        Tower &tw = tr->tower[t];
        for (int l = 0; l <= kMaxLayers; ++l) tw.x[l].release();
        for (int l = 0; l < kMaxLayers; ++l) { tw.a[l].release(); tw.bn[l].release(); }
        DevBuf *tb[] = {&tw.red, &tw.red_part, &tw.d_act[0], &tw.d_act[1], &tw.d_pad, &tw.w_flip, &tw.tc_act, &tw.tc_wimg,
                        &tw.wg_scratch, &tw.grads};
        for (DevBuf *b : tb) b->release();
    }
    delete tr;
    return STEREO_OK;
}

size_t stereo_train_param_count(const stereo_trainer *tr) { return tr ? tr->n_params : 0; }
size_t stereo_train_stat_count(const stereo_trainer *tr) { return tr ? tr->n_stats : 0; }
int stereo_train_patch_size(const stereo_trainer *tr) { return tr ? tr->patch : 0; }

int stereo_train_set_state(stereo_trainer *tr, const float *params, const float *stats, const float *adam_m,
                           const float *adam_v, long long step)
{
    STEREO_REQUIRE(tr && params && stats, "null pointer");
    cudaStream_t s = tr->ctx->stream;
    STEREO_CUDA_CHECK(cudaMemcpyAsync(tr->params.ptr, params, tr->n_params * sizeof(float), cudaMemcpyHostToDevice, s));
    STEREO_CUDA_CHECK(cudaMemcpyAsync(tr->stats.ptr, stats, tr->n_stats * sizeof(float), cudaMemcpyHostToDevice, s));
    if (adam_m) STEREO_CUDA_CHECK(cudaMemcpyAsync(tr->adam_m.ptr, adam_m, tr->n_params * sizeof(float), cudaMemcpyHostToDevice, s));
    else STEREO_CUDA_CHECK(cudaMemsetAsync(tr->adam_m.ptr, 0, tr->n_params * sizeof(float), s));
    if (adam_v) STEREO_CUDA_CHECK(cudaMemcpyAsync(tr->adam_v.ptr, adam_v, tr->n_params * sizeof(float), cudaMemcpyHostToDevice, s));
    else STEREO_CUDA_CHECK(cudaMemsetAsync(tr->adam_v.ptr, 0, tr->n_params * sizeof(float), s));
    tr->step = step;
    return finish(tr->ctx);
}

int stereo_train_get_state(stereo_trainer *tr, float *params, float *stats, float *adam_m, float *adam_v, long long *step)
{
    STEREO_REQUIRE(tr != nullptr, "null pointer");
    cudaStream_t s = tr->ctx->stream;
    if (params) STEREO_CUDA_CHECK(cudaMemcpyAsync(params, tr->params.ptr, tr->n_params * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (stats) STEREO_CUDA_CHECK(cudaMemcpyAsync(stats, tr->stats.ptr, tr->n_stats * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (adam_m) STEREO_CUDA_CHECK(cudaMemcpyAsync(adam_m, tr->adam_m.ptr, tr->n_params * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (adam_v) STEREO_CUDA_CHECK(cudaMemcpyAsync(adam_v, tr->adam_v.ptr, tr->n_params * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (step) *step = tr->step;
    return finish(tr->ctx);
}

// Device address of the flat gradient array (n_params floats): a data-parallel host all-reduces it in
// place between stereo_train_forward_backward and stereo_train_apply.
int stereo_train_grads_dev(stereo_trainer *tr, float **grads)
{
    STEREO_REQUIRE(tr && grads, "null pointer");
    *grads = tr->grads.as<float>();
    return STEREO_OK;
}

// Host pointers: left [B][p][p][3], right [B][p][p + D - 1][3], targets [B][D] (fp32, already normalised
// patches as data_handler.py:60-85 produces them).  Computes loss (mean over the batch), the number of
// patches whose arg max is within 3 of the label's (train.py:117-119) and the gradients (kept on the device;
// copied to `grads_out` if not null).  Moving BatchNorm statistics are updated by both tower passes.
int stereo_train_forward_backward(stereo_trainer *tr, const float *left, const float *right, const float *targets,
                                  float *loss, int *correct3, float *grads_out)
{
    STEREO_REQUIRE(tr && left && right && targets, "null pointer");
    stereo_ctx *ctx = tr->ctx;
    STEREO_CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    const int B = tr->batch, D = tr->D, p = tr->patch, pw = p + D - 1;
    STEREO_CUDA_CHECK(cudaMemcpyAsync(tr->tower[0].x[0].ptr, left, (size_t)B * p * p * 3 * sizeof(float), cudaMemcpyHostToDevice, s));
    STEREO_CUDA_CHECK(cudaMemcpyAsync(tr->tower[1].x[0].ptr, right, (size_t)B * p * pw * 3 * sizeof(float), cudaMemcpyHostToDevice, s));
    STEREO_CUDA_CHECK(cudaMemcpyAsync(tr->targets.ptr, targets, (size_t)B * D * sizeof(float), cudaMemcpyHostToDevice, s));
    // The step is ~150 small launches on two streams with fixed shapes and buffers: from the second call on (every buffer
    // has its size by then) it is captured once into a CUDA graph and replayed.  STEREO_TRAIN_GRAPH=0 and kernel-class
    // profiling keep the plain launches.
    static const bool graphs_on = [] { const char *e = getenv("STEREO_TRAIN_GRAPH"); return !(e && e[0] == '0'); }();
    if (!graphs_on || ctx->prof.on || !tr->warm) {
        STEREO_TRY(enqueue_step(tr));
        tr->warm = true;
    } else {
        if (!tr->graph_exec) {
            const long long launches0 = ctx->launches;
            STEREO_CUDA_CHECK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
            const int rc = enqueue_step(tr);
            cudaGraph_t graph = nullptr;
            const cudaError_t ce = cudaStreamEndCapture(s, &graph);
            tr->graph_launches = ctx->launches - launches0;
            ctx->launches = launches0;
            if (rc != STEREO_OK || ce != cudaSuccess) {
                if (graph) cudaGraphDestroy(graph);
                STEREO_TRY(rc);
                STEREO_CUDA_CHECK(ce);
            }
            const cudaError_t ie = cudaGraphInstantiate(&tr->graph_exec, graph, 0);
            cudaGraphDestroy(graph);
            STEREO_CUDA_CHECK(ie);
        }
        STEREO_CUDA_CHECK(cudaGraphLaunch(tr->graph_exec, s));
        ctx->launches += tr->graph_launches;
    }
    float host[2] = {0.f, 0.f};
    STEREO_CUDA_CHECK(cudaMemcpyAsync(host, tr->loss_buf.as<float>() + 2 * B, sizeof(host), cudaMemcpyDeviceToHost, s));
    if (grads_out)
        STEREO_CUDA_CHECK(cudaMemcpyAsync(grads_out, tr->grads.ptr, tr->n_params * sizeof(float), cudaMemcpyDeviceToHost, s));
    STEREO_TRY(finish(ctx));
    if (loss) *loss = host[0];
    if (correct3) *correct3 = (int)(host[1] + 0.5f);
    return STEREO_OK;
}

// One Adam step (beta1 0.9, beta2 0.999, epsilon 1e-7: tf.keras.optimizers.Adam defaults, train.py:80) with the
// gradients of the last forward_backward call (or `grads_dev`, a device array, if not null).
int stereo_train_apply(stereo_trainer *tr, const float *grads_dev, float lr)
{
    STEREO_REQUIRE(tr != nullptr, "null pointer");
    stereo_ctx *ctx = tr->ctx;
    STEREO_CUDA_CHECK(cudaSetDevice(ctx->device));
    tr->step += 1;
    const double b1 = 0.9, b2 = 0.999;
    const float lr_t = (float)((double)lr * sqrt(1.0 - pow(b2, (double)tr->step)) / (1.0 - pow(b1, (double)tr->step)));
    const float *g = grads_dev ? grads_dev : tr->grads.as<float>();
    train_adam_kernel<<<(unsigned)((tr->n_params + 255) / 256), 256, 0, ctx->stream>>>(
        tr->params.as<float>(), g, tr->adam_m.as<float>(), tr->adam_v.as<float>(), tr->n_params, lr_t, (float)b1, (float)b2, 1e-7f);
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return finish(ctx);
}

}  // extern "C"
