// adcensus.cbca (smooth/nyu.lua:37-40) in ONE pass over the volume, for arm lengths up to L1 = 5.
//
// The separable form in post.cu writes the row-segment sums (hsum, fp32) and lengths (hcnt, uint8) of every voxel
// to HBM and reads them back nine rows at a time: 1.84 GB of DRAM traffic and ~260 instructions per voxel for
// 0.85 GB of algorithmic bytes.  Here a thread owns one disparity of TWO neighbouring pixels and walks down a band
// of image rows: for every row it forms the two row-segment sums from 10 loaded costs (as cbca_h4_kernel does) and
// keeps the last nine rows of (sum, length) in registers; the output of row y - 4 is the sum over the rows its
// vertical arms admit (as cbca_v4_kernel does).  Same additions in the same order as the two-pass kernels, so the
// results are bit-identical to them; nothing but the volume is read and nothing but the result written.
//
// Arms are repacked once per call into one 32-bit record per pixel: bit b of byte 0 says whether the pixel's
// horizontal arms reach column offset -4..-1, +1..+4 (b = 0..7), byte 1 the same for row offsets.  Relative to the
// pixel, the matched pixel's arms "x1c[xm] - d*dir - x" are simply ITS relative arms, so the intersection of the two
// support intervals is the bitwise AND of two records, the segment length a population count, and every
// conditional addition one bit test - no arm arithmetic is left in the inner loops.
#include <algorithm>

#include "common.cuh"

namespace stereo {

namespace {

constexpr int kFR = 4;                   // window half width = L1 - 1
constexpr int kFB = 2;                   // pixels per thread (4: 118 registers, two CTAs per SM, slower; 1: slower still)
constexpr int kFN = kFB + 2 * kFR;       // costs loaded per row
constexpr int kFW = 2 * kFR + 1;         // rows kept

// bit of offset jj (-4..-1, 1..4) inside a mask byte
__host__ __device__ constexpr uint32_t off_bit(int jj) { return 1u << (jj < 0 ? jj + kFR : jj + kFR - 1); }

__global__ void cbca_pack_arms_kernel(const short *__restrict__ arms, int h, int w, int pitch, uint32_t *__restrict__ rec)
{
    const int total = h * pitch;
    const int hw = h * w;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
        const int y = idx / pitch, x = idx - y * pitch;
        uint32_t r = 0;                                      // padding columns: the pixel alone
        if (x < w) {
            const int pix = y * w + x;
            // exclusive arm ends relative to the pixel: offsets strictly between them belong to the support region
            const int l = arms[pix] - x, rt = arms[hw + pix] - x, u = arms[2 * hw + pix] - y, dn = arms[3 * hw + pix] - y;
#pragma unroll
            for (int k = 1; k <= kFR; ++k) {
                if (-k > l) r |= off_bit(-k);
                if (k < rt) r |= off_bit(k);
                if (-k > u) r |= off_bit(-k) << 8;
                if (k < dn) r |= off_bit(k) << 8;
            }
        }
        rec[idx] = r;
    }
}

// sum += v if (mask & bit): one bit test and one predicated add
__device__ __forceinline__ void add_if(float &sum, float v, uint32_t mask, uint32_t bit)
{
    asm("{\n\t.reg .pred p;\n\t.reg .b32 t;\n\tand.b32 t, %2, %3;\n\tsetp.ne.u32 p, t, 0;\n\t@p add.rn.f32 %0, %0, %1;\n\t}"
        : "+f"(sum) : "f"(v), "r"(mask), "r"(bit));
}
__device__ __forceinline__ void add_if(float &sum, int &cnt, float v, int c, uint32_t mask, uint32_t bit)
{
    asm("{\n\t.reg .pred p;\n\t.reg .b32 t;\n\tand.b32 t, %4, %5;\n\tsetp.ne.u32 p, t, 0;\n\t@p add.rn.f32 %0, %0, %2;\n\t"
        "@p add.s32 %1, %1, %3;\n\t}"
        : "+f"(sum), "+r"(cnt) : "f"(v), "r"(c), "r"(mask), "r"(bit));
}

// The round-1 form of the kernel: every row is loaded and used at once, the arm records of an output row are
// re-read when the row is written.  Kept as the cross-check of the default below (STEREO_CBCA_KERNEL=1).
__global__ __launch_bounds__(256, 3) void cbca_fused_r1_kernel(const uint32_t *__restrict__ rec0, const uint32_t *__restrict__ rec1,
                                                               int pitch, const float *__restrict__ vol, float *__restrict__ out,
                                                               int h, int w, int D, int direction, int rows_per_band)
{
    const int d = threadIdx.x;
    if (d >= D) return;
    const int xb = blockIdx.x * kFB;
    const int y0 = blockIdx.y * rows_per_band, y1 = min(h, y0 + rows_per_band);
    const int shift = d * direction;

    int offs[kFN];                           // element offsets of the 10 columns inside a row (clamped: loads stay in the row)
#pragma unroll
    for (int j = 0; j < kFN; ++j) offs[j] = min(max(xb - kFR + j, 0), w - 1) * D + d;
    bool pix_ok[kFB], match_ok[kFB];
    int xm[kFB];
#pragma unroll
    for (int i = 0; i < kFB; ++i) {
        const int x = xb + i;
        pix_ok[i] = x < w;
        match_ok[i] = (unsigned)(x + shift) < (unsigned)w;
        xm[i] = min(max(x + shift, 0), w - 1);
    }

    float hs[kFW][kFB];
    int hc[kFW][kFB];
#pragma unroll
    for (int u = 0; u < kFW; ++u)
#pragma unroll
        for (int i = 0; i < kFB; ++i) { hs[u][i] = 0.f; hc[u][i] = 0; }
    const int row_elems = w * D;

    for (int base = y0 - kFR; base < y1 + kFR; base += kFW) {
#pragma unroll
        for (int u = 0; u < kFW; ++u) {
            const int r = base + u;
            if (r >= y1 + kFR) break;
            if (r >= 0 && r < h) {
                // ---- row-segment sums of image row r (cbca_h4_kernel), into window slot u ----
                const int vbase = r * row_elems, rbase = r * pitch;       // 32-bit element indices: one IMAD.WIDE per load
                float v[kFN];
#pragma unroll
                for (int j = 0; j < kFN; ++j) v[j] = __ldg(vol + (vbase + offs[j]));
                uint32_t own_r[kFB];
#pragma unroll
                for (int i = 0; i < kFB; ++i) own_r[i] = __ldg(rec0 + (rbase + xb + i));
#pragma unroll
                for (int i = 0; i < kFB; ++i) {
                    const uint32_t cm = own_r[i] & __ldg(rec1 + (rbase + xm[i]));   // both arms reach the column
                    float sum = 0.f;
#pragma unroll
                    for (int jj = -kFR; jj <= kFR; ++jj) {           // columns x-R .. x+R, ascending
                        if (jj == 0) sum = __fadd_rn(sum, v[kFR + i]);
                        else add_if(sum, v[kFR + i + jj], cm, off_bit(jj));
                    }
                    // unmatched disparities are copied through: keep the cost itself, length 0
                    hs[u][i] = match_ok[i] ? sum : v[kFR + i];
                    hc[u][i] = match_ok[i] ? __popc(cm & 0xffu) + 1 : 0;
                }
            }
            const int y = r - kFR;
            if (y >= y0 && y < y1) {
                // ---- output row y: rows y-R .. y+R sit in slots u+1 .. u+9 (mod 9) (cbca_v4_kernel) ----
                const int rbase = y * pitch, obase = y * row_elems;
                uint32_t own_r[kFB];
#pragma unroll
                for (int i = 0; i < kFB; ++i) own_r[i] = __ldg(rec0 + (rbase + xb + i));
#pragma unroll
                for (int i = 0; i < kFB; ++i) {
                    if (!pix_ok[i]) continue;
                    const uint32_t cm = (own_r[i] & __ldg(rec1 + (rbase + xm[i]))) >> 8;   // both arms reach the row
                    float sum = 0.f;
                    int cnt = 0;
#pragma unroll
                    for (int jj = -kFR; jj <= kFR; ++jj) {           // rows y-R .. y+R, ascending
                        const int slot = (u + 1 + jj + kFR) % kFW;
                        if (jj == 0) {
                            sum = __fadd_rn(sum, hs[slot][i]);
                            cnt += hc[slot][i];
                        } else {
                            add_if(sum, cnt, hs[slot][i], hc[slot][i], cm, off_bit(jj));
                        }
                    }
                    const float centre = hs[(u + 1 + kFR) % kFW][i];
                    out[obase + offs[kFR + i]] = match_ok[i] ? __fdiv_rn(sum, (float)cnt) : centre;
                }
            }
        }
    }
}

// The default form.  What the ncu source page of the round-1 form showed (profiles/r02_ncu_summary.md): 44 % of the
// stall samples were waits for global loads, 30 % of all samples on ONE instruction per unrolled copy - the arm
// records of the output row, re-read right before their use - and half of the executed instructions were integer
// address arithmetic (one IMAD.WIDE per load, D a run-time value).  Here
//   * the costs and records of row r + 1 are requested right after the row sums of row r have consumed the
//     registers of the previous request (software pipeline): their latency runs under the column sums, the
//     division and the store of output row r - 4;
//   * the column masks of a row are kept from the moment its records are read for the row sums (a ring of nine
//     16-bit entries in five registers) - the output phase loads nothing;
//   * blocks whose 10-column window lies inside the image (all but five block columns) address it as
//     row pointer + j * D with D a compile-time constant for the common ranges, so that the ten loads are one
//     pointer and ten immediates (EDGE = false; the registers of the ten clamped offsets pay for the mask ring);
//     edge blocks keep the clamped offsets (EDGE = true, a launch of its own).
// Same additions in the same order as the round-1 form: bit-identical (tests/test_gpu_variants.py).
// DINC / Din_rt: floats between the [d] rows of neighbouring pixels of the INPUT volume (the output always has D):
// the SGM result of a range that is not a multiple of 4 comes with rows padded to one (post_api.cu).
template <int DC, int DINC, bool EDGE>
__global__ __launch_bounds__(256, EDGE ? 2 : 3) void cbca_fused_kernel(const uint32_t *__restrict__ rec0, const uint32_t *__restrict__ rec1,
                                                                       int pitch, const float *__restrict__ vol, float *__restrict__ out,
                                                                       int h, int w, int D_rt, int Din_rt, int direction, int rows_per_band,
                                                                       int xb_first, int n_left, int xb_right)
{
    const int D = DC ? DC : D_rt;
    const int Din = DINC ? DINC : Din_rt;
    const int d = threadIdx.x;
    if (d >= D) return;
    // interior launch: consecutive blocks from xb_first; edge launch: n_left blocks from pixel 0, the rest from xb_right
    const int xb = EDGE ? ((int)blockIdx.x < n_left ? (int)blockIdx.x * kFB : xb_right + ((int)blockIdx.x - n_left) * kFB)
                        : xb_first + (int)blockIdx.x * kFB;
    const int y0 = blockIdx.y * rows_per_band, y1 = min(h, y0 + rows_per_band);
    const int shift = d * direction;

    int offs[EDGE ? kFN : 1];                // EDGE: element offsets of the 10 columns, clamped into the row
    if (EDGE) {
#pragma unroll
        for (int j = 0; j < kFN; ++j) offs[j] = min(max(xb - kFR + j, 0), w - 1) * Din + d;
    }
    const int col0 = (xb - kFR) * Din + d;   // !EDGE: input offset of column xb - R; column j is + j * Din
    bool pix_ok[kFB], match_ok[kFB];
    int xm[kFB];
#pragma unroll
    for (int i = 0; i < kFB; ++i) {
        const int x = xb + i;
        pix_ok[i] = x < w;
        match_ok[i] = (unsigned)(x + shift) < (unsigned)w;
        xm[i] = min(max(x + shift, 0), w - 1);
    }

    float hs[kFW][kFB];
    int hc[kFW][kFB];
    uint32_t vring[(kFW + 1) / 2];           // column masks of the last nine rows: entry u = two pixels x 8 bits
#pragma unroll
    for (int u = 0; u < kFW; ++u)
#pragma unroll
        for (int i = 0; i < kFB; ++i) { hs[u][i] = 0.f; hc[u][i] = 0; }
#pragma unroll
    for (int u = 0; u < (kFW + 1) / 2; ++u) vring[u] = 0;
    const int row_elems = w * D, in_row_elems = w * Din;

    // costs and arm records of the row whose sums come next, requested one iteration ahead
    float v[kFN];
    uint32_t own_n[kFB], oth_n[kFB];
    auto request_row = [&](int r) {
        if (EDGE) {
            const int vbase = r * in_row_elems;
#pragma unroll
            for (int j = 0; j < kFN; ++j) v[j] = __ldg(vol + (vbase + offs[j]));
        } else {
            const float *prow = vol + ((size_t)r * in_row_elems + col0);
#pragma unroll
            for (int j = 0; j < kFN; ++j) v[j] = __ldg(prow + j * Din);
        }
        const int rbase = r * pitch;
#pragma unroll
        for (int i = 0; i < kFB; ++i) {
            own_n[i] = __ldg(rec0 + (rbase + xb + i));
            oth_n[i] = __ldg(rec1 + (rbase + xm[i]));
        }
    };
    const int r_first = max(0, y0 - kFR), r_end = min(h, y1 + kFR);       // rows whose sums are formed: [r_first, r_end)
    request_row(r_first);

    for (int base = y0 - kFR; base < y1 + kFR; base += kFW) {
#pragma unroll
        for (int u = 0; u < kFW; ++u) {
            const int r = base + u;
            if (r >= y1 + kFR) break;
            if (r >= 0 && r < h) {
                // ---- row-segment sums of image row r (cbca_h4_kernel), into window slot u ----
                uint32_t vm = 0;
#pragma unroll
                for (int i = 0; i < kFB; ++i) {
                    const uint32_t cm = own_n[i] & oth_n[i];         // both arms reach the column (byte 0) / the row (byte 1)
                    vm |= ((cm >> 8) & 0xffu) << (8 * i);
                    float sum = 0.f;
#pragma unroll
                    for (int jj = -kFR; jj <= kFR; ++jj) {           // columns x-R .. x+R, ascending
                        if (jj == 0) sum = __fadd_rn(sum, v[kFR + i]);
                        else add_if(sum, v[kFR + i + jj], cm, off_bit(jj));
                    }
                    // unmatched disparities are copied through: keep the cost itself, length 0
                    hs[u][i] = match_ok[i] ? sum : v[kFR + i];
                    hc[u][i] = match_ok[i] ? __popc(cm & 0xffu) + 1 : 0;
                }
                // the row's column masks wait in the ring until the row is written, four iterations on
                vring[u / 2] = (u & 1) ? ((vring[u / 2] & 0xffffu) | (vm << 16)) : ((vring[u / 2] & 0xffff0000u) | vm);
                if (r + 1 < r_end) request_row(r + 1);
            }
            const int y = r - kFR;
            if (y >= y0 && y < y1) {
                // ---- output row y: rows y-R .. y+R sit in slots u+1 .. u+9 (mod 9) (cbca_v4_kernel) ----
                const int uy = (u + kFW - kFR) % kFW;                // ring entry of row y (a constant after unrolling)
                const uint32_t vm = (vring[uy / 2] >> (16 * (uy & 1))) & 0xffffu;
                const int obase = y * row_elems;
#pragma unroll
                for (int i = 0; i < kFB; ++i) {
                    if (!pix_ok[i]) continue;
                    const uint32_t cm = (vm >> (8 * i)) & 0xffu;     // both arms reach the row
                    float sum = 0.f;
                    int cnt = 0;
#pragma unroll
                    for (int jj = -kFR; jj <= kFR; ++jj) {           // rows y-R .. y+R, ascending
                        const int slot = (u + 1 + jj + kFR) % kFW;
                        if (jj == 0) {
                            sum = __fadd_rn(sum, hs[slot][i]);
                            cnt += hc[slot][i];
                        } else {
                            add_if(sum, cnt, hs[slot][i], hc[slot][i], cm, off_bit(jj));
                        }
                    }
                    const float centre = hs[(u + 1 + kFR) % kFW][i];
                    const int oidx = obase + (xb + i) * D + d;
                    out[oidx] = match_ok[i] ? __fdiv_rn(sum, (float)cnt) : centre;
                }
            }
        }
    }
}

}  // namespace

// x0c / x1c: int16 arms [4][h][w] of the two views; rec: scratch for 2 * h * pitch records
int launch_cbca_fused(stereo_ctx *ctx, const short *x0c, const short *x1c, const float *vol, float *out, int h, int w,
                      int D, int direction, int L1, int in_pitch)
{
    STEREO_REQUIRE((size_t)h * w * (size_t)std::max(D, in_pitch) < (size_t)1 << 31, "volume too large for 32-bit indexing");
    STEREO_REQUIRE(in_pitch == 0 || in_pitch >= D, "input pitch %d smaller than the disparity range %d", in_pitch, D);
    STEREO_REQUIRE(D <= 256 && L1 - 1 <= kFR, "fused CBCA: D <= 256 and L1 <= %d, got D=%d L1=%d", kFR + 1, D, L1);
    const int pitch = (w + 3) / 4 * 4 + 4;
    DevBuf &buf = ctx->cbca_rec[direction > 0];
    STEREO_TRY(buf.reserve((size_t)2 * h * pitch * sizeof(uint32_t)));
    uint32_t *rec0 = buf.as<uint32_t>(), *rec1 = rec0 + (size_t)h * pitch;
    ProfScope prof(ctx, STEREO_K_CBCA);
    const int pack_blocks = std::min((h * pitch + 255) / 256, 148 * 8);
    cbca_pack_arms_kernel<<<pack_blocks, 256, 0, ctx->stream>>>(x0c, h, w, pitch, rec0);
    cbca_pack_arms_kernel<<<pack_blocks, 256, 0, ctx->stream>>>(x1c, h, w, pitch, rec1);
    const int nt = (D + 31) / 32 * 32;
    const int bands = std::max(1, (h + 62) / 63);
    const int rows_per_band = (h + bands - 1) / bands;
    // STEREO_CBCA_KERNEL=1 keeps the round-1 form as the cross-check
    static const bool round1 = [] { const char *e = getenv("STEREO_CBCA_KERNEL"); return e && atoi(e) == 1; }();
    const int nbx = (w + kFB - 1) / kFB;
    if (round1 && (in_pitch == 0 || in_pitch == D)) {
        cbca_fused_r1_kernel<<<dim3(nbx, bands), nt, 0, ctx->stream>>>(rec0, rec1, pitch, vol, out, h, w, D, direction, rows_per_band);
    } else {
        // interior blocks: the window xb - R .. xb + 1 + R lies inside the image
        const int n_left = std::min(nbx, (kFR + kFB - 1) / kFB);                       // blocks with xb < R
        int first_right = n_left;                                                      // first block with xb + kFB + R > w
        while (first_right < nbx && first_right * kFB + kFB + kFR <= w) ++first_right;
        const int n_int = first_right - n_left, n_edge = nbx - n_int;
        const int Din = in_pitch > 0 ? in_pitch : D;
        auto launch = [&](auto interior, auto edge) {
            if (n_int > 0)
                interior<<<dim3(n_int, bands), nt, 0, ctx->stream>>>(rec0, rec1, pitch, vol, out, h, w, D, Din, direction,
                                                                     rows_per_band, n_left * kFB, 0, 0);
            // the few edge block columns get short bands: with the interior's 63-row bands their 24 blocks ran for
            // 80 us on 24 SMs behind a 520 us interior launch (ncu launch list); the extra halo rows cost nothing
            // next to that
            const int edge_rows = 12, edge_bands = (h + edge_rows - 1) / edge_rows;
            if (n_edge > 0)
                edge<<<dim3(n_edge, edge_bands), nt, 0, ctx->stream>>>(rec0, rec1, pitch, vol, out, h, w, D, Din, direction,
                                                                       edge_rows, 0, n_left, first_right * kFB);
        };
        // disparity ranges known at compile time: the benchmark's 228, the reference's default 201 (whose SGM result
        // comes with rows of 204 floats); any other at run time
        if (D == 228 && Din == 228) launch(cbca_fused_kernel<228, 228, false>, cbca_fused_kernel<0, 0, true>);
        else if (D == 201 && Din == 201) launch(cbca_fused_kernel<201, 201, false>, cbca_fused_kernel<0, 0, true>);
        else if (D == 201 && Din == 204) launch(cbca_fused_kernel<201, 204, false>, cbca_fused_kernel<0, 0, true>);
        else launch(cbca_fused_kernel<0, 0, false>, cbca_fused_kernel<0, 0, true>);
        ctx->launches += (n_int > 0 && n_edge > 0) ? 1 : 0;
    }
    ctx->launches += 3;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
