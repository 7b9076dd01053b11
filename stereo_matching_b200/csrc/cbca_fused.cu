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

// VARIANT 0: the row loop loads a row and uses it at once (round 1).  VARIANT 1 (default): software pipeline - the
// ten costs and the four arm records of row r + 1 are requested right after the row sums of row r have consumed
// the registers of the previous request, so their latency runs under the column sums, the division and the store
// of output row r - 4 instead of in front of the next row sums.  Same additions in the same order.
template <int MIN_BLOCKS, int VARIANT>
__global__ __launch_bounds__(256, MIN_BLOCKS) void cbca_fused_kernel(const uint32_t *__restrict__ rec0, const uint32_t *__restrict__ rec1,
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

    // costs and arm records of the row whose sums come next (VARIANT 1: requested one iteration ahead)
    float v[kFN];
    uint32_t own_n[kFB], oth_n[kFB];
    auto request_row = [&](int r) {
        const int vbase = r * row_elems, rbase = r * pitch;       // 32-bit element indices: one IMAD.WIDE per load
#pragma unroll
        for (int j = 0; j < kFN; ++j) v[j] = __ldg(vol + (vbase + offs[j]));
#pragma unroll
        for (int i = 0; i < kFB; ++i) {
            own_n[i] = __ldg(rec0 + (rbase + xb + i));
            oth_n[i] = __ldg(rec1 + (rbase + xm[i]));
        }
    };
    const int r_first = max(0, y0 - kFR), r_end = min(h, y1 + kFR);       // rows whose sums are formed: [r_first, r_end)
    if (VARIANT == 1) request_row(r_first);

    for (int base = y0 - kFR; base < y1 + kFR; base += kFW) {
#pragma unroll
        for (int u = 0; u < kFW; ++u) {
            const int r = base + u;
            if (r >= y1 + kFR) break;
            if (r >= 0 && r < h) {
                // ---- row-segment sums of image row r (cbca_h4_kernel), into window slot u ----
                if (VARIANT == 0) request_row(r);
#pragma unroll
                for (int i = 0; i < kFB; ++i) {
                    const uint32_t cm = own_n[i] & oth_n[i];         // both arms reach the column
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
                if (VARIANT == 1 && r + 1 < r_end) request_row(r + 1);
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

}  // namespace

// x0c / x1c: int16 arms [4][h][w] of the two views; rec: scratch for 2 * h * pitch records
int launch_cbca_fused(stereo_ctx *ctx, const short *x0c, const short *x1c, const float *vol, float *out, int h, int w,
                      int D, int direction, int L1)
{
    STEREO_REQUIRE((size_t)h * w * D < (size_t)1 << 31, "volume too large for 32-bit indexing");
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
    // STEREO_CBCA_KERNEL=1 keeps the round-1 form (three blocks per SM, every row loaded and used at once) as the
    // cross-check; STEREO_CBCA_KERNEL=2: that form compiled for four blocks per SM (64 registers, a few spills)
    static const int variant = [] { const char *e = getenv("STEREO_CBCA_KERNEL"); return e ? atoi(e) : 0; }();
    const dim3 grid((w + kFB - 1) / kFB, bands);
    if (variant == 1) cbca_fused_kernel<3, 0><<<grid, nt, 0, ctx->stream>>>(rec0, rec1, pitch, vol, out, h, w, D, direction, rows_per_band);
    else if (variant == 2) cbca_fused_kernel<4, 0><<<grid, nt, 0, ctx->stream>>>(rec0, rec1, pitch, vol, out, h, w, D, direction, rows_per_band);
    else if (variant == 3) cbca_fused_kernel<4, 1><<<grid, nt, 0, ctx->stream>>>(rec0, rec1, pitch, vol, out, h, w, D, direction, rows_per_band);
    else cbca_fused_kernel<3, 1><<<grid, nt, 0, ctx->stream>>>(rec0, rec1, pitch, vol, out, h, w, D, direction, rows_per_band);
    ctx->launches += 3;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
