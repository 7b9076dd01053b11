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
//
// RING = true (aligned volumes, the default): the costs do not come through per-thread global loads - 12 of them in
// flight per thread were all that hid the DRAM latency at 24 warps per SM - but through a ring of shared-memory row
// slots filled by bulk copies: the 10 pixel columns a block needs from one image row are ONE contiguous run of
// 10 x D floats ([x][D] layout), one cp.async.bulk per row, up to eight rows ahead, completion on an mbarrier.  A warp
// reads its 10 values of a row from shared memory (conflict-free: consecutive disparities), releases the slot, and
// does the same additions as before.  One elected thread of warp 0 refills a slot as soon as every warp has released
// it; no producer warp (it would cost a third of the occupancy at 80 registers).
#include <algorithm>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {
using namespace tc;

namespace {

constexpr int kRingMaxSlots = 8;

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

// order-preserving integer key of a float (-0 counted as +0): a < b  <=>  key(a) < key(b) as unsigned
__device__ __forceinline__ uint32_t order_key(float v)
{
    const uint32_t u = __float_as_uint(__fadd_rn(v, 0.f));
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

constexpr int kFMaxBandRows = 64;
constexpr int kFMaxWarps = 8;

// RING: costs through the shared-memory ring of row slots.  MODE 0: the volume only; 1: + moments and / or arg MAX
// map of the output (first CBCA of the chain); 2: + arg MIN map (second CBCA).  The row loop is unrolled nine
// times (the register window must not move), so everything inside it is kept small: slot and phase of the ring
// are running counters (no divisions), the extras are compiled in only where they are used - an earlier version
// with all of it inlined nine times was four times the size and ran out of the instruction cache.
template <bool RING, int MODE>
__global__ __launch_bounds__(256) void cbca_fused_kernel(const uint32_t *__restrict__ rec0, const uint32_t *__restrict__ rec1,
                                                         int pitch, const float *__restrict__ vol, float *__restrict__ out,
                                                         int h, int w, int D, int direction, int rows_per_band,
                                                         double *__restrict__ moments, uint8_t *__restrict__ wta,
                                                         int take_min, int n_slots)
{
    // the lanes past D of the last warp shadow disparity D - 1 (same loads, nothing stored or counted) so that
    // warp-wide operations see whole warps
    const bool active = (int)threadIdx.x < D;
    const int d = active ? (int)threadIdx.x : D - 1;
    const int xb = blockIdx.x * kFB;
    const int y0 = blockIdx.y * rows_per_band, y1 = min(h, y0 + rows_per_band);
    const int shift = d * direction;

    // EXTRAS: per (row of the band, pixel, warp) the packed (value key, disparity) of the warp's best disparity:
    // plain stores, no barrier on the row loop; merged over the warps at the end of the band
    constexpr bool EXTRAS = MODE != 0;
    __shared__ unsigned long long s_best[EXTRAS ? kFMaxBandRows : 1][kFB][kFMaxWarps];
    __shared__ double s_red[2][kFMaxWarps];
    double m_s = 0.0, m_ss = 0.0;
    const bool want_wta = EXTRAS && wta != nullptr, want_moments = MODE == 1 && moments != nullptr;

    int offs[kFN];                           // element offsets of the 10 columns inside a row (clamped: loads stay in the row)
#pragma unroll
    for (int j = 0; j < kFN; ++j) offs[j] = min(max(xb - kFR + j, 0), w - 1) * D + d;

    // ---- ring of row slots (RING): rows r_first .. r_last-1 of columns xlo .. xhi-1, in order ----
    extern __shared__ __align__(128) uint8_t ring[];
    const int r_first = max(0, y0 - kFR), n_rows = min(h, y1 + kFR) - r_first;
    const int xlo = max(0, xb - kFR), xhi = min(w, xb + kFB + kFR);
    const uint32_t slot_bytes = (uint32_t)kFN * D * 4, run_bytes = (uint32_t)(xhi - xlo) * D * 4;
    const uint32_t s_ring = smem_u32(ring);
    const float *ring_src = vol + ((size_t)r_first * w + xlo) * D;      // row i of the block: + i * w * D
    const uint32_t bar_full = s_ring + (uint32_t)n_slots * slot_bytes, bar_empty = bar_full + 8 * kRingMaxSlots;
    // running state of the ring: slot and phase of the row being read, rows issued so far (thread 0)
    uint32_t cur_slot = 0, cur_phase = 0;
    int rows_issued = min(n_slots, n_rows);
    if (RING) {
        if (threadIdx.x == 0) {
            for (int i = 0; i < n_slots; ++i) {
                mbar_init(bar_full + 8 * i, 1);
                mbar_init(bar_empty + 8 * i, blockDim.x >> 5);
            }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (threadIdx.x == 0)
            for (int i = 0; i < rows_issued; ++i) {
                mbar_expect_tx(bar_full + 8 * i, run_bytes);
                bulk_g2s(s_ring + (uint32_t)i * slot_bytes, ring_src + (size_t)i * w * D, run_bytes, bar_full + 8 * i);
            }
        ring_src += (size_t)rows_issued * w * D;           // next row to issue
    }
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
                if (RING) {
                    mbar_wait(bar_full + 8 * cur_slot, cur_phase);
                    const float *srow = reinterpret_cast<const float *>(ring + cur_slot * slot_bytes) - xlo * D;
#pragma unroll
                    for (int j = 0; j < kFN; ++j) v[j] = srow[offs[j]];
                    __syncwarp();
                    if ((threadIdx.x & 31) == 0) mbar_arrive(bar_empty + 8 * cur_slot);
                    // one thread refills the slot as soon as every warp has released it (row rows_issued)
                    if (threadIdx.x == 0 && rows_issued < n_rows) {
                        mbar_wait(bar_empty + 8 * cur_slot, cur_phase);
                        mbar_expect_tx(bar_full + 8 * cur_slot, run_bytes);
                        bulk_g2s(s_ring + cur_slot * slot_bytes, ring_src, run_bytes, bar_full + 8 * cur_slot);
                        ring_src += (size_t)w * D;
                        ++rows_issued;
                    }
                    if (++cur_slot == (uint32_t)n_slots) { cur_slot = 0; cur_phase ^= 1; }
                } else {
#pragma unroll
                    for (int j = 0; j < kFN; ++j) v[j] = __ldg(vol + (vbase + offs[j]));
                }
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
                    const float res = match_ok[i] ? __fdiv_rn(sum, (float)cnt) : centre;
                    if (active) out[obase + offs[kFR + i]] = res;
                    if (EXTRAS) {
                        if (want_moments && active) {
                            const double rd = (double)res;
                            m_s += rd;
                            m_ss = fma(rd, rd, m_ss);
                        }
                        if (want_wta) {
                            // best disparity of the warp's 32: one redux on the keys (arg max: inverted keys), the
                            // lowest lane holding the winner is the lowest disparity
                            uint32_t key = order_key(res);
                            if (MODE == 1) key = ~key;
                            if (!active) key = 0xffffffffu;
                            const uint32_t mn = __reduce_min_sync(0xffffffffu, key);
                            const unsigned winners = __ballot_sync(0xffffffffu, key == mn);
                            if ((int)(threadIdx.x & 31) == __ffs(winners) - 1)
                                s_best[y - y0][i][threadIdx.x >> 5] = ((unsigned long long)mn << 32) | (unsigned)d;
                        }
                    }
                }
            }
        }
    }
    if (EXTRAS && want_wta) {
        __syncthreads();
        const int n_warps = blockDim.x >> 5;
        for (int i = threadIdx.x; i < (y1 - y0) * kFB; i += blockDim.x) {
            const int row = i / kFB, px = i - row * kFB;
            if (xb + px >= w) continue;
            unsigned long long best = s_best[row][px][0];
            for (int k = 1; k < n_warps; ++k) best = min(best, s_best[row][px][k]);      // lowest key, then lowest disparity
            wta[(size_t)(y0 + row) * w + xb + px] = (uint8_t)(best & 0xffu);
        }
    }
    if (EXTRAS && want_moments) {
        // block partials in a fixed order (deterministic): lanes by a shuffle tree, warps one after the other
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m_s += __shfl_down_sync(0xffffffffu, m_s, o);
            m_ss += __shfl_down_sync(0xffffffffu, m_ss, o);
        }
        if ((threadIdx.x & 31) == 0) { s_red[0][threadIdx.x >> 5] = m_s; s_red[1][threadIdx.x >> 5] = m_ss; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double a = 0.0, b = 0.0;
            for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { a += s_red[0][i]; b += s_red[1][i]; }
            const size_t bid = (size_t)blockIdx.y * gridDim.x + blockIdx.x;
            moments[2 * bid] = a;
            moments[2 * bid + 1] = b;
        }
    }
}

}  // namespace

static void cbca_fused_grid(int h, int w, int *bands, int *rows_per_band)
{
    *bands = std::max(1, (h + 62) / 63);
    *rows_per_band = (h + *bands - 1) / *bands;
}

// number of (sum, sum of squares) partial pairs launch_cbca_fused writes into CbcaExtras::moments
int cbca_fused_blocks(int h, int w)
{
    int bands, rows;
    cbca_fused_grid(h, w, &bands, &rows);
    return ((w + kFB - 1) / kFB) * bands;
}

// x0c / x1c: int16 arms [4][h][w] of the two views; rec: scratch for 2 * h * pitch records
int launch_cbca_fused(stereo_ctx *ctx, const short *x0c, const short *x1c, const float *vol, float *out, int h, int w,
                      int D, int direction, int L1, const CbcaExtras &ex)
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
    int bands, rows_per_band;
    cbca_fused_grid(h, w, &bands, &rows_per_band);
    static_assert(63 <= kFMaxBandRows, "band rows must fit the arg-min table");
    // bulk copies need 16-byte aligned runs: D % 4 == 0 and an aligned volume; STEREO_CBCA_KERNEL=1 keeps the
    // per-thread global loads (the cross-check; also what D = 201 takes)
    static const bool no_ring = [] { const char *e = getenv("STEREO_CBCA_KERNEL"); return e && atoi(e) == 1; }();
    const bool ring = !no_ring && D % 4 == 0 && ((uintptr_t)vol % 16 == 0);
    const dim3 grid((w + kFB - 1) / kFB, bands);
    const bool extras = ex.moments != nullptr || ex.wta != nullptr;
    // as many row slots as three resident blocks per SM leave room for
    const size_t slot_bytes = (size_t)kFN * D * 4;
    const int n_slots = ring ? (int)std::min<size_t>(kRingMaxSlots, std::max<size_t>(2, ((extras ? 64 : 72) * 1024) / slot_bytes)) : 1;
    const size_t smem = ring ? n_slots * slot_bytes + 2 * 8 * kRingMaxSlots : 0;
    auto launch = [&](auto kern) -> int {
        if (smem > 48 * 1024) STEREO_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<grid, nt, smem, ctx->stream>>>(rec0, rec1, pitch, vol, out, h, w, D, direction, rows_per_band, ex.moments,
                                              ex.wta, ex.take_min, n_slots);
        return STEREO_OK;
    };
    // mode 1: moments and / or arg max; mode 2: arg min only
    STEREO_REQUIRE(!(ex.moments && ex.wta && ex.take_min), "moments come with the arg MAX map (first CBCA of the chain)");
    const int mode = !extras ? 0 : ((ex.wta && ex.take_min) ? 2 : 1);
    if (ring) STEREO_TRY(mode == 0 ? launch(cbca_fused_kernel<true, 0>) : mode == 1 ? launch(cbca_fused_kernel<true, 1>) : launch(cbca_fused_kernel<true, 2>));
    else STEREO_TRY(mode == 0 ? launch(cbca_fused_kernel<false, 0>) : mode == 1 ? launch(cbca_fused_kernel<false, 1>) : launch(cbca_fused_kernel<false, 2>));
    if (ex.moment_blocks) *ex.moment_blocks = ((w + kFB - 1) / kFB) * bands;
    ctx->launches += 3;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
