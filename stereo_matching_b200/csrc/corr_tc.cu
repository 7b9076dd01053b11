// Left/right correlation cost volumes and fused WTA on the tensor cores
// (/root/reference/inference.py:139-183).
//
// For one image row the reference's D-iteration loop is a banded Gram matrix
// S[xl][xr] = <L[xl], R[xr]>, 0 <= xl - xr < D.  A tile is 128 consecutive pixels of one
// side: M = 128 "own" feature vectors against N = 384 >= 128 + D - 1 candidates of the other
// view, K = 64 channels - two tcgen05.mma blocks (N = 256 + 128) per K-step into 384 TMEM
// columns, bf16 (hi, lo) split operands with fp32 accumulation like the trunk.
//   left  tile: own = L[x0 .. x0+127], others = R[x0-(D-1) .. ],  d = xi + D-1 - u
//   right tile: own = R[x0 .. x0+127], others = L[x0 .. ],        d = u - xi
// Features are read from the chunk-planar bf16 layout the trunk writes; the buffers carry
// zero margins left and right of every row, so candidates outside the image contribute
// exactly 0.0 - the zero-filled slots of inference.py:140-141.
//
// Epilogue.  tcgen05.ld hands lane i of an epilogue warp one row of the tile, 32 columns at
// a time.  Fused WTA: the thread scans its own row and keeps the arg max - the volume never
// exists.  Materialising: each warp parks the 32x32 block in shared memory and writes it out
// row by row, so that every warp store is one 128-byte run of a [x][d] output row.
//
// The kernel is persistent (one CTA per SM, tiles round-robin); the TMA producer refills the
// operand buffers for tile t+1 while the epilogue warps are still draining tile t from TMEM.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {

namespace tcc {

constexpr int kTileM = 128;
constexpr int kWinN = 384;
constexpr int kARun = kTileM * 16, kAPlane = 8 * kARun;      // 16 KB per plane
constexpr int kBRun = kWinN * 16, kBPlane = 8 * kBRun;       // 48 KB per plane
constexpr int kSmemA = 0, kSmemB = 2 * kAPlane;
constexpr int kSmemStage = kSmemB + 2 * kBPlane;             // one [32][33] float block per epilogue warp
constexpr int kEpiWarps = 16;                                // four per TMEM lane quarter
constexpr int kParts = kEpiWarps / 4;                        // column parts of a tile row
constexpr int kStagePitch = 36;                              // floats; 16-byte stores, conflict-free both ways
constexpr int kChunksPerWarp = (kWinN / 32) / kParts;
constexpr int kSmemWta = kSmemStage + kEpiWarps * 32 * kStagePitch * 4;   // [2 tiles][kParts][128] (value, d)
constexpr int kSmemBars = kSmemWta + 2 * kParts * 128 * 8;
constexpr int kSmemTotal = kSmemBars + 64;
constexpr uint32_t kTmemCols = 512;

struct Params {
    const __nv_bfloat16 *l_hi, *l_lo, *r_hi, *r_lo;   // [n][h][8][wf][8], pixel x at column ml + x
    int n, h, w, wf, ml, D, planes;
    float *left_vol, *right_vol;                       // [n][h][w][D] or null
    uint8_t *left_disp, *right_disp;                   // [n][h][w] or null
    int tiles_x, ntiles;
};

// One epilogue warp drains `kChunksPerWarp` 32-column chunks of its 32 TMEM lanes.
template <int SIDE, bool VOL, bool WTA>
__device__ __forceinline__ void drain_tile(const Params &p, uint32_t tmem_row, float *stage, int lane, int q,
                                           int c_begin, int img, int y, int x0, float &best, int &best_d)
{
    const int xi = q * 32 + lane;
    const int D = p.D;
    best = -INFINITY;
    best_d = 0x7fffffff;
    float *vol = SIDE == 0 ? p.left_vol : p.right_vol;
    const size_t row_pix = ((size_t)img * p.h + y) * p.w;
    const int rows_ok = min(32, p.w - (x0 + q * 32));            // rows of this warp that are inside the image
#pragma unroll 1
    for (int c = c_begin; c < c_begin + kChunksPerWarp; ++c) {
        uint32_t v[32];
        tc::tmem_ld_32x32(tmem_row + c * 32, v);
        if (WTA) {
            // left: d falls as u rises, so ">=" keeps the lowest d among equals; right: ">" does
            const int dbase = SIDE == 0 ? xi + D - 1 - c * 32 : c * 32 - xi;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const int d = SIDE == 0 ? dbase - i : dbase + i;
                const float f = __uint_as_float(v[i]);
                const bool in = (unsigned)d < (unsigned)D;
                const bool take = in && (SIDE == 0 ? f >= best : f > best);
                best = take ? f : best;
                best_d = take ? d : best_d;
            }
        }
        if (VOL) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
                reinterpret_cast<uint4 *>(stage + lane * kStagePitch)[i] = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            __syncwarp();
            // row r of the block goes to out[x0 + q*32 + r][d00 +- r]: one 128-byte run per warp store.
            // Offsets are 32-bit and advance by a constant per row (one image row of a volume is < 2^31 floats).
            const int d00 = SIDE == 0 ? q * 32 + D - 1 - (c * 32 + lane) : (c * 32 + lane) - q * 32;
            const int dstep = SIDE == 0 ? 1 : -1;
            const int stride = D + dstep;
            float *rowbase = vol + row_pix * D;
            int off = (x0 + q * 32) * D + d00;
            int dd = d00;
#pragma unroll 1
            for (int r0 = 0; r0 < 32; r0 += 8) {
                float t[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) t[j] = stage[(r0 + j) * kStagePitch + lane];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (r0 + j < rows_ok && (unsigned)dd < (unsigned)D) rowbase[off] = t[j];
                    off += stride;
                    dd += dstep;
                }
            }
            __syncwarp();
        }
    }
}

template <bool VOL, bool WTA>
__global__ __launch_bounds__(128 + 32 * kEpiWarps, 1) void corr_tc_kernel(Params p)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t s_base = tc::smem_u32(smem);
    const uint32_t s_a = s_base + kSmemA, s_b = s_base + kSmemB;
    const uint32_t bar_full = s_base + kSmemBars, bar_empty = bar_full + 8;
    const uint32_t bar_acc_full = bar_full + 16, bar_acc_empty = bar_full + 32;     // [2] each
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + kSmemBars + 48);

    if (tid == 0) {
        tc::mbar_init(bar_full, 1);
        tc::mbar_init(bar_empty, 1);
        for (int i = 0; i < 2; ++i) {
            tc::mbar_init(bar_acc_full + 8 * i, 1);
            tc::mbar_init(bar_acc_empty + 8 * i, kEpiWarps / 2);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc::smem_u32(tmem_slot)),
                     "r"(kTmemCols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const size_t img_elems = (size_t)p.h * 8 * p.wf * 8;

    auto decode = [&](int t, int &img, int &side, int &y, int &x0) {
        const int tx = t % p.tiles_x;
        int rest = t / p.tiles_x;
        y = rest % p.h;
        rest /= p.h;
        side = rest & 1;
        img = rest >> 1;
        x0 = tx * kTileM;
    };

    if (warp == 0) {
        if (lane == 0) {
            // ---- TMA producer ----
            int k = 0;
            for (int t = blockIdx.x; t < p.ntiles; t += gridDim.x, ++k) {
                int img, side, y, x0;
                decode(t, img, side, y, x0);
                tc::mbar_wait(bar_empty, (k & 1) ^ 1);
                tc::mbar_expect_tx(bar_full, (uint32_t)p.planes * (kAPlane + kBPlane));
                const int a_col = p.ml + x0;
                const int b_col = side == 0 ? p.ml + x0 - (p.D - 1) : p.ml + x0;
                const size_t row0 = (size_t)img * img_elems + (size_t)y * 8 * p.wf * 8;
                for (int pl = 0; pl < p.planes; ++pl) {
                    const __nv_bfloat16 *own = side == 0 ? (pl ? p.l_lo : p.l_hi) : (pl ? p.r_lo : p.r_hi);
                    const __nv_bfloat16 *oth = side == 0 ? (pl ? p.r_lo : p.r_hi) : (pl ? p.l_lo : p.l_hi);
                    for (int cc = 0; cc < 8; ++cc) {
                        const size_t chunk = row0 + (size_t)cc * p.wf * 8;
                        tc::bulk_g2s(s_a + pl * kAPlane + cc * kARun, own + chunk + (size_t)a_col * 8, kARun, bar_full);
                        tc::bulk_g2s(s_b + pl * kBPlane + cc * kBRun, oth + chunk + (size_t)b_col * 8, kBRun, bar_full);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ---- MMA issuer ----
            constexpr uint32_t idesc_half = tc::make_idesc(kTileM, kWinN / 2);
            const int n_combo = p.planes == 2 ? 3 : 1;
            int k = 0;
            for (int t = blockIdx.x; t < p.ntiles; t += gridDim.x, ++k) {
                int img, side, y, x0;
                decode(t, img, side, y, x0);
                tc::mbar_wait(bar_full, k & 1);
                // two column blocks of 192 with their own full/empty barriers: block 1 of this tile
                // is computed while block 0 is drained, block 0 of the next tile while block 1 is
#pragma unroll 1
                for (int nb = 0; nb < 2; ++nb) {
                    tc::mbar_wait(bar_acc_empty + 8 * nb, (k & 1) ^ 1);
                    tc::tc_fence_after();
                    uint32_t acc = 0;
                    for (int combo = 0; combo < n_combo; ++combo) {
                        // L_lo*R_hi, then L_hi*R_lo, then L_hi*R_hi on both sides, so that the left
                        // and the right volume accumulate the same terms in the same order
                        const int lo_first = side == 0 ? 0 : 1;
                        const int apl = (n_combo == 3 && combo == lo_first) ? 1 : 0;
                        const int bpl = (n_combo == 3 && combo == 1 - lo_first) ? 1 : 0;
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            uint64_t ad = tc::smem_desc(s_a + apl * kAPlane + 2 * ks * kARun, kARun, 128);
                            uint64_t bd = tc::smem_desc(s_b + bpl * kBPlane + 2 * ks * kBRun + nb * (kWinN / 2) * 16, kBRun, 128);
                            tc::umma_bf16(tmem_base + nb * (kWinN / 2), ad, bd, idesc_half, acc);
                            acc = 1;
                        }
                    }
                    tc::tc_commit(bar_acc_full + 8 * nb);     // this column block is complete
                }
                tc::tc_commit(bar_empty);        // operands may be overwritten
            }
        }
    } else if (warp >= 4) {
        // ---- epilogue: warp (q, part) drains kChunksPerWarp 32-column chunks of TMEM lanes 32q.. ----
        const int q = warp & 3;
        const int half = (warp - 4) >> 2;            // column part of this warp
        float *stage = reinterpret_cast<float *>(smem + kSmemStage) + (warp - 4) * 32 * kStagePitch;
        float2 *wta_s = reinterpret_cast<float2 *>(smem + kSmemWta);
        const uint32_t tmem_row = tmem_base + ((uint32_t)(q * 32) << 16);
        int k = 0;
        for (int t = blockIdx.x; t < p.ntiles; t += gridDim.x, ++k) {
            int img, side, y, x0;
            decode(t, img, side, y, x0);
            const int nb = half / (kParts / 2);          // column block this warp drains
            tc::mbar_wait(bar_acc_full + 8 * nb, k & 1);
            tc::tc_fence_after();
            float best;
            int best_d;
            if (side == 0) drain_tile<0, VOL, WTA>(p, tmem_row, stage, lane, q, half * kChunksPerWarp, img, y, x0, best, best_d);
            else drain_tile<1, VOL, WTA>(p, tmem_row, stage, lane, q, half * kChunksPerWarp, img, y, x0, best, best_d);
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(bar_acc_empty + 8 * nb);
            if (WTA) {
                // merge the two halves of each row: larger value wins, equal values -> lower disparity
                float2 *slot = wta_s + (k & 1) * (kParts * 128);
                slot[half * 128 + q * 32 + lane] = make_float2(best, __int_as_float(best_d));
                asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
                if (half == 0) {
#pragma unroll
                    for (int part = 1; part < kParts; ++part) {
                        const float2 o = slot[part * 128 + q * 32 + lane];
                        const int od = __float_as_int(o.y);
                        if (o.x > best || (o.x == best && od < best_d)) { best = o.x; best_d = od; }
                    }
                    const int x = x0 + q * 32 + lane;
                    if (x < p.w)
                        (side == 0 ? p.left_disp : p.right_disp)[((size_t)img * p.h + y) * p.w + x] = (uint8_t)(best_d & 255);
                }
            }
        }
    }

    tc::tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc::tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
    }
}

// fp32 NHWC features -> chunk-planar bf16 (hi, lo) with margins (for the entry points that take
// feature maps from the caller; the fused pipeline gets this layout straight from the trunk)
__global__ void nhwc_to_planar_kernel(const float *__restrict__ in, int n, int h, int w, int wf, int ml,
                                      __nv_bfloat16 *__restrict__ hi, __nv_bfloat16 *__restrict__ lo)
{
    size_t total = (size_t)n * h * 8 * w;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int x = (int)(i % w);
        size_t rest = i / w;
        int cc = (int)(rest & 7);
        size_t row = rest >> 3;                       // img * h + y
        const float4 *src = reinterpret_cast<const float4 *>(in + (row * w + x) * 64 + cc * 8);
        float4 a = __ldg(src), b = __ldg(src + 1);
        uint32_t h4[4], l4[4];
        tc::split_bf16x2(a.x, a.y, h4[0], l4[0]);
        tc::split_bf16x2(a.z, a.w, h4[1], l4[1]);
        tc::split_bf16x2(b.x, b.y, h4[2], l4[2]);
        tc::split_bf16x2(b.z, b.w, h4[3], l4[3]);
        size_t off = ((row * 8 + cc) * wf + ml + x) * 8;
        *reinterpret_cast<uint4 *>(hi + off) = make_uint4(h4[0], h4[1], h4[2], h4[3]);
        if (lo) *reinterpret_cast<uint4 *>(lo + off) = make_uint4(l4[0], l4[1], l4[2], l4[3]);
    }
}

}  // namespace tcc

int corr_tc_max_disp() { return tcc::kWinN - tcc::kTileM + 1; }

void corr_tc_geometry(int w, int D, int *wf, int *ml)
{
    // left margin: what the row-sweep kernel wants (corr_sweep.cu), or D-1 rounded up if that is more;
    // right margin: a whole candidate window past the last 128-pixel block
    const int need = (D - 1 + 7) / 8 * 8;
    *ml = need > corr_sweep_margin() ? need : corr_sweep_margin();
    const int nb = (w + tcc::kTileM - 1) / tcc::kTileM;
    const int cols = nb * tcc::kTileM + (tcc::kWinN - tcc::kTileM);
    // rows are laid out for whole PAIRS of 128-pixel blocks plus two more blocks (the trunk's CTA pairs store
    // 256-pixel runs, conv_tc3.cu)
    const int pair_cols = tcc::kTileM * (2 * ((nb + 1) / 2) + 2);
    *wf = *ml + (cols > pair_cols ? cols : pair_cols);
}

int launch_nhwc_to_planar(stereo_ctx *ctx, const float *in, int n, int h, int w, int wf, int ml, void *hi, void *lo)
{
    ProfScope prof(ctx, STEREO_K_PREP);      // layout conversion, not part of the correlation kernel's time
    tcc::nhwc_to_planar_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(
        in, n, h, w, wf, ml, static_cast<__nv_bfloat16 *>(hi), static_cast<__nv_bfloat16 *>(lo));
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

bool corr_both_enabled()
{
    static const bool on = [] {
        const char *k = getenv("STEREO_CORR_KERNEL"), *s = getenv("STEREO_CORR_SWEEPS");
        return !(k && k[0] == '1') && !(s && s[0] == '2');
    }();
    return on;
}

int launch_corr_tc(stereo_ctx *ctx, const void *l_hi, const void *l_lo, const void *r_hi, const void *r_lo, int n,
                   int h, int w, int wf, int ml, int D, int planes, float *left_vol, float *right_vol,
                   uint8_t *left_disp, uint8_t *right_disp)
{
    STEREO_REQUIRE(D >= 1 && D <= corr_tc_max_disp(), "tensor-core correlation supports disp_range <= %d, got %d",
                   corr_tc_max_disp(), D);
    STEREO_REQUIRE(ml >= D - 1 && wf >= ml + (w + tcc::kTileM - 1) / tcc::kTileM * tcc::kTileM + tcc::kWinN - tcc::kTileM,
                   "feature margins too small");
    bool vol = left_vol != nullptr, wta = left_disp != nullptr;
    STEREO_REQUIRE(vol || wta, "nothing to compute");
    // default: the row-sweep kernel (corr_sweep.cu); STEREO_CORR_KERNEL=1 keeps the per-tile kernel below as the
    // cross-check, which also serves rows whose margins are too small for the sweep
    static const bool per_tile = [] { const char *e = getenv("STEREO_CORR_KERNEL"); return e && e[0] == '1'; }();
    // fused WTA of both maps: one sweep per row serves both (corr_both.cu); STEREO_CORR_SWEEPS=2 keeps the two sweeps
    if (corr_both_enabled() && !vol && left_disp && right_disp && ml >= corr_sweep_margin())
        return launch_corr_both(ctx, l_hi, l_lo, r_hi, r_lo, n, h, w, wf, ml, D, planes, left_disp, right_disp);
    if (!per_tile && ml >= corr_sweep_margin())
        return launch_corr_sweep(ctx, l_hi, l_lo, r_hi, r_lo, n, h, w, wf, ml, D, planes, left_vol, right_vol, left_disp,
                                 right_disp);
    tcc::Params p;
    p.l_hi = static_cast<const __nv_bfloat16 *>(l_hi);
    p.l_lo = static_cast<const __nv_bfloat16 *>(l_lo);
    p.r_hi = static_cast<const __nv_bfloat16 *>(r_hi);
    p.r_lo = static_cast<const __nv_bfloat16 *>(r_lo);
    p.n = n; p.h = h; p.w = w; p.wf = wf; p.ml = ml; p.D = D; p.planes = planes;
    p.left_vol = left_vol; p.right_vol = right_vol; p.left_disp = left_disp; p.right_disp = right_disp;
    p.tiles_x = (w + tcc::kTileM - 1) / tcc::kTileM;
    p.ntiles = n * 2 * h * p.tiles_x;
    int grid = p.ntiles < ctx->sm_count ? p.ntiles : ctx->sm_count;
    auto launch = [&](auto kern) -> int {
        STEREO_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tcc::kSmemTotal));
        ProfScope prof(ctx, STEREO_K_CORR);
        kern<<<grid, 128 + 32 * tcc::kEpiWarps, tcc::kSmemTotal, ctx->stream>>>(p);
        ctx->launches += 1;
        STEREO_CUDA_CHECK(cudaGetLastError());
        return STEREO_OK;
    };
    if (vol && wta) return launch(tcc::corr_tc_kernel<true, true>);
    if (vol) return launch(tcc::corr_tc_kernel<true, false>);
    return launch(tcc::corr_tc_kernel<false, true>);
}

}  // namespace stereo
