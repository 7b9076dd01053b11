// Fused correlation + WTA, sweeping along image rows (/root/reference/inference.py:139-183).
//
// corr_tc.cu treats every 128-pixel tile on its own and reloads the whole 384-column candidate
// window (128 KB of operands per tile): it is bound by that reload, which cannot overlap the MMAs
// because only one operand set fits into shared memory.  Neighbouring tiles of one image row share
// two thirds of their window, so here a CTA sweeps along a row: the candidates live in a ring of
// four 128-pixel blocks, every tile adds ONE new block (and its own 128 pixels, double-buffered),
// i.e. 64 KB per tile, loaded by TMA two tiles ahead of the MMAs.
//
//   left  sweep: own = L block k, candidates = R blocks k-2, k-1, k,   d = xi + 256 - u
//   right sweep: own = R block k, candidates = L blocks k,   k+1, k+2, d = u - xi
//   (xi = pixel inside the own block, u = column inside the 384-wide window)
//
// Per tile the issuer runs three M=128 x N=128 accumulations (one per candidate block, K = 64,
// bf16 hi/lo split = 12 MMAs each) into a ring of four 128-column tensor-memory blocks; 16 epilogue
// warps drain each block as soon as its MMAs commit (one 32-column chunk per warp, chunks outside
// the disparity band skipped), keep the running arg max in registers and merge the four column
// parts through shared memory at the end of the tile.  The volume never exists.
#include <cuda_bf16.h>

#include <cstdlib>
#include <type_traits>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {

namespace tcs {

constexpr int kBlk = 128;                       // pixels per block
constexpr int kRun = kBlk * 16;                 // one channel chunk: 2 KB
constexpr int kPlane = 8 * kRun;                // 16 KB
constexpr int kBlock = 2 * kPlane;              // hi + lo: 32 KB
constexpr int kAccBlocks = 4;
constexpr int kStageRows = 8, kStagePitch = 36;  // materialising epilogue: [8 rows][32 + 4] floats per warp
// Shared-memory layout.  The materialising variant gives up one of the three own-block slots for the
// staging buffers in which the epilogue warps transpose 8x32 blocks before they store them.
// BULK (materialising, 16-byte aligned volumes with D % 4 == 0): the staging area is four [16 pixels][D] fp32 tiles,
// one per group of epilogue warps; a tile is ONE contiguous run of the [x][d] volume and leaves for HBM as a
// single bulk copy.  The candidate ring gives up a slot for them (three: just the blocks a tile multiplies).
constexpr int kBulkStageBytes = 16 * 256 * 4;
template <bool VOL, bool BULK = false>
struct Layout {
    static constexpr int kASlots = VOL ? 2 : 3;
    static constexpr int kBSlots = BULK ? 3 : 4;
    static constexpr int kA = 0;
    static constexpr int kB = kA + kASlots * kBlock;
    static constexpr int kWtaV = kB + kBSlots * kBlock;          // [3 parts][128] best value of column parts 1-3
    static constexpr int kWtaD = kWtaV + 3 * 128 * 4;            // [3 parts][128] its disparity (16 bit)
    static constexpr int kStage = kWtaD + 3 * 128 * 2;
    static constexpr int kBars = kStage + (BULK ? 4 * kBulkStageBytes : (VOL ? 16 * kStageRows * kStagePitch * 4 : 0));
    static constexpr int kTotal = kBars + 256;
    static_assert(kTotal <= 232448, "shared memory budget of one CTA");
};
constexpr int kEpiWarps = 16;
constexpr int kThreads = 128 + 32 * kEpiWarps;
constexpr uint32_t kTmemCols = 512;
constexpr int kMargin = 256;                    // zero columns the feature rows need left of pixel 0

struct Params {
    const __nv_bfloat16 *l_hi, *l_lo, *r_hi, *r_lo;   // [n][h][8][wf][8], pixel x at column ml + x
    int n, h, w, wf, ml, D, planes;
    uint8_t *left_disp, *right_disp;                   // [n][h][w] or null
    float *left_vol, *right_vol;                       // [n][h][w][D] or null
    int nb, tps, segs, nitems;                         // blocks per row, tiles per segment, segments per row
};

struct Item {
    int img, side, y, k0, k1;
};

__device__ __forceinline__ Item decode(const Params &p, int item)
{
    Item it;
    // left and right sweep of one row are neighbouring items: they run at the same time on
    // neighbouring CTAs and the second reader of every feature block finds it in L2
    const int seg = item % p.segs;
    int rest = item / p.segs;
    it.side = rest & 1;
    rest >>= 1;
    it.y = rest % p.h;
    it.img = rest / p.h;
    it.k0 = seg * p.tps;
    it.k1 = min(p.nb, it.k0 + p.tps);
    return it;
}

// Running arg max over the 32 columns a lane holds of one accumulator chunk.  SIDE 0 (left volume): the disparity
// falls as the column rises, ">=" keeps the lowest disparity among equals; SIDE 1: it rises, ">" does.  CHECK = false:
// the caller knows that every disparity of the chunk, for every lane of the warp, lies inside [0, D) (true for most
// chunks of the band) - no range test per element.
template <int SIDE, bool CHECK>
__device__ __forceinline__ void scan_chunk(const uint32_t (&v)[32], int dbase, int D, float &best, int &best_d)
{
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        const int d = SIDE == 0 ? dbase - i : dbase + i;
        const float f = __uint_as_float(v[i]);
        const bool in = !CHECK || (unsigned)d < (unsigned)D;
        const bool take = in && (SIDE == 0 ? f >= best : f > best);
        best = take ? f : best;
        best_d = take ? d : best_d;
    }
}

// The same 32 values into row `trow` of a [pixel][d] staging tile (bulk-store variant); `mine`: this lane's row
// belongs to the half of the group that is being staged.
template <int SIDE, bool CHECK>
__device__ __forceinline__ void stage_chunk(const uint32_t (&v)[32], int dbase, int D, float *trow, bool mine)
{
    if (!mine) return;
    float *tp = trow + dbase;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        const int d = SIDE == 0 ? dbase - i : dbase + i;
        if (!CHECK || (unsigned)d < (unsigned)D) tp[SIDE == 0 ? -i : i] = __uint_as_float(v[i]);
    }
}

// DC: the disparity range as a compile-time constant (0: run-time p.D).  The materialising epilogue stores eight
// rows of a [x][d] tile per round, one row D +- 1 floats after the other: with D known the eight stores are one
// pointer and eight immediates, and the range tests compare against a constant - the round-2 ncu source page had
// 40 % of this kernel's instructions in 64-bit address arithmetic and predicates of these stores.
// BULK: see Layout.  Every group of four epilogue warps (32 pixels of the tile) owns a staging tile of 16 pixels: it
// copies the upper 16 rows of its three accumulator blocks there (conflict-free 4-byte shared stores: consecutive
// lanes are D +- 1 words apart), one thread of the group sends the tile off with cp.async.bulk, then the same for the
// lower 16 rows (the accumulators are simply read twice).  One store instruction per 16 x D voxels instead of one
// per 32: the per-store address arithmetic that made the round-1 epilogue instruction bound is gone, and HBM sees
// 14 KB runs.  Measured alternatives: the groups taking turns on one 32-row tile (0.55 ms per pair), on two
// (0.39 ms, like the per-row stores): only four of the sixteen warps work at a time.
template <bool VOL, bool WTA, int DC, bool BULK = false>
__global__ __launch_bounds__(kThreads, 1) void corr_sweep_kernel(Params p)
{
    static_assert(!BULK || VOL, "the bulk store belongs to the materialising variant");
    using L = Layout<VOL, BULK>;
    constexpr int kASlots = L::kASlots;
    constexpr int kBSlots = L::kBSlots;
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t s_base = tc::smem_u32(smem);
    const uint32_t s_a = s_base + L::kA, s_b = s_base + L::kB;
    const uint32_t bar_a_full = s_base + L::kBars;          // [3]
    const uint32_t bar_a_empty = bar_a_full + 24;            // [3]
    const uint32_t bar_b_full = bar_a_full + 48;             // [4]
    const uint32_t bar_b_empty = bar_a_full + 80;            // [4]
    const uint32_t bar_acc_full = bar_a_full + 112;          // [4]
    const uint32_t bar_acc_empty = bar_a_full + 144;         // [4]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + L::kBars + 176);

    if (tid == 0) {
        for (int i = 0; i < kASlots; ++i) {
            tc::mbar_init(bar_a_full + 8 * i, 1);
            tc::mbar_init(bar_a_empty + 8 * i, 1);
        }
        for (int i = 0; i < 4; ++i) {
            tc::mbar_init(bar_b_full + 8 * i, 1);
            tc::mbar_init(bar_b_empty + 8 * i, 1);
            tc::mbar_init(bar_acc_full + 8 * i, 1);
            tc::mbar_init(bar_acc_empty + 8 * i, kEpiWarps);
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

    if (warp == 0) {
        if (lane == 0) {
            // ---- TMA producer: candidate blocks two tiles ahead, own blocks one tile ahead ----
            uint32_t E = 0, T = 0;                   // running candidate-block and tile counters
            for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
                const Item it = decode(p, item);
                const size_t row0 = (size_t)it.img * img_elems + (size_t)it.y * 8 * p.wf * 8;
                const __nv_bfloat16 *own[2] = {it.side == 0 ? p.l_hi : p.r_hi, it.side == 0 ? p.l_lo : p.r_lo};
                const __nv_bfloat16 *oth[2] = {it.side == 0 ? p.r_hi : p.l_hi, it.side == 0 ? p.r_lo : p.l_lo};
                auto load_block = [&](uint32_t dst, const __nv_bfloat16 *const *src, int col, uint32_t bar) {
                    tc::mbar_expect_tx(bar, (uint32_t)p.planes * kPlane);
                    for (int pl = 0; pl < p.planes; ++pl)
                        for (int cc = 0; cc < 8; ++cc)
                            tc::bulk_g2s(dst + pl * kPlane + cc * kRun,
                                         src[pl] + row0 + ((size_t)cc * p.wf + col) * 8, kRun, bar);
                };
                auto load_entry = [&](int j) {       // j-th candidate block of this row
                    const uint32_t slot = E % kBSlots, ph = (E / kBSlots) & 1;
                    ++E;
                    tc::mbar_wait(bar_b_empty + 8 * slot, ph ^ 1);
                    load_block(s_b + slot * kBlock, oth, p.ml + kBlk * (it.side == 0 ? j - 2 : j), bar_b_full + 8 * slot);
                };
                auto load_own = [&](int k) {
                    const uint32_t slot = T % kASlots, ph = (T / kASlots) & 1;
                    ++T;
                    tc::mbar_wait(bar_a_empty + 8 * slot, ph ^ 1);
                    load_block(s_a + slot * kBlock, own, p.ml + kBlk * k, bar_a_full + 8 * slot);
                };
                load_own(it.k0);
                load_entry(it.k0);
                load_entry(it.k0 + 1);
                for (int k = it.k0; k < it.k1; ++k) {
                    load_entry(k + 2);
                    if (k + 1 < it.k1) load_own(k + 1);
                }
            }
        }
    } else if (warp == 1) {
        {
            // ---- MMA issuer: the whole warp runs the loop (converged), one elected lane issues ----
            constexpr uint32_t idesc = tc::make_idesc(kBlk, kBlk);
            const uint32_t d_hi = tc::desc_hi(128);
            const int n_mma = p.planes == 2 ? 12 : 4;
            uint32_t E0 = 0, T = 0, G = 0;           // first candidate block of the item, tiles, accumulator blocks
            bool prewaited = false;                  // barriers of the coming block already waited for
            for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
                const Item it = decode(p, item);
                // L_lo*R_hi, then L_hi*R_lo, then L_hi*R_hi on both sides: the left and the right
                // volume accumulate the same terms in the same order
                const int lo_first = it.side == 0 ? 0 : 1;
                for (int k = it.k0; k < it.k1; ++k) {
                    const uint32_t as = T % kASlots;
                    tc::mbar_wait(bar_a_full + 8 * as, (T / kASlots) & 1);
                    const uint32_t e_first = E0 + (uint32_t)(k - it.k0);
#pragma unroll 1
                    for (int o = 0; o < 3; ++o, ++G) {
                        const uint32_t e = e_first + o, bs = e % kBSlots;
                        const uint32_t ab = G % kAccBlocks;
                        if (!prewaited) {
                            if (k == it.k0 || o == 2) tc::mbar_wait(bar_b_full + 8 * bs, (e / kBSlots) & 1);
                            tc::mbar_wait(bar_acc_empty + 8 * ab, ((G / kAccBlocks) & 1) ^ 1);
                        }
                        prewaited = false;
                        tc::tc_fence_after();
#pragma unroll
                        for (int m = 0; m < 12; ++m) {
                            if (m < n_mma) {
                                const int combo = m >> 2, ks = m & 3;
                                const int apl = (n_mma == 12 && combo == lo_first) ? 1 : 0;
                                const int bpl = (n_mma == 12 && combo == 1 - lo_first) ? 1 : 0;
                                const uint32_t a_lo = tc::desc_lo(s_a + as * kBlock + apl * kPlane, kRun);
                                const uint32_t b_lo = tc::desc_lo(s_b + bs * kBlock + bpl * kPlane, kRun);
                                if (tc::elect_one())
                                    tc::umma_bf16_parts(tmem_base + ab * kBlk, a_lo + ((2 * ks * kRun) >> 4), d_hi,
                                                        b_lo + ((2 * ks * kRun) >> 4), d_hi, idesc, m > 0);
                            }
                            // With the last MMAs of the block still to issue and several queued in the tensor pipe,
                            // wait for what the next block of this tile needs: its accumulator and its candidates.
                            if (m == 8 && o < 2) {
                                if (k == it.k0 || o == 1) {
                                    const uint32_t en = e + 1;
                                    tc::mbar_wait(bar_b_full + 8 * (en % kBSlots), (en / kBSlots) & 1);
                                }
                                tc::mbar_wait(bar_acc_empty + 8 * ((G + 1) % kAccBlocks), (((G + 1) / kAccBlocks) & 1) ^ 1);
                                prewaited = true;
                            }
                        }
                        if (tc::elect_one()) {
                            tc::tc_commit(bar_acc_full + 8 * ab);
                            // the oldest candidate block has had its last use (at the end of a row all three have)
                            if (o == 0 || k == it.k1 - 1) tc::tc_commit(bar_b_empty + 8 * bs);
                        }
                    }
                    if (tc::elect_one()) tc::tc_commit(bar_a_empty + 8 * as);
                    ++T;
                }
                E0 += (uint32_t)(it.k1 - it.k0) + 2;
            }
        }
    } else if (warp >= 4) {
        // ---- epilogue: warp (q, part) drains columns 32*part.. of every accumulator block, lanes 32q.. ----
        const int q = warp & 3, part = (warp - 4) >> 2;
        const int xi = q * 32 + lane;
        const int D = DC ? DC : p.D;
        float *wta_v = reinterpret_cast<float *>(smem + L::kWtaV);
        uint16_t *wta_d = reinterpret_cast<uint16_t *>(smem + L::kWtaD);
        float *stage = reinterpret_cast<float *>(smem + L::kStage) + (warp - 4) * kStageRows * kStagePitch;
        uint32_t T = 0, G = 0;
        for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
            const Item it = decode(p, item);
            for (int k = it.k0; k < it.k1; ++k, ++T) {
                float best = -INFINITY;
                int best_d = 0x7fffffff;
                if (BULK) {
                    float *vol = it.side == 0 ? p.left_vol : p.right_vol;
                    float *rowbase = vol + (((size_t)it.img * p.h + it.y) * p.w) * (size_t)D;
                    float *tile = reinterpret_cast<float *>(smem + L::kStage + q * kBulkStageBytes);
                    const bool store_thread = part == 0 && lane == 0;       // bulk groups are tracked per thread
#pragma unroll 1
                    for (int half = 0; half < 2; ++half) {
                        // the tile is free once the copy engine has read the previous copy out of it
                        if (store_thread) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                        asm volatile("bar.sync %0, 128;" ::"r"(3 + q) : "memory");
                        const bool mine = (lane >> 4) == half;
                        float *trow = tile + (lane & 15) * D;
#pragma unroll 1
                        for (int o = 0; o < 3; ++o) {
                            const uint32_t Go = G + o, ab = Go % kAccBlocks;
                            tc::mbar_wait(bar_acc_full + 8 * ab, (Go / kAccBlocks) & 1);
                            tc::tc_fence_after();
                            const int u0 = 128 * o + 32 * part;
                            const int dlo = it.side == 0 ? 32 * q + kMargin - u0 - 31 : u0 - 32 * q - 31;
                            if (dlo + 62 >= 0 && dlo <= D - 1) {
                                uint32_t v[32];
                                tc::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + ab * kBlk + part * 32, v);
                                // row `lane` of the group, disparity of column i: dbase -+ i; chunks whose disparities all
                                // lie inside the range (for every lane) skip the per-element range tests
                                const bool all_in = dlo >= 0 && dlo + 62 <= D - 1;
                                const bool scan = WTA && half == 0;
                                if (it.side == 0) {
                                    const int dbase = xi + kMargin - u0;
                                    if (all_in) {
                                        stage_chunk<0, false>(v, dbase, D, trow, mine);
                                        if (scan) scan_chunk<0, false>(v, dbase, D, best, best_d);
                                    } else {
                                        stage_chunk<0, true>(v, dbase, D, trow, mine);
                                        if (scan) scan_chunk<0, true>(v, dbase, D, best, best_d);
                                    }
                                } else {
                                    const int dbase = u0 - xi;
                                    if (all_in) {
                                        stage_chunk<1, false>(v, dbase, D, trow, mine);
                                        if (scan) scan_chunk<1, false>(v, dbase, D, best, best_d);
                                    } else {
                                        stage_chunk<1, true>(v, dbase, D, trow, mine);
                                        if (scan) scan_chunk<1, true>(v, dbase, D, best, best_d);
                                    }
                                }
                            }
                        }
                        // generic-proxy writes of this thread become visible to the copy engine
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        asm volatile("bar.sync %0, 128;" ::"r"(3 + q) : "memory");
                        if (store_thread) {
                            const int xq = kBlk * k + q * 32 + 16 * half;
                            const int rows_ok = min(16, p.w - xq);
                            if (rows_ok > 0)
                                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(rowbase + (size_t)xq * D),
                                             "r"(tc::smem_u32(tile)), "r"((uint32_t)(rows_ok * D * 4))
                                             : "memory");
                            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                        }
                    }
                    // every group has read its rows: the three accumulator blocks go back to the issuer
                    tc::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
#pragma unroll
                        for (int o = 0; o < 3; ++o) tc::mbar_arrive(bar_acc_empty + 8 * ((G + o) % kAccBlocks));
                    }
                    G += 3;
                } else {
#pragma unroll 1
                for (int o = 0; o < 3; ++o, ++G) {
                    const uint32_t ab = G % kAccBlocks;
                    tc::mbar_wait(bar_acc_full + 8 * ab, (G / kAccBlocks) & 1);
                    tc::tc_fence_after();
                    const int u0 = 128 * o + 32 * part;                  // first window column of this chunk
                    // disparities this warp can see in the chunk (warp-uniform): skip it when none is in [0, D)
                    const int dlo = it.side == 0 ? 32 * q + kMargin - u0 - 31 : u0 - 32 * q - 31;
                    if (dlo + 62 >= 0 && dlo <= D - 1) {
                        uint32_t v[32];
                        tc::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + ab * kBlk + part * 32, v);
                        if (WTA) {
                            const bool all_in = dlo >= 0 && dlo + 62 <= D - 1;
                            if (it.side == 0) {
                                const int dbase = xi + kMargin - u0;
                                if (all_in) scan_chunk<0, false>(v, dbase, D, best, best_d);
                                else scan_chunk<0, true>(v, dbase, D, best, best_d);
                            } else {
                                const int dbase = u0 - xi;
                                if (all_in) scan_chunk<1, false>(v, dbase, D, best, best_d);
                                else scan_chunk<1, true>(v, dbase, D, best, best_d);
                            }
                        }
                        if (VOL) {
                            // Eight rows at a time go through the warp's staging block, so that every warp store
                            // is one 128-byte run of a [x][d] output row: row r of the chunk is pixel
                            // x0 + 32q + r, its column j is disparity d00 + r - j (left) or d00 - r + j (right).
                            float *vol = it.side == 0 ? p.left_vol : p.right_vol;
                            float *rowbase = vol + (((size_t)it.img * p.h + it.y) * p.w) * (size_t)D;
                            const int xq = kBlk * k + q * 32;
                            const int rows_ok = min(32, p.w - xq);
                            // disparity of this lane's column in row 0 of the chunk; it moves by dstep per row
                            const int dd0 = it.side == 0 ? q * 32 + kMargin - (u0 + lane) : (u0 + lane) - q * 32;
                            auto store_rounds = [&](auto dstep_c) {
                                constexpr int dstep = decltype(dstep_c)::value;
                                const int stride = D + dstep;               // floats from a row's slot to the next row's
                                float *p0 = rowbase + (xq * D + dd0);       // 32-bit offset: a row of a volume is < 2^31 floats
#pragma unroll 1
                                for (int r0 = 0; r0 < 32; r0 += kStageRows) {
                                    if (lane >= r0 && lane < r0 + kStageRows) {
#pragma unroll
                                        for (int i = 0; i < 8; ++i)
                                            reinterpret_cast<uint4 *>(stage + (lane - r0) * kStagePitch)[i] =
                                                make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                                    }
                                    __syncwarp();
                                    float t[kStageRows];
#pragma unroll
                                    for (int j = 0; j < kStageRows; ++j) t[j] = stage[j * kStagePitch + lane];
                                    float *pr = p0 + r0 * stride;
                                    const int dr = dd0 + r0 * dstep;
                                    if (r0 + kStageRows <= rows_ok) {
#pragma unroll
                                        for (int j = 0; j < kStageRows; ++j)
                                            if ((unsigned)(dr + j * dstep) < (unsigned)D) pr[j * stride] = t[j];
                                    } else {
#pragma unroll
                                        for (int j = 0; j < kStageRows; ++j)
                                            if (r0 + j < rows_ok && (unsigned)(dr + j * dstep) < (unsigned)D) pr[j * stride] = t[j];
                                    }
                                    __syncwarp();
                                }
                            };
                            if (it.side == 0) store_rounds(std::integral_constant<int, 1>());
                            else store_rounds(std::integral_constant<int, -1>());
                        }
                    }
                    tc::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive(bar_acc_empty + 8 * ab);
                }
                }
                if (!WTA) continue;
                // merge the four column parts of each row: larger value wins, equal values -> lower disparity.
                // The rows of a group of four warps (same q) are merged inside the group: named barriers 7 + q ("parts
                // 1-3 have written") and 11 + q ("part 0 has read, the buffer can be rewritten"), 128 threads each - a
                // barrier over all sixteen warps put the groups in lockstep once per tile, which serialised the
                // fill / drain phases of the bulk-store variant (0.38 ms with the fused WTA against 0.27 ms without).
                if (part != 0) {
                    if (T > 0) asm volatile("bar.sync %0, 128;" ::"r"(11 + q) : "memory");
                    wta_v[(part - 1) * 128 + xi] = best;
                    wta_d[(part - 1) * 128 + xi] = (uint16_t)(best_d == 0x7fffffff ? 0xffff : best_d);
                    asm volatile("bar.arrive %0, 128;" ::"r"(7 + q) : "memory");
                } else {
                    asm volatile("bar.sync %0, 128;" ::"r"(7 + q) : "memory");
#pragma unroll
                    for (int pt = 0; pt < 3; ++pt) {
                        const float ov = wta_v[pt * 128 + xi];
                        const int od = wta_d[pt * 128 + xi];
                        if (ov > best || (ov == best && od < best_d)) { best = ov; best_d = od; }
                    }
                    asm volatile("bar.arrive %0, 128;" ::"r"(11 + q) : "memory");
                    const int x = kBlk * k + xi;
                    if (x < p.w)
                        (it.side == 0 ? p.left_disp : p.right_disp)[((size_t)it.img * p.h + it.y) * p.w + x] =
                            (uint8_t)(best_d & 255);
                }
            }
        }
        if (WTA && part != 0 && T > 0) asm volatile("bar.sync %0, 128;" ::"r"(11 + q) : "memory");
        // the bulk stores this thread issued are complete (not only read out of shared memory) before the block ends
        if (BULK && part == 0 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }

    tc::tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc::tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
    }
}

}  // namespace tcs

int corr_sweep_margin() { return tcs::kMargin; }

int launch_corr_sweep(stereo_ctx *ctx, const void *l_hi, const void *l_lo, const void *r_hi, const void *r_lo, int n,
                      int h, int w, int wf, int ml, int D, int planes, float *left_vol, float *right_vol,
                      uint8_t *left_disp, uint8_t *right_disp)
{
    const int nb = (w + tcs::kBlk - 1) / tcs::kBlk;
    STEREO_REQUIRE(D >= 1 && D - 1 <= tcs::kMargin, "row-sweep correlation supports disp_range <= %d, got %d",
                   tcs::kMargin + 1, D);
    STEREO_REQUIRE(ml >= tcs::kMargin && wf >= ml + tcs::kBlk * (nb + 2), "feature margins too small");
    tcs::Params p;
    p.l_hi = static_cast<const __nv_bfloat16 *>(l_hi);
    p.l_lo = static_cast<const __nv_bfloat16 *>(l_lo);
    p.r_hi = static_cast<const __nv_bfloat16 *>(r_hi);
    p.r_lo = static_cast<const __nv_bfloat16 *>(r_lo);
    p.n = n; p.h = h; p.w = w; p.wf = wf; p.ml = ml; p.D = D; p.planes = planes;
    p.left_disp = left_disp; p.right_disp = right_disp;
    p.left_vol = left_vol; p.right_vol = right_vol;
    p.nb = nb;
    // rows may be cut into segments when that balances the persistent CTAs better: cost model =
    // rounds x (tiles per segment + half a tile of pipeline fill)
    const int rows = n * 2 * h;
    int best_segs = 1;
    double best_cost = 1e30;
    for (int segs = 1; segs <= 4 && segs <= nb; ++segs) {
        const int tps = (nb + segs - 1) / segs;
        const int real_segs = (nb + tps - 1) / tps;
        const long items = (long)rows * real_segs;
        const double cost = (double)((items + ctx->sm_count - 1) / ctx->sm_count) * (tps + 0.5);
        if (cost < best_cost) { best_cost = cost; best_segs = real_segs; }
    }
    if (const char *e = getenv("STEREO_CORR_SEGS")) {      // tuning override
        const int v = atoi(e);
        if (v >= 1 && v <= nb) best_segs = v;
    }
    p.tps = (nb + best_segs - 1) / best_segs;
    p.segs = (nb + p.tps - 1) / p.tps;
    p.nitems = rows * p.segs;
    const int grid = p.nitems < ctx->sm_count ? p.nitems : ctx->sm_count;
    const bool vol = left_vol != nullptr, wta = left_disp != nullptr;
    auto launch = [&](auto kern, int smem_bytes) -> int {
        STEREO_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
        ProfScope prof(ctx, STEREO_K_CORR);
        kern<<<grid, tcs::kThreads, smem_bytes, ctx->stream>>>(p);
        ctx->launches += 1;
        STEREO_CUDA_CHECK(cudaGetLastError());
        return STEREO_OK;
    };
    // Materialising variants.  16-byte aligned volumes with D % 4 == 0 (D <= 256) leave through one bulk copy per
    // 32-pixel group (STEREO_CORR_STORE=1 keeps the per-row stores as the cross-check); the rest take the per-row
    // stores, with the disparity range as a compile-time constant for the reference's default 201.
    static const bool no_bulk = [] { const char *e = getenv("STEREO_CORR_STORE"); return e && atoi(e) == 1; }();
    const bool bulk = vol && !no_bulk && D % 4 == 0 && D * 64 <= tcs::kBulkStageBytes &&      // 16 rows x D x 4 bytes
                      ((uintptr_t)left_vol % 16 == 0) && ((uintptr_t)right_vol % 16 == 0);
    if (vol && wta) {
        if (bulk) return launch(tcs::corr_sweep_kernel<true, true, 0, true>, tcs::Layout<true, true>::kTotal);
        if (D == 228) return launch(tcs::corr_sweep_kernel<true, true, 228>, tcs::Layout<true>::kTotal);
        if (D == 201) return launch(tcs::corr_sweep_kernel<true, true, 201>, tcs::Layout<true>::kTotal);
        return launch(tcs::corr_sweep_kernel<true, true, 0>, tcs::Layout<true>::kTotal);
    }
    if (vol) {
        if (bulk) return launch(tcs::corr_sweep_kernel<true, false, 0, true>, tcs::Layout<true, true>::kTotal);
        if (D == 228) return launch(tcs::corr_sweep_kernel<true, false, 228>, tcs::Layout<true>::kTotal);
        if (D == 201) return launch(tcs::corr_sweep_kernel<true, false, 201>, tcs::Layout<true>::kTotal);
        return launch(tcs::corr_sweep_kernel<true, false, 0>, tcs::Layout<true>::kTotal);
    }
    STEREO_TRY(launch(tcs::corr_sweep_kernel<false, true, 0>, tcs::Layout<false>::kTotal));
    return STEREO_OK;
}

}  // namespace stereo
