// Fused correlation + WTA of BOTH volumes from one sweep (/root/reference/inference.py:139-183).
//
// corr_sweep.cu computes the product of a row's features twice: own = L for the left volume, own = R for the right
// one.  The two volumes hold the same numbers, right[x][d] = left[x + d][d] (inference.py:160-183), so here a CTA
// runs the LEFT sweep only (own = L block k, candidates = R blocks k-2, k-1, k; S[xi][u], d = xi + 256 - u) and
// reads every accumulator chunk twice:
//   along its rows    -> arg max over u for pixel xl = 128 k + xi            : the left disparity (as in corr_sweep)
//   along its columns -> arg max over xi for pixel xr = 128 (k - 2) + u      : the right disparity
// A right pixel meets its candidates in three consecutive tiles (its block is the window's third, second, first
// block); warp (q, part) always sees the same 32 columns of a block, so lane i keeps the running best of column i
// for the three blocks in flight in registers.  When a block leaves the window the four row groups q write their
// partial result to shared memory and warp 3 merges them and stores the disparities.
//
// Columns run across the lanes of a warp (tcgen05.ld hands every lane one accumulator row), so the column scan goes
// through shared memory: 16 columns of the chunk at a time are written row by row into a per-warp buffer and read
// back column by column (scan_cols).  Ties resolve as in the reference's argmax: the lowest disparity, i.e. the last
// column of a row and the first row of a column.  Slots of the right volume the reference leaves zero (x + d >= w)
// take part in its arg max: rows up to the end of the last block are zero feature rows (the product is 0.0f), the
// rows beyond enter as one 0.0f candidate when the block is flushed.
//
// Measured on one B200 (375 x 1242, D = 228, batch of 16 pairs, bf16x3): 102 us per pair against 121 us for the two
// sweeps of corr_sweep.cu on the same box (scripts/corr_fused_timing.py); 68 us without the column scan.
#include <cuda_bf16.h>

#include <cstdlib>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {

namespace tcb {

constexpr int kBlk = 128;                       // pixels per block
constexpr int kRun = kBlk * 16;                 // one channel chunk: 2 KB
constexpr int kPlane = 8 * kRun;                // 16 KB
constexpr int kBlock = 2 * kPlane;              // hi + lo: 32 KB
constexpr int kAccBlocks = 4;
// Measured first for the column maxima: one redux.sync.max.f32 + ballot per column - 139 us per pair, the reductions
// run at about one per eight cycles and SM.  Eight columns at a time through shared memory (16 KB, leaves room for a
// third own-block slot): 113 us; sixteen (32 KB, two own-block slots): 102 us.  (The sweep without any column scan -
// a build that existed only for this measurement - took 68 us.)
#ifndef STEREO_BOTH_NC
#define STEREO_BOTH_NC 16
#endif
constexpr int kNC = STEREO_BOTH_NC;             // columns of a chunk that go through shared memory at a time
static_assert(kNC == 8 || kNC == 16, "half or quarter chunks");
constexpr int kASlots = kNC == 16 ? 2 : 3, kBSlots = 3;
constexpr int kXposeWarp = 32 * kNC * 4;        // [32 rows][kNC columns] fp32 per epilogue warp
struct Layout {
    static constexpr int kA = 0;
    static constexpr int kB = kA + kASlots * kBlock;
    static constexpr int kWtaV = kB + kBSlots * kBlock;    // [3 parts][128] best value of column parts 1-3 (left map)
    static constexpr int kWtaD = kWtaV + 3 * 128 * 4;      // [3 parts][128] its disparity (16 bit)
    static constexpr int kColV = kWtaD + 3 * 128 * 2;      // [4 slots][4 q][128] best value of a right pixel over the rows of group q
    static constexpr int kColD = kColV + 4 * 4 * 128 * 4;  // [4 slots][4 q][128] its disparity
    static constexpr int kXpose = kColD + 4 * 4 * 128 * 4;
    static constexpr int kBars = kXpose + 16 * kXposeWarp;
    static constexpr int kTotal = kBars + 256;
    static_assert(kTotal <= 232448 && kColV % 16 == 0 && kBars % 8 == 0, "shared memory budget of one CTA");
};
constexpr int kEpiWarps = 16;
constexpr int kThreads = 128 + 32 * kEpiWarps;
constexpr uint32_t kTmemCols = 512;
constexpr int kMargin = 256;
constexpr int kNoDisp = 0x7fffffff;

struct Params {
    const __nv_bfloat16 *l_hi, *l_lo, *r_hi, *r_lo;   // [n][h][8][wf][8], pixel x at column ml + x
    int n, h, w, wf, ml, D, planes;
    uint8_t *left_disp, *right_disp;                   // [n][h][w]
    int nb, nitems;                                    // blocks per row, rows
    size_t img_stride;                                 // elements between consecutive images of one side
};

// Row scan (left map), as in corr_sweep.cu: running arg max of the lane's pixel over the 32 columns of the chunk,
// column i = disparity dbase - i (">=": the disparity falls as the column rises, equal values keep the lowest).
// Measured alternatives, all slower on the GPU although they execute fewer ALU-pipe instructions: four independent
// chains of eight columns; maximum first (a tree of three-input max) and the position by equality, with the index
// move as a SEL or as a predicated IMAD on the FMA pipe (108 us per pair against 102 us).
template <bool CHECK>
__device__ __forceinline__ void scan_rows(const uint32_t (&v)[32], int dbase, int D, float &best, int &best_d)
{
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        const int d = dbase - i;
        const float f = __uint_as_float(v[i]);
        const bool take = (!CHECK || (unsigned)d < (unsigned)D) && f >= best;
        best = take ? f : best;
        best_d = take ? d : best_d;
    }
}

// Column scan through shared memory, kNC columns of the chunk at a time.  `xp`: the warp's [32 rows][kNC columns]
// buffer; the 16-byte pieces of a row are stored at piece ^ f(row) (kNC = 16: f = (row >> 1) & 3, kNC = 8:
// f = (row >> 2) & 1), which keeps the 128-bit stores of a quarter warp and the 32-bit loads of the whole warp free of
// bank conflicts.  Per pass every lane writes its row, then lane (col = lane % kNC, hr = lane / kNC) scans rows hr,
// hr + 32 / kNC, ... of column kNC * pass + col (disparities rise with the row: ">" keeps the lowest among equals) and
// the lanes of a column merge.  Lane i leaves with column i of the chunk: its maximum and the disparity
// dcol0 + row - column of the lowest row that holds it.
template <bool CHECK>
__device__ __forceinline__ void scan_cols(const uint32_t (&v)[32], float *xp, int lane, int dcol0, int D, float &cmax,
                                          int &cd)
{
    constexpr int kLpc = 32 / kNC, kPieces = kNC / 4;          // lanes per column, 16-byte pieces per row
    const int col = lane % kNC, hr = lane / kNC;
    uint4 *wrow = reinterpret_cast<uint4 *>(xp + lane * kNC);
    const int wsw = kNC == 16 ? (lane >> 1) & 3 : (lane >> 2) & 1;
    const float *rbase = xp + hr * kNC + (col & 3);
    const int rpiece = col >> 2;
#pragma unroll
    for (int pass = 0; pass < 32 / kNC; ++pass) {
#pragma unroll
        for (int i4 = 0; i4 < kPieces; ++i4)
            wrow[i4 ^ wsw] = make_uint4(v[kNC * pass + 4 * i4], v[kNC * pass + 4 * i4 + 1], v[kNC * pass + 4 * i4 + 2],
                                        v[kNC * pass + 4 * i4 + 3]);
        __syncwarp();
        float bv = -INFINITY;
        int br = 0;
        const int d0 = dcol0 + hr - (kNC * pass + col);        // disparity of row hr in this lane's column
#pragma unroll
        for (int rr = 0; rr < kNC; ++rr) {                     // 32 / kLpc rows
            const float f = rbase[32 * rr + 4 * (rpiece ^ (rr & (kPieces - 1)))];
            const bool take = (!CHECK || (unsigned)(d0 + kLpc * rr) < (unsigned)D) && f > bv;
            bv = take ? f : bv;
            br = take ? rr : br;
        }
        int bd = d0 + kLpc * br;
#pragma unroll
        for (int sh = kNC; sh < 32; sh <<= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, sh);
            const int od = __shfl_xor_sync(0xffffffffu, bd, sh);
            if (ov > bv || (ov == bv && od < bd)) { bv = ov; bd = od; }
        }
        if (hr == pass) { cmax = bv; cd = bd; }
        __syncwarp();
    }
}

__global__ __launch_bounds__(kThreads, 1) void corr_both_kernel(Params p)
{
    using L = Layout;
    constexpr int kA = L::kA, kB = L::kB, kWtaV = L::kWtaV, kWtaD = L::kWtaD, kColV = L::kColV, kColD = L::kColD,
                  kBars = L::kBars;
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t s_base = tc::smem_u32(smem);
    const uint32_t s_a = s_base + kA, s_b = s_base + kB;
    const uint32_t bar_a_full = s_base + kBars;              // [kASlots <= 3]
    const uint32_t bar_a_empty = bar_a_full + 24;            // [kASlots]
    const uint32_t bar_b_full = bar_a_full + 48;             // [4]
    const uint32_t bar_b_empty = bar_a_full + 80;            // [4]
    const uint32_t bar_acc_full = bar_a_full + 112;          // [4]
    const uint32_t bar_acc_empty = bar_a_full + 144;         // [4]
    // [4] the 16 warps have written their column results of tile T (barrier T & 3: warp 3 may lag up to three tiles
    // behind the epilogue, which waits for slot e - 4 of the column results before it writes slot e)
    const uint32_t bar_col_done = bar_a_full + 176;
    const uint32_t bar_col_free = bar_a_full + 208;          // [4] warp 3 has read a slot of column results
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + kBars + 240);

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
            tc::mbar_init(bar_col_free + 8 * i, 1);
            tc::mbar_init(bar_col_done + 8 * i, kEpiWarps);
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
    float *col_v = reinterpret_cast<float *>(smem + kColV);
    int *col_d = reinterpret_cast<int *>(smem + kColD);

    if (warp == 0) {
        if (lane == 0) {
            // ---- TMA producer: candidate blocks two tiles ahead, own blocks one tile ahead ----
            uint32_t E = 0, T = 0;                   // running candidate-block and tile counters
            for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
                const int img = item / p.h;                               // item = img * h + y
                const size_t row0 = (size_t)img * p.img_stride + (size_t)(item - img * p.h) * 8 * p.wf * 8;
                const __nv_bfloat16 *own[2] = {p.l_hi, p.l_lo};
                const __nv_bfloat16 *oth[2] = {p.r_hi, p.r_lo};
                auto load_block = [&](uint32_t dst, const __nv_bfloat16 *const *src, int col, uint32_t bar) {
                    tc::mbar_expect_tx(bar, (uint32_t)p.planes * kPlane);
                    for (int pl = 0; pl < p.planes; ++pl)
                        for (int cc = 0; cc < 8; ++cc)
                            tc::bulk_g2s(dst + pl * kPlane + cc * kRun,
                                         src[pl] + row0 + ((size_t)cc * p.wf + col) * 8, kRun, bar);
                };
                auto load_entry = [&](int j) {       // j-th candidate block of this row = R block j - 2
                    const uint32_t slot = E % kBSlots, ph = (E / kBSlots) & 1;
                    ++E;
                    tc::mbar_wait(bar_b_empty + 8 * slot, ph ^ 1);
                    load_block(s_b + slot * kBlock, oth, p.ml + kBlk * (j - 2), bar_b_full + 8 * slot);
                };
                auto load_own = [&](int k) {
                    const uint32_t slot = T % kASlots, ph = (T / kASlots) & 1;
                    ++T;
                    tc::mbar_wait(bar_a_empty + 8 * slot, ph ^ 1);
                    load_block(s_a + slot * kBlock, own, p.ml + kBlk * k, bar_a_full + 8 * slot);
                };
                load_own(0);
                load_entry(0);
                load_entry(1);
                for (int k = 0; k < p.nb; ++k) {
                    load_entry(k + 2);
                    if (k + 1 < p.nb) load_own(k + 1);
                }
            }
        }
    } else if (warp == 1) {
        // ---- MMA issuer: one thread waits, issues and commits.  (With the whole warp in the loop and an elected lane
        // per MMA the issue sequence took ~28 instructions per MMA, ~140 cycles against the 64 the tensor pipe needs
        // for a 128 x 128 x 16 MMA: the issuer, not the pipe, bounded the kernel - ncu round 2, tensor pipe 36 % busy.)
        if (lane == 0) {
            constexpr uint32_t idesc = tc::make_idesc(kBlk, kBlk);
            constexpr uint32_t kStep = (2 * kRun) >> 4, kLo = kPlane >> 4;      // descriptor units of 16 bytes
            const uint32_t d_hi = tc::desc_hi(128);
            const int n_mma = p.planes == 2 ? 12 : 4;
            uint32_t E0 = 0, T = 0, G = 0;           // first candidate block of the item, tiles, accumulator blocks
            bool prewaited = false;                  // barriers of the coming block already waited for
            for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
                for (int k = 0; k < p.nb; ++k) {
                    const uint32_t as = T % kASlots;
                    tc::mbar_wait(bar_a_full + 8 * as, (T / kASlots) & 1);
                    const uint32_t a0 = tc::desc_lo(s_a + as * kBlock, kRun);
                    const uint32_t e_first = E0 + (uint32_t)k;
#pragma unroll 1
                    for (int o = 0; o < 3; ++o, ++G) {
                        const uint32_t e = e_first + o, bs = e % kBSlots;
                        const uint32_t ab = G % kAccBlocks;
                        if (!prewaited) {
                            if (k == 0 || o == 2) tc::mbar_wait(bar_b_full + 8 * bs, (e / kBSlots) & 1);
                            tc::mbar_wait(bar_acc_empty + 8 * ab, ((G / kAccBlocks) & 1) ^ 1);
                        }
                        prewaited = false;
                        tc::tc_fence_after();
                        const uint32_t b0 = tc::desc_lo(s_b + bs * kBlock, kRun);
                        const uint32_t acc = tmem_base + ab * kBlk;
                        // the same order of terms as corr_sweep.cu's left sweep: L_lo*R_hi, L_hi*R_lo, L_hi*R_hi
#pragma unroll
                        for (int m = 0; m < 12; ++m) {
                            if (m < n_mma) {
                                const int combo = m >> 2, ks = m & 3;
                                const uint32_t apl = (n_mma == 12 && combo == 0) ? kLo : 0;
                                const uint32_t bpl = (n_mma == 12 && combo == 1) ? kLo : 0;
                                tc::umma_bf16_parts(acc, a0 + apl + ks * kStep, d_hi, b0 + bpl + ks * kStep, d_hi, idesc, m > 0);
                            }
                            // With the last MMAs of the block still to issue and several queued in the tensor pipe,
                            // wait for what the next block of this tile needs: its accumulator and its candidates.
                            if (m == 8 && o < 2) {
                                if (k == 0 || o == 1) {
                                    const uint32_t en = e + 1;
                                    tc::mbar_wait(bar_b_full + 8 * (en % kBSlots), (en / kBSlots) & 1);
                                }
                                tc::mbar_wait(bar_acc_empty + 8 * ((G + 1) % kAccBlocks), (((G + 1) / kAccBlocks) & 1) ^ 1);
                                prewaited = true;
                            }
                        }
                        tc::tc_commit(bar_acc_full + 8 * ab);
                        // the oldest candidate block has had its last use (at the end of a row all three have)
                        if (o == 0 || k == p.nb - 1) tc::tc_commit(bar_b_empty + 8 * bs);
                    }
                    tc::tc_commit(bar_a_empty + 8 * as);
                    ++T;
                }
                E0 += (uint32_t)p.nb + 2;
            }
        }
    } else if (warp == 3) {
        // ---- right map: merge the four row groups of every block that has left the window ----
        uint32_t E0 = 0, T = 0;
        for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
            uint8_t *out = p.right_disp + (size_t)item * p.w;
            for (int k = 0; k < p.nb; ++k, ++T) {
                tc::mbar_wait(bar_col_done + 8 * (T & 3), (T >> 2) & 1);
                const int nflush = k == p.nb - 1 ? 3 : 1;
                for (int f = 0; f < nflush; ++f) {
                    const uint32_t e = E0 + (uint32_t)k + f, slot = e & 3;
                    const int j = k - 2 + f;                      // R block
                    if (j >= 0) {
#pragma unroll
                        for (int m = 0; m < 4; ++m) {
                            const int c = lane + 32 * m, xr = kBlk * j + c;
                            float bv = -INFINITY;
                            int bd = kNoDisp;
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const float ov = col_v[(slot * 4 + q) * 128 + c];
                                const int od = col_d[(slot * 4 + q) * 128 + c];
                                if (ov > bv || (ov == bv && od < bd)) { bv = ov; bd = od; }
                            }
                            // rows beyond the last block: right[x][d] stays 0.0f there (x + d >= w)
                            const int d0 = kBlk * p.nb - xr;
                            if (d0 < p.D && (0.0f > bv || (0.0f == bv && d0 < bd))) bd = d0;
                            if (xr < p.w) out[xr] = (uint8_t)(bd & 255);
                        }
                    }
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive(bar_col_free + 8 * slot);
                }
            }
            E0 += (uint32_t)p.nb + 2;
        }
    } else if (warp >= 4) {
        // ---- epilogue: warp (q, part) drains columns 32*part.. of every accumulator block, lanes 32q.. ----
        const int q = warp & 3, part = (warp - 4) >> 2;
        const int xi = q * 32 + lane;
        const int D = p.D;
        float *wta_v = reinterpret_cast<float *>(smem + kWtaV);
        uint16_t *wta_d = reinterpret_cast<uint16_t *>(smem + kWtaD);
        uint32_t T = 0, G = 0, E0 = 0;
        for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
            // running best of this lane's column in the window's three blocks (oldest first)
            float c0v = -INFINITY, c1v = -INFINITY, c2v = -INFINITY;
            int c0d = kNoDisp, c1d = kNoDisp, c2d = kNoDisp;
            for (int k = 0; k < p.nb; ++k, ++T) {
                float best = -INFINITY;
                int best_d = kNoDisp;
#pragma unroll 1
                for (int o = 0; o < 3; ++o, ++G) {
                    const uint32_t ab = G % kAccBlocks;
                    tc::mbar_wait(bar_acc_full + 8 * ab, (G / kAccBlocks) & 1);
                    tc::tc_fence_after();
                    const int u0 = 128 * o + 32 * part;                  // first window column of this chunk
                    // disparities this warp can see in the chunk (warp-uniform): skip it when none is in [0, D)
                    const int dlo = 32 * q + kMargin - u0 - 31;
                    if (dlo + 62 >= 0 && dlo <= D - 1) {
                        uint32_t v[32];
                        tc::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + ab * kBlk + part * 32, v);
                        // the values are in registers: the issuer may overwrite the block
                        tc::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) tc::mbar_arrive(bar_acc_empty + 8 * ab);
                        const int dbase = xi + kMargin - u0;
                        const bool all_in = dlo >= 0 && dlo + 62 <= D - 1;
                        float cmax = -INFINITY;
                        int cd = kNoDisp;               // column `lane` of the chunk: its maximum over the rows, its disparity
                        if (all_in) scan_rows<false>(v, dbase, D, best, best_d);
                        else scan_rows<true>(v, dbase, D, best, best_d);
                        {
                            float *xp = reinterpret_cast<float *>(smem + L::kXpose + (warp - 4) * kXposeWarp);
                            if (all_in) scan_cols<false>(v, xp, lane, 32 * q + kMargin - u0, D, cmax, cd);
                            else scan_cols<true>(v, xp, lane, 32 * q + kMargin - u0, D, cmax, cd);
                        }
                        // later tiles bring larger disparities: only a larger value replaces the running best
                        if (o == 0) { if (cmax > c0v) { c0v = cmax; c0d = cd; } }
                        else if (o == 1) { if (cmax > c1v) { c1v = cmax; c1d = cd; } }
                        else { if (cmax > c2v) { c2v = cmax; c2d = cd; } }
                    } else {
                        tc::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) tc::mbar_arrive(bar_acc_empty + 8 * ab);
                    }
                }
                // ---- right map: the window's first block is complete (at the end of the row all three are) ----
                {
                    const int nflush = k == p.nb - 1 ? 3 : 1;
                    for (int f = 0; f < nflush; ++f) {
                        const uint32_t e = E0 + (uint32_t)k + f, slot = e & 3;
                        if (e >= 4) tc::mbar_wait(bar_col_free + 8 * slot, ((e >> 2) - 1) & 1);
                        const float cv = f == 0 ? c0v : (f == 1 ? c1v : c2v);
                        const int cdd = f == 0 ? c0d : (f == 1 ? c1d : c2d);
                        col_v[(slot * 4 + q) * 128 + part * 32 + lane] = cv;
                        col_d[(slot * 4 + q) * 128 + part * 32 + lane] = cdd;
                    }
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive(bar_col_done + 8 * (T & 3));
                    c0v = c1v; c0d = c1d;
                    c1v = c2v; c1d = c2d;
                    c2v = -INFINITY; c2d = kNoDisp;
                }
                // ---- left map: merge the four column parts of each row inside the group (corr_sweep.cu) ----
                if (part != 0) {
                    if (T > 0) asm volatile("bar.sync %0, 128;" ::"r"(11 + q) : "memory");
                    wta_v[(part - 1) * 128 + xi] = best;
                    wta_d[(part - 1) * 128 + xi] = (uint16_t)(best_d == kNoDisp ? 0xffff : best_d);
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
                    if (x < p.w) p.left_disp[(size_t)item * p.w + x] = (uint8_t)(best_d & 255);
                }
            }
            E0 += (uint32_t)p.nb + 2;
        }
        if (part != 0 && T > 0) asm volatile("bar.sync %0, 128;" ::"r"(11 + q) : "memory");
    }

    tc::tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc::tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
    }
}

}  // namespace tcb

int launch_corr_both(stereo_ctx *ctx, const void *l_hi, const void *l_lo, const void *r_hi, const void *r_lo, int n,
                     int h, int w, int wf, int ml, int D, int planes, uint8_t *left_disp, uint8_t *right_disp,
                     size_t img_stride)
{
    const int nb = (w + tcb::kBlk - 1) / tcb::kBlk;
    STEREO_REQUIRE(D >= 1 && D - 1 <= tcb::kMargin, "one-sweep correlation supports disp_range <= %d, got %d",
                   tcb::kMargin + 1, D);
    STEREO_REQUIRE(ml >= tcb::kMargin && wf >= ml + tcb::kBlk * nb, "feature margins too small");
    STEREO_REQUIRE(left_disp && right_disp, "the one-sweep kernel writes both maps");
    tcb::Params p;
    p.l_hi = static_cast<const __nv_bfloat16 *>(l_hi);
    p.l_lo = static_cast<const __nv_bfloat16 *>(l_lo);
    p.r_hi = static_cast<const __nv_bfloat16 *>(r_hi);
    p.r_lo = static_cast<const __nv_bfloat16 *>(r_lo);
    p.n = n; p.h = h; p.w = w; p.wf = wf; p.ml = ml; p.D = D; p.planes = planes;
    p.left_disp = left_disp; p.right_disp = right_disp;
    p.nb = nb;
    p.nitems = n * h;
    p.img_stride = img_stride ? img_stride : (size_t)h * 8 * wf * 8;
    const int grid = p.nitems < ctx->sm_count ? p.nitems : ctx->sm_count;
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(tcb::corr_both_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, tcb::Layout::kTotal));
    {
        ProfScope prof(ctx, STEREO_K_CORR);
        tcb::corr_both_kernel<<<grid, tcb::kThreads, tcb::Layout::kTotal, ctx->stream>>>(p);
    }
    ctx->launches += 1;
    STEREO_CUDA_CHECK(cudaGetLastError());
    return STEREO_OK;
}

}  // namespace stereo
