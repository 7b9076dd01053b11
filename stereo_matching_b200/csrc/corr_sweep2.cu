// Fused correlation + WTA on CTA pairs (/root/reference/inference.py:139-183).
//
// Same row sweep as corr_sweep.cu, but two CTAs (one thread-block cluster, tcgen05 cta_group::2) work on
// one image row together.  Measured on B200 the single-CTA kernel is limited by what an M=128 x N=128
// MMA has to fetch from shared memory (8 KB for 64 cycles of math); a CTA pair runs M=256 x N=256 MMAs
// in which every CTA fetches its own 128 rows of A and only HALF of B - the design point at which the
// operand fetch and the math take the same time.
//
// A pair-tile is 256 consecutive pixels of one side: CTA c owns block 2t + c (128 pixels).  The 512
// candidate columns are four 128-pixel blocks of the other view; an N = 256 MMA takes its first 128
// columns from CTA 0's shared memory and the other 128 from CTA 1's, so CTA 0 keeps the even and CTA 1
// the odd candidate blocks, each in a ring of four, and every tile costs each CTA one new candidate
// block and one own block (64 KB), loaded by TMA two tiles ahead.
//
//   left  sweep: candidates = R blocks 2t-2 .. 2t+1,  d = 128c + xi + 256 - u
//   right sweep: candidates = L blocks 2t   .. 2t+3,  d = u - 128c - xi
//   (xi = pixel inside the CTA's own block, u = column of the 512-wide window)
//
// The accumulators fill tensor memory (2 x 256 columns x 128 lanes per CTA): while the pair multiplies
// one half of the window, the 16 epilogue warps of each CTA drain the other half - tcgen05.ld, running
// arg max in registers, chunks outside the disparity band skipped - and the four column parts of a row
// are merged through shared memory at the end of the tile.  The volume never exists.
//
// Cross-CTA synchronisation as in conv_tc3.cu: CTA 1 forwards "block landed" to the leader with remote
// mbarrier arrives, the leader's tcgen05.commit is multicast to both CTAs, and both epilogues release
// the accumulator halves on the leader's barriers.
#include <cuda_bf16.h>

#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "tc_common.cuh"

namespace stereo {

namespace tcs2 {

using namespace tc;

constexpr int kBlk = 128;                       // pixels per block
constexpr int kRun = kBlk * 16;                 // one channel chunk: 2 KB
constexpr int kPlane = 8 * kRun;                // 16 KB
constexpr int kBlock = 2 * kPlane;              // hi + lo: 32 KB
constexpr int kASlots = 3, kBSlots = 4;
constexpr int kSmemA = 0;
constexpr int kSmemB = kSmemA + kASlots * kBlock;
constexpr int kSmemWtaV = kSmemB + kBSlots * kBlock;         // [3 parts][128] best value of column parts 1-3
constexpr int kSmemWtaD = kSmemWtaV + 3 * 128 * 4;           // [3 parts][128] its disparity (16 bit)
constexpr int kSmemBars = kSmemWtaD + 3 * 128 * 2;
constexpr int kSmemTotal = kSmemBars + 256;
static_assert(kSmemTotal <= 232448, "shared memory budget of one CTA");
constexpr int kEpiWarps = 16;
constexpr int kThreads = 128 + 32 * kEpiWarps;
constexpr uint32_t kTmemCols = 512;
constexpr int kHalf = 256;                      // accumulator columns of one MMA
constexpr int kMargin = 256;                    // zero columns the feature rows need left of pixel 0

struct Params {
    const __nv_bfloat16 *l_hi, *l_lo, *r_hi, *r_lo;   // [n][h][8][wf][8], pixel x at column ml + x
    int n, h, w, wf, ml, D, planes;
    uint8_t *left_disp, *right_disp;                   // [n][h][w]
    long long *stamps;                                 // STEREO_CORR_DEBUG=16: clock stamps of pair 0
    int nt, nitems;                                    // pair-tiles per row, rows x sides
};

// Arg max over one 32-column chunk of a row.  d = dbase - i (SIDE 0, left volume) or dbase + i (SIDE 1).
// Equal values must keep the lowest d: with d falling along the chunk ">=" does that, with d rising ">".
// CHECK = false when every lane's 32 disparities are known to lie in [0, D).
template <int SIDE, bool CHECK>
__device__ __forceinline__ void scan_chunk(const uint32_t (&v)[32], int dbase, int D, float &best, int &best_d)
{
    // the running maximum is a chain of max instructions only; the index follows it off the critical path
    int bi = -1;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        const float f = __uint_as_float(v[i]);
        const bool valid = !CHECK || (unsigned)(SIDE == 0 ? dbase - i : dbase + i) < (unsigned)D;
        const bool take = valid && (SIDE == 0 ? f >= best : f > best);
        bi = take ? i : bi;
        best = fmaxf(best, valid ? f : -INFINITY);
    }
    if (bi >= 0) best_d = SIDE == 0 ? dbase - bi : dbase + bi;
}

__global__ __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1) void corr_sweep2_kernel(Params p)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int c = (int)cluster_rank();               // 0 = leader (issues the MMAs)
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
    const uint32_t s_base = smem_u32(smem);
    const uint32_t s_a = s_base + kSmemA, s_b = s_base + kSmemB;
    const uint32_t bar_a_full = s_base + kSmemBars;          // [3] own block landed
    const uint32_t bar_peer_a = bar_a_full + 24;             // [3] (leader) CTA 1's own block landed
    const uint32_t bar_a_empty = bar_a_full + 48;            // [3] own block consumed (multicast commit)
    const uint32_t bar_b_full = bar_a_full + 72;             // [4] candidate block landed
    const uint32_t bar_peer_b = bar_a_full + 104;            // [4] (leader) CTA 1's candidate block landed
    const uint32_t bar_b_empty = bar_a_full + 136;           // [4] candidate block consumed (multicast commit)
    const uint32_t bar_acc_full = bar_a_full + 168;          // [2] window half complete (multicast commit)
    const uint32_t bar_acc_empty = bar_a_full + 184;         // [2] (leader) both epilogues have drained the half
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + kSmemBars + 208);

    if (tid == 0) {
        for (int i = 0; i < kASlots; ++i) {
            mbar_init(bar_a_full + 8 * i, 1);
            mbar_init(bar_peer_a + 8 * i, 1);
            mbar_init(bar_a_empty + 8 * i, 1);
        }
        for (int i = 0; i < kBSlots; ++i) {
            mbar_init(bar_b_full + 8 * i, 1);
            mbar_init(bar_peer_b + 8 * i, 1);
            mbar_init(bar_b_empty + 8 * i, 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar_acc_full + 8 * i, 1);
            mbar_init(bar_acc_empty + 8 * i, 2 * kEpiWarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(kTmemCols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                              // the peer's barriers exist before anyone arrives remotely
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const size_t img_elems = (size_t)p.h * 8 * p.wf * 8;
    // item -> (image, row, side); left and right sweep of a row are neighbours, so they run at the same
    // time on neighbouring pairs and the second reader of every feature block finds it in L2
    auto decode = [&](int item, int &img, int &y, int &side) {
        side = item & 1;
        const int rest = item >> 1;
        y = rest % p.h;
        img = rest / p.h;
    };

    if (warp == 0) {
        if (lane == 0) {
            // ---- TMA producer: candidate blocks two tiles ahead, own blocks two tiles ahead ----
            uint32_t E = 0, T = 0;                   // running candidate-block and tile counters
            for (int item = pair; item < p.nitems; item += npairs) {
                int img, y, side;
                decode(item, img, y, side);
                const size_t row0 = (size_t)img * img_elems + (size_t)y * 8 * p.wf * 8;
                const __nv_bfloat16 *own[2] = {side == 0 ? p.l_hi : p.r_hi, side == 0 ? p.l_lo : p.r_lo};
                const __nv_bfloat16 *oth[2] = {side == 0 ? p.r_hi : p.l_hi, side == 0 ? p.r_lo : p.l_lo};
                auto load_block = [&](uint32_t dst, const __nv_bfloat16 *const *src, int col, uint32_t bar) {
                    mbar_expect_tx(bar, (uint32_t)p.planes * kPlane);
                    for (int pl = 0; pl < p.planes; ++pl)
                        for (int cc = 0; cc < 8; ++cc)
                            bulk_g2s(dst + pl * kPlane + cc * kRun, src[pl] + row0 + ((size_t)cc * p.wf + col) * 8, kRun, bar);
                };
                auto load_entry = [&](int e) {       // e-th candidate block of this CTA in this row
                    const uint32_t slot = E % kBSlots, ph = (E / kBSlots) & 1;
                    ++E;
                    mbar_wait(bar_b_empty + 8 * slot, ph ^ 1);
                    load_block(s_b + slot * kBlock, oth, p.ml + kBlk * (2 * e + c - (side == 0 ? 2 : 0)), bar_b_full + 8 * slot);
                };
                auto load_own = [&](int t) {
                    const uint32_t slot = T % kASlots, ph = (T / kASlots) & 1;
                    ++T;
                    mbar_wait(bar_a_empty + 8 * slot, ph ^ 1);
                    load_block(s_a + slot * kBlock, own, p.ml + kBlk * (2 * t + c), bar_a_full + 8 * slot);
                };
                load_own(0);
                load_entry(0);
                load_entry(1);
                for (int t = 0; t < p.nt; ++t) {
                    if (t + 2 <= p.nt) load_entry(t + 2);
                    if (t + 1 < p.nt) load_own(t + 1);
                }
            }
        }
    } else if (warp == 3) {
        if (lane == 0 && c == 1) {
            // ---- CTA 1: tell the leader when this CTA's blocks have landed (same order as the producer) ----
            uint32_t E = 0, T = 0;
            auto fwd_entry = [&]() {
                const uint32_t slot = E % kBSlots, ph = (E / kBSlots) & 1;
                ++E;
                mbar_wait(bar_b_full + 8 * slot, ph);
                mbar_arrive_remote(bar_peer_b + 8 * slot, 0);
            };
            auto fwd_own = [&]() {
                const uint32_t slot = T % kASlots, ph = (T / kASlots) & 1;
                ++T;
                mbar_wait(bar_a_full + 8 * slot, ph);
                mbar_arrive_remote(bar_peer_a + 8 * slot, 0);
            };
            for (int item = pair; item < p.nitems; item += npairs) {
                fwd_own();
                fwd_entry();
                fwd_entry();
                for (int t = 0; t < p.nt; ++t) {
                    if (t + 2 <= p.nt) fwd_entry();
                    if (t + 1 < p.nt) fwd_own();
                }
            }
        }
    } else if (warp == 1) {
        if (c == 0) {
            // ---- MMA issuer (leader CTA; the whole warp runs the loop, one elected lane issues) ----
            constexpr uint32_t idesc = make_idesc(2 * kBlk, kHalf);
            const uint32_t d_hi = desc_hi(128);
            const int n_combo = p.planes == 2 ? 3 : 1;
            uint32_t E0 = 0, T = 0, G = 0;           // first candidate block of the row, tiles, window halves
            for (int item = pair; item < p.nitems; item += npairs) {
                int img, y, side;
                decode(item, img, y, side);
                // L_lo*R_hi, then L_hi*R_lo, then L_hi*R_hi on both sides: the left and the right
                // volume accumulate the same terms in the same order
                const int lo_first = side == 0 ? 0 : 1;
                for (int t = 0; t < p.nt; ++t, ++T) {
                    const uint32_t as = T % kASlots;
                    mbar_wait(bar_a_full + 8 * as, (T / kASlots) & 1);
                    mbar_wait(bar_peer_a + 8 * as, (T / kASlots) & 1);
#pragma unroll 1
                    for (int h = 0; h < 2; ++h, ++G) {
                        const uint32_t e = E0 + (uint32_t)(t + h), bs = e % kBSlots;
                        if (t == 0 || h == 1) {
                            mbar_wait(bar_b_full + 8 * bs, (e / kBSlots) & 1);
                            mbar_wait(bar_peer_b + 8 * bs, (e / kBSlots) & 1);
                        }
                        const uint32_t ab = G & 1;
                        mbar_wait(bar_acc_empty + 8 * ab, ((G >> 1) & 1) ^ 1);
                        if (p.stamps && pair == 0 && G < 64 && lane == 0) p.stamps[G * 8 + 0] = clock64();
                        tc_fence_after();
                        uint32_t acc = 0;
                        for (int combo = 0; combo < n_combo; ++combo) {
                            const int apl = (n_combo == 3 && combo == lo_first) ? 1 : 0;
                            const int bpl = (n_combo == 3 && combo == 1 - lo_first) ? 1 : 0;
                            const uint32_t a_lo = desc_lo(s_a + as * kBlock + apl * kPlane, kRun);
                            const uint32_t b_lo = desc_lo(s_b + bs * kBlock + bpl * kPlane, kRun);
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks) {
                                if (elect_one())
                                    umma_pair(tmem_base + ab * kHalf, a_lo + ((2 * ks * kRun) >> 4), d_hi,
                                              b_lo + ((2 * ks * kRun) >> 4), d_hi, idesc, acc);
                                acc = 1;
                            }
                        }
                        if (elect_one()) {
                            tc_commit_pair(bar_acc_full + 8 * ab);
                            // the older candidate block has had its last use (at the end of a row both have)
                            if (h == 0 || t == p.nt - 1) tc_commit_pair(bar_b_empty + 8 * bs);
                        }
                        if (p.stamps && pair == 0 && G < 64 && lane == 0) p.stamps[G * 8 + 1] = clock64();
                    }
                    if (elect_one()) tc_commit_pair(bar_a_empty + 8 * as);
                }
                E0 += (uint32_t)p.nt + 1;
            }
        }
    } else if (warp >= 4) {
        // ---- epilogue (both CTAs): warp (q, part) drains chunks 2*part, 2*part+1 of every window half ----
        const int q = warp & 3, part = (warp - 4) >> 2;
        const int xi = q * 32 + lane;
        const int D = p.D;
        float *wta_v = reinterpret_cast<float *>(smem + kSmemWtaV);
        uint16_t *wta_d = reinterpret_cast<uint16_t *>(smem + kSmemWtaD);
        uint32_t T = 0, G = 0;
        for (int item = pair; item < p.nitems; item += npairs) {
            int img, y, side;
            decode(item, img, y, side);
            for (int t = 0; t < p.nt; ++t, ++T) {
                float best = -INFINITY;
                int best_d = 0x7fffffff;
#pragma unroll 1
                for (int h = 0; h < 2; ++h, ++G) {
                    const uint32_t ab = G & 1;
                    mbar_wait(bar_acc_full + 8 * ab, (G >> 1) & 1);
                    const bool st = p.stamps && pair == 0 && c == 0 && warp == 4 && lane == 0 && G < 64;
                    if (st) p.stamps[G * 8 + 2] = clock64();
                    tc_fence_after();
#pragma unroll 1
                    for (int j = 0; j < 2; ++j) {
                        const int cb = 2 * part + j;
                        const int u0 = kHalf * h + 32 * cb;                  // first window column of this chunk
                        // disparities this warp can see in the chunk (warp-uniform): skip it when none is in [0, D)
                        const int dlo = side == 0 ? 128 * c + 32 * q + kMargin - u0 - 31 : u0 - 128 * c - 32 * q - 31;
                        if (dlo + 62 < 0 || dlo > D - 1) continue;
                        uint32_t v[32];
                        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + ab * kHalf + cb * 32, v);
                        if (st) p.stamps[G * 8 + 3 + 2 * j] = clock64();
                        const bool inside = dlo >= 0 && dlo + 62 <= D - 1;   // whole chunk in the band
                        if (side == 0) {
                            const int dbase = 128 * c + xi + kMargin - u0;
                            if (inside) scan_chunk<0, false>(v, dbase, D, best, best_d);
                            else scan_chunk<0, true>(v, dbase, D, best, best_d);
                        } else {
                            const int dbase = u0 - 128 * c - xi;
                            if (inside) scan_chunk<1, false>(v, dbase, D, best, best_d);
                            else scan_chunk<1, true>(v, dbase, D, best, best_d);
                        }
                        if (st) p.stamps[G * 8 + 4 + 2 * j] = clock64() + (best_d & 0);
                    }
                    if (st) p.stamps[G * 8 + 7] = clock64();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        if (c == 0) mbar_arrive(bar_acc_empty + 8 * ab);
                        else mbar_arrive_remote(bar_acc_empty + 8 * ab, 0);
                    }
                }
                // merge the four column parts of each row: larger value wins, equal values -> lower disparity.
                // Barrier 1: parts 1-3 have written; barrier 2: part 0 has read (so the buffer can be rewritten).
                if (part != 0) {
                    if (T > 0) asm volatile("bar.sync 2, %0;" ::"n"(32 * kEpiWarps) : "memory");
                    wta_v[(part - 1) * 128 + xi] = best;
                    wta_d[(part - 1) * 128 + xi] = (uint16_t)(best_d == 0x7fffffff ? 0xffff : best_d);
                    asm volatile("bar.arrive 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
                } else {
                    asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
#pragma unroll
                    for (int pt = 0; pt < 3; ++pt) {
                        const float ov = wta_v[pt * 128 + xi];
                        const int od = wta_d[pt * 128 + xi];
                        if (ov > best || (ov == best && od < best_d)) { best = ov; best_d = od; }
                    }
                    asm volatile("bar.arrive 2, %0;" ::"n"(32 * kEpiWarps) : "memory");
                    const int x = kBlk * (2 * t + c) + xi;
                    if (x < p.w)
                        (side == 0 ? p.left_disp : p.right_disp)[((size_t)img * p.h + y) * p.w + x] = (uint8_t)(best_d & 255);
                }
            }
        }
        if (part != 0 && T > 0) asm volatile("bar.sync 2, %0;" ::"n"(32 * kEpiWarps) : "memory");
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();      // neither CTA frees tensor memory while the pair's MMAs may still touch it
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
    }
}

}  // namespace tcs2

// columns a feature row needs right of its last 128-pixel block (the right sweep of the last pair-tile
// reads candidate blocks up to 2 * nt + 1)
int corr_sweep2_row_cols(int w)
{
    const int nb = (w + tcs2::kBlk - 1) / tcs2::kBlk, nt = (nb + 1) / 2;
    return tcs2::kBlk * (2 * nt + 2);
}

int launch_corr_sweep2(stereo_ctx *ctx, const void *l_hi, const void *l_lo, const void *r_hi, const void *r_lo, int n,
                       int h, int w, int wf, int ml, int D, int planes, uint8_t *left_disp, uint8_t *right_disp)
{
    STEREO_REQUIRE(D >= 1 && D - 1 <= tcs2::kMargin, "row-sweep correlation supports disp_range <= %d, got %d",
                   tcs2::kMargin + 1, D);
    STEREO_REQUIRE(ml >= tcs2::kMargin && wf >= ml + corr_sweep2_row_cols(w), "feature margins too small");
    tcs2::Params p;
    p.l_hi = static_cast<const __nv_bfloat16 *>(l_hi);
    p.l_lo = static_cast<const __nv_bfloat16 *>(l_lo);
    p.r_hi = static_cast<const __nv_bfloat16 *>(r_hi);
    p.r_lo = static_cast<const __nv_bfloat16 *>(r_lo);
    p.n = n; p.h = h; p.w = w; p.wf = wf; p.ml = ml; p.D = D; p.planes = planes;
    p.left_disp = left_disp; p.right_disp = right_disp;
    const int nb = (w + tcs2::kBlk - 1) / tcs2::kBlk;
    p.nt = (nb + 1) / 2;
    p.nitems = n * h * 2;
    int pairs = ctx->sm_count / 2;
    if (pairs > p.nitems) pairs = p.nitems;
    STEREO_CUDA_CHECK(cudaFuncSetAttribute(tcs2::corr_sweep2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           tcs2::kSmemTotal));
    p.stamps = nullptr;
    const char *dbg = getenv("STEREO_CORR_DEBUG");
    const bool stamps = dbg && (atoi(dbg) & 16);
    if (stamps) {
        STEREO_CUDA_CHECK(cudaMalloc(&p.stamps, 64 * 8 * sizeof(long long)));
        STEREO_CUDA_CHECK(cudaMemsetAsync(p.stamps, 0, 64 * 8 * sizeof(long long), ctx->stream));
    }
    {
        ProfScope prof(ctx, STEREO_K_CORR);
        tcs2::corr_sweep2_kernel<<<2 * pairs, tcs2::kThreads, tcs2::kSmemTotal, ctx->stream>>>(p);
        ctx->launches += 1;
        STEREO_CUDA_CHECK(cudaGetLastError());
    }
    if (stamps) {
        // issuer: granted, committed; epilogue warp 4: seen, ld0, scan0, ld1, scan1, done (0 = chunk skipped)
        static int dumps = 0;
        long long h[64 * 8];
        STEREO_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
        STEREO_CUDA_CHECK(cudaMemcpy(h, p.stamps, sizeof(h), cudaMemcpyDeviceToHost));
        cudaFree(p.stamps);
        if (dumps++ == 2)
            for (int g = 0; g < 64; ++g) {
                fprintf(stderr, "half %2d", g);
                for (int k = 0; k < 8; ++k) fprintf(stderr, " %6lld", h[g * 8 + k] ? h[g * 8 + k] - h[0] : 0);
                fprintf(stderr, "\n");
            }
    }
    return STEREO_OK;
}

}  // namespace stereo
