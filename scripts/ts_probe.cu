// Probe: does tcgen05.cp.128x256b + tcgen05.mma with the A operand in tensor memory reproduce the
// shared-memory (SS) MMA?  One CTA, M = 128, N = 128, K = 64 bf16, K-major no-swizzle operands in the
// [k/8][row][8] layout the kernels of this repo use.  Build and run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -I stereo_matching_b200/csrc scripts/ts_probe.cu -o /tmp/ts_probe
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <vector>

#include "tc_common.cuh"

using namespace stereo::tc;

__device__ __forceinline__ void umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t acc)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}

__global__ void probe(const __nv_bfloat16 *a, const __nv_bfloat16 *b, float *d_ss, float *d_ts)
{
    extern __shared__ __align__(128) uint8_t smem[];
    __nv_bfloat16 *sa = reinterpret_cast<__nv_bfloat16 *>(smem);               // [8][128][8]
    __nv_bfloat16 *sb = reinterpret_cast<__nv_bfloat16 *>(smem + 16384);       // [8][128][8]
    const uint32_t bar = smem_u32(smem + 32768);
    uint32_t *slot = reinterpret_cast<uint32_t *>(smem + 32768 + 16);
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 8 * 128 * 8; i += blockDim.x) { sa[i] = a[i]; sb[i] = b[i]; }
    if (tid == 0) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(512u));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot;
    const uint32_t idesc = make_idesc(128, 128);
    const uint32_t s_a = smem_u32(sa), s_b = smem_u32(sb);
    if (tid == 0) {
        // SS: D at columns 0..127
        for (int ks = 0; ks < 4; ++ks)
            umma_bf16(tmem, smem_desc(s_a + 2 * ks * 2048, 2048, 128), smem_desc(s_b + 2 * ks * 2048, 2048, 128), idesc, ks > 0);
        // TS: stage A at columns 256.. (8 columns per K=16 slab), D at columns 128..255
        for (int ks = 0; ks < 4; ++ks) {
            const uint64_t ad = smem_desc(s_a + 2 * ks * 2048, 2048, 128);
            asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(tmem + 256 + 8 * ks), "l"(ad) : "memory");
        }
        for (int ks = 0; ks < 4; ++ks)
            umma_ts(tmem + 128, tmem + 256 + 8 * ks, smem_desc(s_b + 2 * ks * 2048, 2048, 128), idesc, ks > 0);
        tc_commit(bar);
    }
    mbar_wait(bar, 0);
    tc_fence_after();
    if (warp < 4) {
        const int row = tid;
        for (int c = 0; c < 4; ++c) {
            uint32_t v[32];
            tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + c * 32, v);
            for (int i = 0; i < 32; ++i) d_ss[row * 128 + c * 32 + i] = __uint_as_float(v[i]);
            tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + 128 + c * 32, v);
            for (int i = 0; i < 32; ++i) d_ts[row * 128 + c * 32 + i] = __uint_as_float(v[i]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u));
}

int main()
{
    std::vector<__nv_bfloat16> ha(8 * 128 * 8), hb(8 * 128 * 8);
    std::vector<float> fa(128 * 64), fb(128 * 64);
    for (int m = 0; m < 128; ++m)
        for (int k = 0; k < 64; ++k) {
            fa[m * 64 + k] = (float)((m * 3 + k * 5) % 17 - 8);
            fb[m * 64 + k] = (float)((m * 7 + k) % 13 - 6);
            ha[((k / 8) * 128 + m) * 8 + k % 8] = __float2bfloat16(fa[m * 64 + k]);
            hb[((k / 8) * 128 + m) * 8 + k % 8] = __float2bfloat16(fb[m * 64 + k]);
        }
    __nv_bfloat16 *da, *db;
    float *dss, *dts;
    cudaMalloc(&da, ha.size() * 2); cudaMalloc(&db, hb.size() * 2);
    cudaMalloc(&dss, 128 * 128 * 4); cudaMalloc(&dts, 128 * 128 * 4);
    cudaMemcpy(da, ha.data(), ha.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(db, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice);
    cudaMemset(dss, 0, 128 * 128 * 4); cudaMemset(dts, 0, 128 * 128 * 4);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 40000);
    probe<<<1, 128, 40000>>>(da, db, dss, dts);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    std::vector<float> ss(128 * 128), ts(128 * 128);
    cudaMemcpy(ss.data(), dss, ss.size() * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(ts.data(), dts, ts.size() * 4, cudaMemcpyDeviceToHost);
    int bad_ss = 0, bad_ts = 0;
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 128; ++n) {
            float ref = 0;
            for (int k = 0; k < 64; ++k) ref += fa[m * 64 + k] * fb[n * 64 + k];
            bad_ss += ss[m * 128 + n] != ref;
            bad_ts += ts[m * 128 + n] != ref;
        }
    printf("SS mismatches %d, TS mismatches %d of %d\n", bad_ss, bad_ts, 128 * 128);
    for (int m = 0; m < 3; ++m) printf("row %d: ss %g %g %g  ts %g %g %g\n", m, ss[m * 128], ss[m * 128 + 1], ss[m * 128 + 2],
                                       ts[m * 128], ts[m * 128 + 1], ts[m * 128 + 2]);
    return 0;
}
