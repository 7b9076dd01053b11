// Exhaustive check of stereo::div_exact against IEEE division: every one of the 2^32 float bit patterns as the
// dividend, for a list of divisors (fixed ones, then pseudo-random ones in the range volume std values take).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/_build/div_check scripts/div_check.cu && scripts/_build/div_check [n_random]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../stereo_matching_b200/csrc/exact_div.cuh"

__global__ void check_kernel(float s, unsigned long long *bad)
{
    const stereo::ConstDivisor c = stereo::make_divisor(s);
    unsigned long long local = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < (1ull << 32);
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const float x = __uint_as_float((unsigned)i);
        const float a = stereo::div_exact(x, c), b = __fdiv_rn(x, s);
        const bool same = __float_as_uint(a) == __float_as_uint(b) || (a != a && b != b);
        local += same ? 0 : 1;
    }
    if (local) atomicAdd(bad, local);
}

int main(int argc, char **argv)
{
    const int n_random = argc > 1 ? atoi(argv[1]) : 64;
    unsigned long long *bad, host = 0, total = 0;
    cudaMalloc(&bad, sizeof(*bad));
    float fixed[] = {1.f, 3.f, 0.1f, 1.9999999f, 1.0000001f, 7.25f, 123.456f, 1e-3f, 6.02e5f, 0x1.fffffep+0f, 0x1.000002p+3f};
    unsigned long long seed = 88172645463325252ull;
    const int n_fixed = sizeof(fixed) / sizeof(fixed[0]);
    for (int k = 0; k < n_fixed + n_random; ++k) {
        float s;
        if (k < n_fixed) s = fixed[k];
        else {
            seed ^= seed << 13; seed ^= seed >> 7; seed ^= seed << 17;
            unsigned mant = (unsigned)(seed >> 20) & 0x7fffffu, ex = 117u + (unsigned)((seed >> 50) % 20);   // 2^-10 .. 2^9
            unsigned bits = (ex << 23) | mant;
            memcpy(&s, &bits, 4);
        }
        cudaMemset(bad, 0, sizeof(*bad));
        check_kernel<<<148 * 16, 256>>>(s, bad);
        cudaMemcpy(&host, bad, sizeof(host), cudaMemcpyDeviceToHost);
        total += host;
        if (host) printf("divisor %.9g (%a): %llu mismatches\n", s, s, host);
    }
    printf("{\"divisors\": %d, \"dividends_each\": 4294967296, \"mismatches\": %llu}\n", n_fixed + n_random, total);
    return total != 0;
}
