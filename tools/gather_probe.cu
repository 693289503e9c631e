/*
 * gather_probe.cu -- what the B200 gives a kernel made of random 8-byte gathers (one 32-byte sector each), which is
 * what PART 1 (k_vote) is made of: deep-table entry, sax entry, SA value.  Prints gathers/s for table sizes from
 * L2-resident to human-scale, for U independent loads in flight per thread, and for the three settings of
 * cudaLimitMaxL2FetchGranularity.  This is the denominator of north_star's "gather-bandwidth roofline".
 *   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/gather_probe tools/gather_probe.cu && tools/gather_probe
 */
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t mix(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}

/* U independent random loads per round; DEP = the next round's addresses depend on this round's data (a chain of
 * rounds, like table entry -> sax entry), else they only depend on the counter */
template <int U, bool DEP>
__global__ void __launch_bounds__(128) k_gather(const unsigned long long *__restrict__ tab, uint64_t mask, int rounds,
                                                unsigned long long *out) {
    uint64_t s = mix(blockIdx.x * 128ull + threadIdx.x + 1);
    unsigned long long acc = 0;
    for (int r = 0; r < rounds; r++) {
        unsigned long long v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = __ldg(tab + (mix(s + u * 0x9E3779B97F4A7C15ULL) & mask));
#pragma unroll
        for (int u = 0; u < U; u++) acc += v[u];
        s = mix(s + 0x1234567ull + (DEP ? acc : 0ull));
    }
    if (acc == 0x12345) out[0] = acc;
}

template <int U, bool DEP>
static void run(const unsigned long long *tab, uint64_t entries, unsigned long long *out, int sm, const char *tag, int bps = 0) {
    int occ = 0;
    const size_t smem = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_gather<U, DEP>, 128, smem));
    if (bps > 0 && bps < occ) occ = bps;          /* fewer resident blocks per SM: the grid is all that is resident */
    const int grid = sm * occ;
    const int rounds = 256 / U * 4;
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    k_gather<U, DEP><<<grid, 128, smem>>>(tab, entries - 1, rounds, out);
    CK(cudaEventRecord(a));
    k_gather<U, DEP><<<grid, 128, smem>>>(tab, entries - 1, rounds, out);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, a, b));
    const double n = (double)grid * 128 * rounds * U;
    printf("  %-22s U=%d dep=%d warps/SM=%2d : %7.2f G gathers/s  (%6.0f GB/s at 32 B, %6.0f GB/s at 64 B)  %.3f ms\n", tag, U,
           (int)DEP, occ * 4, n / ms / 1e6, n * 32 / ms / 1e6, n * 64 / ms / 1e6, ms);
}

int main(int argc, char **argv) {
    int dev = 0, sm = 0;
    CK(cudaSetDevice(dev));
    CK(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
    size_t max_gb = argc > 1 ? (size_t)atoi(argv[1]) : 32;
    const size_t sizes_mb[] = {32, 64, 128, 512, 8192, 32768};
    unsigned long long *out;
    CK(cudaMalloc(&out, 64));
    for (int gran : {0, 32}) {
        if (gran) CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, gran));
        size_t got = 0;
        CK(cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity));
        printf("cudaLimitMaxL2FetchGranularity %s -> %zu\n", gran ? "set" : "default", got);
        for (size_t mb : sizes_mb) {
            if (mb > max_gb * 1024) continue;
            /* power-of-two entry count not above the size */
            uint64_t entries = 1;
            while (entries * 2 * 8 <= mb * 1048576ull) entries *= 2;
            unsigned long long *tab;
            if (cudaMalloc(&tab, entries * 8) != cudaSuccess) { printf("  %zu MB: no memory\n", mb); continue; }
            CK(cudaMemset(tab, 1, entries * 8));
            char tag[64];
            snprintf(tag, sizeof tag, "table %zu MB", (size_t)(entries * 8 >> 20));
            run<1, true>(tab, entries, out, sm, tag);
            run<2, true>(tab, entries, out, sm, tag);
            run<4, true>(tab, entries, out, sm, tag);
            run<8, false>(tab, entries, out, sm, tag);
            /* k_vote's occupancy: 6 blocks of 4 warps per SM */
            run<1, true>(tab, entries, out, sm, tag, 6);
            run<2, true>(tab, entries, out, sm, tag, 6);
            run<4, true>(tab, entries, out, sm, tag, 6);
            CK(cudaFree(tab));
        }
    }
    return 0;
}
