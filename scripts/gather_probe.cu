// Micro-benchmark behind DESIGN.md's seed-lookup roofline: DRAM bytes moved per random read of 4 / 16 / 32 bytes
// from a 4 GiB table (one read per thread, addresses from a hash), to be run under
//   ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum --clock-control none ./gather_probe
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_probe scripts/gather_probe.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
__device__ __forceinline__ uint64_t mix(uint64_t x) { x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33; return x; }
template <int BYTES>
__global__ void k_gather(const uint8_t *__restrict__ tab, uint64_t n_slots, uint64_t n, uint32_t *__restrict__ out) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint8_t *p = tab + (mix(i) % n_slots) * BYTES;
    uint32_t acc;
    if (BYTES == 4) acc = *(const uint32_t *) p;
    else if (BYTES == 16) { uint4 v = *(const uint4 *) p; acc = v.x ^ v.y ^ v.z ^ v.w; }
    else { uint4 a = ((const uint4 *) p)[0], b = ((const uint4 *) p)[1]; acc = a.x ^ a.w ^ b.x ^ b.w; }
    out[i] = acc;
}
int main(int argc, char **argv) {
    if (argc > 1) { size_t g = atoi(argv[1]); cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, g); size_t got = 0; cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity); printf("set L2 fetch granularity %zu: %s, now %zu\n", g, cudaGetErrorString(e), got); }
    const uint64_t bytes = 4ull << 30, n = 28000000;
    uint8_t *tab; uint32_t *out;
    cudaMalloc(&tab, bytes); cudaMalloc(&out, n * 4); cudaMemset(tab, 1, bytes);
    for (int rep = 0; rep < 2; rep++) {
        k_gather<4><<<(n + 255) / 256, 256>>>(tab, bytes / 4, n, out);
        k_gather<16><<<(n + 255) / 256, 256>>>(tab, bytes / 16, n, out);
        k_gather<32><<<(n + 255) / 256, 256>>>(tab, bytes / 32, n, out);
    }
    cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
