// cub_sort_baseline - BASELINE ONLY (never on the product path): cub::DeviceRadixSort::SortKeys on
// 64-bit keys, timed with CUDA events, next to our K8 radix sort (launch_sort_u64) on the same data.
//   cub_sort_baseline [log2 sizes...]      default 20 22 24 26
// One JSON line per size.
#include "kernels/kernels.h"
#include <cub/device/device_radix_sort.cuh>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                                  \
    do {                                                                                       \
        cudaError_t e_ = (x);                                                                  \
        if (e_ != cudaSuccess) {                                                               \
            fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); \
            exit(1);                                                                           \
        }                                                                                      \
    } while (0)

__global__ void gen_kernel(uint64_t *k, uint64_t n, uint64_t seed) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t x = (i + seed) * 0x9E3779B97F4A7C15ull;   // splitmix64
        x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
        x ^= x >> 27; x *= 0x94D049BB133111EBull;
        k[i] = x ^ (x >> 31);
    }
}

__global__ void check_kernel(const uint64_t *a, const uint64_t *b, uint64_t n, unsigned long long *bad) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        if (a[i] != b[i] || (i && a[i - 1] > a[i])) atomicAdd(bad, 1ull);
}

int main(int argc, char **argv) {
    std::vector<int> sizes;
    for (int i = 1; i < argc; ++i) sizes.push_back(atoi(argv[i]));
    if (sizes.empty()) sizes = {20, 22, 24, 26};
    const int iters = 5;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int lg : sizes) {
        const uint64_t n = 1ull << lg;
        uint64_t *src, *a, *b, *c;
        unsigned long long *bad;
        CK(cudaMalloc(&src, n * 8));
        CK(cudaMalloc(&a, n * 8));
        CK(cudaMalloc(&b, n * 8));
        CK(cudaMalloc(&c, n * 8));
        CK(cudaMalloc(&bad, 8));
        CK(cudaMemset(bad, 0, 8));
        gen_kernel<<<1024, 256>>>(src, n, 12345);
        // CUB keys-only
        size_t tmp_bytes = 0;
        CK(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, src, a, n));
        void *tmp;
        CK(cudaMalloc(&tmp, tmp_bytes));
        float cub_ms = 1e30f, ours_ms = 1e30f;
        for (int it = 0; it < iters + 1; ++it) {
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(e0));
            CK(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, src, a, n));
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (it && ms < cub_ms) cub_ms = ms;
        }
        // ours (in place on a copy; the copy is outside the timed region)
        int ctas = 0;
        const size_t sb = ocmk::sort_scratch_bytes(n, &ctas);
        void *scratch;
        CK(cudaMalloc(&scratch, sb));
        for (int it = 0; it < iters + 1; ++it) {
            CK(cudaMemcpy(b, src, n * 8, cudaMemcpyDeviceToDevice));
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(e0));
            CK(ocmk::launch_sort_u64(b, c, n, scratch, sb, 0, nullptr));
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (it && ms < ours_ms) ours_ms = ms;
        }
        check_kernel<<<1024, 256>>>(a, b, n, bad);
        unsigned long long h_bad = 0;
        CK(cudaMemcpy(&h_bad, bad, 8, cudaMemcpyDeviceToHost));
        printf("{\"log2_keys\": %d, \"cub_keys_only_ms\": %.4f, \"ours_ms\": %.4f, \"ours_over_cub\": %.3f, "
               "\"cub_Gkeys_s\": %.2f, \"ours_Gkeys_s\": %.2f, \"mismatches\": %llu}\n",
               lg, cub_ms, ours_ms, cub_ms / ours_ms, n / cub_ms / 1e6, n / ours_ms / 1e6, h_bad);
        fflush(stdout);
        cudaFree(src); cudaFree(a); cudaFree(b); cudaFree(c); cudaFree(bad); cudaFree(tmp); cudaFree(scratch);
    }
    return 0;
}
