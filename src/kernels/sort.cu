// K8 - device building block of the out-of-core sort pipeline (BASELINE.json
// config 4): stable LSD radix sort of 64-bit keys, 8 bits per pass.
// No counterpart in the reference (it ships no compute at all).
// Per pass: (1) per-CTA digit histograms into global scratch, (2) one-CTA
// exclusive scan in digit-major order, (3) stable scatter - every CTA walks its
// contiguous chunk in order, 256 keys per round, ranking equal digits with warp
// match masks and a warp-sequential hand-over of the running digit offsets.
// The partition exchange between GPUs is K1/K4 (oncilla_b200/models/sort.py).
#include "kernels.h"
#include "ptx.cuh"
#include <algorithm>

namespace ocmk {

constexpr int kSortThreads = 256;
constexpr int kRadix = 256;

__device__ __forceinline__ uint32_t digit_of(uint64_t k, int shift) { return (uint32_t)(k >> shift) & 0xffu; }

// counts[d * n_ctas + cta]
__global__ void __launch_bounds__(kSortThreads) ocm_sort_hist_kernel(const uint64_t *keys, uint64_t n, int shift,
                                                                     uint32_t *counts, uint64_t chunk) {
    __shared__ uint32_t h[kRadix];
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint64_t lo = (uint64_t)blockIdx.x * chunk, hi = lo + chunk < n ? lo + chunk : n;
    for (uint64_t i = lo + threadIdx.x; i < hi; i += kSortThreads) atomicAdd(&h[digit_of(keys[i], shift)], 1u);
    __syncthreads();
    counts[(uint64_t)threadIdx.x * gridDim.x + blockIdx.x] = h[threadIdx.x];
}

// exclusive scan of m 32-bit counts into 64-bit offsets (single CTA)
__global__ void __launch_bounds__(1024) ocm_sort_scan_kernel(const uint32_t *counts, uint64_t *offsets, uint64_t m) {
    __shared__ uint64_t part[1024];
    const uint64_t per = (m + 1023) / 1024;
    const uint64_t lo = (uint64_t)threadIdx.x * per, hi = lo + per < m ? lo + per : m;
    uint64_t s = 0;
    for (uint64_t i = lo; i < hi; ++i) s += counts[i];
    part[threadIdx.x] = s;
    __syncthreads();
    // Hillis-Steele over the 1024 partial sums
    for (int o = 1; o < 1024; o <<= 1) {
        uint64_t v = threadIdx.x >= (unsigned)o ? part[threadIdx.x - o] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    uint64_t run = threadIdx.x ? part[threadIdx.x - 1] : 0;
    for (uint64_t i = lo; i < hi; ++i) {
        offsets[i] = run;
        run += counts[i];
    }
}

__global__ void __launch_bounds__(kSortThreads) ocm_sort_scatter_kernel(const uint64_t *in, uint64_t *out, uint64_t n,
                                                                        int shift, const uint64_t *offsets,
                                                                        uint64_t chunk) {
    __shared__ uint64_t off[kRadix];
    off[threadIdx.x] = offsets[(uint64_t)threadIdx.x * gridDim.x + blockIdx.x];
    __syncthreads();
    const uint64_t lo = (uint64_t)blockIdx.x * chunk, hi = lo + chunk < n ? lo + chunk : n;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (uint64_t base = lo; base < hi; base += kSortThreads) {
        const uint64_t i = base + threadIdx.x;
        const bool valid = i < hi;
        const uint64_t k = valid ? in[i] : 0;
        const uint32_t d = digit_of(k, shift);
        // lanes of this warp holding the same digit (invalid lanes get a private class)
        const uint32_t peers = __match_any_sync(0xffffffffu, valid ? d : 0x100u + lane);
        const uint32_t rank = __popc(peers & ((1u << lane) - 1));
        const bool leader = rank == 0;
        uint64_t dst = 0;
        for (int w = 0; w < kSortThreads / 32; ++w) {
            if (warp == w) {
                uint64_t b = 0;
                if (leader && valid) {
                    b = off[d];
                    off[d] = b + __popc(peers);
                }
                b = __shfl_sync(0xffffffffu, b, __ffs(peers) - 1);
                dst = b + rank;
            }
            __syncthreads();
        }
        if (valid) out[dst] = k;
    }
}

size_t sort_scratch_bytes(uint64_t n, int *ctas_out) {
    int ctas = (int)std::min<uint64_t>(std::max<uint64_t>((n + 8191) / 8192, 1), (uint64_t)num_sms() * 4);
    if (ctas_out) *ctas_out = ctas;
    return (size_t)kRadix * ctas * (sizeof(uint32_t) + sizeof(uint64_t));
}

// keys and tmp: n elements each.  scratch: sort_scratch_bytes(n) bytes.
cudaError_t launch_sort_u64(uint64_t *keys, uint64_t *tmp, uint64_t n, void *scratch, size_t scratch_bytes,
                            cudaStream_t stream, unsigned long long *launches) {
    if (n < 2) return cudaSuccess;
    int ctas = 0;
    if (scratch_bytes < sort_scratch_bytes(n, &ctas)) return cudaErrorInvalidValue;
    uint32_t *counts = static_cast<uint32_t *>(scratch);
    uint64_t *offsets = reinterpret_cast<uint64_t *>(static_cast<uint8_t *>(scratch) + (size_t)kRadix * ctas * sizeof(uint32_t));
    const uint64_t chunk = ((n + ctas - 1) / ctas + kSortThreads - 1) / kSortThreads * kSortThreads;
    uint64_t *a = keys, *b = tmp;
    for (int pass = 0; pass < 8; ++pass) {
        const int shift = pass * 8;
        ocm_sort_hist_kernel<<<ctas, kSortThreads, 0, stream>>>(a, n, shift, counts, chunk);
        ocm_sort_scan_kernel<<<1, 1024, 0, stream>>>(counts, offsets, (uint64_t)kRadix * ctas);
        ocm_sort_scatter_kernel<<<ctas, kSortThreads, 0, stream>>>(a, b, n, shift, offsets, chunk);
        std::swap(a, b);
        if (launches) *launches += 3;
    }
    return cudaGetLastError();
}

} // namespace ocmk
