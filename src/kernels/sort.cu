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

// ---------------------------------------------------------------------------------------------
// ocmk_sort_u64_noalloc - the same radix sort for callers that must not allocate.
//
// Who needs it: (1) an application attached to a cluster whose daemons have already claimed
// nearly all of the HBM for their pools - a cudaMalloc for sort scratch may simply fail there,
// or succeed and starve the next pool growth; (2) a caller in the middle of a latency-critical
// section (copies in flight on other streams): cudaMalloc / cudaFree may synchronise the whole
// device, which is exactly what the rest of this runtime avoids on its allocation path.
// The regular entry point ocmk_sort_u64 grows a cudaMalloc'ed scratch buffer on demand; this
// one keeps its digit histograms and scanned offsets in the fixed array below instead.
//
// Limits and rules:
//   * at most OCMK_SORT_NOALLOC_MAX_KEYS (2^24) keys per call; larger n is refused with
//     -OCM_ERANGE and a log line, nothing is launched;
//   * the scratch array is 384 KiB and lives on every GPU that loads this library, from load
//     to process exit;
//   * the array's name is only a host-side stand-in in host code; its device address is looked
//     up once per device (cudaGetSymbolAddress, with that device current) the first time the
//     entry point runs on that device.  Make that first call before starting any stream capture;
//   * all calls on a device share the one array, which keeps whatever the previous call left in
//     it.  Nothing in it has to start at zero: every pass rewrites each histogram and offset
//     entry before reading it, so there is nothing to clear between calls;
//   * because the array is shared, two calls whose kernels overlap on the same GPU (different
//     streams, or different host threads) race on it and produce garbage - serialise them;
//   * no cudaMalloc, cudaFree or other allocating call is made here; keys and tmp (n elements
//     each) are the caller's.
// Returns 0 or a negative enum ocm_err.
// ---------------------------------------------------------------------------------------------
#include "../../include/oncillamem.h"
#include "../common/log.hpp"
#include <atomic>

static constexpr int      kNoAllocMaxCtas = 128;
static constexpr uint64_t kNoAllocMaxKeys = 1ull << 24;
static constexpr size_t   kNoAllocScratchBytes = (size_t)ocmk::kRadix * kNoAllocMaxCtas * (sizeof(uint32_t) + sizeof(uint64_t));
// the scratch itself: fixed size, one copy per GPU that loads the library
static __device__ __align__(256) unsigned char g_sort_scratch_static[kNoAllocScratchBytes];
// its device address per CUDA device, filled in by the first call on that device
static std::atomic<void *> g_sort_scratch_addr[64];

extern "C" int ocmk_sort_u64_noalloc(uint64_t *keys, uint64_t *tmp, uint64_t n, void *stream) {
    using namespace ocmk;
    if (n < 2) return 0;
    if (!keys || !tmp) return -OCM_EINVAL;
    if (n > kNoAllocMaxKeys) {
        LOGE("ocmk_sort_u64_noalloc: %llu keys exceed the fixed-scratch limit of %llu (OCMK_SORT_NOALLOC_MAX_KEYS); "
             "use ocmk_sort_u64 or sort in pieces",
             (unsigned long long)n, (unsigned long long)kNoAllocMaxKeys);
        return -OCM_ERANGE;
    }
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return -OCM_ENOGPU;
    void *scratch = g_sort_scratch_addr[dev].load(std::memory_order_acquire);
    if (!scratch) {
        // one-time set-up for this device (it is the current one): resolve the array's device address
        cudaError_t e = cudaGetSymbolAddress(&scratch, g_sort_scratch_static);
        if (e != cudaSuccess || !scratch) {
            LOGE("ocmk_sort_u64_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            cudaGetLastError();
            return -OCM_ECUDA;
        }
        g_sort_scratch_addr[dev].store(scratch, std::memory_order_release);
    }
    const int ctas = (int)std::min<uint64_t>(std::max<uint64_t>((n + 8191) / 8192, 1), (uint64_t)kNoAllocMaxCtas);
    uint32_t *counts = static_cast<uint32_t *>(scratch);
    uint64_t *offsets = reinterpret_cast<uint64_t *>(static_cast<uint8_t *>(scratch) + (size_t)kRadix * ctas * sizeof(uint32_t));
    const uint64_t chunk = ((n + ctas - 1) / ctas + kSortThreads - 1) / kSortThreads * kSortThreads;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    uint64_t *a = keys, *b = tmp;
    for (int pass = 0; pass < 8; ++pass) {
        const int shift = pass * 8;
        ocm_sort_hist_kernel<<<ctas, kSortThreads, 0, st>>>(a, n, shift, counts, chunk);
        ocm_sort_scan_kernel<<<1, 1024, 0, st>>>(counts, offsets, (uint64_t)kRadix * ctas);
        ocm_sort_scatter_kernel<<<ctas, kSortThreads, 0, st>>>(a, b, n, shift, offsets, chunk);
        std::swap(a, b);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        LOGE("ocmk_sort_u64_noalloc: launch failed: %s", cudaGetErrorString(e));
        return -OCM_ECUDA;
    }
    return 0;
}
