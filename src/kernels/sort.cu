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

// Exclusive scan of the digit-major count matrix in two small steps (a single-CTA scan of all
// 256 x ctas entries took 253 us per pass - more than the histogram and the scatter together):
//   rows:   CTA d scans row d (one entry per scatter CTA) -> offsets[d][cta] relative to the row, row_total[d]
//   digits: one CTA scans the 256 row totals -> digit_base[d]
// The scatter kernel starts bucket d of CTA c at digit_base[d] + offsets[d][c].
__device__ __forceinline__ uint64_t block_exclusive_scan_256(uint64_t v, uint64_t *total, uint64_t *warp_sums) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint64_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint64_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    uint64_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        const uint64_t s = warp_sums[w];
        if (w < warp) base += s;
        tot += s;
    }
    __syncthreads();
    *total = tot;
    return base + inc - v;
}

__global__ void __launch_bounds__(256) ocm_sort_scan_rows_kernel(const uint32_t *counts, uint64_t *offsets,
                                                                 uint64_t *row_total, int ctas) {
    __shared__ uint64_t warp_sums[8];
    const uint32_t *row = counts + (uint64_t)blockIdx.x * ctas;
    uint64_t *orow = offsets + (uint64_t)blockIdx.x * ctas;
    uint64_t carry = 0;
    for (int base = 0; base < ctas; base += 256) {
        const int i = base + threadIdx.x;
        const uint64_t v = i < ctas ? row[i] : 0;
        uint64_t total;
        const uint64_t ex = block_exclusive_scan_256(v, &total, warp_sums);
        if (i < ctas) orow[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) row_total[blockIdx.x] = carry;
}

__global__ void __launch_bounds__(256) ocm_sort_scan_digits_kernel(const uint64_t *row_total, uint64_t *digit_base) {
    __shared__ uint64_t warp_sums[8];
    uint64_t total;
    digit_base[threadIdx.x] = block_exclusive_scan_256(row_total[threadIdx.x], &total, warp_sums);
}

__global__ void __launch_bounds__(kSortThreads) ocm_sort_scatter_kernel(const uint64_t *in, uint64_t *out, uint64_t n,
                                                                        int shift, const uint64_t *offsets,
                                                                        const uint64_t *digit_base, uint64_t chunk) {
    // One round = kKeysPerThread x 256 keys, taken in index order as "slots" (sub-round j, warp w):
    // slot s = j * 8 + w holds the keys  base + j*256 + w*32 + lane.
    //   1. every warp counts its keys per digit (match masks), leaders write cnt[slot][digit]
    //   2. thread d turns cnt[.][d] into an exclusive prefix over the 32 slots and advances off[d]
    //   3. every key goes to off_old[d] + prefix[slot][d] + rank inside its warp
    // Three barriers per 1024 keys (the first version handed the running offsets from warp to warp and
    // paid eight barriers per 256 keys).  cnt is double-buffered so that clearing overlaps the scatter.
    constexpr int kKeysPerThread = 4;
    constexpr int kWarps = kSortThreads / 32;
    constexpr int kSlots = kKeysPerThread * kWarps;
    __shared__ uint64_t off[2][kRadix];
    __shared__ uint16_t cnt[2][kSlots][kRadix];
    off[0][threadIdx.x] = digit_base[threadIdx.x] + offsets[(uint64_t)threadIdx.x * gridDim.x + blockIdx.x];
    for (int i = threadIdx.x; i < kSlots * kRadix; i += kSortThreads) (&cnt[0][0][0])[i] = 0;
    __syncthreads();
    const uint64_t lo = (uint64_t)blockIdx.x * chunk, hi = lo + chunk < n ? lo + chunk : n;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int buf = 0;
    for (uint64_t base = lo; base < hi; base += (uint64_t)kSortThreads * kKeysPerThread, buf ^= 1) {
        uint64_t k[kKeysPerThread];
        uint32_t d[kKeysPerThread], rank[kKeysPerThread];
        bool valid[kKeysPerThread];
#pragma unroll
        for (int j = 0; j < kKeysPerThread; ++j) {
            const uint64_t i = base + (uint64_t)j * kSortThreads + threadIdx.x;
            valid[j] = i < hi;
            k[j] = valid[j] ? in[i] : 0;
            d[j] = digit_of(k[j], shift);
            // lanes of this warp holding the same digit (invalid lanes get a private class)
            const uint32_t peers = __match_any_sync(0xffffffffu, valid[j] ? d[j] : 0x100u + lane);
            rank[j] = __popc(peers & ((1u << lane) - 1));
            if (valid[j] && rank[j] == 0) cnt[buf][j * kWarps + warp][d[j]] = (uint16_t)__popc(peers);
        }
        __syncthreads();
        {   // thread t owns digit t
            const int t = threadIdx.x;
            uint32_t run = 0;
#pragma unroll 8
            for (int s = 0; s < kSlots; ++s) {
                const uint32_t c = cnt[buf][s][t];
                cnt[buf][s][t] = (uint16_t)run;
                run += c;
            }
            off[buf ^ 1][t] = off[buf][t] + run;
        }
        // clear the other buffer for the next round (nobody reads it any more: the barrier above is past its scatter)
        for (int i = threadIdx.x; i < kSlots * kRadix; i += kSortThreads) (&cnt[buf ^ 1][0][0])[i] = 0;
        __syncthreads();
#pragma unroll
        for (int j = 0; j < kKeysPerThread; ++j)
            if (valid[j]) out[off[buf][d[j]] + cnt[buf][j * kWarps + warp][d[j]] + rank[j]] = k[j];
        __syncthreads();
    }
}

// scratch layout: counts u32[256][ctas] | offsets u64[256][ctas] | row_total u64[256] | digit_base u64[256]
constexpr size_t sort_scratch_for(int ctas) {
    return (size_t)kRadix * ctas * (sizeof(uint32_t) + sizeof(uint64_t)) + 2 * (size_t)kRadix * sizeof(uint64_t);
}

size_t sort_scratch_bytes(uint64_t n, int *ctas_out) {
    int ctas = (int)std::min<uint64_t>(std::max<uint64_t>((n + 8191) / 8192, 1), (uint64_t)num_sms() * 4);
    if (ctas_out) *ctas_out = ctas;
    return sort_scratch_for(ctas);
}

// The eight passes (hist, two scan steps, scatter), shared by the regular and the allocation-free entry point.
static void sort_passes(uint64_t *keys, uint64_t *tmp, uint64_t n, int ctas, void *scratch, cudaStream_t stream) {
    uint32_t *counts = static_cast<uint32_t *>(scratch);
    uint64_t *offsets = reinterpret_cast<uint64_t *>(static_cast<uint8_t *>(scratch) + (size_t)kRadix * ctas * sizeof(uint32_t));
    uint64_t *row_total = offsets + (size_t)kRadix * ctas;
    uint64_t *digit_base = row_total + kRadix;
    const uint64_t chunk = ((n + ctas - 1) / ctas + kSortThreads - 1) / kSortThreads * kSortThreads;
    uint64_t *a = keys, *b = tmp;
    for (int pass = 0; pass < 8; ++pass) {
        const int shift = pass * 8;
        ocm_sort_hist_kernel<<<ctas, kSortThreads, 0, stream>>>(a, n, shift, counts, chunk);
        ocm_sort_scan_rows_kernel<<<kRadix, 256, 0, stream>>>(counts, offsets, row_total, ctas);
        ocm_sort_scan_digits_kernel<<<1, 256, 0, stream>>>(row_total, digit_base);
        ocm_sort_scatter_kernel<<<ctas, kSortThreads, 0, stream>>>(a, b, n, shift, offsets, digit_base, chunk);
        std::swap(a, b);
    }
}

// keys and tmp: n elements each.  scratch: sort_scratch_bytes(n) bytes.
cudaError_t launch_sort_u64(uint64_t *keys, uint64_t *tmp, uint64_t n, void *scratch, size_t scratch_bytes,
                            cudaStream_t stream, unsigned long long *launches) {
    if (n < 2) return cudaSuccess;
    int ctas = 0;
    if (scratch_bytes < sort_scratch_bytes(n, &ctas)) return cudaErrorInvalidValue;
    sort_passes(keys, tmp, n, ctas, scratch, stream);
    if (launches) *launches += 8 * 4;
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
//   * the scratch array is 388 KiB and lives on every GPU that loads this library, from load
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
static constexpr size_t   kNoAllocScratchBytes = ocmk::sort_scratch_for(kNoAllocMaxCtas);   // 388 KiB
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
    sort_passes(keys, tmp, n, ctas, scratch, static_cast<cudaStream_t>(stream));
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        LOGE("ocmk_sort_u64_noalloc: launch failed: %s", cudaGetErrorString(e));
        return -OCM_ECUDA;
    }
    return 0;
}
