// K8 - device building block of the out-of-core sort pipeline (BASELINE.json
// config 4): stable LSD radix sort of 64-bit keys, 8 bits per pass.
// No counterpart in the reference (it ships no compute at all).
// Per pass: (1) per-CTA digit histograms into global scratch, (2) two-step
// exclusive scan in digit-major order, (3) stable scatter - every CTA walks its
// contiguous chunk in tiles, ranks equal digits with warp match masks, sorts the
// tile by digit in shared memory and stores whole digit runs.
// The partition exchange between GPUs is K1/K4 (oncilla_b200/models/sort.py,
// oncilla_b200/models/dsort.py).
#include "kernels.h"
#include "ptx.cuh"
#include <algorithm>
#include <cstdlib>

namespace ocmk {

constexpr int kSortThreads = 256;
constexpr int kRadix = 256;

__device__ __forceinline__ uint32_t digit_of(uint64_t k, int shift) { return (uint32_t)(k >> shift) & 0xffu; }

// Lanes of the warp whose 8-bit digit equals mine, from eight ballots.  __match_any_sync gives the same
// mask in one instruction, but MATCH.ANY occupies the SM's single address-divergence unit for ~64 cycles
// per warp: with it the scatter kernel ran that pipe at 80 % and nothing else mattered (profiles/README.md 7f).
__device__ __forceinline__ uint32_t match_digit8(uint32_t d) {
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        const bool bit = (d >> b) & 1u;
        const uint32_t m = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? m : ~m;
    }
    return peers;
}

// counts[d * n_ctas + cta]
__global__ void __launch_bounds__(kSortThreads) ocm_sort_hist_kernel(const uint64_t *keys, uint64_t n, int shift,
                                                                     uint32_t *counts, uint64_t chunk,
                                                                     unsigned int *scan_ticket) {
    __shared__ uint32_t h[kRadix];
    h[threadIdx.x] = 0;
    if (blockIdx.x == 0 && threadIdx.x == 0) *scan_ticket = 0;   // for the scan kernel that follows in the stream
    __syncthreads();
    const uint64_t lo = (uint64_t)blockIdx.x * chunk, hi = lo + chunk < n ? lo + chunk : n;
    if ((reinterpret_cast<uintptr_t>(keys + lo) & 15) != 0) {   // odd base pointer: one key per load
        for (uint64_t i = lo + threadIdx.x; i < hi; i += kSortThreads) atomicAdd(&h[digit_of(keys[i], shift)], 1u);
        __syncthreads();
        counts[(uint64_t)threadIdx.x * gridDim.x + blockIdx.x] = h[threadIdx.x];
        return;
    }
    // 16-byte aligned chunk: two keys per load, four loads in flight
    const uint64_t pairs = hi > lo ? (hi - lo) / 2 : 0;
    const ulonglong2 *kp = reinterpret_cast<const ulonglong2 *>(keys + lo);
    uint64_t p = threadIdx.x;
    for (; p + 3 * kSortThreads < pairs; p += 4 * kSortThreads) {
        ulonglong2 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = kp[p + u * kSortThreads];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            atomicAdd(&h[digit_of(v[u].x, shift)], 1u);
            atomicAdd(&h[digit_of(v[u].y, shift)], 1u);
        }
    }
    for (; p < pairs; p += kSortThreads) {
        const ulonglong2 v = kp[p];
        atomicAdd(&h[digit_of(v.x, shift)], 1u);
        atomicAdd(&h[digit_of(v.y, shift)], 1u);
    }
    if (threadIdx.x == 0 && hi > lo && ((hi - lo) & 1)) atomicAdd(&h[digit_of(keys[hi - 1], shift)], 1u);
    __syncthreads();
    counts[(uint64_t)threadIdx.x * gridDim.x + blockIdx.x] = h[threadIdx.x];
}

// Exclusive scan of the digit-major count matrix in two small steps inside one kernel (a single-CTA
// scan of all 256 x ctas entries took 253 us per pass - more than the histogram and the scatter together):
//   rows:   CTA d scans row d (one entry per scatter CTA) -> offsets[d][cta] relative to the row, row_total[d]
//   digits: the CTA that finishes last (ticket) scans the 256 row totals -> digit_base[d]
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

// zero_counts (may be null): the count matrix the FOLLOWING scatter kernel accumulates the next pass's histogram
// into (fused mode, small inputs); row d is cleared by CTA d here, one launch ahead of its use.
__global__ void __launch_bounds__(256) ocm_sort_scan_kernel(const uint32_t *counts, uint64_t *offsets, uint64_t *row_total,
                                                            uint64_t *digit_base, int ctas, unsigned int *ticket,
                                                            uint32_t *zero_counts) {
    __shared__ uint64_t warp_sums[8];
    __shared__ bool last;
    const uint32_t *row = counts + (uint64_t)blockIdx.x * ctas;
    uint64_t *orow = offsets + (uint64_t)blockIdx.x * ctas;
    if (zero_counts)
        for (int i = threadIdx.x; i < ctas; i += 256) zero_counts[(uint64_t)blockIdx.x * ctas + i] = 0;
    uint64_t carry = 0;
    for (int base = 0; base < ctas; base += 256) {
        const int i = base + threadIdx.x;
        const uint64_t v = i < ctas ? row[i] : 0;
        uint64_t total;
        const uint64_t ex = block_exclusive_scan_256(v, &total, warp_sums);
        if (i < ctas) orow[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) {
        row_total[blockIdx.x] = carry;
        __threadfence();                                       // my total is visible before my ticket
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();                                           // every other CTA's total is visible now
    uint64_t total;
    digit_base[threadIdx.x] = block_exclusive_scan_256(__ldcg(row_total + threadIdx.x), &total, warp_sums);
    if (threadIdx.x == 0) *ticket = 0;   // ready for the next pass (fused passes have no histogram kernel to do it)
}

// Stable scatter.  A CTA walks its chunk in tiles of kSortThreads * IPT keys; warp w owns the w-th
// contiguous 32*IPT keys of the tile (item j of lane l = key w*32*IPT + j*32 + l, i.e. index order).
//   1. ranking: per item, lanes with equal digits find each other (eight ballots); the lowest one
//      advances the warp's running counter cnt[w][digit] in shared memory and broadcasts the old value
//      -> rank of every key among the keys of its digit inside its warp, no block barrier involved.
//      The ballots of all items are issued first (independent of each other), the counter hand-over
//      follows as a second, sequential sweep.
//   2. thread d turns cnt[.][d] into a prefix over the warps; a block scan over the digit totals gives
//      the tile-local start of every digit (tprefix) and advances the CTA's global cursors (off)
//   3. keys are first placed in shared memory at their tile-local sorted position, then written out in
//      that order: consecutive threads store consecutive addresses inside each digit run, so the global
//      stores are whole sectors instead of 8-byte fragments (avg run = tile/256 keys)
// The next tile's keys are requested as soon as the registers are free (after the placement), so the
// load latency hides behind the write-out.  Four barriers per tile; every warp clears its own counter row.
// (History: version 1 handed the running offsets from warp to warp, eight barriers per 256 keys; version 2
// ranked 1024 keys with three barriers but scattered straight from registers, 8 bytes per sector; version 3
// used __match_any_sync, which saturates the SM's address-divergence unit.)
template <int IPT>
__global__ void __launch_bounds__(kSortThreads, IPT <= 8 ? 4 : 2)
    ocm_sort_scatter_kernel(const uint64_t *in, uint64_t *out, uint64_t n, int shift, const uint64_t *offsets,
                            const uint64_t *digit_base, uint64_t chunk, uint32_t *next_counts, int chunk_shift) {
    constexpr int kWarps = kSortThreads / 32;
    constexpr int kTile = kSortThreads * IPT;
    __shared__ uint64_t skeys[kTile];
    __shared__ uint64_t off[kRadix];      // next global position of digit d for this CTA
    __shared__ uint64_t gpos[kRadix];     // global position of tile-local index 0 as seen from digit d
    __shared__ uint32_t cnt[kWarps][kRadix];
    __shared__ uint32_t tprefix[kRadix];
    __shared__ uint32_t wsum[kWarps];
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const uint32_t lt_mask = (1u << lane) - 1;
    off[t] = digit_base[t] + offsets[(uint64_t)t * gridDim.x + blockIdx.x];
#pragma unroll
    for (int i = 0; i < kRadix / 32; ++i) cnt[warp][lane + 32 * i] = 0;
    const uint64_t lo = (uint64_t)blockIdx.x * chunk, hi = lo + chunk < n ? lo + chunk : n;
    const uint64_t woff = (uint64_t)warp * 32 * IPT + lane;   // my first key inside a tile
    uint64_t k[IPT];
#pragma unroll
    for (int j = 0; j < IPT; ++j) k[j] = lo + woff + j * 32 < hi ? in[lo + woff + j * 32] : ~0ull;
    __syncwarp();
    for (uint64_t base = lo; base < hi; base += kTile) {
        // items of this lane that exist (only the last tile of a chunk is short)
        const uint64_t left = hi - base;
        const int nv = left >= (uint64_t)kTile ? IPT
                       : (left > woff ? (int)((left - woff + 31) / 32 < (uint64_t)IPT ? (left - woff + 31) / 32 : IPT) : 0);
        uint32_t pr[IPT];   // peers, then rank
#pragma unroll
        for (int j = 0; j < IPT; ++j) {
            // lanes of this warp holding the same digit; lanes past the end of the chunk stand alone
            const uint32_t vmask = __ballot_sync(0xffffffffu, j < nv);
            const uint32_t same = match_digit8(digit_of(k[j], shift)) & vmask;
            pr[j] = j < nv ? same : (1u << lane);
        }
#pragma unroll
        for (int j = 0; j < IPT; ++j) {
            const uint32_t peers = pr[j];
            const int leader = __ffs(peers) - 1;
            uint32_t before = 0;
            if (j < nv && lane == leader) {
                const uint32_t d = digit_of(k[j], shift);
                before = cnt[warp][d];
                cnt[warp][d] = before + __popc(peers);
            }
            __syncwarp();   // the next item's leader may be another lane reading this counter
            pr[j] = __shfl_sync(0xffffffffu, before, leader) + __popc(peers & lt_mask);
        }
        __syncthreads();
        // thread t owns digit t: prefix over the warps, then a block scan over the digits
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
            const uint32_t c = cnt[w][t];
            cnt[w][t] = run;
            run += c;
        }
        uint32_t inc = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        uint32_t ex = inc - run;
#pragma unroll
        for (int w = 0; w < kWarps; ++w)
            if (w < warp) ex += wsum[w];
        tprefix[t] = ex;
        gpos[t] = off[t] - ex;
        off[t] += run;
        __syncthreads();
#pragma unroll
        for (int j = 0; j < IPT; ++j)
            if (j < nv) {
                const uint32_t d = digit_of(k[j], shift);
                skeys[tprefix[d] + cnt[warp][d] + pr[j]] = k[j];
            }
        // the registers are free: request the next tile now, clear my warp's counter row for it
        const uint64_t nb = base + kTile + woff;
#pragma unroll
        for (int j = 0; j < IPT; ++j) k[j] = nb + j * 32 < hi ? in[nb + j * 32] : ~0ull;
        __syncwarp();
#pragma unroll
        for (int i = 0; i < kRadix / 32; ++i) cnt[warp][lane + 32 * i] = 0;
        __syncthreads();
        const uint32_t tile_n = left < (uint64_t)kTile ? (uint32_t)left : (uint32_t)kTile;
#pragma unroll
        for (int j = 0; j < IPT; ++j) {
            const uint32_t i = t + j * kSortThreads;
            if (i < tile_n) {
                const uint64_t kk = skeys[i];
                const uint64_t pos = gpos[digit_of(kk, shift)] + i;
                out[pos] = kk;
                // fused mode (small inputs, chunk = 1 << chunk_shift): count the key for the next pass's histogram
                // of the CTA that will own position pos - saves that pass's histogram launch
                if (next_counts)
                    atomicAdd(&next_counts[(uint64_t)digit_of(kk, shift + 8) * gridDim.x + (pos >> chunk_shift)], 1u);
            }
        }
    }
}

// scratch layout: counts u32[256][ctas] | offsets u64[256][ctas] | row_total u64[256] | digit_base u64[256] | ticket
// (the ticket word is reset by each pass's histogram kernel: nothing in the scratch has to start at zero)
constexpr size_t sort_scratch_for(int ctas) {
    return (size_t)kRadix * ctas * (sizeof(uint32_t) + sizeof(uint64_t)) + 2 * (size_t)kRadix * sizeof(uint64_t) + 8;
}

// Small inputs are launch-bound (24 launches of ~5.5 us): up to this many keys the scatter kernel of pass p also
// builds the histogram of pass p+1 with global atomics (one tile of 2048 keys per CTA, so the CTA that will own an
// output position is position >> 11), which removes seven of the eight histogram launches.  For large inputs
// 2^26 atomics per pass would cost more than the separate histogram kernel's extra read.
constexpr uint64_t kSortFuseMaxKeys = 1ull << 19;
constexpr int kSortFuseChunkShift = 11;
static bool sort_fused(uint64_t n) {
    static const bool enabled = [] {
        const char *e = getenv("OCM_SORT_FUSE");   // 0 disables the fused small-input mode
        return !(e && atoi(e) == 0);
    }();
    return enabled && n <= kSortFuseMaxKeys;
}

size_t sort_scratch_bytes(uint64_t n, int *ctas_out) {
    // one scatter tile (2048 keys) per CTA until the grid is four CTAs per SM, longer chunks after that
    int ctas = (int)std::min<uint64_t>(std::max<uint64_t>((n + 2047) / 2048, 1), (uint64_t)num_sms() * 4);
    if (ctas_out) *ctas_out = ctas;
    // fused mode keeps a second count matrix behind the regular layout
    return sort_scratch_for(ctas) + (sort_fused(n) ? (size_t)kRadix * ctas * sizeof(uint32_t) : 0);
}

// The eight passes (hist, scan, scatter), shared by the regular and the allocation-free entry point.
// fuse: see kSortFuseMaxKeys; needs ctas == ceil(n / 2048) and the second count matrix in the scratch.
static void sort_passes(uint64_t *keys, uint64_t *tmp, uint64_t n, int ctas, void *scratch, cudaStream_t stream, bool fuse) {
    uint32_t *counts = static_cast<uint32_t *>(scratch);
    uint64_t *offsets = reinterpret_cast<uint64_t *>(static_cast<uint8_t *>(scratch) + (size_t)kRadix * ctas * sizeof(uint32_t));
    uint64_t *row_total = offsets + (size_t)kRadix * ctas;
    uint64_t *digit_base = row_total + kRadix;
    unsigned int *ticket = reinterpret_cast<unsigned int *>(digit_base + kRadix);
    uint32_t *counts2 = reinterpret_cast<uint32_t *>(reinterpret_cast<uint8_t *>(ticket) + 8);   // fused mode only
    const uint64_t chunk = fuse ? (1ull << kSortFuseChunkShift)
                                : ((n + ctas - 1) / ctas + kSortThreads - 1) / kSortThreads * kSortThreads;
    uint64_t *a = keys, *b = tmp;
    uint32_t *cur = counts, *nxt = counts2;
    static const int ipt = [] {
        const char *e = getenv("OCM_SORT_IPT");   // keys per thread per tile of the scatter: 8 (default) or 16
        return e && atoi(e) == 16 ? 16 : 8;
    }();
    for (int pass = 0; pass < 8; ++pass) {
        const int shift = pass * 8;
        const bool feed_next = fuse && pass < 7;   // this pass's scatter counts the digits of the next pass
        if (!fuse || pass == 0) ocm_sort_hist_kernel<<<ctas, kSortThreads, 0, stream>>>(a, n, shift, cur, chunk, ticket);
        ocm_sort_scan_kernel<<<kRadix, 256, 0, stream>>>(cur, offsets, row_total, digit_base, ctas, ticket,
                                                        feed_next ? nxt : nullptr);
        if (ipt == 16)
            ocm_sort_scatter_kernel<16><<<ctas, kSortThreads, 0, stream>>>(a, b, n, shift, offsets, digit_base, chunk,
                                                                           feed_next ? nxt : nullptr, kSortFuseChunkShift);
        else
            ocm_sort_scatter_kernel<8><<<ctas, kSortThreads, 0, stream>>>(a, b, n, shift, offsets, digit_base, chunk,
                                                                          feed_next ? nxt : nullptr, kSortFuseChunkShift);
        std::swap(a, b);
        if (fuse) std::swap(cur, nxt);
    }
}

// keys and tmp: n elements each.  scratch: sort_scratch_bytes(n) bytes.
cudaError_t launch_sort_u64(uint64_t *keys, uint64_t *tmp, uint64_t n, void *scratch, size_t scratch_bytes,
                            cudaStream_t stream, unsigned long long *launches) {
    if (n < 2) return cudaSuccess;
    int ctas = 0;
    if (scratch_bytes < sort_scratch_bytes(n, &ctas)) return cudaErrorInvalidValue;
    // the fused small-input mode assumes one 2048-key tile per CTA (true while n/2048 <= 4 CTAs per SM)
    const bool fuse = sort_fused(n) && (uint64_t)ctas == (n + 2047) / 2048;
    sort_passes(keys, tmp, n, ctas, scratch, stream, fuse);
    if (launches) *launches += fuse ? 3 + 7 * 2 : 8 * 3;
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
//     entry before reading it and its histogram kernel resets the scan ticket word, so there is
//     nothing to clear between calls;
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
    sort_passes(keys, tmp, n, ctas, scratch, static_cast<cudaStream_t>(stream), /*fuse=*/false);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        LOGE("ocmk_sort_u64_noalloc: launch failed: %s", cudaGetErrorString(e));
        // the cached address belongs to a context that may be gone (cudaDeviceReset): look it up again next time
        g_sort_scratch_addr[dev].store(nullptr, std::memory_order_release);
        return -OCM_ECUDA;
    }
    ocmk::extra_launches().fetch_add(24, std::memory_order_relaxed);   // 8 passes x (histogram, scan, scatter)
    return 0;
}
