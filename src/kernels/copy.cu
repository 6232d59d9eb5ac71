// K1/K2/K3 - contiguous one-sided put / get / third-party copy for sm_100a.
//
// Replaces the reference's data plane: ib_write/ib_read + ib_poll (reference
// src/rdma.c:46-85,241-302), extoll_write/extoll_read with 2 x 8 MiB overlap
// (src/extoll.c:40-173) and the staging memcpy/cudaMemcpy hops in ocm_copy
// (src/lib.c:523-656).  Here one kernel moves the bytes end to end:
//
//  * bulk path: every warp of a persistent CTA runs an independent ring of
//    `stages` tiles.  One elected lane issues cp.async.bulk global->shared
//    (TMA engine, mbarrier complete_tx) and, when a tile has landed,
//    cp.async.bulk shared->global to the destination - and optionally to a
//    second destination, so a two-hop "window + far side" ocm_copy reads the
//    source once.  Source and destination are generic addresses: local HBM,
//    peer HBM over NVLink, or pinned host memory.  No SM registers touch the
//    payload.
//  * 256-bit vector path (default): 32 bytes per thread per access with the
//    sm_100 LDG/STG.256, optional second destination.  Measured faster than the
//    bulk path for pure copies (no shared-memory round trip), see
//    profiles/README.md; the bulk path is selected with COPY_BULK.
//  * narrow vector path: 16/8/4/1-byte LDG/STG grid-stride loop for transfers
//    whose pointers are not congruent mod 32.
//
// Completion: the last CTA to finish (elected through a device-local ticket
// word) adds 1 to the destination's signal word with release semantics at
// system scope, after its own system fence.
#include "kernels.h"
#include "ptx.cuh"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace ocmk {

constexpr int kMaxAgents = 8;
constexpr int kMaxStages = 8;

struct CopyArgs {
    const uint8_t *src;
    uint8_t       *dst;
    uint8_t       *dst2;
    uint64_t       bytes;
    uint32_t       tile;
    uint32_t       stages;
    uint64_t      *signal;
    unsigned      *ticket;
    uint64_t      *done_flag;     // optional host-visible completion word
    uint64_t       done_value;
};

__global__ void __launch_bounds__(kMaxAgents * 32, 1) ocm_copy_bulk_kernel(CopyArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[kMaxAgents][kMaxStages];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int agents = blockDim.x >> 5;
    const uint32_t tile = a.tile, stages = a.stages;

    if (lane == 0) {
        for (uint32_t s = 0; s < stages; ++s) mbar_init(&full[warp][s], 1);
        mbar_fence_init();
    }
    __syncthreads();

    const uint64_t bulk_bytes = a.bytes & ~15ull;
    const uint64_t n_tiles = (bulk_bytes + tile - 1) / tile;
    const uint64_t ag = (uint64_t)blockIdx.x * agents + warp;
    const uint64_t total = (uint64_t)gridDim.x * agents;

    if (lane == 0 && ag < n_tiles) {
        uint8_t *my = smem + (size_t)warp * stages * tile;
        const uint64_t n_my = (n_tiles - ag + total - 1) / total;
        auto tile_off = [&](uint64_t j) { return (ag + j * total) * (uint64_t)tile; };
        auto tile_len = [&](uint64_t j) {
            uint64_t off = tile_off(j);
            return (uint32_t)(bulk_bytes - off < tile ? bulk_bytes - off : tile);
        };
        auto issue_load = [&](uint64_t j) {
            uint32_t s = (uint32_t)(j % stages);
            uint32_t len = tile_len(j);
            mbar_arrive_expect_tx(&full[warp][s], len);
            bulk_g2s(my + (size_t)s * tile, a.src + tile_off(j), len, &full[warp][s]);
        };
        const uint64_t npro = n_my < stages ? n_my : stages;
        for (uint64_t j = 0; j < npro; ++j) issue_load(j);
        for (uint64_t i = 0; i < n_my; ++i) {
            uint32_t s = (uint32_t)(i % stages);
            mbar_wait(&full[warp][s], (uint32_t)((i / stages) & 1));
            uint64_t off = tile_off(i);
            uint32_t len = tile_len(i);
            bulk_s2g(a.dst + off, my + (size_t)s * tile, len);
            if (a.dst2) bulk_s2g(a.dst2 + off, my + (size_t)s * tile, len);
            bulk_commit();
            // the store of tile i-1 has finished reading its stage once at most
            // one group (tile i) is still pending: refill that stage
            if (i >= 1 && i - 1 + stages < n_my) {
                bulk_wait_read<1>();
                issue_load(i - 1 + stages);
            }
        }
        bulk_wait_all<0>();
        fence_proxy_async_all();
        __threadfence_system();
    }
    // tail (< 16 bytes) - plain byte stores by the first CTA
    if (blockIdx.x == 0 && threadIdx.x < (a.bytes & 15ull)) {
        uint64_t o = bulk_bytes + threadIdx.x;
        uint8_t v = a.src[o];
        a.dst[o] = v;
        if (a.dst2) a.dst2[o] = v;
    }
    __syncthreads();
    if (threadIdx.x == 0) finish_and_signal(a.signal, a.ticket, a.done_flag, a.done_value);
}

// Vector path.  V = access width in bytes (16, 8, 4 or 1): the widest power of
// two dividing every pointer.
template <typename VT>
__global__ void __launch_bounds__(256) ocm_copy_vec_kernel(CopyArgs a) {
    constexpr uint64_t V = sizeof(VT);
    const uint64_t nvec = a.bytes / V;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const VT *__restrict__ s = reinterpret_cast<const VT *>(a.src);
    VT *__restrict__ d = reinterpret_cast<VT *>(a.dst);
    VT *__restrict__ d2 = reinterpret_cast<VT *>(a.dst2);
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    // 4 independent loads in flight per thread
    for (; i + 3 * stride < nvec; i += 4 * stride) {
        VT v0 = s[i], v1 = s[i + stride], v2 = s[i + 2 * stride], v3 = s[i + 3 * stride];
        d[i] = v0; d[i + stride] = v1; d[i + 2 * stride] = v2; d[i + 3 * stride] = v3;
        if (d2) { d2[i] = v0; d2[i + stride] = v1; d2[i + 2 * stride] = v2; d2[i + 3 * stride] = v3; }
    }
    for (; i < nvec; i += stride) {
        VT v = s[i];
        d[i] = v;
        if (d2) d2[i] = v;
    }
    const uint64_t done = nvec * V;
    if (blockIdx.x == 0 && threadIdx.x < a.bytes - done) {
        uint8_t v = a.src[done + threadIdx.x];
        a.dst[done + threadIdx.x] = v;
        if (a.dst2) a.dst2[done + threadIdx.x] = v;
    }
    __syncthreads();
    if (threadIdx.x == 0) finish_and_signal(a.signal, a.ticket, a.done_flag, a.done_value);
}

// 256-bit vector path (LDG.E.ENL2.256 / STG.E.ENL2.256, new on sm_100): the measured winner for pure
// copies on B200 - it matches cudaMemcpy inside one HBM and over NVLink and has ~2 us less fixed cost
// than staging through shared memory (profiles/kbench_*.md).  All pointers must be congruent mod 32;
// up to 31 head bytes and the tail are moved bytewise by the first CTA.
struct V32 {
    uint32_t w[8];
};
__device__ __forceinline__ V32 ld_v32(const uint8_t *p) {
    V32 v;
    asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3]), "=r"(v.w[4]), "=r"(v.w[5]), "=r"(v.w[6]),
                   "=r"(v.w[7])
                 : "l"(p));
    return v;
}
__device__ __forceinline__ void st_v32(uint8_t *p, const V32 &v) {
    asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v.w[0]), "r"(v.w[1]),
                 "r"(v.w[2]), "r"(v.w[3]), "r"(v.w[4]), "r"(v.w[5]), "r"(v.w[6]), "r"(v.w[7])
                 : "memory");
}

template <bool DUAL>
__global__ void __launch_bounds__(256) ocm_copy_v32_kernel(CopyArgs a) {
    uint64_t head = (32 - ((uintptr_t)a.dst & 31)) & 31;
    if (head > a.bytes) head = a.bytes;
    const uint64_t n = (a.bytes - head) / 32;
    const uint8_t *s = a.src + head;
    uint8_t *d = a.dst + head;
    uint8_t *d2 = DUAL ? a.dst2 + head : nullptr;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + stride < n; i += 2 * stride) {
        V32 v0 = ld_v32(s + i * 32), v1 = ld_v32(s + (i + stride) * 32);
        st_v32(d + i * 32, v0);
        st_v32(d + (i + stride) * 32, v1);
        if (DUAL) {
            st_v32(d2 + i * 32, v0);
            st_v32(d2 + (i + stride) * 32, v1);
        }
    }
    for (; i < n; i += stride) {
        V32 v = ld_v32(s + i * 32);
        st_v32(d + i * 32, v);
        if (DUAL) st_v32(d2 + i * 32, v);
    }
    if (blockIdx.x == 0) {
        const uint64_t tail0 = head + n * 32;
        for (uint64_t o = threadIdx.x; o < head + (a.bytes - tail0); o += blockDim.x) {
            const uint64_t k = o < head ? o : tail0 + (o - head);
            const uint8_t v = a.src[k];
            a.dst[k] = v;
            if (DUAL) a.dst2[k] = v;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) finish_and_signal(a.signal, a.ticket, a.done_flag, a.done_value);
}

// ---- batched copies: one launch, many segments ------------------------------------------------------------------
struct BatchArgs {
    CopySeg  seg[kMaxBatch];
    uint32_t cta_first[kMaxBatch + 1];   // CTAs [cta_first[i], cta_first[i+1]) work on segment i
    uint32_t n;
    uint32_t signal_add;
    uint64_t *signal;
    unsigned *ticket;
    uint64_t *done_flag;
    uint64_t  done_value;
};

__global__ void __launch_bounds__(256) ocm_copy_batch_kernel(const __grid_constant__ BatchArgs a) {
    // which segment is mine (the table lives in constant-bank kernel parameters: a short scan is cheap)
    uint32_t i = 0;
    while (i + 1 < a.n && blockIdx.x >= a.cta_first[i + 1]) ++i;
    const uint32_t part = blockIdx.x - a.cta_first[i], parts = a.cta_first[i + 1] - a.cta_first[i];
    const uint8_t *src = static_cast<const uint8_t *>(a.seg[i].src);
    uint8_t *dst = static_cast<uint8_t *>(a.seg[i].dst);
    const uint64_t bytes = a.seg[i].bytes;
    const uintptr_t r = (uintptr_t)dst & 31;
    if (((uintptr_t)src & 31) == r) {
        uint64_t head = (32 - r) & 31;
        if (head > bytes) head = bytes;
        const uint64_t n = (bytes - head) / 32;
        const uint64_t stride = (uint64_t)parts * blockDim.x;
        uint64_t k = (uint64_t)part * blockDim.x + threadIdx.x;
        for (; k + stride < n; k += 2 * stride) {
            V32 v0 = ld_v32(src + head + k * 32), v1 = ld_v32(src + head + (k + stride) * 32);
            st_v32(dst + head + k * 32, v0);
            st_v32(dst + head + (k + stride) * 32, v1);
        }
        for (; k < n; k += stride) st_v32(dst + head + k * 32, ld_v32(src + head + k * 32));
        if (part == 0) {
            const uint64_t tail0 = head + n * 32;
            for (uint64_t o = threadIdx.x; o < head + (bytes - tail0); o += blockDim.x) {
                const uint64_t b = o < head ? o : tail0 + (o - head);
                dst[b] = src[b];
            }
        }
    } else {
        for (uint64_t b = (uint64_t)part * blockDim.x + threadIdx.x; b < bytes; b += (uint64_t)parts * blockDim.x) dst[b] = src[b];
    }
    __syncthreads();
    if (threadIdx.x == 0 && (a.signal || a.done_flag)) {
        bool last = true;
        if (gridDim.x > 1) {
            unsigned t;
            asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(t) : "l"(a.ticket) : "memory");
            last = (t == gridDim.x - 1);
            if (last) atomicExch(a.ticket, 0u);
        }
        if (last) {
            asm volatile("fence.acq_rel.sys;" ::: "memory");
            if (a.signal) red_release_sys_add(a.signal, (uint64_t)a.signal_add);
            if (a.done_flag) st_release_sys(a.done_flag, a.done_value);
        }
    }
}

__global__ void ocm_signal_kernel(uint64_t *signal) {
    __threadfence_system();
    red_release_sys_add(signal, 1ull);
}

// Completion of a LARGE copy: published by a one-thread kernel that follows the copy kernel on the stream.
// The kernel boundary orders every byte; this avoids a release in each of the ~2400 CTAs of the copy
// (measured 8-14 us on a 128 us two-destination copy; the extra launch costs ~1 us on the device).
__global__ void ocm_complete_kernel(uint64_t *signal, uint64_t *done_flag, uint64_t done_value) {
    asm volatile("fence.acq_rel.sys;" ::: "memory");
    if (signal) red_release_sys_add(signal, 1ull);
    if (done_flag) st_release_sys(done_flag, done_value);
}

// ------------------------------------------------------------------ host ----
std::atomic<unsigned long long> &extra_launches() {
    static std::atomic<unsigned long long> n{0};
    return n;
}

int num_sms(int device) {
    static int cached[64] = {0};
    int dev = device;
    if (dev < 0 && cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev >= 0 && dev < 64 && cached[dev]) return cached[dev];
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    if (dev >= 0 && dev < 64) cached[dev] = n;
    return n;
}

static long env_long(const char *k, long dflt) {
    const char *e = getenv(k);
    if (!e || !*e) return dflt;
    char *end = nullptr;
    long v = strtol(e, &end, 0);
    if (end && (*end == 'k' || *end == 'K')) v <<= 10;
    if (end && (*end == 'm' || *end == 'M')) v <<= 20;
    return v;
}

const CopyTuning &default_copy_tuning() {
    static CopyTuning t = [] {
        CopyTuning c;
        if (const char *m = getenv("OCM_COPY_MODE")) {
            if (!strcmp(m, "bulk")) c.mode = COPY_BULK;
            else if (!strcmp(m, "vec")) c.mode = COPY_VEC;
        }
        c.ctas = (int)env_long("OCM_COPY_CTAS", c.ctas);
        c.agents = (int)env_long("OCM_COPY_AGENTS", c.agents);
        c.stages = (int)env_long("OCM_COPY_STAGES", c.stages);
        c.tile_bytes = (int)env_long("OCM_COPY_TILE", c.tile_bytes);
        c.bulk_min = (size_t)env_long("OCM_BULK_MIN", (long)c.bulk_min);
        return c;
    }();
    return t;
}

static cudaError_t ensure_bulk_smem(size_t bytes) {
    static size_t configured[64] = {0};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 64 && configured[dev] >= bytes) return cudaSuccess;
    e = cudaFuncSetAttribute(ocm_copy_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e == cudaSuccess && dev < 64) configured[dev] = bytes;
    return e;
}

cudaError_t launch_copy(void *dst, void *dst2, const void *src, size_t bytes, uint64_t *signal, unsigned *ticket,
                        cudaStream_t stream, const CopyTuning *tuning, unsigned long long *launches,
                        const CopyNotify *notify) {
    if (bytes == 0) {
        if (signal) {
            ocm_signal_kernel<<<1, 1, 0, stream>>>(signal);
            if (launches) ++*launches;
        }
        return cudaGetLastError();
    }
    const CopyTuning &t = tuning ? *tuning : default_copy_tuning();
    CopyArgs a;
    a.src = static_cast<const uint8_t *>(src);
    a.dst = static_cast<uint8_t *>(dst);
    a.dst2 = static_cast<uint8_t *>(dst2);
    a.bytes = bytes;
    a.signal = signal;
    a.ticket = ticket;
    a.done_flag = notify ? notify->flag : nullptr;
    a.done_value = notify ? notify->value : 0;
    const uintptr_t align_or = (uintptr_t)src | (uintptr_t)dst | (uintptr_t)dst2;
    const int sms = num_sms();

    // AUTO: the 256-bit vector kernel whenever the pointers are congruent mod 32 (pool blocks always are);
    // the TMA kernel only on request (OCM_COPY_MODE=bulk / mode=COPY_BULK); narrower vectors otherwise.
    bool bulk = t.mode == COPY_BULK && (align_or & 15) == 0 && bytes >= 16;
    const uintptr_t r32 = (uintptr_t)dst & 31;
    const bool v32 = !bulk && t.mode != COPY_VEC && ((uintptr_t)src & 31) == r32 && (!dst2 || ((uintptr_t)dst2 & 31) == r32);

    if (v32) {
        a.tile = 0;
        a.stages = 0;
        const uint64_t n32 = bytes / 32;
        const uint64_t want = (n32 + 511) / 512;           // 2 x 32 B per thread per iteration
        // two destinations write twice as much: 16 CTAs/SM (two waves) measured +9 % over 4/SM inside one HBM
        // (kbench dual: 2113 vs 1935 GB/s at 256 MiB = 6.3 TB/s of traffic); one destination is flat beyond 4/SM
        int cap = t.ctas > 0 ? t.ctas : sms * (dst2 ? 16 : 4);
        int ctas = (int)std::min<uint64_t>(std::max<uint64_t>(want, 1), (uint64_t)cap);
        if (ctas > 1 && signal && !ticket) return cudaErrorInvalidValue;
        // many CTAs: completion moves to a follow-up one-thread kernel (see ocm_complete_kernel)
        const bool split = (a.signal || a.done_flag) && ctas > sms * 4;
        uint64_t *sig = a.signal, *flag = a.done_flag;
        if (split) a.signal = a.done_flag = nullptr;
        if (dst2) ocm_copy_v32_kernel<true><<<ctas, 256, 0, stream>>>(a);
        else ocm_copy_v32_kernel<false><<<ctas, 256, 0, stream>>>(a);
        if (split) {
            ocm_complete_kernel<<<1, 1, 0, stream>>>(sig, flag, a.done_value);
            if (launches) ++*launches;
        }
    } else if (bulk) {
        int agents = std::min(std::max(t.agents, 1), kMaxAgents);
        int stages = std::min(std::max(t.stages, 2), kMaxStages);
        size_t tile = ((size_t)std::max(t.tile_bytes, 1024) + 127) & ~(size_t)127;
        while ((size_t)agents * stages * tile > (size_t)(224 << 10) && stages > 2) --stages;
        while ((size_t)agents * stages * tile > (size_t)(224 << 10) && agents > 1) --agents;
        while ((size_t)agents * stages * tile > (size_t)(224 << 10)) tile >>= 1;
        a.tile = (uint32_t)tile;
        a.stages = (uint32_t)stages;
        uint64_t n_tiles = ((bytes & ~15ull) + tile - 1) / tile;
        uint64_t want = (n_tiles + agents - 1) / agents;
        int ctas = t.ctas > 0 ? t.ctas : sms;
        if ((uint64_t)ctas > want) ctas = (int)want;
        if (ctas < 1) ctas = 1;
        if (ctas > 1 && signal && !ticket) return cudaErrorInvalidValue;
        size_t smem = (size_t)agents * stages * tile;
        cudaError_t e = ensure_bulk_smem(224 << 10);
        if (e != cudaSuccess) return e;
        ocm_copy_bulk_kernel<<<ctas, agents * 32, smem, stream>>>(a);
    } else {
        a.tile = 0;
        a.stages = 0;
        const int threads = 256;
        int width = 16;
        while (width > 1 && (align_or & (uintptr_t)(width - 1))) width >>= 1;
        if (width == 2) width = 1;
        uint64_t nvec = bytes / (uint64_t)width;
        uint64_t want = (nvec + (uint64_t)threads * 4 - 1) / ((uint64_t)threads * 4);
        int cap = t.ctas > 0 ? t.ctas : sms * 4;
        int ctas = (int)std::min<uint64_t>(std::max<uint64_t>(want, 1), (uint64_t)cap);
        if (ctas > 1 && signal && !ticket) return cudaErrorInvalidValue;
        switch (width) {
        case 16: ocm_copy_vec_kernel<uint4><<<ctas, threads, 0, stream>>>(a); break;
        case 8: ocm_copy_vec_kernel<uint2><<<ctas, threads, 0, stream>>>(a); break;
        case 4: ocm_copy_vec_kernel<uint32_t><<<ctas, threads, 0, stream>>>(a); break;
        default: ocm_copy_vec_kernel<uint8_t><<<ctas, threads, 0, stream>>>(a); break;
        }
    }
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_copy_batch(const CopySeg *segs, int n, uint64_t *signal, unsigned *ticket, cudaStream_t stream,
                              unsigned long long *launches, const CopyNotify *notify) {
    if (n <= 0) return cudaSuccess;
    if ((signal || notify) && !ticket) return cudaErrorInvalidValue;
    const int sms = num_sms();
    for (int base = 0; base < n; base += kMaxBatch) {
        const int m = std::min(kMaxBatch, n - base);
        const bool final_launch = base + m >= n;
        BatchArgs a;
        memset(&a, 0, sizeof a);
        a.n = (uint32_t)m;
        // CTAs per segment: one per 16 KiB, the whole launch capped at four CTAs per SM
        uint64_t total = 0;
        for (int i = 0; i < m; ++i) total += segs[base + i].bytes;
        const uint64_t cap = (uint64_t)sms * 4;
        uint32_t next = 0;
        for (int i = 0; i < m; ++i) {
            a.seg[i] = segs[base + i];
            uint64_t want = std::max<uint64_t>(1, (a.seg[i].bytes + 16383) / 16384);
            if (total > cap * 16384) want = std::max<uint64_t>(1, a.seg[i].bytes * cap / total);
            a.cta_first[i] = next;
            next += (uint32_t)want;
        }
        a.cta_first[m] = next;
        // the completion of a long list is published by its last launch (stream order covers the earlier ones)
        a.signal = final_launch ? signal : nullptr;
        a.signal_add = (uint32_t)n;
        a.ticket = ticket;
        a.done_flag = final_launch && notify ? notify->flag : nullptr;
        a.done_value = notify ? notify->value : 0;
        ocm_copy_batch_kernel<<<next, 256, 0, stream>>>(a);
        if (launches) ++*launches;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

cudaError_t launch_signal(uint64_t *signal, cudaStream_t stream) {
    ocm_signal_kernel<<<1, 1, 0, stream>>>(signal);
    return cudaGetLastError();
}

} // namespace ocmk
