// K6 - one-to-many put (no counterpart in the reference: it has neither
// multicast nor collectives).
//
// (a) Cluster-shared staging: a thread-block cluster of C CTAs (C = number of
//     replicas, <= 8) moves each tile ONCE out of HBM.  Every CTA bulk-loads
//     1/C of the tile and multicasts it into the shared memory of all C CTAs
//     (cp.async.bulk ... .multicast::cluster), so each CTA ends up with the
//     whole tile after issuing 1/C of the reads; CTA r then bulk-stores the
//     tile to replica r (a peer GPU over NVLink).  Stage recycling is a
//     cluster-wide empty barrier (remote mbarrier arrives through DSMEM).
//     This form is taken only when the caller knows the source is in the
//     launching GPU's own memory; for any other source every CTA bulk-loads
//     the whole tile itself (template argument MC below).
// (b) NVLS: with a multicast virtual address the NVSwitch replicates a single
//     multimem.st stream to every bound GPU.
#include "kernels.h"
#include "ptx.cuh"
#include <algorithm>

namespace ocmk {

constexpr int kRepMaxDst = 8;
constexpr int kRepStages = 4;
constexpr uint32_t kRepTile = 32768;

struct RepArgs {
    const uint8_t *src;
    uint8_t       *dst[kRepMaxDst];
    uint64_t      *signal[kRepMaxDst];
    uint64_t       bytes;
    unsigned      *ticket;
    int            n_dst;
};

// MC: the cluster shares each tile with multicast loads.  Only for a source in THIS GPU's own memory: TMA multicast
// loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU.
// Any other source (a peer's memory, pinned host memory, unknown) takes MC = false: every CTA loads the whole tile
// itself with plain bulk copies and no CTA depends on another one's loads.
template <bool MC>
__global__ void __launch_bounds__(128, 1) ocm_put_replicated_kernel(RepArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[kRepStages];
    __shared__ __align__(8) uint64_t empty[kRepStages];
    const uint32_t C = cluster_nctarank();
    const uint32_t r = cluster_ctarank();
    if (threadIdx.x == 0) {
        for (int s = 0; s < kRepStages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], C);
        }
        mbar_fence_init();
    }
    cluster_sync_all();   // barriers of every CTA are initialised before any remote traffic

    const uint64_t bulk_bytes = a.bytes & ~15ull;
    const uint64_t n_tiles = (bulk_bytes + kRepTile - 1) / kRepTile;
    const uint64_t cluster_id = blockIdx.x / C, n_clusters = gridDim.x / C;
    const uint16_t mask_all = (uint16_t)((1u << C) - 1);

    if (threadIdx.x == 0 && cluster_id < n_tiles) {
        const uint64_t n_my = (n_tiles - cluster_id + n_clusters - 1) / n_clusters;
        auto tile_off = [&](uint64_t j) { return (cluster_id + j * n_clusters) * (uint64_t)kRepTile; };
        auto tile_len = [&](uint64_t j) {
            uint64_t off = tile_off(j);
            return (uint32_t)(bulk_bytes - off < kRepTile ? bulk_bytes - off : kRepTile);
        };
        auto issue_load = [&](uint64_t j) {
            const uint32_t s = (uint32_t)(j % kRepStages);
            const uint32_t len = tile_len(j);
            // every CTA will receive the whole tile, a slice from each cluster member
            mbar_arrive_expect_tx(&full[s], len);
            if (!MC) {
                bulk_g2s(smem + (size_t)s * kRepTile, a.src + tile_off(j), len, &full[s]);
                return;
            }
            const uint32_t slice = ((len + C - 1) / C + 15) & ~15u;
            const uint32_t off = r * slice;
            if (off < len) {
                const uint32_t n = len - off < slice ? len - off : slice;
                bulk_g2s_multicast(smem + (size_t)s * kRepTile + off, a.src + tile_off(j) + off, n, &full[s], mask_all);
            }
        };
        const uint64_t npro = n_my < (uint64_t)kRepStages ? n_my : (uint64_t)kRepStages;
        for (uint64_t j = 0; j < npro; ++j) issue_load(j);
        for (uint64_t i = 0; i < n_my; ++i) {
            const uint32_t s = (uint32_t)(i % kRepStages);
            mbar_wait(&full[s], (uint32_t)((i / kRepStages) & 1));
            bulk_s2g(a.dst[r] + tile_off(i), smem + (size_t)s * kRepTile, tile_len(i));
            bulk_commit();
            if (i >= 1) {
                // store i-1 has drained its stage: tell the whole cluster, then refill it
                bulk_wait_read<1>();
                const uint32_t ps = (uint32_t)((i - 1) % kRepStages);
                if (MC)
                    for (uint32_t c = 0; c < C; ++c) mbar_arrive_remote(&empty[ps], c);
                const uint64_t j = i - 1 + kRepStages;
                if (j < n_my) {
                    // shared tiles: every CTA of the cluster must have drained the stage; private tiles: ours has
                    if (MC) mbar_wait(&empty[ps], (uint32_t)(((j / kRepStages) - 1) & 1));
                    issue_load(j);
                }
            }
        }
        bulk_wait_all<0>();
        fence_proxy_async_all();
        __threadfence_system();
    }
    // tail bytes: CTA r of the first cluster writes replica r
    if (cluster_id == 0 && threadIdx.x < (a.bytes & 15ull)) {
        uint64_t o = bulk_bytes + threadIdx.x;
        a.dst[r][o] = a.src[o];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // GPU-scope release on the ticket, one system fence in the CTA that draws the last one (ptx.cuh,
        // finish_and_signal): cumulativity carries every CTA's stores to the signals that follow
        bool last = true;
        if (gridDim.x > 1) {
            unsigned t = 0;
            if (a.ticket) asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(t) : "l"(a.ticket) : "memory");
            last = a.ticket && t == gridDim.x - 1;
            if (last) atomicExch(a.ticket, 0u);
        }
        if (last) {
            asm volatile("fence.acq_rel.sys;" ::: "memory");
            for (int d = 0; d < a.n_dst; ++d)
                if (a.signal[d]) red_release_sys_add(a.signal[d], 1ull);
        }
    }
    cluster_sync_all();   // nobody leaves while peers may still touch its shared memory
}

// (b) NVLS multimem store stream
__global__ void __launch_bounds__(256) ocm_put_multimem_kernel(uint8_t *mc_dst, const uint8_t *src, uint64_t bytes,
                                                               uint64_t *mc_signal, unsigned *ticket) {
    const uint64_t n16 = bytes / 16;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
        uint4 v = ldg_stream(src + i * 16);
        multimem_st_v4(mc_dst + i * 16, v);
    }
    __syncthreads();
    if (threadIdx.x == 0 && mc_signal) {
        bool last = true;
        if (gridDim.x > 1) {
            unsigned t = 0;
            if (ticket) asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(t) : "l"(ticket) : "memory");
            last = ticket && t == gridDim.x - 1;
            if (last) atomicExch(ticket, 0u);
        }
        if (last) {
            // every replica's completion word moves at once: one reduction on the word's multicast address
            asm volatile("fence.acq_rel.sys;" ::: "memory");
            multimem_red_release_sys_add(mc_signal, 1ull);
        }
    }
}

// +1 on the completion word of a multicast-backed allocation on every GPU, ordered after prior work on the stream
// (for puts that could not take the multimem.st path: odd alignment, unaligned tails).
__global__ void ocm_signal_multimem_kernel(uint64_t *mc_signal) {
    asm volatile("fence.acq_rel.sys;" ::: "memory");
    multimem_red_release_sys_add(mc_signal, 1ull);
}
cudaError_t launch_signal_multimem(uint64_t *mc_signal, cudaStream_t stream) {
    ocm_signal_multimem_kernel<<<1, 1, 0, stream>>>(mc_signal);
    return cudaGetLastError();
}

cudaError_t launch_put_replicated(void *const *dsts, int n_dst, const void *src, size_t bytes,
                                  uint64_t *const *signals, unsigned *ticket, cudaStream_t stream, bool src_on_this_gpu) {
    if (n_dst < 1 || n_dst > kRepMaxDst) return cudaErrorInvalidValue;
    if (!bytes) return cudaSuccess;
    uintptr_t al = (uintptr_t)src;
    for (int i = 0; i < n_dst; ++i) al |= (uintptr_t)dsts[i];
    if (al & 15) return cudaErrorMisalignedAddress;
    RepArgs a;
    a.src = static_cast<const uint8_t *>(src);
    for (int i = 0; i < kRepMaxDst; ++i) {
        a.dst[i] = i < n_dst ? static_cast<uint8_t *>(dsts[i]) : nullptr;
        a.signal[i] = (i < n_dst && signals) ? signals[i] : nullptr;
    }
    a.bytes = bytes;
    a.ticket = ticket;
    a.n_dst = n_dst;
    const uint64_t n_tiles = std::max<uint64_t>(((bytes & ~15ull) + kRepTile - 1) / kRepTile, 1);
    int clusters = (int)std::min<uint64_t>(n_tiles, (uint64_t)std::max(1, num_sms() / n_dst));
    const size_t smem = (size_t)kRepStages * kRepTile;
    static bool attr[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 64 && !attr[dev]) {
        cudaError_t e = cudaFuncSetAttribute(ocm_put_replicated_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(ocm_put_replicated_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        attr[dev] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(clusters * n_dst));
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)n_dst;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return src_on_this_gpu ? cudaLaunchKernelEx(&cfg, ocm_put_replicated_kernel<true>, a)
                           : cudaLaunchKernelEx(&cfg, ocm_put_replicated_kernel<false>, a);
}

// (c) NVLS gather-reduce: dst[i] = sum over all GPUs bound to the multicast object of their replica's
// element i (fp32).  One multimem.ld_reduce per 16 bytes: the NVSwitch fetches and adds the replicas, the
// reading GPU receives one result - the "get" counterpart of the multimem.st put, and the first half of an
// in-switch all-reduce.  Two loads in flight per thread.
__global__ void __launch_bounds__(256) ocm_get_reduce_f32_kernel(float4 *dst, const uint8_t *mc_src, uint64_t n16) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + stride < n16; i += 2 * stride) {
        const float4 a = multimem_ld_reduce_add_f32x4(mc_src + i * 16);
        const float4 b = multimem_ld_reduce_add_f32x4(mc_src + (i + stride) * 16);
        dst[i] = a;
        dst[i + stride] = b;
    }
    if (i < n16) dst[i] = multimem_ld_reduce_add_f32x4(mc_src + i * 16);
}

cudaError_t launch_get_reduce_f32(void *dst, const void *mc_src, size_t bytes, cudaStream_t stream) {
    if (((uintptr_t)dst | (uintptr_t)mc_src | bytes) & 15) return cudaErrorMisalignedAddress;
    if (!bytes) return cudaSuccess;
    int ctas = (int)std::min<uint64_t>(std::max<uint64_t>((bytes / 16 + 511) / 512, 1), (uint64_t)num_sms() * 4);
    ocm_get_reduce_f32_kernel<<<ctas, 256, 0, stream>>>(static_cast<float4 *>(dst), static_cast<const uint8_t *>(mc_src),
                                                       bytes / 16);
    return cudaGetLastError();
}

cudaError_t launch_put_multimem(void *mc_dst, const void *src, size_t bytes, uint64_t *mc_signal, unsigned *ticket,
                                cudaStream_t stream) {
    if (((uintptr_t)mc_dst | (uintptr_t)src | bytes) & 15) return cudaErrorMisalignedAddress;
    if (!bytes) return cudaSuccess;
    int ctas = (int)std::min<uint64_t>(std::max<uint64_t>((bytes / 16 + 1023) / 1024, 1), (uint64_t)num_sms() * 2);
    ocm_put_multimem_kernel<<<ctas, 256, 0, stream>>>(static_cast<uint8_t *>(mc_dst), static_cast<const uint8_t *>(src),
                                                     bytes, mc_signal, ticket);
    return cudaGetLastError();
}

} // namespace ocmk
