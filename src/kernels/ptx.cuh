// Thin inline-PTX wrappers for sm_100a used by the copy kernels:
// mbarrier, bulk async copies (TMA engine, SASS UBLKCP), proxy fences,
// system-scope release/acquire, multimem (NVLS) and cluster helpers.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ocmk {

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---- mbarrier --------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok)
                 : "r"(smem_u32(bar)), "r"(parity)
                 : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}

// ---- bulk async copy (1-D TMA) -----------------------------------------------
// global -> shared, completion on an mbarrier.  16-byte aligned, size % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// global -> shared of every CTA in cta_mask (cluster multicast); each receiving
// CTA's mbarrier at the same smem offset gets the complete_tx.
__device__ __forceinline__ void bulk_g2s_multicast(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar,
                                                   uint16_t cta_mask) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster "
                 "[%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "h"(cta_mask)
                 : "memory");
}
// shared -> global, tracked by the issuing thread's bulk group.
__device__ __forceinline__ void bulk_s2g(void *gdst, const void *smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N> __device__ __forceinline__ void bulk_wait_all() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// ---- system-scope signalling ---------------------------------------------------
__device__ __forceinline__ void red_release_sys_add(uint64_t *addr, uint64_t v) {
    asm volatile("red.release.sys.global.add.u64 [%0], %1;" ::"l"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t *addr) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(addr) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t *addr) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys_u32(uint32_t *addr, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void st_release_sys(uint64_t *addr, uint64_t v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(addr), "l"(v) : "memory");
}

// ---- streaming 16-byte global access (no L1 allocation) ------------------------
__device__ __forceinline__ uint4 ldg_stream(const void *p) {
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void stg_stream(void *p, const uint4 &v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}

// ---- NVLS multimem (multicast virtual addresses only) --------------------------
__device__ __forceinline__ void multimem_st_v4(void *mc_ptr, const uint4 &v) {
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc_ptr), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ void multimem_red_release_sys_add(uint64_t *mc_addr, uint64_t v) {
    asm volatile("multimem.red.release.sys.global.add.u64 [%0], %1;" ::"l"(mc_addr), "l"(v) : "memory");
}
__device__ __forceinline__ float4 multimem_ld_reduce_add_f32x4(const void *mc_ptr) {
    float4 r;
    asm volatile("multimem.ld_reduce.weak.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(mc_ptr)
                 : "memory");
    return r;
}

// ---- cluster helpers ---------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive (count 1) on the mbarrier at the same smem offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint64_t *bar, uint32_t cta) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(bar)), "r"(cta));
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}

// Elect the last CTA and publish completion.  Called by one thread per CTA
// after a CTA-wide barrier that follows the CTA's last store.
// `done_flag` (optional) is a host-visible word that receives `done_value` (release, system scope) at the same
// point: a blocked host thread can spin on it instead of paying kernel tear-down + stream synchronisation.
__device__ __forceinline__ void finish_and_signal(uint64_t *signal, unsigned *ticket, uint64_t *done_flag = nullptr,
                                                  uint64_t done_value = 0) {
    if (signal == nullptr && done_flag == nullptr) return;   // nobody to notify: kernel completion orders the bytes
    bool last = true;
    if (gridDim.x > 1) {
        if (ticket == nullptr) return; // nobody to notify
        // Release at GPU scope on the ticket (ordered after the CTA's stores by the preceding barrier); only
        // the last CTA pays the system-scope fence.  A system fence in EVERY CTA cost 11 % of a 256 MiB
        // two-destination copy (2368 CTAs, profiles/README.md); cumulativity makes the single one sufficient.
        unsigned t;
        asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(t) : "l"(ticket) : "memory");
        last = (t == gridDim.x - 1);
        if (last) atomicExch(ticket, 0u); // ready for the next launch
    }
    if (last) {
        asm volatile("fence.acq_rel.sys;" ::: "memory");
        if (signal) red_release_sys_add(signal, 1ull);
        if (done_flag) st_release_sys(done_flag, done_value);
    }
}

} // namespace ocmk
