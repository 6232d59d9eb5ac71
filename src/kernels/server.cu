// H2 - resident copy server.
//
// The reference's small-message latency is structural: every blocking ocm_copy_onesided posts one work request
// and blocks on a completion-channel event (reference src/rdma.c:254-302), and its daemon polls its mailbox
// every 500 us (src/main.c:105-129).  On a GPU the structural cost of a small one-sided copy is the kernel launch
// (~4-5 us of host + device work for 64 bytes of payload).  This kernel removes it: one CTA stays resident
// and polls a descriptor ring in pinned host memory; the host posts a 64-byte descriptor with a release store and
// spins on a completion counter the kernel advances - no launch, no stream synchronisation on the path.
//
//  * descriptor i (0-based) is valid once ring[i % R].seq == i + 1 (system-scope acquire by thread 0);
//  * the whole CTA moves the bytes (256-bit LDG/STG when the pointers are congruent mod 32, bytes otherwise),
//    source and destinations being generic addresses as in copy.cu: local HBM, peer HBM over NVLink, pinned host;
//  * completion: CTA barrier, fence.acq_rel.sys, optional +1 on the destination's signal word, then
//    ctl->done = i + 1 (release, system scope) - descriptors complete in order;
//  * the kernel is resident on demand, not forever: after `idle_ns` without work (or `max_ns` of life) it leaves,
//    because a kernel that never ends would block cudaDeviceSynchronize / cudaFree in the application for ever.
//    Leaving races with a host thread posting work, so both sides run a Dekker hand-shake over ctl->state:
//        kernel: state = EXITING; fence.sc.sys; look at the ring once more -> found ? state = RUNNING : state = EXITED
//        host  : publish descriptor; seq_cst fence; read state -> EXITING ? wait : EXITED ? relaunch : nothing
//    at least one side sees the other's store, so no descriptor is ever left without a kernel.
#include "kernels.h"
#include "ptx.cuh"

namespace ocmk {

__device__ __forceinline__ uint64_t globaltimer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void fence_sc_sys() { asm volatile("fence.sc.sys;" ::: "memory"); }
// Polling load: relaxed at system scope (never served from L1).  The acquire that orders the descriptor's fields
// after its sequence word is a single fence once the poll has hit, not part of every poll: measured 2.8 us per
// ld.acquire.sys of pinned host memory on the B200 VMs, most of it the fence.
__device__ __forceinline__ uint64_t ld_relaxed_sys(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}


struct V32s {
    uint32_t w[8];
};
__device__ __forceinline__ V32s srv_ld32(const uint8_t *p) {
    V32s v;
    asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3]), "=r"(v.w[4]), "=r"(v.w[5]), "=r"(v.w[6]), "=r"(v.w[7])
                 : "l"(p));
    return v;
}
__device__ __forceinline__ void srv_st32(uint8_t *p, const V32s &v) {
    asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v.w[0]), "r"(v.w[1]),
                 "r"(v.w[2]), "r"(v.w[3]), "r"(v.w[4]), "r"(v.w[5]), "r"(v.w[6]), "r"(v.w[7])
                 : "memory");
}

// One CTA copies `bytes` from src to dst (and dst2).
__device__ void cta_copy(uint8_t *dst, uint8_t *dst2, const uint8_t *src, uint64_t bytes) {
    const unsigned tid = threadIdx.x, nt = blockDim.x;
    const uintptr_t r = (uintptr_t)dst & 31;
    const bool wide = ((uintptr_t)src & 31) == r && (!dst2 || ((uintptr_t)dst2 & 31) == r);
    if (!wide) {
        for (uint64_t i = tid; i < bytes; i += nt) {
            const uint8_t v = src[i];
            dst[i] = v;
            if (dst2) dst2[i] = v;
        }
        return;
    }
    uint64_t head = (32 - r) & 31;
    if (head > bytes) head = bytes;
    const uint64_t n = (bytes - head) / 32;
    const uint8_t *s = src + head;
    uint8_t *d = dst + head, *d2 = dst2 ? dst2 + head : nullptr;
    uint64_t i = tid;
    for (; i + nt < n; i += 2 * nt) {
        V32s v0 = srv_ld32(s + i * 32), v1 = srv_ld32(s + (i + nt) * 32);
        srv_st32(d + i * 32, v0);
        srv_st32(d + (i + nt) * 32, v1);
        if (d2) {
            srv_st32(d2 + i * 32, v0);
            srv_st32(d2 + (i + nt) * 32, v1);
        }
    }
    for (; i < n; i += nt) {
        V32s v = srv_ld32(s + i * 32);
        srv_st32(d + i * 32, v);
        if (d2) srv_st32(d2 + i * 32, v);
    }
    const uint64_t tail0 = head + n * 32;
    for (uint64_t o = tid; o < head + (bytes - tail0); o += nt) {
        const uint64_t k = o < head ? o : tail0 + (o - head);
        const uint8_t v = src[k];
        dst[k] = v;
        if (dst2) dst2[k] = v;
    }
}

__global__ void __launch_bounds__(512, 1) ocm_copy_server_kernel(ServerDesc *ring, uint32_t ring_size, ServerCtl *ctl,
                                                                 uint64_t start_seq, uint64_t idle_ns, uint64_t max_ns) {
    __shared__ ServerDesc cur;
    __shared__ int action;      // 1: descriptor in `cur`, 2: leave
    uint64_t seq = start_seq;
    uint64_t t_start = 0, t_last = 0;
    uint64_t acc_ops = 0, acc_copy = 0, acc_fence = 0, acc_polls = 0, acc_poll_ns = 0, t_seen = 0;   // thread 0 only
    if (threadIdx.x == 0) t_start = t_last = globaltimer_ns();
    for (;;) {
        if (threadIdx.x == 0) {
            ServerDesc *slot = &ring[seq % ring_size];
            action = 0;
            for (;;) {
                if (ld_relaxed_sys(&slot->seq) == seq + 1) {
                    action = 1;
                    break;
                }
                const uint64_t now = globaltimer_ns();
                ++acc_polls;
                if (now - t_last > idle_ns || now - t_start > max_ns) {
                    st_release_sys_u32(&ctl->state, SERVER_EXITING);
                    fence_sc_sys();
                    if (ld_acquire_sys(&slot->seq) == seq + 1) {     // work arrived while we were leaving: stay
                        st_release_sys_u32(&ctl->state, SERVER_RUNNING);
                        t_start = now;
                        action = 1;
                        break;
                    }
                    acc_poll_ns += now - t_last;
                    ctl->t_ops += acc_ops;
                    ctl->t_copy_ns += acc_copy;
                    ctl->t_fence_ns += acc_fence;
                    ctl->t_polls += acc_polls;
                    ctl->t_poll_ns += acc_poll_ns;
                    st_release_sys_u32(&ctl->state, SERVER_EXITED);
                    action = 2;
                    break;
                }
            }
            if (action == 1) {
                asm volatile("fence.acq_rel.sys;" ::: "memory");   // the descriptor's fields (and the data they name) after its seq
                t_seen = globaltimer_ns();
                acc_poll_ns += t_seen - t_last;
                cur.src = slot->src;
                cur.dst = slot->dst;
                cur.dst2 = slot->dst2;
                cur.bytes = slot->bytes;
                cur.signal = slot->signal;
            }
        }
        __syncthreads();
        if (action == 2) return;
        cta_copy(static_cast<uint8_t *>(cur.dst), static_cast<uint8_t *>(cur.dst2), static_cast<const uint8_t *>(cur.src),
                 cur.bytes);
        __syncthreads();
        if (threadIdx.x == 0) {
            const uint64_t t_copied = globaltimer_ns();
            // a system-scope release covers the CTA's stores (ordered before it by the barrier).  The host side counts
            // the completion signal itself for server copies (lib/copy.cpp), so the red is the rare case; it needs its
            // own release so that the counter can never be seen before the signal.
            if (cur.signal) red_release_sys_add(cur.signal, 1ull);
            st_release_sys(&ctl->done, seq + 1);
            t_last = globaltimer_ns();
            ++acc_ops;
            acc_copy += t_copied - t_seen;
            acc_fence += t_last - t_copied;
        }
        ++seq;
        __syncthreads();
    }
}

cudaError_t launch_copy_server(ServerDesc *ring, uint32_t ring_size, ServerCtl *ctl, uint64_t start_seq, uint64_t idle_ns,
                               uint64_t max_ns, cudaStream_t stream) {
    ocm_copy_server_kernel<<<1, 512, 0, stream>>>(ring, ring_size, ctl, start_seq, idle_ns, max_ns);
    extra_launches().fetch_add(1, std::memory_order_relaxed);
    return cudaGetLastError();
}

} // namespace ocmk
