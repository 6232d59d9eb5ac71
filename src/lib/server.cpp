// Host side of the resident copy server (H2, src/kernels/server.cu): descriptor ring, posting, completion wait and
// the launch / re-launch hand-shake.  Used by BLOCKING small copies only (ocm_copy, ocm_copy_onesided,
// ocm_copy_in/out on the library stream): the caller waits for the bytes anyway, so the only thing that changes
// is that no kernel is launched and no stream is synchronised for them.
#include "ctx.hpp"
#include "../common/log.hpp"
#include "../kernels/kernels.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#include <time.h>

namespace ocm {

static uint64_t srv_now_ns() {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (uint64_t)ts.tv_sec * 1000000000ull + (uint64_t)ts.tv_nsec;
}

static uint64_t env_u64(const char *k, uint64_t dflt) {
    const char *e = getenv(k);
    return (e && *e) ? strtoull(e, nullptr, 0) : dflt;
}

void server_init() {
    Ctx &c = ctx();
    CopyServer &s = c.server;
    s.enabled = false;
    if (c.device < 0 || !c.done_word) return;
    if (const char *e = getenv("OCM_SERVER"))
        if (!strcmp(e, "0") || !strcmp(e, "off")) return;
    s.max_bytes = env_u64("OCM_SERVER_MAX", 128 << 10);
    s.idle_ns = env_u64("OCM_SERVER_IDLE_US", 100) * 1000;
    s.max_ns = env_u64("OCM_SERVER_LIFE_US", 20000) * 1000;
    if (!s.max_bytes) return;
    DeviceGuard dg(c.device);
    void *p = nullptr;
    const size_t bytes = sizeof(ocmk::ServerDesc) * kServerRing + sizeof(ocmk::ServerCtl);
    if (cudaHostAlloc(&p, bytes, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        return;
    }
    memset(p, 0, bytes);
    cudaStream_t st = nullptr;
    if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) {
        cudaGetLastError();
        cudaFreeHost(p);
        return;
    }
    s.ring = static_cast<ocmk::ServerDesc *>(p);
    s.ctl = reinterpret_cast<ocmk::ServerCtl *>(s.ring + kServerRing);
    s.stream = st;
    s.next_seq = 0;
    s.enabled = true;
}

void server_shutdown() {
    Ctx &c = ctx();
    CopyServer &s = c.server;
    if (!s.ring) return;
    s.enabled = false;
    DeviceGuard dg(c.device);
    // a resident kernel leaves by itself once idle (at most idle_ns from now)
    if (s.stream) {
        cudaStreamSynchronize((cudaStream_t)s.stream);
        cudaStreamDestroy((cudaStream_t)s.stream);
    }
    if (getenv("OCM_SERVER_TRACE") && s.ctl->t_ops) {
        const ocmk::ServerCtl &k = *s.ctl;
        fprintf(stderr, "{\"kind\": \"copy_server_trace\", \"ops\": %llu, \"kernel_starts\": %llu, \"copy_us_avg\": %.2f, "
                        "\"fence_and_done_us_avg\": %.2f, \"ring_polls\": %llu, \"us_per_poll\": %.2f}\n",
                (unsigned long long)k.t_ops, (unsigned long long)s.n_launches.load(), k.t_copy_ns * 1e-3 / k.t_ops,
                k.t_fence_ns * 1e-3 / k.t_ops, (unsigned long long)k.t_polls, k.t_polls ? k.t_poll_ns * 1e-3 / k.t_polls : 0.0);
    }
    cudaFreeHost(s.ring);
    s.ring = nullptr;
    s.ctl = nullptr;
    s.stream = nullptr;
}

bool server_eligible(uint64_t bytes) {
    Ctx &c = ctx();
    return c.server.enabled && bytes <= c.server.max_bytes && !c.lib_async_pending();
}

bool server_post(const void *src, void *dst, void *dst2, uint64_t bytes, uint64_t *signal, uint64_t *seq_out) {
    Ctx &c = ctx();
    CopyServer &s = c.server;
    std::lock_guard<std::mutex> g(s.mu);
    if (!s.enabled) return false;
    const uint64_t seq = s.next_seq;
    volatile uint64_t *done = &s.ctl->done;
    // ring full: wait for the oldest descriptor (bounded: every descriptor is at most max_bytes of work)
    for (uint64_t t0 = 0; seq - *done >= kServerRing;) {
        if (!t0) t0 = srv_now_ns();
        if (srv_now_ns() - t0 > 2000000000ull) return false;
        __builtin_ia32_pause();
    }
    ocmk::ServerDesc &d = s.ring[seq % kServerRing];
    d.src = src;
    d.dst = dst;
    d.dst2 = dst2;
    d.bytes = bytes;
    d.signal = signal;
    __atomic_store_n(&d.seq, seq + 1, __ATOMIC_RELEASE);
    s.next_seq = seq + 1;
    // Dekker hand-shake with a kernel that may be leaving right now (server.cu): our descriptor is published, now
    // look at its state.  RUNNING: it will find the descriptor.  EXITING: wait until it has decided.
    std::atomic_thread_fence(std::memory_order_seq_cst);
    volatile uint32_t *state = &s.ctl->state;
    uint32_t st = *state;
    for (uint64_t t0 = srv_now_ns(); st == ocmk::SERVER_EXITING; st = *state) {
        if (srv_now_ns() - t0 > 2000000000ull) break;
        __builtin_ia32_pause();
    }
    if (st == ocmk::SERVER_EXITED) {
        std::atomic_thread_fence(std::memory_order_acquire);
        const uint64_t start = *done;          // a leaving kernel has completed everything before `start`
        __atomic_store_n(&s.ctl->state, (uint32_t)ocmk::SERVER_RUNNING, __ATOMIC_SEQ_CST);
        DeviceGuard dg(c.device);
        cudaError_t e = ocmk::launch_copy_server(s.ring, kServerRing, s.ctl, start, s.idle_ns, s.max_ns, (cudaStream_t)s.stream);
        if (e != cudaSuccess) {
            LOGE("copy server launch: %s", cudaGetErrorString(e));
            cudaGetLastError();
            __atomic_store_n(&s.ctl->state, (uint32_t)ocmk::SERVER_EXITED, __ATOMIC_SEQ_CST);
            // take the descriptor back: the caller falls back to an ordinary launch
            __atomic_store_n(&d.seq, (uint64_t)0, __ATOMIC_RELEASE);
            s.next_seq = seq;
            s.enabled = false;
            return false;
        }
        s.n_launches.fetch_add(1, std::memory_order_relaxed);
    }
    s.n_ops.fetch_add(1, std::memory_order_relaxed);
    *seq_out = seq;
    return true;
}

int server_wait(uint64_t seq) {
    Ctx &c = ctx();
    CopyServer &s = c.server;
    if (!s.ctl) return OCM_ECUDA;
    volatile uint64_t *done = &s.ctl->done;
    uint64_t t0 = 0;
    for (unsigned spin = 0; *done < seq + 1; ++spin) {
        __builtin_ia32_pause();
        if ((spin & 1023) != 1023) continue;
        if (!t0) t0 = srv_now_ns();
        if (srv_now_ns() - t0 > 5000000000ull) {
            DeviceGuard dg(c.device);
            cudaError_t e = cudaStreamQuery((cudaStream_t)s.stream);
            LOGE("copy server: descriptor %llu not completed after 5 s (stream: %s)", (unsigned long long)seq,
                 cudaGetErrorString(e));
            cudaGetLastError();
            return OCM_ECUDA;
        }
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    return OCM_OK;
}

} // namespace ocm

extern "C" int ocm_server_stats(uint64_t *ops, uint64_t *launches) {
    ocm::Ctx &c = ocm::ctx();
    if (ops) *ops = c.server.n_ops.load();
    if (launches) *launches = c.server.n_launches.load();
    return c.server.enabled ? 1 : 0;
}
