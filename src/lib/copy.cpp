// Copy engine of libocm: ocm_copy / ocm_copy_onesided / ocm_copy_in / ocm_copy_out
// and their stream-ordered and strided variants.
//
// The reference implements a sparse kind matrix with serial blocking hops
// (memcpy -> ib_write -> ib_poll, cudaMemcpy -> ib_write, ...; reference
// src/lib.c:502-723) and BUG()s on remote<->remote, silently skips
// GPU->GPU.  Here every allocation side is a generic address in this process,
// so any (source side, destination side) combination is one primitive:
//   GPU<->GPU (same or peer device)     our sm_100a copy kernel over NVLink
//   window + far side in one pass        the same kernel with two destinations
//   pinned host -> {GPU window, far GPU} zero-copy read fused into that kernel
//   plain host<->GPU                     cudaMemcpyAsync on the stream (host tier only)
//   host<->host                          memcpy
// Offsets: src_offset/dest_offset drive the first hop, src_offset_2/dest_offset_2
// the second; when the second hop starts where the first one landed
// the two hops are fused.
#include "ctx.hpp"
#include "trace.hpp"
#include "../common/log.hpp"
#include "../kernels/kernels.h"
#include <algorithm>
#include <cstring>
#include <cuda_runtime.h>

using namespace ocm;

namespace ocm {

static bool kernel_addressable(const Side &s) {
    if (s.tier == TIER_GPU) return true;
    return s.backing == SIDE_POOL; // pool host memory is pinned + mapped in this process
}

// Blocking calls on the library stream: the kernel itself tells a host-visible word when its last byte is
// visible, and the caller spins on that word instead of paying kernel tear-down + cudaStreamSynchronize
// (~4 us of a ~13 us small put).  Only used when the whole call was exactly one kernel launch.
struct SyncNotify {
    bool armed = false;
    uint64_t value = 0;
    int gpu_ops = 0;        // kernels / async memcpys enqueued by the current call
    int notified = 0;       // ... of which carry the completion word
    int server_ops = 0;     // copies handed to the resident copy server (H2) by the current call
    bool server_waited = false;
    uint64_t server_seq = 0;
    std::atomic<uint64_t> *server_sig = nullptr;   // completion counter the HOST bumps once the server reports done
    uint64_t async_snap = 0;    // Ctx::lib_async_issued when the call began
    ocmk::CopyNotify cn{nullptr, 0};
};
static thread_local SyncNotify t_sn;

// Only kernels that already publish a completion signal carry the word: for those the system fence is paid
// anyway (measured: a get without signal gets 2 us SLOWER with the extra fence + host store, a put 1 us faster).
// Wait for the copy this call handed to the resident server, then publish its completion signal.  The kernel's
// done word is released at system scope after its bytes, so counting on the CPU side (the allocation's CPU word:
// one writing agent per word) keeps the order "bytes, then signal" and saves the kernel a second PCIe round trip.
static int server_drain() {
    if (!t_sn.server_ops || t_sn.server_waited) return OCM_OK;
    t_sn.server_waited = true;
    const int e = server_wait(t_sn.server_seq);
    if (!e && t_sn.server_sig) t_sn.server_sig->fetch_add(1, std::memory_order_release);
    return e;
}

static const ocmk::CopyNotify *note_kernel(bool has_signal) {
    ++t_sn.gpu_ops;
    if (!t_sn.armed || !has_signal) return nullptr;
    ++t_sn.notified;
    return &t_sn.cn;
}
static void note_other_gpu_op(int n = 1) { t_sn.gpu_ops += n; }

// Kernels enqueued by the current side_copy call on this thread; added to the process-wide counter when it returns.
static thread_local unsigned long long launches_ = 0;
struct LaunchFlush {
    ~LaunchFlush() {
        if (launches_) ctx().kernel_launches.fetch_add(launches_, std::memory_order_relaxed);
        launches_ = 0;
    }
};

static int cuda_err(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return OCM_OK;
    LOGE("%s: %s", what, cudaGetErrorString(e));
    cudaGetLastError();
    return OCM_ECUDA;
}

// ---- pull-mode puts ---------------------------------------------------------------------------
// SM-issued stores over NVLink top out near 714 GB/s on B200 while SM-issued loads reach ~780 GB/s
// (profiles/README.md).  For a large put into a peer GPU's pool this process therefore launches the copy
// kernel ON THE OWNER GPU (it can: all GPUs of the box are visible), which pulls the bytes with loads.
// Stream order is kept with two events: user stream -> owner-GPU stream -> user stream.
static std::mutex g_peer_mu;

static PeerExec *get_peer_exec(int dev) {
    Ctx &c = ctx();
    if (dev < 0 || dev >= kMaxGpus || dev == c.device) return nullptr;
    std::lock_guard<std::mutex> g(g_peer_mu);
    PeerExec &p = c.peer_exec[dev];
    if (!p.tried) {
        p.tried = true;
        p.dev = dev;
        bool ok = true;
        {
            DeviceGuard dg(dev);
            cudaStream_t s = nullptr;
            ok = cudaFree(0) == cudaSuccess && cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) == cudaSuccess;
            p.stream = s;
            for (int i = 0; ok && i < 16; ++i) {
                cudaEvent_t e;
                ok = cudaEventCreateWithFlags(&e, cudaEventDisableTiming) == cudaSuccess;
                p.ev_out[i] = e;
            }
            if (ok) {
                ok = cudaMalloc((void **)&p.tickets, 16 * sizeof(unsigned)) == cudaSuccess &&
                     cudaMemset(p.tickets, 0, 16 * sizeof(unsigned)) == cudaSuccess;
            }
        }
        {
            DeviceGuard dg(c.device);
            for (int i = 0; ok && i < 16; ++i) {
                cudaEvent_t e;
                ok = cudaEventCreateWithFlags(&e, cudaEventDisableTiming) == cudaSuccess;
                p.ev_in[i] = e;
            }
        }
        if (!ok) {
            cudaGetLastError();
            LOGW("cannot drive GPU %d from this process: large puts stay in push mode", dev);
        }
        p.ok = ok;
    }
    return p.ok ? &p : nullptr;
}

static bool ensure_access(RankView *v, int dev) {
    Ctx &c = ctx();
    if (!v || !v->gpu.ptr) return false;
    if (dev == c.device || (v->access_mask & (1ull << dev))) return true;
    std::lock_guard<std::mutex> g(g_peer_mu);
    std::string err;
    if (!gpu::pool_grant_access(&v->gpu, dev, &err)) {
        LOGW("%s", err.c_str());
        return false;
    }
    v->access_mask |= 1ull << dev;
    return true;
}

// Returns OCM_EUNSUPPORTED when the transfer does not qualify (caller falls back to push mode).
static int try_pull_put(const Side &dst, uint64_t dst_off, const Side *dst2, uint64_t dst2_off, const Side &src,
                        uint64_t src_off, uint64_t bytes, uint64_t *signal, cudaStream_t st) {
    Ctx &c = ctx();
    if (!c.pull_min || bytes < c.pull_min) return OCM_EUNSUPPORTED;
    if (src.tier != TIER_GPU || dst.tier != TIER_GPU || src.backing != SIDE_POOL || dst.backing != SIDE_POOL)
        return OCM_EUNSUPPORTED;
    if (dst2 && dst2->tier != TIER_GPU) return OCM_EUNSUPPORTED;
    RankView *dv = get_view(dst.rank, true), *sv = get_view(src.rank, true);
    if (!dv || !sv || !dv->gpu.ptr || !sv->gpu.ptr) return OCM_EUNSUPPORTED;
    const int qdev = dv->hdr->gpu_index;
    if (qdev < 0 || qdev == c.device || sv->hdr->gpu_index != c.device) return OCM_EUNSUPPORTED;
    if (dv->gpu.mode != GPU_SHARE_VMM_FD || sv->gpu.mode != GPU_SHARE_VMM_FD) return OCM_EUNSUPPORTED;
    PeerExec *pe = get_peer_exec(qdev);
    if (!pe || !ensure_access(dv, qdev) || !ensure_access(sv, qdev)) return OCM_EUNSUPPORTED;

    unsigned slot;
    {
        std::lock_guard<std::mutex> g(g_peer_mu);
        slot = pe->cursor++ % 16;
    }
    cudaError_t e;
    {
        DeviceGuard dg(c.device);
        e = cudaEventRecord((cudaEvent_t)pe->ev_in[slot], st);
    }
    if (e == cudaSuccess) {
        DeviceGuard dg(qdev);
        e = cudaStreamWaitEvent((cudaStream_t)pe->stream, (cudaEvent_t)pe->ev_in[slot], 0);
        if (e == cudaSuccess)
            e = ocmk::launch_copy((char *)dst.ptr + dst_off, nullptr, (const char *)src.ptr + src_off, bytes, signal,
                                  pe->tickets + slot, (cudaStream_t)pe->stream, nullptr, &launches_);
        if (e == cudaSuccess) e = cudaEventRecord((cudaEvent_t)pe->ev_out[slot], (cudaStream_t)pe->stream);
    }
    if (e == cudaSuccess) {
        DeviceGuard dg(c.device);
        if (dst2)   // the window is local: its copy runs here, concurrently with the pull
            e = ocmk::launch_copy((char *)dst2->ptr + dst2_off, nullptr, (const char *)src.ptr + src_off, bytes, nullptr,
                                  next_ticket(), st, nullptr, &launches_);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(st, (cudaEvent_t)pe->ev_out[slot], 0);
    }
    if (e != cudaSuccess) return cuda_err(e, "pull-mode put");
    c.n_pulled.fetch_add(1, std::memory_order_relaxed);
    return OCM_OK;
}

static int single_copy(const Side &dst, uint64_t dst_off, const Side &src, uint64_t src_off, uint64_t bytes,
                       cudaStream_t st, bool *used_stream, uint64_t *signal, bool *signalled) {
    Ctx &c = ctx();
    char *d = (char *)dst.ptr + dst_off;
    const char *s = (const char *)src.ptr + src_off;
    if (dst.tier == TIER_HOST && src.tier == TIER_HOST) {
        // a previous hop on the stream may still be producing the source
        if (*used_stream) {
            cudaStreamSynchronize(st);
        }
        memcpy(d, s, bytes);
        return OCM_OK;
    }
    if (c.device < 0) return OCM_ENOGPU;
    *used_stream = true;
    if (dst.tier == TIER_GPU && src.tier == TIER_GPU) {
        *signalled = signal != nullptr;
        return cuda_err(ocmk::launch_copy(d, nullptr, s, bytes, signal, next_ticket(), st, nullptr, &launches_,
                                          note_kernel(signal != nullptr)),
                        "copy kernel");
    }
    // host tier <-> GPU: copy engine (H1).  A block of a host pool may lie in several separately pinned pieces of the
    // pool (ctx.hpp host_reg_chunk) and one cudaMemcpyAsync must stay inside one registration: cut at the boundaries.
    note_other_gpu_op();
    const Side &h = dst.tier == TIER_HOST ? dst : src;
    const bool pool_host = (h.backing == SIDE_POOL || h.backing == SIDE_ATTACHED) &&
                           !(h.hdr && h.hdr->host_share_mode == GPU_SHARE_VMM_FD);   // a VMM pool is one mapping
    const uint64_t chunk = host_reg_chunk(pool_host && h.hdr ? h.hdr->pool[TIER_HOST].bytes : 0);
    const cudaMemcpyKind kind = dst.tier == TIER_GPU ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
    uint64_t pos = h.offset + (dst.tier == TIER_HOST ? dst_off : src_off);      // position inside the pool
    for (uint64_t done = 0; done < bytes;) {
        uint64_t n = bytes - done;
        if (pool_host) n = std::min<uint64_t>(n, chunk - pos % chunk);
        int err = cuda_err(cudaMemcpyAsync(d + done, s + done, n, kind, st), "cudaMemcpyAsync");
        if (err) return err;
        done += n;
        pos += n;
    }
    return OCM_OK;
}

// [off, off+n) lies inside the side (written so that huge offsets cannot wrap around)
static inline bool fits(const Side &s, uint64_t off, uint64_t n) { return n <= s.bytes && off <= s.bytes - n; }

int side_copy(const Side &dst, uint64_t dst_off, const Side *dst2, uint64_t dst2_off, const Side &src, uint64_t src_off,
              uint64_t bytes, const Side *sig, void *stream, bool *used_stream) {
    Ctx &c = ctx();
    LaunchFlush flush_launches;
    // only the first hop of a blocking call may go through the copy server; anything after it waits for it first
    if (int e = server_drain()) return e;
    if (!dst.ptr || !src.ptr || (dst2 && !dst2->ptr)) return OCM_EINVAL;
    if (!fits(src, src_off, bytes) || !fits(dst, dst_off, bytes) || (dst2 && !fits(*dst2, dst2_off, bytes)))
        return OCM_ERANGE;
    c.n_copies.fetch_add(1, std::memory_order_relaxed);
    c.copy_bytes.fetch_add(bytes, std::memory_order_relaxed);
    // completion counters of the side named as destination: one word for our GPU, one for CPUs
    uint64_t *signal = sig ? sig->gpu_signal() : nullptr;
    std::atomic<uint64_t> *signal_cpu = sig ? sig->sig_cpu : nullptr;
    cudaStream_t st = (cudaStream_t)(stream ? stream : c.stream);

    const bool all_host = src.tier == TIER_HOST && dst.tier == TIER_HOST && (!dst2 || dst2->tier == TIER_HOST);
    if (all_host) {
        if (*used_stream) {
            DeviceGuard dg(c.device);
            cudaStreamSynchronize(st);
        }
        const char *s = (const char *)src.ptr + src_off;
        if (!dst2) {
            memcpy((char *)dst.ptr + dst_off, s, bytes);
        } else {
            // two destinations: feed both from one pass over the source, a cache-sized piece at a time, so the
            // second copy reads from L2 instead of DRAM (the reference does two full passes, src/lib.c:530-531)
            constexpr uint64_t kPiece = 256 << 10;
            for (uint64_t o = 0; o < bytes; o += kPiece) {
                const uint64_t n = std::min<uint64_t>(kPiece, bytes - o);
                memcpy((char *)dst.ptr + dst_off + o, s + o, n);
                memcpy((char *)dst2->ptr + dst2_off + o, s + o, n);
            }
        }
        if (signal_cpu) signal_cpu->fetch_add(1, std::memory_order_release);
        return OCM_OK;
    }
    if (c.device < 0) return OCM_ENOGPU;
    DeviceGuard dg(c.device);
    bool signalled = false;
    int err = try_pull_put(dst, dst_off, dst2, dst2_off, src, src_off, bytes, signal, st);
    if (err == OCM_OK) {
        *used_stream = true;
        note_other_gpu_op(2);
        return OCM_OK;
    }
    if (err != OCM_EUNSUPPORTED) return err;
    err = OCM_OK;

    // fused path: one kernel reads the source once and feeds every destination
    const bool gpu_dst = dst.tier == TIER_GPU || (dst2 && dst2->tier == TIER_GPU);
    const bool fused = kernel_addressable(src) && kernel_addressable(dst) && (!dst2 || kernel_addressable(*dst2)) &&
                       ((src.tier == TIER_GPU && dst.tier == TIER_GPU && (!dst2 || dst2->tier == TIER_GPU)) ||
                        (dst2 && gpu_dst));
    uint64_t srv_seq = 0;
    if (fused && t_sn.armed && !stream && !*used_stream && t_sn.gpu_ops == 0 && t_sn.server_ops == 0 && server_eligible(bytes) &&
        server_post((const char *)src.ptr + src_off, (char *)dst.ptr + dst_off, dst2 ? (char *)dst2->ptr + dst2_off : nullptr,
                    bytes, signal_cpu ? nullptr : signal, &srv_seq)) {
        // H2: a blocking small copy - the resident kernel moves it, nothing is launched (finish_sync waits for it)
        t_sn.server_ops = 1;
        t_sn.server_seq = srv_seq;
        t_sn.server_sig = signal ? signal_cpu : nullptr;
        return OCM_OK;
    }
    if (fused) {
        *used_stream = true;
        err = cuda_err(ocmk::launch_copy((char *)dst.ptr + dst_off, dst2 ? (char *)dst2->ptr + dst2_off : nullptr,
                                         (const char *)src.ptr + src_off, bytes, signal, next_ticket(), st, nullptr,
                                         &launches_, note_kernel(signal != nullptr)),
                       "copy kernel");
        signalled = signal != nullptr;
    } else {
        err = single_copy(dst, dst_off, src, src_off, bytes, st, used_stream, dst2 ? nullptr : signal, &signalled);
        if (!err && dst2)
            err = single_copy(*dst2, dst2_off, dst, dst_off, bytes, st, used_stream, signal, &signalled);
    }
    if (!err && signal && !signalled) {
        if (*used_stream) {
            err = cuda_err(ocmk::launch_signal(signal, st), "signal kernel");
            note_other_gpu_op();
            ++c.kernel_launches;
        } else if (signal_cpu) {
            signal_cpu->fetch_add(1, std::memory_order_release);
        }
    }
    return err;
}

static bool valid(ocm_alloc_t a) { return a && a->magic == kAllocMagic; }

static bool stale(ocm_alloc_t a) { return side_stale(a->remote); }

// The far side of a pair (single slab kinds only).
static const Side *far_side(ocm_alloc_t a) { return (a->paired && a->remote.valid()) ? &a->remote : nullptr; }

static void mark_pending(ocm_alloc_t a, void *stream, bool used) {
    if (used && !stream) a->lib_stream_pending.store(true, std::memory_order_relaxed);
}

// Multi-slab kinds.  STRIPED: the far side is one logical range of N * stripe_bytes, block-distributed over
// the ranks (the "allocations composed of partitioned distributed allocations" the reference only plans,
// ref src/lib.c:42-45).  REPLICATED: every rank holds a full copy; a put reaches all of them in one pass
// (cluster-multicast kernel K6), a get reads the nearest replica.
// True only when the bytes are KNOWN to live in the HBM of the GPU this context launches on: a block of a pool whose
// table names this GPU, or the library's own cudaMalloc on it, and the driver agrees about the pointer.  The
// replicated put shares its source tiles across a cluster with TMA multicast only then; TMA multicast loads here
// read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU.
static bool source_on_this_gpu(const Side &s, uint64_t off) {
    Ctx &c = ctx();
    if (c.device < 0 || s.tier != TIER_GPU) return false;
    const bool by_table = (s.backing == SIDE_POOL && s.hdr && s.hdr->gpu_index == c.device && s.rank == c.rank) ||
                          s.backing == SIDE_CUDAMALLOC;
    if (!by_table) return false;
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, (const char *)s.ptr + off) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeDevice && at.device == c.device;
}

static int stripes_xfer(ocm_alloc_t a, bool to_far, const Side &local, uint64_t local_off, uint64_t far_off, uint64_t n,
                        void *stream, bool *used) {
    Ctx &c = ctx();
    const size_t N = a->stripes.size();
    if (!N || !local.ptr) return OCM_EINVAL;
    if (!fits(local, local_off, n)) return OCM_ERANGE;
    if (a->kind == OCM_REPLICATED_GPU) {
        if (n > a->stripe_bytes || far_off > a->stripe_bytes - n) return OCM_ERANGE;
        if (!to_far) {
            const Side &near = a->stripes[(size_t)c.rank < N ? (size_t)c.rank : 0];
            return side_copy(local, local_off, nullptr, 0, near, far_off, n, nullptr, stream, used);
        }
        const bool aligned = (((uintptr_t)local.ptr + local_off) | far_off) % 16 == 0;
        if (a->mc && aligned && n % 16 == 0 && local.tier == TIER_GPU && c.device >= 0) {
            // NVLS: one multimem.st stream, the switch replicates it into every GPU's slice
            DeviceGuard dg(c.device);
            // ... and the allocation's completion word on every GPU (multimem.red on the word's multicast address)
            cudaError_t e = ocmk::launch_put_multimem((char *)a->mc_ptr + far_off, (const char *)local.ptr + local_off, n,
                                                      a->mc_sig, next_ticket(), (cudaStream_t)(stream ? stream : c.stream));
            ++c.kernel_launches;
            note_other_gpu_op();
            c.n_copies.fetch_add(1, std::memory_order_relaxed);
            c.copy_bytes.fetch_add(n * N, std::memory_order_relaxed);
            *used = true;
            return cuda_err(e, "multimem put");
        }
        if (N <= 8 && aligned && local.tier == TIER_GPU && c.device >= 0) {
            void *dsts[8];
            uint64_t *sigs[8];
            for (size_t r = 0; r < N; ++r) {
                dsts[r] = (char *)a->stripes[r].ptr + far_off;
                sigs[r] = a->stripes[r].gpu_signal();
            }
            DeviceGuard dg(c.device);
            cudaError_t e = ocmk::launch_put_replicated(dsts, (int)N, (const char *)local.ptr + local_off, n, sigs,
                                                        next_ticket(), (cudaStream_t)(stream ? stream : c.stream),
                                                        source_on_this_gpu(local, local_off));
            ++c.kernel_launches;
            note_other_gpu_op();
            c.n_copies.fetch_add(1, std::memory_order_relaxed);
            c.copy_bytes.fetch_add(n * N, std::memory_order_relaxed);
            *used = true;
            if (e == cudaSuccess && a->mc && a->mc_sig) {
                e = ocmk::launch_signal_multimem(a->mc_sig, (cudaStream_t)(stream ? stream : c.stream));
                ++c.kernel_launches;
            }
            return cuda_err(e, "replicated put");
        }
        for (size_t r = 0; r < N; ++r) {
            int err = side_copy(a->stripes[r], far_off, nullptr, 0, local, local_off, n, &a->stripes[r], stream, used);
            if (err) return err;
        }
        if (a->mc && a->mc_sig && c.device >= 0) {
            // per-replica copies of a multicast-backed allocation: its completion word still moves once, everywhere
            DeviceGuard dg(c.device);
            cudaError_t e = ocmk::launch_signal_multimem(a->mc_sig, (cudaStream_t)(stream ? stream : c.stream));
            ++c.kernel_launches;
            note_other_gpu_op();
            *used = true;
            return cuda_err(e, "multicast signal");
        }
        return OCM_OK;
    }
    // striped
    if (n > a->stripe_bytes * N || far_off > a->stripe_bytes * N - n) return OCM_ERANGE;
    while (n) {
        const size_t r = (size_t)(far_off / a->stripe_bytes);
        const uint64_t in = far_off % a->stripe_bytes;
        const uint64_t piece = std::min<uint64_t>(n, a->stripe_bytes - in);
        int err = to_far ? side_copy(a->stripes[r], in, nullptr, 0, local, local_off, piece, &a->stripes[r], stream, used)
                         : side_copy(local, local_off, nullptr, 0, a->stripes[r], in, piece, nullptr, stream, used);
        if (err) return err;
        far_off += piece;
        local_off += piece;
        n -= piece;
    }
    return OCM_OK;
}

static int do_copy(ocm_alloc_t dst, ocm_alloc_t src, const struct ocm_params *p, void *stream, bool *used) {
    if (!valid(dst) || !valid(src) || !p) return OCM_EINVAL;
    if (stale(dst) || stale(src)) return OCM_ESTALE;
    if (!p->op_flag) std::swap(dst, src);   // "read": data flows the other way (src/lib.c:510-515)
    if (!dst->stripes.empty() && !src->stripes.empty()) return OCM_EUNSUPPORTED;
    if (!dst->stripes.empty() || !src->stripes.empty()) {
        // one hop between a local side and the multi-slab far side: src_offset/dest_offset index source/destination
        ocm_alloc_t multi = dst->stripes.empty() ? src : dst, other = dst->stripes.empty() ? dst : src;
        if (far_side(other) || !other->local.valid()) return OCM_EUNSUPPORTED;
        const bool to_far = multi == dst;
        int err = stripes_xfer(multi, to_far, other->local, to_far ? p->src_offset : p->dest_offset,
                               to_far ? p->dest_offset : p->src_offset, p->bytes, stream, used);
        mark_pending(dst, stream, *used);
        mark_pending(src, stream, *used);
        return err;
    }
    const Side *dfar = far_side(dst), *sfar = far_side(src);
    const uint64_t n = p->bytes;
    int err;
    if (!dfar && !sfar) {
        // local -> local (host/GPU in any combination; GPU->GPU is a no-op in the reference)
        err = side_copy(dst->local, p->dest_offset, nullptr, 0, src->local, p->src_offset, n, &dst->local, stream, used);
    } else if (dfar && !sfar) {
        // local -> pair: hop 1 into the window, hop 2 window -> far side
        if (!dst->local.valid()) {
            // window-less pair: a single hop, src_offset/dest_offset index source/destination
            err = side_copy(*dfar, p->dest_offset, nullptr, 0, src->local, p->src_offset, n, dfar, stream, used);
        } else if (p->src_offset_2 == p->dest_offset) {
            err = side_copy(*dfar, p->dest_offset_2, &dst->local, p->dest_offset, src->local, p->src_offset, n,
                            dfar, stream, used);
        } else if (!fits(dst->local, p->src_offset_2, n) || !fits(*dfar, p->dest_offset_2, n)) {
            err = OCM_ERANGE;   // checked up front so that a bad second hop does not leave the first one done
        } else {
            err = side_copy(dst->local, p->dest_offset, nullptr, 0, src->local, p->src_offset, n, nullptr, stream, used);
            if (!err)
                err = side_copy(*dfar, p->dest_offset_2, nullptr, 0, dst->local, p->src_offset_2, n, dfar, stream, used);
        }
    } else if (!dfar && sfar) {
        // pair -> local: hop 1 far side -> window, hop 2 window -> destination
        if (!src->local.valid()) {
            err = side_copy(dst->local, p->dest_offset, nullptr, 0, *sfar, p->src_offset, n, &dst->local, stream, used);
        } else if (p->src_offset_2 == p->src_offset) {
            err = side_copy(dst->local, p->dest_offset_2, &src->local, p->src_offset, *sfar, p->dest_offset, n,
                            &dst->local, stream, used);
        } else if (!fits(src->local, p->src_offset_2, n) || !fits(dst->local, p->dest_offset_2, n)) {
            err = OCM_ERANGE;
        } else {
            err = side_copy(src->local, p->src_offset, nullptr, 0, *sfar, p->dest_offset, n, nullptr, stream, used);
            if (!err)
                err = side_copy(dst->local, p->dest_offset_2, nullptr, 0, src->local, p->src_offset_2, n, &dst->local,
                                stream, used);
        }
    } else {
        // pair -> pair: third-party copy far side -> far side (BUG(1) in the reference, src/lib.c:591-594)
        err = side_copy(*dfar, p->dest_offset, nullptr, 0, *sfar, p->src_offset, n, dfar, stream, used);
    }
    mark_pending(dst, stream, *used);
    mark_pending(src, stream, *used);
    return err;
}

static int do_onesided(ocm_alloc_t a, const struct ocm_params *p, void *stream, bool *used) {
    if (!valid(a) || !p) return OCM_EINVAL;
    if (!a->stripes.empty()) {
        if (!a->local.valid() || p->bytes > a->local.bytes) return a->local.valid() ? OCM_ERANGE : OCM_EINVAL;
        int err = stripes_xfer(a, p->op_flag != 0, a->local, p->src_offset, p->dest_offset, p->bytes, stream, used);
        mark_pending(a, stream, *used);
        return err;
    }
    const Side *far = far_side(a);
    if (!far || !a->local.valid()) {
        LOGW("one-sided copy needs a paired allocation with a window");
        return OCM_EINVAL;
    }
    if (p->bytes > a->local.bytes) return OCM_ERANGE;   // as the reference (src/lib.c:679,701)
    int err;
    if (p->op_flag) err = side_copy(*far, p->dest_offset, nullptr, 0, a->local, p->src_offset, p->bytes, far, stream, used);
    else err = side_copy(a->local, p->src_offset, nullptr, 0, *far, p->dest_offset, p->bytes, nullptr, stream, used);
    mark_pending(a, stream, *used);
    return err;
}

// The calling thread's completion slot (Ctx::done_word): claimed on first use, returned when the thread exits.
struct DoneSlot {
    int slot = -1;
    uint64_t *epoch = nullptr;      // the done_word array the slot belongs to (a re-init after ocm_tini makes a new one)
    ~DoneSlot() { release(); }
    void release() {
        Ctx &c = ctx();
        if (slot >= 0 && epoch && epoch == c.done_word) c.done_slots_free.fetch_or(1ull << slot, std::memory_order_release);
        slot = -1;
        epoch = nullptr;
    }
    int get() {
        Ctx &c = ctx();
        if (slot >= 0 && epoch == c.done_word) return slot;
        slot = -1;
        uint64_t fr = c.done_slots_free.load(std::memory_order_acquire);
        while (fr) {
            const int s = __builtin_ctzll(fr);
            if (c.done_slots_free.compare_exchange_weak(fr, fr & ~(1ull << s), std::memory_order_acq_rel)) {
                slot = s;
                epoch = c.done_word;
                break;
            }
        }
        return slot;   // -1: more than kDoneSlots threads are blocking at once; this one synchronises its stream
    }
};
static thread_local DoneSlot t_done_slot;

// Arm the completion word for the blocking call that follows on this thread.
static void arm_sync_notify() {
    Ctx &c = ctx();
    t_sn = SyncNotify();
    t_sn.async_snap = c.lib_async_issued.load(std::memory_order_acquire);
    if (!c.done_word || c.device < 0) return;
    const int s = t_done_slot.get();
    if (s < 0) return;
    t_sn.armed = true;
    t_sn.value = ++c.done_seq[s];
    t_sn.cn.flag = c.done_word + (size_t)s * kDoneSlotWords;
    t_sn.cn.value = t_sn.value;
}

static int finish_sync(int err, bool used, void *stream) {
    Ctx &c = ctx();
    if (const int e = server_drain())
        if (!err) err = e;
    if (used && c.device >= 0) {
        bool done = false;
        if (!err && !stream && t_sn.armed && t_sn.gpu_ops == 1 && t_sn.notified == 1) {
            // exactly one kernel: wait for its completion word (bounded, then fall back to the stream)
            volatile uint64_t *w = t_sn.cn.flag;   // this thread's own slot: only its kernels write it, in order
            struct timespec t0, t1;
            clock_gettime(CLOCK_MONOTONIC, &t0);
            for (unsigned spin = 0; !done; ++spin) {
                if (*w >= t_sn.value) {
                    done = true;
                    break;
                }
                if ((spin & 255) == 255) {
                    clock_gettime(CLOCK_MONOTONIC, &t1);
                    if ((t1.tv_sec - t0.tv_sec) * 1000000000ll + (t1.tv_nsec - t0.tv_nsec) > 2000000ll) break;   // 2 ms
                }
                __builtin_ia32_pause();
            }
            std::atomic_thread_fence(std::memory_order_acquire);
        }
        if (!done) {
            DeviceGuard dg(c.device);
            cudaError_t e = cudaStreamSynchronize((cudaStream_t)(stream ? stream : c.stream));
            if (e != cudaSuccess && !err) err = cuda_err(e, "stream sync");
        }
        if (!stream) c.note_lib_synced(t_sn.async_snap);   // the library stream has drained up to this call
    }
    t_sn = SyncNotify();
    set_last_error(err);
    return err ? -1 : 0;
}

// Raw user pointer -> temporary side.
static Side user_side(void *p, uint64_t bytes) {
    Side s;
    s.backing = SIDE_USER;
    s.ptr = p;
    s.bytes = bytes;
    s.tier = TIER_HOST;
    if (ctx().device >= 0) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, p) == cudaSuccess) {
            if (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged) s.tier = TIER_GPU;
        } else {
            cudaGetLastError();
        }
    }
    return s;
}

} // namespace ocm

extern "C" {

int ocm_copy_async(ocm_alloc_t dst, ocm_alloc_t src, const struct ocm_params *p, void *stream) {
    TraceRange tr_("ocm_copy_async");
    bool used = false;
    if (!stream) ctx().note_lib_async();
    int err = do_copy(dst, src, p, stream, &used);
    set_last_error(err);
    return -err;
}

int ocm_copy(ocm_alloc_t dst, ocm_alloc_t src, ocm_param_t p) {
    TraceRange tr_("ocm_copy");
    bool used = false;
    arm_sync_notify();
    int err = do_copy(dst, src, p, nullptr, &used);
    return finish_sync(err, used, nullptr);
}

int ocm_copy_onesided_async(ocm_alloc_t a, const struct ocm_params *p, void *stream) {
    TraceRange tr_("ocm_copy_onesided_async");
    bool used = false;
    if (!stream) ctx().note_lib_async();
    int err = do_onesided(a, p, stream, &used);
    set_last_error(err);
    return -err;
}

int ocm_copy_onesided(ocm_alloc_t a, ocm_param_t p) {
    TraceRange tr_("ocm_copy_onesided");
    bool used = false;
    arm_sync_notify();
    int err = do_onesided(a, p, nullptr, &used);
    return finish_sync(err, used, nullptr);
}

// Many one-sided transfers on one pair in a single kernel launch (entry i: op_flag 1 = window[src_offset] ->
// far[dest_offset], 0 = the reverse).  Small messages are launch-bound (~2 us of device time each, more on the host);
// batching them is what a resident work queue would do, without the resident kernel.  Stream-ordered like the
// *_async calls; the far side's signal word receives +<number of puts> once every byte of the batch is visible.
int ocm_copy_onesided_batch(ocm_alloc_t a, const struct ocm_params *p, int n, void *stream) {
    TraceRange tr_("ocm_copy_onesided_batch");
    Ctx &c = ctx();
    if (!valid(a) || (!p && n > 0) || n < 0) {
        set_last_error(OCM_EINVAL);
        return -OCM_EINVAL;
    }
    if (n == 0) return 0;
    const Side *far = far_side(a);
    bool used = false;
    int err = OCM_OK;
    const bool one_kernel = a->stripes.empty() && far && a->local.valid() && c.device >= 0 && kernel_addressable(*far) &&
                            kernel_addressable(a->local) && (far->tier == TIER_GPU || a->local.tier == TIER_GPU);
    if (stale(a)) err = OCM_ESTALE;
    else if (!one_kernel) {
        if (!stream) c.note_lib_async();
        for (int i = 0; i < n && !err; ++i) err = do_onesided(a, &p[i], stream, &used);
    } else {
        std::vector<ocmk::CopySeg> segs((size_t)n);
        int puts = 0;
        for (int i = 0; i < n && !err; ++i) {
            if (p[i].bytes > a->local.bytes || !fits(a->local, p[i].src_offset, p[i].bytes) ||
                !fits(*far, p[i].dest_offset, p[i].bytes)) {
                err = OCM_ERANGE;
                break;
            }
            char *w = (char *)a->local.ptr + p[i].src_offset, *f = (char *)far->ptr + p[i].dest_offset;
            segs[i].src = p[i].op_flag ? w : f;
            segs[i].dst = p[i].op_flag ? f : w;
            segs[i].bytes = p[i].bytes;
            puts += p[i].op_flag != 0;
        }
        if (!err) {
            DeviceGuard dg(c.device);
            if (!stream) c.note_lib_async();
            unsigned long long nl = 0;
            // the signal counts puts only; a batch of gets publishes nothing
            uint64_t *sig = puts ? far->gpu_signal() : nullptr;
            cudaError_t e = cudaSuccess;
            if (puts == n || puts == 0) {
                e = ocmk::launch_copy_batch(segs.data(), n, sig, next_ticket(), (cudaStream_t)(stream ? stream : c.stream), &nl);
            } else {
                // mixed directions: puts first (they carry the signal), then the gets
                std::stable_partition(segs.begin(), segs.end(), [&](const ocmk::CopySeg &s) {
                    return (const char *)s.dst >= (const char *)far->ptr && (const char *)s.dst < (const char *)far->ptr + far->bytes;
                });
                e = ocmk::launch_copy_batch(segs.data(), puts, sig, next_ticket(), (cudaStream_t)(stream ? stream : c.stream), &nl);
                if (e == cudaSuccess)
                    e = ocmk::launch_copy_batch(segs.data() + puts, n - puts, nullptr, next_ticket(),
                                                (cudaStream_t)(stream ? stream : c.stream), &nl);
            }
            c.kernel_launches.fetch_add(nl, std::memory_order_relaxed);
            c.n_copies.fetch_add((uint64_t)n, std::memory_order_relaxed);
            for (int i = 0; i < n; ++i) c.copy_bytes.fetch_add(p[i].bytes, std::memory_order_relaxed);
            err = cuda_err(e, "batched copy kernel");
            used = true;
        }
    }
    mark_pending(a, stream, used);
    set_last_error(err);
    return -err;
}

// NVLS gather-reduce (K6, multimem.ld_reduce): window = sum over all ranks' replicas.  Stream-ordered like the
// *_async calls (stream NULL = the library's copy stream; ocm_quiet / ocm_free wait for it).
int ocm_reduce_get_f32(ocm_alloc_t a, const struct ocm_params *p, void *stream) {
    TraceRange tr_("ocm_reduce_get_f32");
    Ctx &c = ctx();
    int err = OCM_OK;
    if (!valid(a) || !p) err = OCM_EINVAL;
    else if (a->kind != OCM_REPLICATED_GPU || !a->mc || !a->mc_ptr || c.device < 0) err = OCM_EUNSUPPORTED;
    else if (!a->local.valid() || a->local.tier != TIER_GPU) err = OCM_EINVAL;
    else if (!fits(a->local, p->src_offset, p->bytes) || p->bytes > a->stripe_bytes ||
             p->dest_offset > a->stripe_bytes - p->bytes)
        err = OCM_ERANGE;
    else if ((p->src_offset | p->dest_offset | p->bytes) & 15) err = OCM_EINVAL;
    if (!err && p->bytes) {
        DeviceGuard dg(c.device);
        if (!stream) c.note_lib_async();
        err = cuda_err(ocmk::launch_get_reduce_f32((char *)a->local.ptr + p->src_offset, (const char *)a->mc_ptr + p->dest_offset,
                                                   p->bytes, (cudaStream_t)(stream ? stream : c.stream)),
                       "multimem reduce");
        ++c.kernel_launches;
        note_other_gpu_op();
        c.n_copies.fetch_add(1, std::memory_order_relaxed);
        c.copy_bytes.fetch_add(p->bytes * a->stripes.size(), std::memory_order_relaxed);
        mark_pending(a, stream, true);
    }
    set_last_error(err);
    return -err;
}

int ocm_quiet(void) {
    Ctx &c = ctx();
    if (c.device < 0 || !c.stream) return 0;
    DeviceGuard dg(c.device);
    const uint64_t snap = c.lib_async_issued.load(std::memory_order_acquire);
    if (cudaStreamSynchronize((cudaStream_t)c.stream) != cudaSuccess) return -OCM_ECUDA;
    c.note_lib_synced(snap);
    return 0;
}

int ocm_copy_in(ocm_alloc_t dst, void *src) {
    if (!valid(dst) || !src || !dst->stripes.empty()) {
        set_last_error(OCM_EINVAL);
        return -1;
    }
    if (stale(dst)) {   // an attached handle whose owner has freed (and maybe recycled) the block
        set_last_error(OCM_ESTALE);
        return -1;
    }
    arm_sync_notify();
    bool used = false;
    int err;
    const Side *far = far_side(dst);
    if (!far) {
        Side u = user_side(src, dst->local.bytes);
        err = side_copy(dst->local, 0, nullptr, 0, u, 0, dst->local.bytes, &dst->local, nullptr, &used);
    } else if (!dst->local.valid()) {
        Side u = user_side(src, far->bytes);
        err = side_copy(*far, 0, nullptr, 0, u, 0, far->bytes, far, nullptr, &used);
    } else {
        uint64_t n = std::min(dst->local.bytes, far->bytes);
        Side u = user_side(src, n);
        // window first, then window -> far side (the user pointer may be pageable host memory)
        err = side_copy(dst->local, 0, nullptr, 0, u, 0, n, nullptr, nullptr, &used);
        if (!err) err = side_copy(*far, 0, nullptr, 0, dst->local, 0, n, far, nullptr, &used);
    }
    return finish_sync(err, used, nullptr);
}

int ocm_copy_out(void *dstp, ocm_alloc_t src) {
    if (!valid(src) || !dstp || !src->stripes.empty()) {
        set_last_error(OCM_EINVAL);
        return -1;
    }
    if (stale(src)) {
        set_last_error(OCM_ESTALE);
        return -1;
    }
    arm_sync_notify();
    bool used = false;
    int err;
    const Side *far = far_side(src);
    if (!far) {
        Side u = user_side(dstp, src->local.bytes);
        err = side_copy(u, 0, nullptr, 0, src->local, 0, src->local.bytes, nullptr, nullptr, &used);
    } else if (!src->local.valid()) {
        Side u = user_side(dstp, far->bytes);
        err = side_copy(u, 0, nullptr, 0, *far, 0, far->bytes, nullptr, nullptr, &used);
    } else {
        uint64_t n = std::min(src->local.bytes, far->bytes);
        Side u = user_side(dstp, n);
        err = side_copy(src->local, 0, nullptr, 0, *far, 0, n, nullptr, nullptr, &used);
        if (!err) err = side_copy(u, 0, nullptr, 0, src->local, 0, n, nullptr, nullptr, &used);
    }
    return finish_sync(err, used, nullptr);
}

} // extern "C"
