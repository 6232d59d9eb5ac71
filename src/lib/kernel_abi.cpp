// extern "C" wrappers around the kernel launchers (include/ocm_kernels.h) and the
// strided entry points of the allocation API.  Scratch needed by the kernels
// (election tickets, sort histograms) is allocated lazily per device, never on
// the per-call path once warm.
#include "ctx.hpp"
#include "../common/log.hpp"
#include "../kernels/kernels.h"
#include "../../include/ocm_kernels.h"
#include <atomic>
#include <cstring>
#include <cuda_runtime.h>
#include <mutex>
#include <vector>

using namespace ocm;

namespace {

std::atomic<unsigned long long> g_abi_launches{0};

struct DevScratch {
    unsigned *tickets = nullptr;
    std::atomic<uint32_t> cursor{0};
    void  *sort_scratch = nullptr;
    size_t sort_bytes = 0;
    std::vector<void *> retired;   // outgrown sort scratch buffers (at most log2 of them; freed with the context)
};
DevScratch g_dev[64];
std::mutex g_dev_mu;

unsigned *abi_ticket() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    DevScratch &d = g_dev[dev];
    if (!d.tickets) {
        std::lock_guard<std::mutex> g(g_dev_mu);
        if (!d.tickets) {
            unsigned *p = nullptr;
            if (cudaMalloc((void **)&p, kNumTickets * sizeof(unsigned)) != cudaSuccess) return nullptr;
            cudaMemset(p, 0, kNumTickets * sizeof(unsigned));
            d.tickets = p;
        }
    }
    return d.tickets + (d.cursor.fetch_add(1, std::memory_order_relaxed) % kNumTickets);
}

int ret(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return 0;
    LOGE("%s: %s", what, cudaGetErrorString(e));
    cudaGetLastError();
    return -OCM_ECUDA;
}

ocmk::StridedDesc to_desc(const struct ocmk_strided *d) {
    ocmk::StridedDesc s;
    s.src = d->src; s.dst = d->dst; s.row_bytes = d->row_bytes;
    for (int i = 0; i < 3; ++i) {
        s.n[i] = d->n[i] ? d->n[i] : 1;
        s.src_stride[i] = d->src_stride[i];
        s.dst_stride[i] = d->dst_stride[i];
    }
    s.src_dtype = d->src_dtype; s.dst_dtype = d->dst_dtype; s.scale_block = d->scale_block; s.scales = d->scales;
    return s;
}

} // namespace

extern "C" {

uint64_t ocmk_launch_count(void) {
    return g_abi_launches.load() + ctx().kernel_launches.load() + ocmk::extra_launches().load();
}

int ocmk_copy(void *dst, void *dst2, const void *src, size_t bytes, uint64_t *signal, void *stream) {
    unsigned long long n = 0;
    cudaError_t e = ocmk::launch_copy(dst, dst2, src, bytes, signal, abi_ticket(), (cudaStream_t)stream, nullptr, &n);
    g_abi_launches += n;
    return ret(e, "ocmk_copy");
}

int ocmk_copy_tuned(void *dst, void *dst2, const void *src, size_t bytes, uint64_t *signal, void *stream, int mode,
                    int ctas, int agents, int stages, int tile_bytes) {
    ocmk::CopyTuning t = ocmk::default_copy_tuning();
    t.mode = mode;
    if (mode != ocmk::COPY_AUTO) t.bulk_min = 0;
    if (ctas > 0) t.ctas = ctas;
    if (agents > 0) t.agents = agents;
    if (stages > 0) t.stages = stages;
    if (tile_bytes > 0) t.tile_bytes = tile_bytes;
    unsigned long long n = 0;
    cudaError_t e = ocmk::launch_copy(dst, dst2, src, bytes, signal, abi_ticket(), (cudaStream_t)stream, &t, &n);
    g_abi_launches += n;
    return ret(e, "ocmk_copy_tuned");
}

int ocmk_fill(void *dst, size_t bytes, uint64_t pattern, void *stream) {
    ++g_abi_launches;
    return ret(ocmk::launch_fill(dst, bytes, pattern, (cudaStream_t)stream), "ocmk_fill");
}
int ocmk_verify(const void *p, size_t bytes, uint64_t pattern, uint64_t *out, void *stream) {
    ++g_abi_launches;
    return ret(ocmk::launch_verify(p, bytes, pattern, (unsigned long long *)out, (cudaStream_t)stream), "ocmk_verify");
}
int ocmk_checksum(const void *p, size_t bytes, uint64_t *out, void *stream) {
    ++g_abi_launches;
    return ret(ocmk::launch_checksum(p, bytes, (unsigned long long *)out, (cudaStream_t)stream), "ocmk_checksum");
}
int ocmk_signal(uint64_t *signal, void *stream) {
    ++g_abi_launches;
    return ret(ocmk::launch_signal(signal, (cudaStream_t)stream), "ocmk_signal");
}

int ocmk_strided_copy(const struct ocmk_strided *d, uint64_t *signal, void *stream) {
    if (!d) return -OCM_EINVAL;
    unsigned long long n = 0;
    cudaError_t e = ocmk::launch_strided(to_desc(d), signal, abi_ticket(), (cudaStream_t)stream, &n);
    g_abi_launches += n;
    return ret(e, "ocmk_strided_copy");
}

int ocmk_put_replicated(void *const *dsts, int n_dst, const void *src, size_t bytes, void *stream) {
    ++g_abi_launches;
    // raw pointers: where src lives is not known here, so the tiles are not shared across the cluster
    return ret(ocmk::launch_put_replicated(dsts, n_dst, src, bytes, nullptr, abi_ticket(), (cudaStream_t)stream, false),
               "ocmk_put_replicated");
}

int ocmk_gups(uint64_t *const *bases, int n_ranks, uint64_t words_per_stripe, uint64_t total_words, uint64_t n_updates,
              uint64_t seed, void *stream) {
    if (n_ranks < 1 || n_ranks > 64) return -OCM_EINVAL;
    ocmk::GupsTable t;
    memset(&t, 0, sizeof t);
    for (int i = 0; i < n_ranks; ++i) t.base[i] = bases[i];
    t.n_ranks = n_ranks;
    t.words_per_stripe = words_per_stripe;
    t.total_words = total_words;
    ++g_abi_launches;
    return ret(ocmk::launch_gups(t, n_updates, seed, (cudaStream_t)stream), "ocmk_gups");
}
uint64_t ocmk_gups_threads(uint64_t) { return ocmk::gups_threads(); }

int ocmk_sort_u64(uint64_t *keys, uint64_t *tmp, uint64_t n, void *stream) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return -OCM_ENOGPU;
    DevScratch &d = g_dev[dev];
    size_t need = ocmk::sort_scratch_bytes(n);
    void *scratch = nullptr;
    size_t scratch_bytes = 0;
    {
        // One scratch buffer per device, grown on demand and shared by every call on that device: sorts whose
        // kernels overlap on one GPU (other streams, other threads) race on it - serialise them, like the
        // allocation-free variant.  An outgrown buffer is parked, not freed: a sort launched earlier on another
        // stream may still be using it.
        std::lock_guard<std::mutex> g(g_dev_mu);
        if (d.sort_bytes < need) {
            void *p = nullptr;
            need = (need + (1u << 20) - 1) & ~(size_t)((1u << 20) - 1);
            if (cudaMalloc(&p, need) != cudaSuccess) {
                cudaGetLastError();
                return -OCM_ENOMEM;
            }
            if (d.sort_scratch) d.retired.push_back(d.sort_scratch);
            d.sort_scratch = p;
            d.sort_bytes = need;
        }
        scratch = d.sort_scratch;
        scratch_bytes = d.sort_bytes;
    }
    unsigned long long nl = 0;
    cudaError_t e = ocmk::launch_sort_u64(keys, tmp, n, scratch, scratch_bytes, (cudaStream_t)stream, &nl);
    g_abi_launches += nl;
    return ret(e, "ocmk_sort_u64");
}

int ocmk_merge_u64(const uint64_t *a, uint64_t na, const uint64_t *b, uint64_t nb, uint64_t *out, void *stream) {
    ocmk::MergePairDesc d{a, b, out, na, nb};
    unsigned long long nl = 0;
    cudaError_t e = ocmk::launch_merge_pairs(&d, 1, (cudaStream_t)stream, &nl);
    g_abi_launches += nl;
    return ret(e, "ocmk_merge_u64");
}

// k sorted runs laid end to end in `keys` (run i = keys[offsets[i] .. offsets[i+1])) become one sorted sequence
// through ceil(log2 k) levels of pairwise merges between `keys` and `tmp`; every level is one launch per 32 pairs.
// Returns 0 if the result is in `keys`, 1 if it is in `tmp`, or a negative ocm_err.
int ocmk_merge_runs_u64(uint64_t *keys, uint64_t *tmp, const uint64_t *offsets, int k, void *stream) {
    if (k < 1 || !keys || !tmp || !offsets) return -OCM_EINVAL;
    std::vector<uint64_t> off(offsets, offsets + k + 1);
    for (int i = 0; i < k; ++i)
        if (off[i] > off[i + 1]) return -OCM_EINVAL;
    uint64_t *src = keys, *dst = tmp;
    int where = 0;
    unsigned long long nl = 0;
    while (off.size() > 2) {
        const int runs = (int)off.size() - 1;
        std::vector<ocmk::MergePairDesc> pairs;
        std::vector<uint64_t> next{off[0]};
        for (int i = 0; i < runs; i += 2) {
            ocmk::MergePairDesc d;
            d.a = src + off[i];
            d.na = off[i + 1] - off[i];
            d.b = i + 1 < runs ? src + off[i + 1] : nullptr;
            d.nb = i + 1 < runs ? off[i + 2] - off[i + 1] : 0;      // an odd run out is merged with nothing: a copy
            d.out = dst + off[i];
            pairs.push_back(d);
            next.push_back(off[std::min(i + 2, runs)]);
        }
        cudaError_t e = ocmk::launch_merge_pairs(pairs.data(), (int)pairs.size(), (cudaStream_t)stream, &nl);
        if (e != cudaSuccess) {
            g_abi_launches += nl;
            return ret(e, "ocmk_merge_runs_u64");
        }
        off.swap(next);
        std::swap(src, dst);
        where ^= 1;
    }
    g_abi_launches += nl;
    return where;
}

// ---- strided copies on allocations -------------------------------------------------
static int strided_sides(const Side &dst, const Side &src, const struct ocm_strided *d, const Side *sig, void *stream) {
    Ctx &c = ctx();
    if (c.device < 0) return OCM_ENOGPU;
    if (!dst.ptr || !src.ptr) return OCM_EINVAL;
    auto addressable = [](const Side &s) { return s.tier == TIER_GPU || s.backing == SIDE_POOL; };
    if (!addressable(dst) || !addressable(src)) return OCM_EUNSUPPORTED;
    const size_t ss = ocmk::dtype_size(d->src_dtype), ds = ocmk::dtype_size(d->dst_dtype);
    if (d->extent[0] == 0 || d->extent[0] % ss) return OCM_EINVAL;
    const uint64_t elems = d->extent[0] / ss;
    uint64_t n[3];
    for (int i = 0; i < 3; ++i) n[i] = d->extent[i + 1] ? d->extent[i + 1] : 1;
    // bounds: last byte touched on each side
    auto span = [&](const int64_t *st, uint64_t row_b) {
        int64_t hi = 0, lo = 0;
        for (int i = 0; i < 3; ++i) {
            int64_t e = (int64_t)(n[i] - 1) * st[i];
            if (e > 0) hi += e; else lo += e;
        }
        return std::pair<int64_t, int64_t>(lo, hi + (int64_t)row_b);
    };
    auto sb = span(d->src_stride, d->extent[0]), db = span(d->dst_stride, elems * ds);
    if ((int64_t)d->src_offset + sb.first < 0 || d->src_offset + (uint64_t)sb.second > src.bytes) return OCM_ERANGE;
    if ((int64_t)d->dst_offset + db.first < 0 || d->dst_offset + (uint64_t)db.second > dst.bytes) return OCM_ERANGE;
    ocmk::StridedDesc k;
    k.src = (const char *)src.ptr + d->src_offset;
    k.dst = (char *)dst.ptr + d->dst_offset;
    k.row_bytes = d->extent[0];
    for (int i = 0; i < 3; ++i) {
        k.n[i] = n[i];
        k.src_stride[i] = d->src_stride[i];
        k.dst_stride[i] = d->dst_stride[i];
    }
    k.src_dtype = d->src_dtype; k.dst_dtype = d->dst_dtype; k.scale_block = d->scale_block;
    k.scales = nullptr;
    if (d->src_dtype != d->dst_dtype && (d->src_dtype == OCM_F8E4M3 || d->dst_dtype == OCM_F8E4M3)) {
        // scales live on the fp8 side of the transfer
        const Side &f8 = d->dst_dtype == OCM_F8E4M3 ? dst : src;
        const uint64_t blocks = (elems + d->scale_block - 1) / (d->scale_block ? d->scale_block : 1);
        if (d->scale_block != 32 && d->scale_block != 128) return OCM_EINVAL;
        if (d->scale_offset % 4 || d->scale_offset + blocks * n[0] * n[1] * n[2] * 4 > f8.bytes) return OCM_ERANGE;
        k.scales = reinterpret_cast<float *>((char *)f8.ptr + d->scale_offset);
    }
    DeviceGuard dg(c.device);
    cudaStream_t st = (cudaStream_t)(stream ? stream : c.stream);
    if (!stream) c.note_lib_async();
    c.n_copies.fetch_add(1, std::memory_order_relaxed);
    c.copy_bytes.fetch_add(d->extent[0] * n[0] * n[1] * n[2], std::memory_order_relaxed);
    unsigned long long nl = 0;
    cudaError_t e = ocmk::launch_strided(k, sig ? sig->gpu_signal() : nullptr, next_ticket(), st, &nl);
    c.kernel_launches.fetch_add(nl, std::memory_order_relaxed);
    if (e == cudaSuccess) return OCM_OK;
    LOGE("strided kernel: %s", cudaGetErrorString(e));
    cudaGetLastError();
    return OCM_ECUDA;
}

int ocm_copy_strided(ocm_alloc_t dst, ocm_alloc_t src, const struct ocm_strided *d, void *stream) {
    if (!dst || !src || !d || dst->magic != kAllocMagic || src->magic != kAllocMagic) return -OCM_EINVAL;
    if (!dst->stripes.empty() || !src->stripes.empty()) return -OCM_EUNSUPPORTED;
    if (side_stale(dst->remote) || side_stale(src->remote)) return -OCM_ESTALE;
    if (!d->op_flag) std::swap(dst, src);
    // the far side is the target when there is one, like a one-hop ocm_copy
    const Side &D = (dst->paired && dst->remote.valid()) ? dst->remote : dst->local;
    const Side &S = (src->paired && src->remote.valid()) ? src->remote : src->local;
    int err = strided_sides(D, S, d, &D, stream);
    if (!err && !stream) {
        dst->lib_stream_pending.store(true);
        src->lib_stream_pending.store(true);
    }
    set_last_error(err);
    return -err;
}

int ocm_copy_onesided_strided(ocm_alloc_t a, const struct ocm_strided *d, void *stream) {
    if (!a || !d || a->magic != kAllocMagic || !a->paired || !a->remote.valid() || !a->local.valid()) return -OCM_EINVAL;
    int err = d->op_flag ? strided_sides(a->remote, a->local, d, &a->remote, stream)
                         : strided_sides(a->local, a->remote, d, nullptr, stream);
    if (!err && !stream) a->lib_stream_pending.store(true);
    set_last_error(err);
    return -err;
}

} // extern "C"
