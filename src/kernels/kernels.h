// Host-side launch interface of the sm_100a kernels.
// Every launcher enqueues on `stream` and returns immediately.  Pointers may
// be local HBM, peer HBM mapped into this process (NVLink), or pinned host
// memory registered with CUDA - the kernels only see generic global addresses.
#pragma once
#include <atomic>
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace ocmk {

// ---- K1/K2/K3: contiguous put / get / third-party copy ----------------------
// AUTO picks the 256-bit vector kernel when alignment allows, else VEC; BULK = TMA staging through smem.
enum CopyMode { COPY_AUTO = 0, COPY_BULK = 1, COPY_VEC = 2 };
struct CopyTuning {
    int mode = COPY_AUTO;
    int ctas = 0;          // 0: derive from size and SM count
    int agents = 4;        // independent TMA pipelines (warps) per CTA, <= 8
    int stages = 3;        // ring depth per agent, <= 8
    int tile_bytes = 16384;
    size_t bulk_min = 128 << 10;   // below this the vector kernel is used
};
// Reads OCM_COPY_MODE / OCM_COPY_CTAS / OCM_COPY_AGENTS / OCM_COPY_STAGES /
// OCM_COPY_TILE / OCM_BULK_MIN once.
const CopyTuning &default_copy_tuning();

// dst2 (optional) receives the same bytes (window + far side in one pass).
// signal (optional): 64-bit counter, +1 with release.sys after the last byte.
// ticket: zero-initialised device-local word used to elect the last CTA
// (required when signal != nullptr and more than one CTA runs).
// *launches (optional) is incremented by the number of kernels enqueued.
// notify (optional): a host-visible word that receives `value` (release.sys) once the last byte is visible,
// so that a blocked host thread can spin on it instead of synchronising the stream.
struct CopyNotify {
    uint64_t *flag;
    uint64_t  value;
};
cudaError_t launch_copy(void *dst, void *dst2, const void *src, size_t bytes, uint64_t *signal, unsigned *ticket,
                        cudaStream_t stream, const CopyTuning *tuning = nullptr, unsigned long long *launches = nullptr,
                        const CopyNotify *notify = nullptr);

// Many contiguous copies in ONE launch (small-message batching): segment i moves seg[i].bytes from seg[i].src to
// seg[i].dst.  Every CTA works on a slice of one segment.  signal (optional) receives +n after the last byte of the
// whole batch.  n <= kMaxBatch per launch; launch_copy_batch loops for longer lists.
constexpr int kMaxBatch = 64;
struct CopySeg {
    const void *src;
    void       *dst;
    uint64_t    bytes;
};
cudaError_t launch_copy_batch(const CopySeg *segs, int n, uint64_t *signal, unsigned *ticket, cudaStream_t stream,
                              unsigned long long *launches = nullptr, const CopyNotify *notify = nullptr);

// ---- K9: fill / verify / checksum --------------------------------------------
cudaError_t launch_fill(void *dst, size_t bytes, uint64_t pattern, cudaStream_t stream);
// *mismatches (device or mapped host, u64) += number of 8-byte words != pattern (+ tail bytes)
cudaError_t launch_verify(const void *p, size_t bytes, uint64_t pattern, unsigned long long *mismatches,
                          cudaStream_t stream);
// *sum (u64) += wrapping sum of all 8-byte words (tail bytes zero-extended)
cudaError_t launch_checksum(const void *p, size_t bytes, unsigned long long *sum, cudaStream_t stream);
// +1 (release.sys) on a signal word, ordered after prior work on the stream
cudaError_t launch_signal(uint64_t *signal, cudaStream_t stream);

// ---- K4/K5: strided pack/unpack with optional cast ----------------------------
enum DType { DT_U8 = 0, DT_F32 = 1, DT_BF16 = 2, DT_F8E4M3 = 3 };
struct StridedDesc {
    const void *src;
    void       *dst;
    uint64_t    row_bytes;          // contiguous bytes per row on the source side
    uint64_t    n[3];               // extents of the 3 outer dimensions (>= 1)
    int64_t     src_stride[3];      // bytes
    int64_t     dst_stride[3];      // bytes
    int         src_dtype, dst_dtype;
    int         scale_block;        // 32 or 128 when fp8 is involved
    float      *scales;             // fp8 side: one fp32 per block, row-major over (rows, blocks)
};
cudaError_t launch_strided(const StridedDesc &d, uint64_t *signal, unsigned *ticket, cudaStream_t stream,
                           unsigned long long *launches = nullptr);
#ifdef __CUDACC__
__host__ __device__
#endif
inline size_t dtype_size(int dt) { return dt == DT_F32 ? 4 : dt == DT_BF16 ? 2 : 1; }

// ---- K6: replicated put ---------------------------------------------------------
// One source, n_dst destinations (same bytes): one CTA of a thread-block cluster per replica.
// src_on_this_gpu = the caller KNOWS src is in the launching GPU's own memory: the cluster then shares each staged
// tile (TMA cluster multicast, the source is read once).  Otherwise every CTA loads the tile itself.
cudaError_t launch_put_replicated(void *const *dsts, int n_dst, const void *src, size_t bytes,
                                  uint64_t *const *signals, unsigned *ticket, cudaStream_t stream,
                                  bool src_on_this_gpu = false);
// NVLS variant: one multimem.st stream to a multicast VA.
cudaError_t launch_put_multimem(void *mc_dst, const void *src, size_t bytes, uint64_t *mc_signal, unsigned *ticket,
                                cudaStream_t stream);
cudaError_t launch_signal_multimem(uint64_t *mc_signal, cudaStream_t stream);
// NVLS gather-reduce: dst = element-wise fp32 sum of every GPU's replica behind the multicast VA (multimem.ld_reduce).
cudaError_t launch_get_reduce_f32(void *dst, const void *mc_src, size_t bytes, cudaStream_t stream);

// ---- K7: GUPS over a striped table ------------------------------------------------
struct GupsTable {
    uint64_t *base[64];     // per-rank slab, mapped into this process
    int       n_ranks;
    uint64_t  words_per_stripe;   // power of two
    uint64_t  total_words;        // power of two
};
cudaError_t launch_gups(const GupsTable &t, uint64_t n_updates, uint64_t seed, cudaStream_t stream);
uint64_t gups_threads();   // threads launch_gups uses on the current device (the oracle needs the same number)
cudaError_t launch_gups_init(const GupsTable &t, int my_rank, cudaStream_t stream);

// ---- K8: radix sort of u64 keys -------------------------------------------------------
// scratch holds the per-CTA digit histograms and their scanned offsets.
size_t sort_scratch_bytes(uint64_t n, int *ctas_out = nullptr);
cudaError_t launch_sort_u64(uint64_t *keys, uint64_t *tmp, uint64_t n, void *scratch, size_t scratch_bytes,
                            cudaStream_t stream, unsigned long long *launches = nullptr);

// ---- H2: resident copy server ----------------------------------------------------------
// Descriptor ring + control block live in pinned, mapped host memory (the host posts, the kernel polls).
struct alignas(64) ServerDesc {
    const void *src;
    void       *dst;
    void       *dst2;       // optional second destination
    uint64_t    bytes;
    uint64_t   *signal;     // optional completion counter of the destination allocation (+1, release.sys)
    uint64_t    pad[2];
    uint64_t    seq;        // written last (release): descriptor i is valid once seq == i + 1
};
enum ServerState : uint32_t { SERVER_EXITED = 0, SERVER_RUNNING = 1, SERVER_EXITING = 2 };
struct ServerCtl {
    alignas(64) uint32_t state;   // ServerState; host: EXITED -> RUNNING (before a launch), kernel: the rest
    alignas(64) uint64_t done;    // descriptors completed so far (kernel, release.sys; in order)
    // phase accounting, added up by the kernel when it leaves (one writer; read with OCM_SERVER_TRACE=1 at ocm_tini):
    // descriptors served, ns spent copying / in the completion fence + stores, empty polls of the ring and ns in them
    alignas(64) uint64_t t_ops;
    uint64_t t_copy_ns, t_fence_ns, t_polls, t_poll_ns;
};
// Starts the one-CTA server on `stream`: it serves descriptors from `start_seq` on and leaves after `idle_ns`
// without work or `max_ns` of life (see server.cu for the exit hand-shake).
cudaError_t launch_copy_server(ServerDesc *ring, uint32_t ring_size, ServerCtl *ctl, uint64_t start_seq, uint64_t idle_ns,
                               uint64_t max_ns, cudaStream_t stream);

// ---- K8: merge of sorted u64 runs -----------------------------------------------------------------------
// out = stable merge of a[0..na) and b[0..nb) (both ascending, unsigned).  One launch handles many pairs.
struct MergePairDesc {
    const uint64_t *a, *b;
    uint64_t       *out;
    uint64_t        na, nb;
};
cudaError_t launch_merge_pairs(const MergePairDesc *pairs, int n_pairs, cudaStream_t stream,
                               unsigned long long *launches = nullptr);

int num_sms(int device = -1);
// Launches made by entry points that live inside the kernel library (ocmk_sort_u64_noalloc, the copy server);
// ocmk_launch_count() adds it to the library's own counters.
std::atomic<unsigned long long> &extra_launches();

} // namespace ocmk
