/*
 * oncillamem.h -- application interface of OCM-B200 (libocm.so)
 *
 * C ABI.  The first block keeps the surface of the reference header
 * (reference inc/oncillamem.h:24-89): same enum values, same struct field
 * order, same twelve functions.  Everything below the "extensions" banner is
 * new: kinds that place memory in a peer GPU's HBM or in another rank's
 * pinned host pool, stream-ordered copies with completion signals, strided
 * and block-scaled-cast copies, striping, multicast replication, statistics.
 *
 * Unlike the reference header this one pulls in no private headers
 * (the reference leaks util/list.h, inc/oncillamem.h:18).
 */
#ifndef ONCILLAMEM_H_
#define ONCILLAMEM_H_

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lib_alloc *ocm_alloc_t;

/* Values 1..7 are the reference's (inc/oncillamem.h:26-35). */
enum ocm_kind {
    OCM_LOCAL_HOST = 1,   /* host memory of the calling rank (pinned pool, malloc fallback) */
    OCM_LOCAL_RMA,        /* accepted, granted as OCM_LOCAL_HOST */
    OCM_REMOTE_RMA,       /* legacy name: granted as a remote pinned-host pair */
    OCM_LOCAL_RDMA,       /* accepted, granted as OCM_LOCAL_HOST */
    OCM_REMOTE_RDMA,      /* legacy name: granted as a remote pinned-host pair */
    OCM_LOCAL_GPU,        /* HBM of the calling rank's GPU */
    OCM_REMOTE_GPU,       /* pair: local HBM window + region in a peer GPU's HBM */
    /* ---- new kinds ---- */
    OCM_REMOTE_HOST,      /* pair: local pinned window + region in a peer rank's pinned DRAM */
    OCM_STRIPED_GPU,      /* pair: local HBM window + region striped over every rank's HBM */
    OCM_REPLICATED_GPU,   /* pair: local HBM window + one replica on every rank (NVLS multicast put) */
    OCM_KIND_MAX
};

/* Same layout as the reference (inc/oncillamem.h:39-48). */
struct ocm_params {
    uint64_t src_offset;
    uint64_t dest_offset;
    uint64_t src_offset_2;
    uint64_t dest_offset_2;
    uint64_t bytes;
    int      op_flag;      /* read = 0, write = 1 */
};
typedef struct ocm_params *ocm_param_t;

/* Same layout as the reference (inc/oncillamem.h:53-58). */
struct ocm_alloc_params {
    uint64_t      local_alloc_bytes;
    uint64_t      rem_alloc_bytes;
    enum ocm_kind kind;
};
typedef struct ocm_alloc_params *ocm_alloc_param_t;

/* ---- reference API (inc/oncillamem.h:69-89) ---- */
int         ocm_init(void);
int         ocm_tini(void);
ocm_alloc_t ocm_alloc(ocm_alloc_param_t alloc_param);
int         ocm_free(ocm_alloc_t a);
/* pointer/length of the local side (for a pair: the local window) */
int         ocm_localbuf(ocm_alloc_t a, void **buf, size_t *len);
bool        ocm_is_remote(ocm_alloc_t a);
enum ocm_kind ocm_alloc_kind(ocm_alloc_t a);
/* size of the remote side; -1 for local kinds */
int         ocm_remote_sz(ocm_alloc_t a, size_t *len);
/* Whole-allocation copies between a raw pointer (host or device) and an
 * allocation.  Stubs in the reference (src/lib.c:491-499); real here.  For a
 * pair the far side is the target and min(local,remote) bytes move. */
int         ocm_copy_out(void *dst, ocm_alloc_t src);
int         ocm_copy_in(ocm_alloc_t dst, void *src);
/* Two-sided copy, see DESIGN.md "copy matrix".  Never modifies *options
 * (the reference flips op_flag in the caller's struct, src/lib.c:511-514). */
int         ocm_copy(ocm_alloc_t dst, ocm_alloc_t src, ocm_param_t options);
/* One-sided put (op_flag=1: window[src_offset] -> remote[dest_offset]) or get
 * (op_flag=0: remote[dest_offset] -> window[src_offset]) on a pair. */
int         ocm_copy_onesided(ocm_alloc_t a, ocm_param_t options);

/* ===================================================================== */
/* ============================ extensions ============================== */
/* ===================================================================== */

#define OCM_API_VERSION 2

/* Error codes returned (negated) by the int-returning extension calls; the
 * reference API keeps its 0 / -1 convention. */
enum ocm_err {
    OCM_OK = 0, OCM_EINVAL = 1, OCM_ENOMEM = 2, OCM_ENODAEMON = 3, OCM_ENOGPU = 4,
    OCM_ERANGE = 5, OCM_ECUDA = 6, OCM_EPROTO = 7, OCM_ESTALE = 8, OCM_EUNSUPPORTED = 9
};
const char *ocm_strerror(int err);
/* last error of the calling thread (0 if the previous call succeeded) */
int         ocm_last_error(void);

/* Placement policies understood by the rank-0 governor. */
enum ocm_policy {
    OCM_POLICY_DEFAULT = 0,   /* neighbour: (rank+1) % N, as reference src/alloc.c:107 */
    OCM_POLICY_NEIGHBOUR,
    OCM_POLICY_EXPLICIT,      /* use ocm_alloc_params_ex.rank */
    OCM_POLICY_LEAST_LOADED,
    OCM_POLICY_ROUND_ROBIN,
    OCM_POLICY_SELF           /* stay on the caller's rank */
};

enum ocm_alloc_flags {
    OCM_ALLOC_SLOW_PATH = 1u << 0,   /* force the daemon/rank-0 message protocol */
    OCM_ALLOC_NO_ZERO   = 1u << 1,   /* caller does not need zero-filled memory */
    OCM_ALLOC_NO_WINDOW = 1u << 2    /* pair without a local window (local_alloc_bytes ignored) */
};

struct ocm_alloc_params_ex {
    struct ocm_alloc_params base;
    int      policy;        /* enum ocm_policy */
    int      rank;          /* target rank for OCM_POLICY_EXPLICIT */
    uint32_t flags;         /* enum ocm_alloc_flags */
    uint32_t reserved;
};
ocm_alloc_t ocm_alloc_ex(const struct ocm_alloc_params_ex *p);

/* Introspection. */
int      ocm_rank(void);                 /* rank of the daemon this process is attached to */
int      ocm_num_ranks(void);
int      ocm_device(void);               /* CUDA device of this process, -1 without a GPU */
int      ocm_remote_rank(ocm_alloc_t a); /* owner rank of the far side, -1 for local kinds */
/* Device- or host-usable address of the far side as mapped into this process
 * (peer HBM or peer pinned DRAM); NULL for local and striped kinds. */
void    *ocm_remote_ptr(ocm_alloc_t a);
uint64_t ocm_alloc_id(ocm_alloc_t a);
/* Multi-slab kinds (OCM_STRIPED_GPU / OCM_REPLICATED_GPU): the slab that lives on `rank`, as mapped into
 * this process, and the bytes per slab.  Striped: logical offset o lives in slab o / slab_bytes. */
void    *ocm_slab_ptr(ocm_alloc_t a, int rank);
uint64_t ocm_slab_bytes(ocm_alloc_t a);
/* Map every rank's tables and pools into this process NOW instead of at the first allocation / attach that needs
 * them (mapping a 100 GiB peer pool takes seconds, once per process; an application with a timed phase calls this
 * during set-up).  tiers: bit 0 = HBM pools, bit 1 = host pools.  Returns the number of ranks whose requested pools
 * are mapped, or a negative ocm_err. */
int      ocm_prefetch_views(unsigned tiers);
/* NVLS: 1 once the daemons have bound the multicast region and this process has mapped it; replicated
 * allocations made after that are multicast-backed (one multimem.st stream reaches every GPU). */
int      ocm_multicast_ready(void);
int      ocm_is_multicast(ocm_alloc_t a);
/* NVLS gather-reduce on a multicast-backed OCM_REPLICATED_GPU allocation:
 *   window[src_offset .. +bytes) = element-wise fp32 SUM over every rank's replica of far[dest_offset .. +bytes)
 * in one multimem.ld_reduce stream (the NVSwitch adds the replicas; the "get" counterpart of the multicast
 * put and the reduce half of an in-switch all-reduce).  Each rank writes its own contribution into its replica
 * (ocm_slab_ptr(a, ocm_rank())); ordering those writes before the reduce is the caller's business.  Offsets
 * and bytes are multiples of 16.  -OCM_EUNSUPPORTED when the allocation is not multicast-backed. */
int      ocm_reduce_get_f32(ocm_alloc_t a, const struct ocm_params *p, void *stream);

/* ---- stream-ordered copies ------------------------------------------ */
/* `stream` is a cudaStream_t (NULL = the library's own copy stream).  The
 * call returns once the work is enqueued; host<->host hops that need the CPU
 * are performed before returning. */
int ocm_copy_async(ocm_alloc_t dst, ocm_alloc_t src, const struct ocm_params *p, void *stream);
int ocm_copy_onesided_async(ocm_alloc_t a, const struct ocm_params *p, void *stream);
/* n one-sided transfers on ONE paired allocation in a single kernel launch (entry i: op_flag 1 = put
 * window[src_offset] -> far[dest_offset], 0 = get).  Small messages are launch-bound; a batch costs one launch.
 * The reference posts one work request and blocks once per message (src/lib.c:669-723).  Stream-ordered like the
 * *_async calls; the far side's signal receives +<number of puts> after the last byte of the batch. */
int ocm_copy_onesided_batch(ocm_alloc_t a, const struct ocm_params *p, int n, void *stream);
/* Wait for everything enqueued on the library's own copy stream. */
int ocm_quiet(void);

/* Completion signal of an allocation: a 64-bit count kept in device-visible
 * pinned memory next to the allocation's table entry.  Every put that names
 * the allocation as destination adds 1 to it with release semantics at system
 * scope after its last byte is visible (replaces ib_poll / rma2 notifications,
 * reference src/rdma.c:267-302, src/extoll.c:145-166).  Internally the count is
 * the sum of several words with one writing agent each (the owner's GPU, CPUs,
 * and for shared blocks one word per rank's GPU): GPUs cannot add atomically to
 * the same host word on boxes without PCIe atomics. */
int ocm_signal_read(ocm_alloc_t a, uint64_t *value);
int ocm_signal_wait(ocm_alloc_t a, uint64_t at_least, int timeout_ms);

/* ---- sharing an allocation between applications --------------------- */
/* In the reference an allocation is private to the process that made it; the owner of the memory only
 * learns about arriving data through the EXTOLL completer notification (reference src/extoll.c:145-170,
 * src/extoll_server.c).  Here any application of the cluster can open a block that lives in a daemon's
 * pool: the owner calls ocm_share() and hands the 32-byte token to its peers by any means (file, pipe,
 * torch.distributed); a peer calls ocm_attach() and gets a window-less pair whose far side IS that
 * block (ocm_remote_ptr, ocm_copy, ocm_copy_strided and the completion signal all work on it).  The
 * signal word is shared too: the owner's ocm_signal_wait() returns once the peers' puts have landed.
 * ocm_free() of an attached handle only drops the handle.  The block belongs to the owner: copies
 * through an attached handle fail with OCM_ESTALE once the owner has freed it, and an owner that frees
 * while peers still write is a use-after-free of the peers' making (as with CUDA IPC handles). */
struct ocm_share_token {
    uint64_t id;          /* allocation id on the rank that holds the block */
    uint64_t bytes;
    int32_t  rank;
    uint32_t region;
    uint32_t generation;
    uint32_t tier;        /* 0 HBM, 1 pinned host, OCM_SHARE_TIER_MULTICAST: a range of the NVLS region */
};
#define OCM_SHARE_TIER_MULTICAST 2u
/* The far side of a pair or the memory of a local kind; -OCM_EUNSUPPORTED for multi-slab kinds and for
 * local kinds that fell back to malloc / cudaMalloc (src/lib.c:226,240 behaviour). */
int ocm_share(ocm_alloc_t a, struct ocm_share_token *out);
int ocm_attach(const struct ocm_share_token *tok, ocm_alloc_t *out);
/* handles other applications currently hold on the block behind `a` (owner or attached handle) */
int ocm_attach_count(ocm_alloc_t a);

/* ---- strided / cast copies (kernels K4/K5) -------------------------- */
enum ocm_dtype { OCM_U8 = 0, OCM_F32 = 1, OCM_BF16 = 2, OCM_F8E4M3 = 3 };

/* Up to 4-D box.  extent[0] is the contiguous run in BYTES of the source
 * dtype; stride[i] (bytes) steps dimension i+1 on each side. */
struct ocm_strided {
    uint64_t src_offset, dst_offset;
    uint64_t extent[4];        /* extent[0] bytes per row, extent[1..3] counts (0 or 1 = unused) */
    int64_t  src_stride[3];
    int64_t  dst_stride[3];    /* in bytes of the DESTINATION dtype */
    int      op_flag;          /* 1: window/src -> far side; 0: far side -> window */
    int      src_dtype, dst_dtype;   /* enum ocm_dtype; equal = plain byte copy */
    int      scale_block;      /* elements per scale for OCM_F8E4M3 (32 or 128) */
    uint64_t scale_offset;     /* where the fp32 scales live on the fp8 side */
};
int ocm_copy_strided(ocm_alloc_t dst, ocm_alloc_t src, const struct ocm_strided *d, void *stream);
int ocm_copy_onesided_strided(ocm_alloc_t a, const struct ocm_strided *d, void *stream);

/* ---- statistics ------------------------------------------------------ */
struct ocm_stats {
    uint64_t allocs_live, allocs_total, frees_total;
    uint64_t bytes_gpu_live, bytes_host_live;
    uint64_t copies, copy_bytes, kernel_launches;
    uint64_t fast_path_allocs, slow_path_allocs;
    uint64_t pool_gpu_bytes, pool_gpu_used, pool_host_bytes, pool_host_used;
    /* granules taken out of each pool (what the blocks really cost: pool_*_used rounds every block up to its size
     * class, these also count the free remainder of slabs) and audits that found a rank's tables inconsistent */
    uint64_t pool_gpu_granule_bytes, pool_host_granule_bytes, table_faults;
};
int ocm_get_stats(struct ocm_stats *out);            /* this process + its rank's pools */
int ocm_get_rank_stats(int rank, struct ocm_stats *out); /* via MSG_STATS to the daemon */
/* Resident copy server (blocking copies up to OCM_SERVER_MAX bytes, default 128 KiB, are moved by a kernel that is
 * already running: no launch, no stream synchronisation).  Returns 1 if it is enabled in this process; *ops = copies
 * it has served, *launches = how often the kernel had to be (re)started.  OCM_SERVER=0 disables it. */
int ocm_server_stats(uint64_t *ops, uint64_t *launches);

#ifdef __cplusplus
}
#endif
#endif /* ONCILLAMEM_H_ */
