// Internal state of libocm: daemon attachment, per-rank mapping cache,
// allocation records.  Counterpart of the globals in reference src/lib.c:36-91.
#pragma once
#include "../common/gpu_rt.hpp"
#include "../common/msg.hpp"
#include "../common/tables.hpp"
#include "../ipc/pmsg.hpp"
#include "../kernels/kernels.h"
#include "../../include/oncillamem.h"
#include <atomic>
#include <cstdlib>
#include <condition_variable>
#include <unordered_map>
#include <mutex>
#include <unordered_set>
#include <vector>

namespace ocm {

constexpr uint32_t kAllocMagic = 0x0C3A110Cu;

// Everything this process knows about one rank: its tables and pools mapped
// into our address space.  Filled lazily, kept until ocm_tini ("mapping cache":
// after the first touch a remote region is plain pointer arithmetic).
struct RankView {
    bool         tried = false, ok = false;
    ShmMapping   tbl_map;
    TableHeader *hdr = nullptr;
    ShmMapping   host_map;
    gpu::PoolMapping host_vmm;              // host pool imported as a pinned VMM allocation (host_map.base aliases it)
    bool         host_registered = false;   // at least one chunk of the host pool is pinned + mapped for our GPU
    std::vector<uint8_t> host_chunk_reg;    // per chunk of the host pool: registered for this process (api.cpp)
    bool         gpu_tried = false;
    gpu::PoolMapping gpu;       // ptr == nullptr if the rank has no HBM pool or we have no GPU
    uint64_t     access_mask = 0;   // devices of this process that may touch `gpu` (bit = CUDA ordinal)
};

// Host half of the resident copy server (H2; kernels/server.cu, lib/server.cpp).
constexpr uint32_t kServerRing = 64;
struct CopyServer {
    bool        enabled = false;
    void       *stream = nullptr;           // the stream the resident kernel lives on
    ocmk::ServerDesc *ring = nullptr;       // pinned + mapped
    ocmk::ServerCtl  *ctl = nullptr;
    std::mutex  mu;                         // posting threads, one at a time
    uint64_t    next_seq = 0;
    uint64_t    max_bytes = 0, idle_ns = 0, max_ns = 0;
    std::atomic<uint64_t> n_ops{0}, n_launches{0};
};

// A second GPU this process launches kernels on: the OWNER of a far side, so that a large put runs as
// loads over NVLink from the destination GPU ("pull") instead of stores from the source GPU ("push").
struct PeerExec {
    bool      tried = false, ok = false;
    int       dev = -1;
    void     *stream = nullptr;
    void     *ev_in[16] = {nullptr}, *ev_out[16] = {nullptr};
    unsigned  cursor = 0;
    unsigned *tickets = nullptr;
};

enum SideBacking : uint8_t { SIDE_NONE = 0, SIDE_POOL = 1, SIDE_HEAP = 2, SIDE_CUDAMALLOC = 3, SIDE_USER = 4,
                             SIDE_MC = 5 /* replica inside the NVLS multicast region */,
                             SIDE_ATTACHED = 6 /* another application's pool block (ocm_attach); never released here */ };

// One contiguous piece of memory addressable from this process.
struct Side {
    SideBacking backing = SIDE_NONE;
    uint8_t     tier = TIER_HOST;
    bool        via_daemon = false;   // allocated through the message protocol
    int         rank = -1;            // owner rank (pool backing)
    uint32_t    region = 0;
    uint32_t    generation = 0;
    uint64_t    id = 0;
    uint64_t    offset = 0;
    uint64_t    bytes = 0;            // usable (requested) bytes
    void       *ptr = nullptr;
    Region     *rec = nullptr;        // table record
    TableHeader *hdr = nullptr;       // the table that holds rec (for the signal total)
    // The completion counters THIS handle adds to.  One writing agent per word (tables.hpp, SigExt): the
    // owner's handle uses Region::signal / signal_cpu, an attached handle the slot of its rank's GPU / the
    // CPU slot of the block's extension record.
    std::atomic<uint64_t> *sig_gpu = nullptr;
    std::atomic<uint64_t> *sig_cpu = nullptr;
    uint64_t *gpu_signal() const { return reinterpret_cast<uint64_t *>(sig_gpu); }
    bool valid() const { return backing != SIDE_NONE; }
};

// Size of the pieces in which a process pins a rank's host pool (api.cpp view_host_ready).  A library copy between
// host and GPU memory must not straddle two pieces: they may be separate registrations (copy.cpp single_copy).
inline uint64_t host_reg_chunk(uint64_t pool_bytes) {
    static const uint64_t env = [] {
        const char *e = getenv("OCM_HOST_REG_CHUNK");
        uint64_t b = e && *e ? strtoull(e, nullptr, 0) : 0;
        if (b && b < (2ull << 20)) b = 2ull << 20;
        return (b + 4095) & ~4095ull;
    }();
    if (env) return env;
    // small pools are pinned whole (one registration: any copy, also one issued by the application itself, may span
    // the pool); large ones in 256 MiB pieces
    return pool_bytes <= (2ull << 30) ? (pool_bytes ? ((pool_bytes + 4095) & ~4095ull) : 4096) : (256ull << 20);
}

// An attached side (ocm_attach) goes stale when the owner frees the block.
inline bool side_stale(const Side &s) {
    return s.backing == SIDE_ATTACHED &&
           !(s.rec->state.load(std::memory_order_acquire) == REGION_LIVE && s.rec->id == s.id &&
             s.rec->generation == s.generation);
}

} // namespace ocm

// The opaque handle of the public API.
struct lib_alloc {
    uint32_t      magic;
    enum ocm_kind kind;          // granted kind
    bool          paired;
    ocm::Side     local;         // local kinds: the memory; pairs: the window
    ocm::Side     remote;        // pairs: the far side
    std::vector<ocm::Side> stripes;   // striped / replicated kinds: one per rank
    uint64_t      stripe_bytes;
    bool          mc;                 // replicated through the NVLS multicast region
    uint64_t      mc_offset;          // offset of this allocation inside the region
    void         *mc_ptr;             // its multicast virtual address
    // completion word of a multicast-backed allocation (tables.hpp, kMcSignalGranules): its multicast address (a
    // multimem.red there bumps the word on every GPU) and the copy of it in the replica nearest to this process
    uint64_t     *mc_sig = nullptr;
    uint64_t     *mc_sig_near = nullptr;
    bool          mc_attached = false;   // opened with ocm_attach: the range belongs to another application
    std::atomic<bool> lib_stream_pending;
};

namespace ocm {

struct Ctx {
    std::mutex   mu;             // init/tini + allocation set
    // Daemon exchanges: mq_mu guards the sequence counter, the table of waiting threads and the "receiver" role
    // (the one waiter that currently reads the mailbox and distributes replies by sequence number).
    struct PendingReply {
        Message  rep;
        uint16_t want_type = 0;
        bool     ready = false;
    };
    std::mutex   mq_mu;
    std::condition_variable mq_cv;
    std::unordered_map<uint32_t, PendingReply *> mq_pending;
    bool         mq_receiving = false;
    bool         inited = false;
    int          rank = 0, n_ranks = 1;
    int          daemon_pid = -1;
    int          device = -1;    // CUDA device of this process (-1: none)
    bool         force_slow = false;
    Mailbox      box;
    std::string  daemon_mq;
    uint32_t     seq = 0;
    ShmMapping   cluster_map;
    ClusterTable *cluster = nullptr;
    RankView     view[kMaxRanks];
    // Views of ranks whose daemon has been restarted since: the mappings stay (handles made before the restart
    // still point into them) but nothing new is placed there.  Released in ocm_tini.
    std::vector<RankView> orphaned;
    std::unordered_set<lib_alloc *> allocs;
    void        *stream = nullptr;      // library copy stream
    unsigned    *tickets = nullptr;     // device-local election words
    std::atomic<uint32_t> ticket_cursor{0};
    // counters
    std::atomic<uint64_t> n_alloc{0}, n_free{0}, n_fast{0}, n_slow{0}, n_copies{0}, copy_bytes{0};
    std::atomic<unsigned long long> kernel_launches{0};
    std::atomic<uint64_t> bytes_live[TIER_COUNT];
    PeerExec     peer_exec[kMaxGpus];
    // Completion words of blocking calls: pinned, device-visible, one 64-byte slot per calling thread (a thread
    // claims a slot for its lifetime).  The kernel of a blocking call stores that thread's next sequence number
    // into the thread's own slot, so one thread's kernel can never satisfy another thread's wait and a slot only
    // ever counts upwards (kernels of one thread run in launch order on the library stream).
    uint64_t    *done_word = nullptr;          // kDoneSlots slots of 8 words each
    std::atomic<uint64_t> done_slots_free{0};  // bit s set: slot s is free
    uint64_t     done_seq[64] = {0};           // last sequence number handed out per slot (owner thread only)
    CopyServer   server;
    // *_async calls on the library stream count up `issued`; waiting for that stream moves `synced` up to the count
    // seen before the wait began.  While synced < issued a blocking copy must queue behind that work on the stream
    // instead of overtaking it through the copy server.
    std::atomic<uint64_t> lib_async_issued{0}, lib_async_synced{0};
    void note_lib_async() { lib_async_issued.fetch_add(1, std::memory_order_acq_rel); }
    // everything counted up to `snap` (taken BEFORE the work that was waited for was enqueued) has completed
    void note_lib_synced(uint64_t snap) {
        uint64_t cur = lib_async_synced.load(std::memory_order_relaxed);
        while (cur < snap && !lib_async_synced.compare_exchange_weak(cur, snap, std::memory_order_acq_rel)) {}
    }
    bool lib_async_pending() const {
        return lib_async_synced.load(std::memory_order_acquire) < lib_async_issued.load(std::memory_order_acquire);
    }
    gpu::McObject mc;            // rank 0's multicast object mapped for our device
    bool         mc_tried = false;
    // Puts at least this large are pulled by the owner GPU (0: never, the default).  Measured on 2xB200:
    // pulling reaches 765 GB/s at 1 GiB where pushing stops at 705, but only while the owner GPU is not
    // busy with ANOTHER process' kernels - with one application per GPU the two contexts time-slice and
    // the aggregate halves (profiles/README.md).  Opt in with OCM_PULL_MIN=<bytes> for single-initiator use.
    uint64_t     pull_min = 0;
    std::atomic<uint64_t> n_pulled{0};
};
constexpr int kNumTickets = 1024;
constexpr int kDoneSlots = 64;
constexpr int kDoneSlotWords = 8;

Ctx &ctx();
void set_last_error(int e);
// Map (once) the tables and host pool of `rank`; with_gpu also maps its HBM pool.
RankView *get_view(int rank, bool with_gpu);
// Exchange one message with the local daemon.  0 or positive ocm_err.
int daemon_call(Message &m, int timeout_ms = 10000);
unsigned *next_ticket();

// resident copy server (server.cpp)
void server_init();
void server_shutdown();
bool server_eligible(uint64_t bytes);
bool server_post(const void *src, void *dst, void *dst2, uint64_t bytes, uint64_t *signal, uint64_t *seq_out);
int  server_wait(uint64_t seq);

// copy engine (copy.cpp)
int side_copy(const Side &dst, uint64_t dst_off, const Side *dst2, uint64_t dst2_off, const Side &src, uint64_t src_off,
              uint64_t bytes, const Side *sig, void *stream, bool *used_stream);
struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    explicit DeviceGuard(int dev);
    ~DeviceGuard();
};

} // namespace ocm
