// Lock-free, device-visible allocation tables.
//
// Every rank owns one shared-memory segment ("/ocm_<ns>tbl_r<rank>") holding a
// header, one PoolDesc per tier (GPU HBM pool, pinned-host pool) and one Region
// record per granule of each pool.  Rank 0 additionally owns the cluster table
// ("/ocm_<ns>cluster") with per-rank capacity accounting.  All mutable words are
// 32/64-bit atomics, so the daemon, every attached application and - once the
// segment is cudaHostRegister'ed - device code can read and update them
// without locks and without a daemon round trip.  ocm_alloc on the fast path
// is therefore a handful of CAS operations and never issues a CUDA driver call
// that could serialise with an in-flight copy (the reference calls
// cudaMalloc/cudaFree per allocation, reference src/lib.c:240,365, and keeps
// rank-0 bookkeeping in a mutex-protected list it never prunes,
// src/alloc.c:133-137, src/mem.c:220-229).
//
// Allocator: one bit per granule in a shared bitmap.  A block is a run of
// granules claimed word by word with compare-and-swap (rolled back if a later
// word is contended); releasing clears the bits, so free space coalesces by
// construction and a pool never fragments into unusable size classes.  Freed
// blocks first pass through a "dirty" Treiber stack (ABA tag in the upper half
// of the head word) so that the owner daemon can zero them before their bits
// are cleared (the reference zero-fills remote buffers at allocation,
// src/alloc.c:171, src/extoll_client.c:46 - here the cost is paid off the
// allocation path).
#pragma once
#include "msg.hpp"
#include <atomic>
#include <cstddef>
#include <cstdint>
#include <string>
#include <utility>
#include <vector>

namespace ocm {

constexpr uint64_t kTableMagic   = 0x4F434D5441424C45ull; // "OCMTABLE"
constexpr uint32_t kTableVersion = 8;
constexpr uint32_t kMaxGranules  = 1u << 18;
constexpr uint64_t kMcGranule     = 64ull << 10;
constexpr uint32_t kMcMaxGranules = 1u << 14;     // multicast region <= 1 GiB
// The first granules of the multicast region hold one 64-bit completion word per granule of the region (word g
// belongs to the allocation that STARTS at granule g).  They are part of the region, i.e. bound to the multicast
// object: a multimem.red on the word's multicast address bumps the word on every GPU at once.
constexpr uint32_t kMcSignalGranules = (uint32_t)((uint64_t)kMcMaxGranules * 8 / kMcGranule);   // 2
enum McState : uint32_t { MC_NONE = 0, MC_CREATED = 1, MC_READY = 2, MC_FAILED = 3 };

enum RegionState : uint32_t { REGION_FREE = 0, REGION_LIVE = 1, REGION_DIRTY = 2, REGION_SCRUBBING = 3 };

struct alignas(64) Region {
    std::atomic<uint32_t> state;
    uint32_t              n_granules;  // length of the block starting here (valid while not FREE)
    std::atomic<uint32_t> next;        // dirty-stack link: granule index + 1, 0 = end
    uint32_t              generation;
    int32_t               owner_pid;
    int32_t               owner_rank;
    uint64_t              bytes;       // requested bytes
    uint64_t              id;          // rem_alloc_id
    std::atomic<uint64_t> signal;      // completion counter: puts issued by the OWNER's GPU add 1
    std::atomic<uint64_t> signal_cpu;  //   "      "     : copies done by CPUs (host <-> host) add 1
    std::atomic<uint32_t> attached;    // handles other applications hold on this block (ocm_attach)
    std::atomic<uint32_t> ext;         // shared blocks: index + 1 of the SigExt record, 0 = none
};
static_assert(sizeof(Region) == 64, "one cache line per region");

// ---- slabs: size classes below the granule --------------------------------------------------------
// A pool's granule grows with the pool (64 KiB for HBM pools up to 16 GiB, 1 MiB at 160 GiB), so a 4 KiB request
// would burn a whole granule.  Requests of at most half a granule are therefore served from SLABS: one granule cut
// into 2^k equal sub-blocks (k = 1..6, sub-blocks of at least 256 bytes), its free sub-blocks a 64-bit mask claimed
// with compare-and-swap.  Slab i always serves class i % kSlabClasses, so a mask bit means the same thing in every
// incarnation of slab i and a stale CAS can only ever claim a block that really is free.  Every sub-block has a
// full Region record of its own (owner, id, generation, completion signals, share/attach state): it is an
// allocation like any other, named by kSlabFlag | slab << 6 | sub wherever a granule index is expected.
// Runs of at least this many granules are placed first fit from the bottom of the pool, shorter ones next fit from a
// rotating start, single granules from the top (tables.cpp granule_alloc; measured in unit_host frag).
constexpr uint32_t kFirstFitMinGranules = 16;
constexpr uint32_t kSlabFlag    = 0x80000000u;
constexpr uint32_t kSlabSubs    = 64;
constexpr uint32_t kSlabClasses = 6;       // class c: 2^(c+1) sub-blocks of granule >> (c+1) bytes
constexpr uint32_t kMaxSlabs    = 1536;    // per tier: 256 slabs per class
enum SlabState : uint32_t { SLAB_FREE = 0, SLAB_INIT = 1, SLAB_ACTIVE = 2, SLAB_RETIRING = 3 };
struct alignas(64) Slab {
    std::atomic<uint32_t> state;
    uint32_t              region;       // backing granule (valid while ACTIVE)
    uint64_t              backing_id;   // allocation id of the backing block
    std::atomic<uint64_t> free_mask;    // bit s: sub-block s is free and clean
    uint64_t              pad[5];
};
static_assert(sizeof(Slab) == 64, "one cache line per slab");
inline bool     region_is_sub(uint32_t region) { return (region & kSlabFlag) != 0; }
inline uint32_t sub_slab(uint32_t region) { return (region & ~kSlabFlag) >> 6; }
inline uint32_t sub_index(uint32_t region) { return region & 63u; }
inline uint32_t slab_class(uint32_t slab) { return slab % kSlabClasses; }
inline uint32_t class_subs(uint32_t cls) { return 2u << cls; }

// Completion counters of a block that is shared between applications (ocm_share / ocm_attach).
// A system-scope reduction from a GPU to host memory is only atomic with respect to the SAME GPU on boxes
// without PCIe atomics (cudaDevAttrHostNativeAtomicSupported = 0 on the B200 HGX boxes: four GPUs adding
// to one pinned word lost 25 % of the updates, with a CPU thread joining in 75 % - bench/atomics_probe.cu).
// So every word has exactly one writing agent: slot r is written by the GPU of rank r, the last slot by
// CPUs (coherent among themselves); the reader adds them up.  The owner's own GPU keeps Region::signal.
constexpr int kSigGpuSlots = kMaxRanks;
constexpr int kSigCpuSlot  = kMaxRanks;
struct alignas(64) SigExt {
    std::atomic<uint64_t> owner;       // 0 = free, else (tier << 32 | region) + 1
    uint64_t              id;          // allocation id of the block it serves
    uint64_t              pad[6];
    std::atomic<uint64_t> slot[kMaxRanks + 8];
};
constexpr uint32_t kSigExtRecords = 1024;

struct PoolDesc {
    uint64_t bytes;            // capacity
    uint64_t granule;          // bytes per granule (power of two)
    uint32_t n_granules;
    uint32_t tier;
    uint64_t regions_offset;   // from segment start
    uint64_t bitmap_offset;    // from segment start; one bit per granule, 1 = taken
    uint32_t bitmap_words;
    uint32_t pad;
    std::atomic<uint64_t> hint;          // lower bound of the lowest free granule: where the first-fit search for long runs starts
    std::atomic<uint64_t> rot_hint;      // where the next-fit search for short runs starts (rotating)
    std::atomic<uint64_t> used_bytes;    // sum of live block bytes
    std::atomic<uint64_t> live;
    std::atomic<uint64_t> total_allocs;
    std::atomic<uint64_t> total_frees;
    std::atomic<uint64_t> dirty_head;    // {tag:32, idx+1:32}
    std::atomic<uint64_t> dirty_bytes;
    // slabs (see above): headers, one Region per sub-block, their own dirty stack, free sub-blocks per class
    uint64_t slabs_offset;               // Slab[n_slabs], from segment start
    uint64_t subs_offset;                // Region[n_slabs * kSlabSubs]
    uint32_t n_slabs;
    uint32_t pad_slab;
    std::atomic<uint64_t> sub_dirty_head;
    std::atomic<uint64_t> sub_live;      // live sub-blocks
    std::atomic<uint64_t> sub_bytes;     // bytes of live sub-blocks (their class size)
    std::atomic<int64_t>  class_free[kSlabClasses];
    std::atomic<uint32_t> class_hint[kSlabClasses];
};

enum GpuShareMode : uint32_t { GPU_SHARE_NONE = 0, GPU_SHARE_VMM_FD = 1, GPU_SHARE_CUDA_IPC = 2 };

struct TableHeader {
    uint64_t magic;
    uint32_t version;
    int32_t  rank;
    int32_t  n_ranks;
    int32_t  daemon_pid;
    int32_t  gpu_index;        // CUDA ordinal of the pool's GPU, -1 if none
    uint32_t gpu_share_mode;   // GpuShareMode
    uint8_t  cuda_ipc_handle[64];
    uint8_t  gpu_uuid[16];
    char     ns[32];
    uint64_t segment_bytes;
    std::atomic<uint32_t> ready;       // set last by the daemon
    std::atomic<uint32_t> scrub_enabled;
    std::atomic<uint32_t> slabs_enabled;   // sub-granule size classes (OCM_NO_SLABS clears it)
    uint32_t host_share_mode;          // how applications map the host pool: GPU_SHARE_NONE = the shared-memory segment,
                                       // GPU_SHARE_VMM_FD = pinned VMM allocation (CUDA host NUMA location), fd from the daemon
    std::atomic<uint64_t> next_id;
    std::atomic<uint64_t> heartbeat_ns; // daemon liveness (CLOCK_MONOTONIC)
    uint64_t mc_tail_offset;           // NVLS: byte offset inside the HBM pool mapping of the multicast slice
    uint64_t mc_bytes;                 //       its length (0: none)
    uint64_t sigext_offset;            // SigExt[kSigExtRecords], from segment start
    uint32_t n_sigext;
    int32_t  host_numa_node;           // NUMA node of a VMM host pool (its CUDA memory location id)
    PoolDesc pool[TIER_COUNT];
};

struct NodeSlot {
    std::atomic<uint32_t> joined;
    int32_t  daemon_pid;
    int32_t  gpu_index;
    int32_t  num_gpu;
    uint64_t pool_bytes[TIER_COUNT];
    std::atomic<uint64_t> used[TIER_COUNT];   // reserved bytes, rank-0 capacity accounting
    std::atomic<uint64_t> live;
    uint64_t gpu_mem[kMaxGpus];
    char     ip[48];
    // Topology reported at join (common/topo.hpp): PCI address and host NUMA node of the rank's GPU (-1: the
    // kernel does not know, e.g. inside a VM), which CUDA ordinals that GPU reaches directly, NodeCaps bits.
    // governor_place keeps HBM far sides off GPUs the requester's GPU cannot reach.
    char     pci_bdf[16];
    int32_t  numa_node;
    uint32_t peer_mask;
    uint32_t caps;
    uint32_t pad_topo;
    // Liveness: the rank's daemon stores CLOCK_MONOTONIC here on every timer tick (20 ms).  0 = has not
    // beaten yet (just joined, or a table without daemons in the unit tests) and counts as alive.
    std::atomic<uint64_t> heartbeat_ns;
};
constexpr uint64_t kRankDeadAfterNs = 1500ull * 1000000ull;

struct ClusterTable {
    uint64_t magic;
    uint32_t version;
    int32_t  n_ranks;
    std::atomic<uint32_t> ready;
    uint32_t pad;
    std::atomic<uint64_t> rr_cursor;
    std::atomic<uint64_t> total_allocs;
    std::atomic<uint64_t> total_frees;
    std::atomic<uint64_t> rejected;
    NodeSlot node[kMaxRanks];
    // NVLS multicast region: the same byte range on every rank's GPU, bound to one multicast object
    // (daemon/daemon.cpp).  Replicated allocations are carved out of it with the bitmap below.
    std::atomic<uint32_t> mc_state;      // McState
    uint32_t mc_n_granules;
    uint64_t mc_bytes;
    std::atomic<uint32_t> mc_added;      // daemons that added their device
    std::atomic<uint32_t> mc_bound;      // daemons that bound their slice
    std::atomic<uint64_t> mc_bitmap[kMcMaxGranules / 64];
    // Owner of the range that STARTS at granule g (0: none), packed by mc_owner_pack: when the owning process
    // dies the daemon of its rank marks the range dead, every daemon zeroes its own GPU's slice and sets its bit
    // in mc_scrubbed; the daemon that completes the set returns the range to the bitmap.
    std::atomic<uint64_t> mc_owner[kMcMaxGranules];
    std::atomic<uint64_t> mc_scrubbed[kMcMaxGranules];
};

enum McOwnerState : uint32_t { MC_RANGE_NONE = 0, MC_RANGE_OWNED = 1, MC_RANGE_DEAD = 2 };
inline uint64_t mc_owner_pack(int pid, uint32_t n_granules, int rank, uint32_t state) {
    return ((uint64_t)(uint32_t)pid << 32) | ((uint64_t)(n_granules & 0xffffu) << 16) | ((uint64_t)(rank & 0xff) << 8) | (state & 0xffu);
}
inline int      mc_owner_pid(uint64_t w) { return (int)(uint32_t)(w >> 32); }
inline uint32_t mc_owner_granules(uint64_t w) { return (uint32_t)(w >> 16) & 0xffffu; }   // kMcMaxGranules = 2^14 fits
inline int      mc_owner_rank(uint64_t w) { return (int)((w >> 8) & 0xff); }
inline uint32_t mc_owner_state(uint64_t w) { return (uint32_t)(w & 0xff); }

// Sub-allocation inside the multicast region (any process).  Returns the byte offset or -1.
int64_t mc_range_alloc(ClusterTable *c, uint64_t bytes, int owner_pid = 0, int owner_rank = 0);
void    mc_range_free(ClusterTable *c, uint64_t offset, uint64_t bytes);
// Daemon of `rank`: mark the ranges of dead owners (pred(pid) true) on that rank as MC_RANGE_DEAD.  Returns how many.
size_t  mc_mark_dead(ClusterTable *c, int rank, bool (*pred)(int pid));
// Any daemon: the dead ranges `rank` has not zeroed yet, as (start granule, granules).
size_t  mc_dead_ranges(ClusterTable *c, int rank, std::vector<std::pair<uint32_t, uint32_t>> *out);
// `rank` has zeroed its slice of the dead range at `start`; frees the range when every rank has.  True if freed.
bool    mc_range_scrubbed(ClusterTable *c, int rank, uint32_t start);

uint64_t default_granule(uint64_t pool_bytes, Tier tier);
inline std::atomic<uint64_t> *table_bitmap(TableHeader *h, int tier) {
    return reinterpret_cast<std::atomic<uint64_t> *>(reinterpret_cast<char *>(h) + h->pool[tier].bitmap_offset);
}

// ---- segment lifecycle (POSIX shm) ---------------------------------------
struct ShmMapping {
    void  *base = nullptr;
    size_t bytes = 0;
    int    fd = -1;
    bool   owner = false;
    std::string name;
};
// Create (O_EXCL unless replace) and size a segment; zero-filled.
bool shm_create(const std::string &name, size_t bytes, bool replace, ShmMapping *out, std::string *err);
bool shm_open_existing(const std::string &name, ShmMapping *out, std::string *err);
void shm_close(ShmMapping *m, bool unlink_it);

size_t table_segment_bytes(uint64_t gpu_pool_bytes, uint64_t host_pool_bytes,
                           uint64_t gpu_granule = 0, uint64_t host_granule = 0);
// Lay out a fresh (zeroed) segment.  Does not set ready.
void table_format(void *base, size_t seg_bytes, int rank, int n_ranks, int daemon_pid,
                  uint64_t gpu_pool_bytes, uint64_t host_pool_bytes,
                  uint64_t gpu_granule = 0, uint64_t host_granule = 0);
inline Region *table_regions(TableHeader *h, int tier) {
    return reinterpret_cast<Region *>(reinterpret_cast<char *>(h) + h->pool[tier].regions_offset);
}
inline Slab *table_slabs(TableHeader *h, int tier) {
    return reinterpret_cast<Slab *>(reinterpret_cast<char *>(h) + h->pool[tier].slabs_offset);
}
inline Region *table_subs(TableHeader *h, int tier) {
    return reinterpret_cast<Region *>(reinterpret_cast<char *>(h) + h->pool[tier].subs_offset);
}
// The record of a block named by a granule index or by a slab sub-block reference; nullptr when out of range.
inline Region *table_region_rec(TableHeader *h, int tier, uint32_t region) {
    const PoolDesc &p = h->pool[tier];
    if (!region_is_sub(region)) return region < p.n_granules ? &table_regions(h, tier)[region] : nullptr;
    const uint32_t s = sub_slab(region);
    if (s >= p.n_slabs || sub_index(region) >= class_subs(slab_class(s))) return nullptr;
    return &table_subs(h, tier)[(size_t)s * kSlabSubs + sub_index(region)];
}
// Byte offset of that block inside the pool (sub-blocks: only meaningful while their slab is active).
inline uint64_t table_region_offset(TableHeader *h, int tier, uint32_t region) {
    const PoolDesc &p = h->pool[tier];
    if (!region_is_sub(region)) return (uint64_t)region * p.granule;
    const uint32_t s = sub_slab(region);
    return (uint64_t)table_slabs(h, tier)[s].region * p.granule + (uint64_t)sub_index(region) * (p.granule >> (slab_class(s) + 1));
}
bool table_validate(const TableHeader *h, size_t mapped_bytes);
inline SigExt *table_sigext(TableHeader *h) {
    return reinterpret_cast<SigExt *>(reinterpret_cast<char *>(h) + h->sigext_offset);
}
// Give a LIVE block its extension record (idempotent).  Returns the record or nullptr when all are taken.
SigExt *sigext_claim(TableHeader *h, int tier, uint32_t region);
// Sum of every completion counter of a block (owner GPU + CPUs + all extension slots).
uint64_t region_signal_total(TableHeader *h, const Region &r);

// ---- lock-free pool operations (any process) ------------------------------
struct BlockRef {
    uint32_t region = 0;       // granule index of the block start
    uint32_t generation = 0;
    uint64_t offset = 0;       // byte offset in the pool
    uint64_t block_bytes = 0;
    uint64_t id = 0;
};
// 0 on success, OCM_ENOMEM / OCM_EINVAL (positive enum ocm_err values) otherwise.
int pool_alloc(TableHeader *h, int tier, uint64_t bytes, int pid, int rank, BlockRef *out);
// Releases a LIVE block whose id matches.  Returns OCM_ESTALE on double free /
// id mismatch.  With scrubbing enabled the block is queued for the daemon.
int pool_free(TableHeader *h, int tier, uint32_t region, uint64_t id);

// ---- daemon-side maintenance ----------------------------------------------
struct DirtyBlock { uint32_t region; uint64_t offset; uint64_t bytes; };
// Detach the whole dirty stack (blocks become SCRUBBING).
size_t pool_take_dirty(TableHeader *h, int tier, std::vector<DirtyBlock> *out);
// Return a scrubbed block to the pool (clears its bits).
void pool_release_scrubbed(TableHeader *h, int tier, uint32_t region);
// Any process that maps the pinned-host pool may help the owner: detach the dirty
// host blocks, zero them and release them.  Returns bytes scrubbed.
uint64_t pool_scrub_host_inline(TableHeader *h, void *host_pool_base);
// True while released blocks are still waiting to be zeroed in this tier.
inline bool pool_has_dirty(const TableHeader *h, int tier) {
    return h->pool[tier].dirty_bytes.load(std::memory_order_acquire) != 0;
}
// Walk carved blocks; for LIVE blocks call pred(owner_pid, owner_rank) and free
// those for which it returns true.  Returns blocks reclaimed; adds their bytes
// to *bytes_out.
size_t pool_reclaim(TableHeader *h, int tier, bool (*pred)(int pid, int rank, void *ctx), void *ctx,
                    uint64_t *bytes_out, std::vector<BlockRef> *reclaimed = nullptr);
// Count LIVE blocks (O(blocks)); consistency checks for tests.  Sub-blocks count, the granules backing slabs do not.
size_t pool_count_live(TableHeader *h, int tier);
// Bytes of granules in use as seen by applications: used_bytes without the free remainder of active slabs.
uint64_t pool_payload_bytes(TableHeader *h, int tier);
// Daemon-side audit: do the counters still agree with the bitmap and the records?  Transient disagreement is
// normal while other processes allocate; a caller that sees the same non-zero result twice with unchanged
// total_allocs/total_frees has found a corrupted table (a misbehaving client).  Returns the number of problems.
struct TableAudit {
    uint64_t bitmap_bytes, used_bytes, dirty_bytes, live_records, live_counter;
    int problems;
};
int pool_audit(TableHeader *h, int tier, TableAudit *out);

// ---- cluster table / governor primitives -----------------------------------
void cluster_format(ClusterTable *c, int n_ranks);
// Reserve capacity on `rank`.  false if it would exceed the pool.
bool cluster_reserve(ClusterTable *c, int rank, int tier, uint64_t bytes);
void cluster_unreserve(ClusterTable *c, int rank, int tier, uint64_t bytes);

} // namespace ocm
