// Wire format shared by the app<->daemon mailboxes and the daemon<->daemon
// sockets.  One versioned POD with fixed-width fields replaces the reference's
// compile-time-variant `struct message` that ships raw host pointers
// (reference inc/msg.h:24-73, inc/alloc.h:32-99).
// Message type / status numbering is the reference's; new types are appended.
#pragma once
#include <cstdint>
#include <cstring>

namespace ocm {

constexpr uint32_t kMsgMagic   = 0x4F434D32u; // "OCM2"
constexpr uint16_t kMsgVersion = 3;
constexpr int      kMaxRanks   = 64;
constexpr int      kMaxGpus    = 8;   // reference inc/alloc.h:30

enum MsgType : uint16_t {
    MSG_INVALID = 0,
    MSG_CONNECT,          // app -> daemon
    MSG_CONNECT_CONFIRM,  // daemon -> app
    MSG_DISCONNECT,       // app -> daemon
    MSG_ADD_NODE,         // ranks > 0 report to rank 0 at boot
    MSG_REQ_ALLOC,        // placement request (to rank 0)
    MSG_DO_ALLOC,         // fulfil on the chosen rank
    MSG_REQ_FREE,         // lib asks for a release
    MSG_DO_FREE,          // owner rank releases the region
    MSG_RELEASE_APP,      // request finished, wake the app
    MSG_ANY,
    MSG_MAX,
    // ---- new ----
    MSG_NOTE_FREE = 16,   // owner/origin -> rank 0: bookkeeping the reference never did (src/mem.c:271-278)
    MSG_STATS,            // counters query
    MSG_PING,
    MSG_SHUTDOWN,         // orderly daemon exit (tests)
    MSG_TYPE_END
};

enum MsgStatus : uint16_t { MSG_NO_STATUS = 0, MSG_REQUEST, MSG_RESPONSE };

// Allocation classes understood by the governor.  1..4 are the reference's
// (inc/alloc.h:32-42); RDMA and RMA both mean "remote pinned host" here.
enum AllocType : uint32_t {
    ALLOC_MEM_INVALID = 0,
    ALLOC_MEM_HOST,       // caller-local host
    ALLOC_MEM_RMA,        // legacy -> remote pinned host
    ALLOC_MEM_RDMA,       // legacy -> remote pinned host
    ALLOC_MEM_GPU,        // caller-local GPU
    ALLOC_MEM_REMOTE_GPU, // peer HBM
    ALLOC_MEM_REMOTE_HOST,// peer pinned DRAM
    ALLOC_MEM_MAX
};

enum Tier : int { TIER_GPU = 0, TIER_HOST = 1, TIER_COUNT = 2 };

struct AllocRequest {
    int32_t  orig_rank;
    int32_t  remote_rank;   // honoured for POLICY_EXPLICIT (a TODO in the reference, inc/alloc.h:49)
    uint64_t bytes;
    uint32_t type;          // AllocType
    uint32_t policy;        // enum ocm_policy
    uint32_t flags;
    uint32_t pad;
};

struct Allocation {
    int32_t  orig_rank;
    int32_t  remote_rank;
    uint64_t rem_alloc_id;  // unique per owner rank
    uint32_t type;          // AllocType actually granted
    uint32_t tier;          // Tier of the far side
    uint64_t bytes;         // requested size
    uint64_t offset;        // byte offset inside the owner's pool
    uint64_t block_bytes;   // size-class block actually reserved
    uint32_t region;        // region-table index (granule of the block start)
    uint32_t generation;
    uint32_t flags;
    uint32_t pad;
};

enum NodeCaps : uint8_t { NODE_CAP_MULTICAST = 1, NODE_CAP_VMM_FD = 2 };
struct NodeConfig {
    char     ip[48];        // ib_ip in the reference (inc/alloc.h:59)
    char     pci_bdf[16];   // PCI address of the rank's GPU ("0000:1b:00.0"), empty without one
    uint64_t ram;           // pinned-host pool bytes
    uint64_t gpu_mem[kMaxGpus];
    int32_t  num_gpu;
    int32_t  gpu_index;     // CUDA ordinal backing this rank, -1 if none
    int32_t  daemon_pid;
    // topology of the rank's GPU (common/topo.hpp): host NUMA node (-1 unknown), the CUDA ordinals it reaches
    // directly (bit per ordinal, kMaxGpus = 8), NodeCaps
    int16_t  numa_node;
    uint8_t  peer_mask;
    uint8_t  caps;
};

struct StatsPayload {
    uint64_t allocs_live, allocs_total, frees_total;
    uint64_t pool_bytes[TIER_COUNT], pool_used[TIER_COUNT];
    uint64_t apps, msgs_in, msgs_out, reclaimed_regions, scrubbed_bytes;
    uint64_t pool_granule_bytes[TIER_COUNT];   // granules taken (slab remainders included)
    uint64_t table_faults;      // audits that found the rank's tables inconsistent (daemon.cpp, audit_tables)
};

struct ConnectInfo {
    int32_t rank, n_ranks;
    int32_t gpu_index;
    int32_t daemon_pid;
    char    ns[32];         // IPC namespace tag (names of shm segments / sockets)
};

struct Message {
    uint32_t magic;
    uint16_t version;
    uint16_t type;      // MsgType
    uint16_t status;    // MsgStatus
    uint16_t hops;
    int32_t  pid;       // app that made the request
    int32_t  rank;      // rank where the request originated
    uint32_t seq;       // request id, echoed in every reply
    int32_t  err;       // 0 or enum ocm_err
    union {
        AllocRequest req;
        Allocation   alloc;
        NodeConfig   node;
        StatsPayload stats;
        ConnectInfo  conn;
        uint8_t      raw[160];
    } u;

    void init(MsgType t, MsgStatus s = MSG_REQUEST) {
        memset(this, 0, sizeof *this);
        magic = kMsgMagic; version = kMsgVersion; type = t; status = s;
    }
    bool valid() const { return magic == kMsgMagic && version == kMsgVersion && type < MSG_TYPE_END; }
};
static_assert(sizeof(Message) == 192, "wire size is part of the protocol");
static_assert(sizeof(NodeConfig) <= 160 && sizeof(Allocation) <= 160, "payload must fit");

inline const char *msg_type_str(uint16_t t) {
    switch (t) {
    case MSG_INVALID: return "MSG_INVALID";
    case MSG_CONNECT: return "MSG_CONNECT";
    case MSG_CONNECT_CONFIRM: return "MSG_CONNECT_CONFIRM";
    case MSG_DISCONNECT: return "MSG_DISCONNECT";
    case MSG_ADD_NODE: return "MSG_ADD_NODE";
    case MSG_REQ_ALLOC: return "MSG_REQ_ALLOC";
    case MSG_DO_ALLOC: return "MSG_DO_ALLOC";
    case MSG_REQ_FREE: return "MSG_REQ_FREE";
    case MSG_DO_FREE: return "MSG_DO_FREE";
    case MSG_RELEASE_APP: return "MSG_RELEASE_APP";
    case MSG_ANY: return "MSG_ANY";
    case MSG_MAX: return "MSG_MAX";
    case MSG_NOTE_FREE: return "MSG_NOTE_FREE";
    case MSG_STATS: return "MSG_STATS";
    case MSG_PING: return "MSG_PING";
    case MSG_SHUTDOWN: return "MSG_SHUTDOWN";
    default: return "INVALID MSG TYPE";
    }
}
inline const char *msg_status_str(uint16_t s) {
    switch (s) {
    case MSG_NO_STATUS: return "MSG_NO_STATUS";
    case MSG_REQUEST: return "MSG_REQUEST";
    case MSG_RESPONSE: return "MSG_RESPONSE";
    default: return "INVALID MSG STATUS";
    }
}
inline const char *alloc_type_str(uint32_t t) {
    static const char *n[] = {"INVALID", "HOST", "RMA", "RDMA", "GPU", "REMOTE_GPU", "REMOTE_HOST"};
    return t < ALLOC_MEM_MAX ? n[t] : "?";
}

} // namespace ocm
