#include "governor.hpp"
#include "log.hpp"
#include "../../include/oncillamem.h"
#include <algorithm>
#include <cstring>
#include <ctime>

namespace ocm {

int governor_joined(const ClusterTable *c) {
    int n = 0;
    for (int r = 0; r < c->n_ranks && r < kMaxRanks; ++r)
        if (c->node[r].joined.load(std::memory_order_acquire)) ++n;
    return n;
}

int governor_add_node(ClusterTable *c, int rank, const NodeConfig &cfg, const uint64_t pool_bytes[TIER_COUNT]) {
    if (rank < 0 || rank >= c->n_ranks || rank >= kMaxRanks) return OCM_EINVAL;
    NodeSlot &n = c->node[rank];
    bool rejoin = n.joined.load(std::memory_order_acquire) != 0;
    // The same daemon asking again keeps its accounting; a NEW daemon process on that rank (restart after a
    // crash) comes with empty pools, so what the old incarnation had handed out no longer counts.
    const bool new_incarnation = rejoin && n.daemon_pid != cfg.daemon_pid;
    if (new_incarnation) rejoin = false;
    n.daemon_pid = cfg.daemon_pid;
    n.gpu_index = cfg.gpu_index;
    n.num_gpu = cfg.num_gpu;
    memcpy(n.gpu_mem, cfg.gpu_mem, sizeof n.gpu_mem);
    memcpy(n.ip, cfg.ip, sizeof n.ip);
    memcpy(n.pci_bdf, cfg.pci_bdf, sizeof n.pci_bdf);
    n.pci_bdf[sizeof n.pci_bdf - 1] = 0;
    n.numa_node = cfg.numa_node;
    n.peer_mask = cfg.peer_mask;
    n.caps = cfg.caps;
    for (int t = 0; t < TIER_COUNT; ++t) n.pool_bytes[t] = pool_bytes[t];
    if (!rejoin) {
        for (int t = 0; t < TIER_COUNT; ++t) n.used[t].store(0, std::memory_order_relaxed);
        n.live.store(0, std::memory_order_relaxed);
    }
    n.heartbeat_ns.store(0, std::memory_order_relaxed);   // alive until the first beat says otherwise
    n.joined.store(1, std::memory_order_release);
    LOGI("rank %d %s: gpu=%d pool gpu=%llu host=%llu", rank,
         new_incarnation ? "re-joined after a restart" : rejoin ? "re-joined" : "joined", cfg.gpu_index,
         (unsigned long long)pool_bytes[TIER_GPU], (unsigned long long)pool_bytes[TIER_HOST]);
    return OCM_OK;
}

void governor_release(ClusterTable *c, int rank, int tier, uint64_t bytes) { cluster_unreserve(c, rank, tier, bytes); }

static uint64_t mono_now_ns() {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (uint64_t)ts.tv_sec * 1000000000ull + (uint64_t)ts.tv_nsec;
}

bool governor_rank_alive(const ClusterTable *c, int rank) {
    if (rank < 0 || rank >= c->n_ranks || !c->node[rank].joined.load(std::memory_order_acquire)) return false;
    const uint64_t beat = c->node[rank].heartbeat_ns.load(std::memory_order_relaxed);
    if (!beat) return true;
    const uint64_t now = mono_now_ns();
    return now < beat || now - beat < kRankDeadAfterNs;
}

bool governor_gpu_reachable(const ClusterTable *c, int from_rank, int to_rank) {
    if (from_rank < 0 || from_rank >= c->n_ranks || to_rank < 0 || to_rank >= c->n_ranks) return false;
    const NodeSlot &f = c->node[from_rank], &t = c->node[to_rank];
    if (t.gpu_index < 0) return false;
    if (from_rank == to_rank || f.gpu_index == t.gpu_index) return true;
    // a requester without a GPU of its own, or one that reported no peer row (older daemon, unit tests): not known
    // to be unreachable - the mapping step decides, as before
    if (f.gpu_index < 0 || f.peer_mask == 0) return true;
    return t.gpu_index < 32 && (f.peer_mask >> t.gpu_index) & 1u;
}

static bool try_rank(ClusterTable *c, int orig, int rank, Tier want, uint64_t bytes, Tier *got, bool *dead) {
    if (rank < 0 || rank >= c->n_ranks) return false;
    if (!c->node[rank].joined.load(std::memory_order_acquire)) return false;
    if (!governor_rank_alive(c, rank)) {   // its daemon stopped beating: place elsewhere (the reference hangs here)
        *dead = true;
        return false;
    }
    // HBM the requester's GPU cannot load from / store to is no use to it: down-tier to that rank's pinned host
    // pool (below) instead of failing later, when the application tries to map the pool
    const bool gpu_ok = want != TIER_GPU || governor_gpu_reachable(c, orig, rank);
    if (gpu_ok && c->node[rank].pool_bytes[want] && cluster_reserve(c, rank, want, bytes)) {
        *got = want;
        return true;
    }
    if (want == TIER_GPU && c->node[rank].pool_bytes[TIER_HOST] && cluster_reserve(c, rank, TIER_HOST, bytes)) {
        *got = TIER_HOST;
        return true;
    }
    return false;
}

int governor_place(ClusterTable *c, const AllocRequest &req, Allocation *out) {
    const int N = c->n_ranks;
    if (req.orig_rank < 0 || req.orig_rank >= N) return OCM_EINVAL;
    if (req.type <= ALLOC_MEM_INVALID || req.type >= ALLOC_MEM_MAX) return OCM_EINVAL;
    const Tier want = alloc_type_tier(req.type);
    const bool remote = alloc_type_is_remote(req.type);
    int order[kMaxRanks];
    int n = 0;

    uint32_t policy = req.policy;
    if (!remote) policy = OCM_POLICY_SELF;
    else if (policy == OCM_POLICY_DEFAULT) policy = OCM_POLICY_NEIGHBOUR;
    switch (policy) {
    case OCM_POLICY_SELF:
        order[n++] = req.orig_rank;
        break;
    case OCM_POLICY_EXPLICIT:
        if (req.remote_rank < 0 || req.remote_rank >= N) return OCM_EINVAL;
        order[n++] = req.remote_rank;
        break;
    case OCM_POLICY_NEIGHBOUR:
        // (orig+1)%N first, then further neighbours, the caller's own rank last.
        for (int d = 1; d <= N; ++d) order[n++] = (req.orig_rank + d) % N;
        break;
    case OCM_POLICY_ROUND_ROBIN: {
        uint64_t start = c->rr_cursor.fetch_add(1, std::memory_order_relaxed);
        for (int d = 0; d < N; ++d) order[n++] = (int)((start + d) % N);
        break;
    }
    case OCM_POLICY_LEAST_LOADED: {
        for (int r = 0; r < N; ++r) order[n++] = r;
        auto load = [&](int r) {
            uint64_t cap = c->node[r].pool_bytes[want];
            if (!cap) return 2.0;
            return (double)c->node[r].used[want].load(std::memory_order_relaxed) / (double)cap;
        };
        std::stable_sort(order, order + n, [&](int a, int b) {
            double la = load(a), lb = load(b);
            if (la != lb) return la < lb;
            // prefer a remote rank over the caller on ties
            return (a == req.orig_rank) < (b == req.orig_rank);
        });
        break;
    }
    default:
        return OCM_EINVAL;
    }

    bool dead = false;
    for (int i = 0; i < n; ++i) {
        Tier got;
        if (!try_rank(c, req.orig_rank, order[i], want, req.bytes, &got, &dead)) continue;
        memset(out, 0, sizeof *out);
        out->orig_rank = req.orig_rank;
        out->remote_rank = order[i];
        out->bytes = req.bytes;
        out->tier = got;
        out->flags = req.flags;
        if (got == want) {
            // legacy RDMA/RMA requests keep their type so the app sees the kind it asked for
            out->type = req.type;
        } else {
            out->type = remote ? ALLOC_MEM_REMOTE_HOST : ALLOC_MEM_HOST;
        }
        return OCM_OK;
    }
    return dead ? OCM_ENODAEMON : OCM_ENOMEM;   // nobody had room - or the only candidates are gone
}

} // namespace ocm
