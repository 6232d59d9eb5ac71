// Allocation governor: decides which rank backs an allocation and keeps
// capacity accounting in the cluster table.  Re-design of the reference's
// rank-0 `alloc_find` (reference src/alloc.c:76-140): same default policy
// (neighbour, (orig+1) % N, src/alloc.c:107,120) plus explicit / least-loaded /
// round-robin / self policies, real capacity checks (absent in the reference,
// src/alloc.c:87-92) and release bookkeeping (absent, src/mem.c:220-229).
// Works on the shared ClusterTable, so rank 0's daemon (slow path) and any
// attached application (fast path) run the very same code on the same state.
#pragma once
#include "msg.hpp"
#include "tables.hpp"

namespace ocm {

// Which tier an AllocType lives in and whether it is placed remotely.
inline bool alloc_type_is_remote(uint32_t t) {
    return t == ALLOC_MEM_RMA || t == ALLOC_MEM_RDMA || t == ALLOC_MEM_REMOTE_GPU || t == ALLOC_MEM_REMOTE_HOST;
}
inline Tier alloc_type_tier(uint32_t t) {
    return (t == ALLOC_MEM_GPU || t == ALLOC_MEM_REMOTE_GPU) ? TIER_GPU : TIER_HOST;
}

// Place `req`; on success fills out->{orig_rank, remote_rank, type, tier, bytes}
// and reserves `bytes` on the chosen rank.  Returns 0 or a positive ocm_err.
// The granted type may differ from the requested one: a GPU request falls back
// to the host tier of the same candidate when that rank has no (or a full) GPU
// pool - the analogue of the reference downgrading on single-node clusters
// (src/alloc.c:82-83).
int governor_place(ClusterTable *c, const AllocRequest &req, Allocation *out);

// Undo the reservation made by governor_place (on failure or on free).
void governor_release(ClusterTable *c, int rank, int tier, uint64_t bytes);

// Register a node (MSG_ADD_NODE; reference src/alloc.c:59-74 BUGs on re-add,
// here a re-join simply refreshes the slot and keeps its accounting).
int governor_add_node(ClusterTable *c, int rank, const NodeConfig &cfg, const uint64_t pool_bytes[TIER_COUNT]);

int governor_joined(const ClusterTable *c);

// True while the rank's daemon keeps its heartbeat in the cluster table fresh (NodeSlot::heartbeat_ns).
// governor_place skips ranks that are not: allocations move on to the next candidate of the policy, and
// return OCM_ENODAEMON instead of OCM_ENOMEM when dead ranks were the only candidates.  The reference has no
// failure detection at all: a request for a dead node blocks in connect() or forever.
bool governor_rank_alive(const ClusterTable *c, int rank);

// Can the GPU of `from_rank` address the HBM pool of `to_rank` directly (same GPU, or a peer its daemon found with
// cudaDeviceCanAccessPeer at start-up)?  governor_place gives unreachable candidates the host tier instead.
bool governor_gpu_reachable(const ClusterTable *c, int from_rank, int to_rank);

} // namespace ocm
