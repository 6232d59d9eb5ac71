// Nodefile / topology description.
// Accepts the reference's format `rank dns ethernet_ip ocm_port rdmacm_port`
// with `#` comments (reference src/nodefile.c:30-37, bin/nodefile.rdma:1-3; the
// .rma variant without the last column, bin/nodefile.rma:1-3) and optional
// trailing key=value columns: gpu=<ordinal> gpu_bytes=<n[KMGT]> host_bytes=<n[KMGT]> numa=<node>.
// Rank discovery: OCM_RANK overrides; otherwise the hostname match of the
// reference (src/nodefile.c:92-103); a one-line file is rank 0.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace ocm {

struct NodeEntry {
    int         rank = -1;
    std::string dns;
    std::string ip_eth;
    int         ocm_port = 0;
    int         rdmacm_port = 0;     // parsed for compatibility, unused
    int         gpu = -2;            // -2: choose rank % device_count, -1: no GPU
    uint64_t    gpu_bytes = 0;       // 0: default
    uint64_t    host_bytes = 0;      // 0: default
    int         numa = -2;           // host NUMA node of the rank's pinned pool; -2: the node of its GPU (sysfs), -1: do not bind
};

struct Nodefile {
    std::vector<NodeEntry> nodes;    // index == rank
    int my_rank = -1;
    int size() const { return (int)nodes.size(); }
};

// Parse a size such as "512M", "16G", "4096".  Returns false on garbage.
bool parse_size(const char *s, uint64_t *out);

// Parse text (testable without files).  err receives a human-readable reason.
bool parse_nodefile_text(const std::string &text, Nodefile *out, std::string *err);
// Parse a file and resolve my_rank.  rank_override < 0: use OCM_RANK / hostname.
bool parse_nodefile(const std::string &path, Nodefile *out, std::string *err, int rank_override = -1);
// Resolve rank for an already parsed file (exposed for tests).
int resolve_rank(const Nodefile &nf, const char *hostname, const char *env_rank);

} // namespace ocm
