// Names of every kernel-visible IPC object (message queues, shared-memory
// segments, unix sockets).  With OCM_NS unset and rank 0 the mailbox names are
// the reference's (`/ocm_mq_daemon`, `/ocm_mq_<pid>`, reference inc/pmsg.h:23-24).
// A rank suffix makes several daemons per host possible, which the reference's
// single global daemon queue forbids; OCM_NS isolates
// concurrent clusters (tests) from each other.
#pragma once
#include <cstdlib>
#include <string>
#include <unistd.h>

namespace ocm {

inline std::string ns_tag() {
    const char *e = getenv("OCM_NS");
    std::string s = e ? e : "";
    if (s.size() > 24) s.resize(24);
    for (auto &c : s)
        if (c == '/') c = '_';
    return s;
}
inline std::string ns_prefix() {
    std::string s = ns_tag();
    return s.empty() ? s : s + "_";
}
inline std::string mq_daemon_name(int rank) {
    std::string s = "/ocm_mq_" + ns_prefix() + "daemon";
    if (rank > 0) s += "_r" + std::to_string(rank);
    return s;
}
inline std::string mq_app_name(int pid) { return "/ocm_mq_" + ns_prefix() + std::to_string(pid); }
inline std::string shm_table_name(int rank) { return "/ocm_" + ns_prefix() + "tbl_r" + std::to_string(rank); }
inline std::string shm_host_pool_name(int rank) { return "/ocm_" + ns_prefix() + "host_r" + std::to_string(rank); }
inline std::string shm_cluster_name() { return "/ocm_" + ns_prefix() + "cluster"; }
// abstract unix socket (leading NUL added by the socket layer)
inline std::string uds_name(int rank) { return "ocm_" + ns_prefix() + "uds_r" + std::to_string(rank); }

} // namespace ocm
