// oncillamem daemon: one per rank (one GPU + a slice of pinned host DRAM).
//
// Re-design of the reference daemon front end + protocol engine (reference
// src/main.c:57-224, src/mem.c:62-535).  Same roles - app registry over pmsg,
// rank-0 master, MSG_ADD_NODE join, REQ_ALLOC -> (rank 0) -> DO_ALLOC -> RELEASE_APP
// and REQ_FREE -> DO_FREE -> RELEASE_APP flows - but:
//   * one epoll loop (mailbox fd, peer sockets, unix fd server, timer, signalfd)
//     instead of a 500 us polling thread, a thread per request and a thread per
//     accepted connection (src/main.c:105-129, src/mem.c:399-433,513-529);
//   * persistent daemon<->daemon connections with request ids instead of one TCP
//     connection per message (src/mem.c:62-111);
//   * the daemon owns the rank's pools (HBM via CUDA VMM, pinned host via shm)
//     and publishes lock-free tables; it scrubs released blocks, reclaims the
//     memory of dead applications (a TODO in the reference, src/main.c:6-7,
//     README:68-69) and tells rank 0 about every release (never done in the
//     reference, src/mem.c:220-229,271-278).
#pragma once
#include "../common/gpu_rt.hpp"
#include "../common/msg.hpp"
#include "../common/nodefile.hpp"
#include "../common/ring.hpp"
#include "../common/tables.hpp"
#include "../common/topo.hpp"
#include "../ipc/pmsg.hpp"
#include <atomic>
#include <map>
#include <string>
#include <thread>
#include <vector>

namespace ocm {

struct DaemonOptions {
    std::string nodefile;
    int      rank_override = -1;
    uint64_t gpu_bytes = 0;     // 0: nodefile / env / default
    uint64_t host_bytes = 0;
    bool     no_gpu = false;
    int      join_timeout_ms = 30000;
};

class Daemon {
  public:
    explicit Daemon(const DaemonOptions &o) : opt_(o) {}
    ~Daemon();
    int  init();          // 0 or negative errno-style code; logs the reason
    int  run();           // event loop until shutdown
    void request_stop() { stop_.store(true); }

  private:
    struct App { int pid; std::string mq; uint64_t connected_ns; };
    struct PeerConn { int fd = -1; std::vector<uint8_t> inbuf; };
    enum Stage { ST_WAIT_PLACE, ST_WAIT_DO_ALLOC, ST_WAIT_DO_FREE, ST_WAIT_STATS, ST_WAIT_JOIN };
    struct Pending { Message orig; Stage stage; uint64_t t0_ns; };
    enum FdKind { FD_MQ, FD_TCP_LISTEN, FD_UDS_LISTEN, FD_PEER_OUT, FD_PEER_IN, FD_TIMER, FD_SIGNAL };

    // set-up
    int  setup_pools();
    int  setup_ipc();
    void join_cluster();
    // event handlers
    void on_mailbox();
    void on_app_msg(Message &m);
    void on_tcp_accept();
    void on_uds_accept();
    void on_peer_readable(int fd, bool outgoing);
    void on_peer_request(int fd, Message &m);
    void on_peer_reply(Message &m);
    void on_timer();
    // protocol steps
    void start_alloc(Message &m);
    void placed(Message &m);                  // placement known, fulfil on owner rank
    void finish_alloc(Message &m, int err);   // reply RELEASE_APP
    void start_free(Message &m);
    void note_free(int owner_rank, int tier, uint64_t bytes);
    int  do_alloc_local(Message &m);          // fills m.u.alloc from own table
    int  do_free_local(const Allocation &a);
    void fill_stats(StatsPayload *s);
    // helpers
    int  peer_fd(int rank);                   // persistent outgoing connection (lazy)
    bool send_peer(int rank, const Message &m);
    void reply_app(const Message &m);
    void drop_app(int pid, bool reclaim);
    void add_fd(int fd, FdKind kind);
    void del_fd(int fd);
    void close_peer(int fd);
    void scrub_loop();
    void reclaim_dead();
    void audit_tables();
    void mc_step();                           // NVLS multicast bring-up state machine (timer driven)
    ClusterTable *shared_cluster();           // rank 0: own table; others: lazily mapped
    uint32_t next_seq() { return ++seq_ ? seq_ : ++seq_; }

    DaemonOptions opt_;
    Nodefile nf_;
    int rank_ = -1, n_ranks_ = 0;
    int gpu_ = -1;
    topo::GpuTopo topo_;                    // what the runtime and sysfs know about gpu_ (peer row, NUMA node, ...)
    int host_numa_ = -1;                    // NUMA node the pinned host pool is bound to (-1: not bound)
    ShmMapping tbl_map_, host_map_, cluster_map_;
    gpu::PoolMapping host_vmm_;             // the host pool as a pinned VMM allocation (host_map_.base aliases its mapping)
    TableHeader *tbl_ = nullptr;
    ClusterTable *cluster_ = nullptr;       // rank 0 only (owner)
    gpu::PoolMapping gpu_pool_;
    Mailbox box_;
    int ep_ = -1, tcp_listen_ = -1, uds_listen_ = -1, timer_fd_ = -1, signal_fd_ = -1;
    std::map<int, FdKind> fds_;
    std::map<int, App> apps_;
    std::vector<int> out_fd_;                // per rank
    std::map<int, PeerConn> conns_;          // by fd (incoming and outgoing)
    std::map<uint32_t, Pending> pending_;
    uint32_t seq_ = 0;
    bool joined_ = false;
    uint64_t join_deadline_ns_ = 0, last_join_try_ns_ = 0, last_reclaim_ns_ = 0, last_audit_ns_ = 0;
    uint64_t audit_suspect_[TIER_COUNT] = {0, 0};   // operation count + 1 at which an audit last disagreed (0: none)
    std::atomic<uint64_t> table_faults_{0};
    std::atomic<bool> stop_{false};
    std::thread scrubber_;
    // counters
    std::atomic<uint64_t> msgs_in_{0}, msgs_out_{0}, reclaimed_{0}, scrubbed_bytes_{0};
    void *scrub_stream_ = nullptr;
    // NVLS multicast: the tail of the HBM pool (mc_bytes_ at mc_tail_off_) is bound to rank 0's object
    gpu::McObject mc_;
    uint64_t mc_bytes_ = 0, mc_tail_off_ = 0;
    bool mc_added_ = false, mc_bound_ = false, mc_done_ = false;
    ShmMapping cluster_peer_map_;
    // Tables of the other ranks of THIS node (shared memory): requests whose owner rank is mapped here are fulfilled
    // without a TCP hop; ranks on other nodes have no segment to open and keep the message path.
    struct PeerTable {
        ShmMapping map;
        TableHeader *hdr = nullptr;
        uint64_t retry_ns = 0;
    };
    PeerTable peer_tbl_[kMaxRanks];
    bool node_local_ = true;                 // OCM_DAEMON_NODE_LOCAL=0: always use the message path
    TableHeader *peer_table(int rank);
    bool inject_alloc_fault();
    // set by the main thread once the multicast region is READY; the scrubber thread zeroes this rank's slice of
    // ranges whose owner died (tables.hpp: mc_owner / mc_scrubbed)
    std::atomic<ClusterTable *> mc_cluster_{nullptr};
    // cluster table for the heartbeat (set once joined).  The scrubber thread beats too, so that a main loop that
    // is busy for a while (multicast bring-up, a long driver call) does not make a live rank look dead.
    std::atomic<ClusterTable *> hb_cluster_{nullptr};
};

uint64_t now_ns();

} // namespace ocm
