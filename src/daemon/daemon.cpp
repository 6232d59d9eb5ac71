#include "daemon.hpp"
#include <sys/mman.h>
#include "../common/governor.hpp"
#include "../common/log.hpp"
#include "../common/names.hpp"
#include "../ipc/sock.hpp"
#include "../kernels/kernels.h"
#include "../../include/oncillamem.h"
#include "uds_proto.hpp"
#include <cerrno>
#include <csignal>
#include <cstring>
#include <sys/epoll.h>
#include <sys/signalfd.h>
#include <sys/socket.h>
#include <sys/sysinfo.h>
#include <sys/timerfd.h>
#include <unistd.h>

namespace ocm {

uint64_t now_ns() {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (uint64_t)ts.tv_sec * 1000000000ull + (uint64_t)ts.tv_nsec;
}

static uint64_t env_size(const char *k, uint64_t dflt) {
    const char *e = getenv(k);
    uint64_t v;
    return (e && parse_size(e, &v)) ? v : dflt;
}

Daemon::~Daemon() {
    stop_.store(true);
    if (scrubber_.joinable()) scrubber_.join();
    for (auto &kv : conns_) sock_close(kv.first);
    sock_close(tcp_listen_);
    sock_close(uds_listen_);
    sock_close(timer_fd_);
    sock_close(signal_fd_);
    sock_close(ep_);
    if (box_.is_open()) box_.close(true);
    if (tbl_) tbl_->ready.store(0);
    if (cluster_) cluster_->mc_state.store(MC_NONE);
    if (mc_.handle) gpu::mc_release(&mc_);
    if (cluster_peer_map_.base) shm_close(&cluster_peer_map_, false);
    for (auto &p : peer_tbl_)
        if (p.map.base) shm_close(&p.map, false);
    if (gpu_pool_.ptr) {
        if (scrub_stream_) cudaStreamDestroy((cudaStream_t)scrub_stream_);
        gpu::pool_release(&gpu_pool_);
    }
    if (host_vmm_.ptr) {
        gpu::pool_release(&host_vmm_);
        host_map_ = ShmMapping();
    }
    if (host_map_.base) {
        gpu::host_unregister(host_map_.base);
        shm_close(&host_map_, true);
    }
    if (tbl_map_.base) {
        gpu::host_unregister(tbl_map_.base);
        shm_close(&tbl_map_, true);
    }
    if (cluster_map_.base) shm_close(&cluster_map_, cluster_map_.owner);
}

void Daemon::add_fd(int fd, FdKind kind) {
    struct epoll_event ev;
    memset(&ev, 0, sizeof ev);
    ev.events = EPOLLIN;
    ev.data.fd = fd;
    if (epoll_ctl(ep_, EPOLL_CTL_ADD, fd, &ev) != 0) LOGE("epoll add fd %d: %s", fd, strerror(errno));
    fds_[fd] = kind;
}

void Daemon::del_fd(int fd) {
    epoll_ctl(ep_, EPOLL_CTL_DEL, fd, nullptr);
    fds_.erase(fd);
}

// ------------------------------------------------------------------ set-up --
int Daemon::setup_pools() {
    const NodeEntry &me = nf_.nodes[rank_];
    uint64_t host_bytes = opt_.host_bytes ? opt_.host_bytes : me.host_bytes ? me.host_bytes : env_size("OCM_HOST_BYTES", 1ull << 30);
    uint64_t gpu_bytes = opt_.gpu_bytes ? opt_.gpu_bytes : me.gpu_bytes ? me.gpu_bytes : env_size("OCM_GPU_BYTES", 0);

    int ndev = opt_.no_gpu ? 0 : gpu::device_count();
    gpu_ = -1;
    if (ndev > 0 && me.gpu != -1) {
        gpu_ = me.gpu >= 0 ? me.gpu : rank_ % ndev;
        if (gpu_ >= ndev) {
            LOGE("nodefile asks for gpu=%d but only %d device(s) are visible", gpu_, ndev);
            return -ENODEV;
        }
        std::string err;
        if (!gpu::set_device(gpu_, &err)) {
            LOGE("%s", err.c_str());
            return -ENODEV;
        }
        if (!gpu_bytes) {
            size_t fr = 0, tot = 0;
            gpu::mem_info(gpu_, &fr, &tot);
            // default: most of the free HBM (a B200 has ~178 GiB free: 160 GiB pooled), leaving room for the contexts
            // and working memory of the applications that share the GPU.  OCM_GPU_BYTES / nodefile gpu_bytes= override.
            // Ranks that share one GPU (more nodefile lines than devices) split it evenly.
            int sharers = 0;
            for (int r = 0; r < n_ranks_; ++r) {
                const int g = nf_.nodes[r].gpu >= 0 ? nf_.nodes[r].gpu : nf_.nodes[r].gpu == -1 ? -1 : r % ndev;
                sharers += g == gpu_;
            }
            if (sharers < 1) sharers = 1;
            gpu_bytes = fr > (24ull << 30) ? fr - std::max<uint64_t>(fr / 10, 16ull << 30) : fr / 2;
            gpu_bytes /= (uint64_t)sharers;
        }
    } else {
        gpu_bytes = 0;
    }

    std::string err;
    // Topology of our GPU: which peers it reaches, multicast, and the host NUMA node it hangs off.  The pinned host
    // pool is first-touched and bound there (nodefile `numa=<n>` overrides, `numa=-1` disables), and the daemon's
    // threads run on that node's CPUs, so the GPU's PCIe reads of the pool stay on one socket.
    if (gpu_ >= 0) topo::discover_gpu(gpu_, &topo_);
    host_numa_ = me.numa >= -1 ? me.numa : topo_.numa_node;
    if (host_numa_ >= 0 && topo::numa_nodes().size() > 1) {
        if (topo::bind_thread_to_node(host_numa_)) LOGI("daemon threads bound to NUMA node %d", host_numa_);
        else LOGW("cannot bind to NUMA node %d (no such node or not permitted)", host_numa_);
    } else if (host_numa_ >= 0 && topo::node_cpus(host_numa_).empty()) {
        host_numa_ = -1;
    }
    // GPU pool first: VMM rounds the size up, the table must describe the real size.
    uint8_t ipc_handle[64] = {0};
    if (gpu_ >= 0 && gpu_bytes) {
        // NVLS: with several ranks a tail of the pool is set aside and later bound to one multicast object
        uint64_t mc_want = 0;
        if (n_ranks_ >= 2 && gpu::mc_supported(gpu_)) {
            mc_want = env_size("OCM_MC_BYTES", 256ull << 20);
            size_t g = gpu::mc_granularity(n_ranks_, mc_want, false);
            if (!g) g = 2 << 20;
            mc_want = std::min<uint64_t>((mc_want + g - 1) / g * g, (uint64_t)kMcMaxGranules * kMcGranule);
        }
        const size_t vg = gpu::vmm_granularity(gpu_);
        gpu_bytes = (gpu_bytes + vg - 1) / vg * vg;
        if (!gpu::pool_create(gpu_, gpu_bytes + mc_want, 0, &gpu_pool_, ipc_handle, &err)) {
            LOGE("GPU pool: %s", err.c_str());
            return -ENOMEM;
        }
        if (mc_want && gpu_pool_.mode == GPU_SHARE_VMM_FD && gpu_pool_.bytes >= gpu_bytes + mc_want) {
            mc_bytes_ = mc_want;
            mc_tail_off_ = gpu_pool_.bytes - mc_want;
            gpu_bytes = mc_tail_off_;
        } else {
            gpu_bytes = gpu_pool_.bytes;
        }
        cudaStream_t s;
        if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) == cudaSuccess) scrub_stream_ = s;
        // fresh pools start zeroed (remote buffers are zero-filled in the reference, src/alloc.c:171)
        ocmk::launch_fill(gpu_pool_.ptr, gpu_pool_.bytes, 0, (cudaStream_t)scrub_stream_);
        cudaStreamSynchronize((cudaStream_t)scrub_stream_);
    }

    size_t seg = table_segment_bytes(gpu_bytes, host_bytes);
    if (!shm_create(shm_table_name(rank_), seg, true, &tbl_map_, &err)) {
        LOGE("%s", err.c_str());
        return -EIO;
    }
    table_format(tbl_map_.base, seg, rank_, n_ranks_, (int)getpid(), gpu_bytes, host_bytes);
    tbl_ = static_cast<TableHeader *>(tbl_map_.base);
    snprintf(tbl_->ns, sizeof tbl_->ns, "%s", ns_tag().c_str());
    tbl_->gpu_index = gpu_;
    tbl_->gpu_share_mode = gpu_pool_.ptr ? gpu_pool_.mode : GPU_SHARE_NONE;
    memcpy(tbl_->cuda_ipc_handle, ipc_handle, 64);
    tbl_->mc_tail_offset = mc_tail_off_;
    tbl_->mc_bytes = mc_bytes_;
    if (gpu_ >= 0) gpu::get_uuid(gpu_, tbl_->gpu_uuid);
    if (getenv("OCM_NO_SCRUB")) tbl_->scrub_enabled.store(0);
    if (getenv("OCM_NO_SLABS")) tbl_->slabs_enabled.store(0);
    if (const char *e = getenv("OCM_DAEMON_NODE_LOCAL")) node_local_ = atoi(e) != 0;

    // Host pool.  With a GPU: one pinned VMM allocation on the GPU's NUMA node, handed to applications by file
    // descriptor (they map it without pinning anything again).  Without one, or if the driver refuses
    // (OCM_HOST_POOL=shm forces it): a shared-memory segment that every process pins for itself.
    const char *hp = getenv("OCM_HOST_POOL");
    if (host_bytes && gpu_ >= 0 && !(hp && !strcmp(hp, "shm"))) {
        const uint64_t want = tbl_->pool[TIER_HOST].bytes;
        if (gpu::host_pool_create(host_numa_, want, gpu_, &host_vmm_, &err) && host_vmm_.bytes >= want) {
            host_map_.base = host_vmm_.ptr;
            host_map_.bytes = want;
            // fresh pools start zeroed; the GPU writes the zeros (PCIe rate; a first CPU pass over the mapping runs
            // at ~2 GB/s, it faults every page into this process)
            ocmk::launch_fill(host_vmm_.ptr, want, 0, (cudaStream_t)scrub_stream_);
            cudaStreamSynchronize((cudaStream_t)scrub_stream_);
            tbl_->host_share_mode = GPU_SHARE_VMM_FD;
            tbl_->host_numa_node = host_vmm_.owner_dev;
            shm_unlink(shm_host_pool_name(rank_).c_str());       // no stale segment of an earlier incarnation
        } else {
            if (host_vmm_.ptr) gpu::pool_release(&host_vmm_);
            LOGW("host pool as a VMM allocation refused (%s): shared memory instead", err.c_str());
        }
    }
    if (host_bytes && !host_vmm_.ptr) {
        if (!shm_create(shm_host_pool_name(rank_), tbl_->pool[TIER_HOST].bytes, true, &host_map_, &err)) {
            LOGE("%s", err.c_str());
            return -ENOMEM;
        }
        // place (NUMA node of the GPU), touch + pin so that the pool is resident and device-visible
        if (host_numa_ >= 0 && !topo::prefer_node_for_range(host_map_.base, host_map_.bytes, host_numa_))
            LOGD("mbind refused: relying on first touch from threads bound to node %d", host_numa_);
        topo::touch_range(host_map_.base, host_map_.bytes, host_numa_, 8);
        if (gpu_ >= 0 && !gpu::host_register(host_map_.base, host_map_.bytes, &err))
            LOGW("host pool not pinned: %s", err.c_str());
    }
    if (gpu_ >= 0 && !gpu::host_register(tbl_map_.base, tbl_map_.bytes, &err))
        LOGW("tables not device-visible: %s", err.c_str());
    LOGI("rank %d pools: gpu=%d %llu B (%s), host %llu B (%s)", rank_, gpu_, (unsigned long long)gpu_bytes,
         gpu_pool_.mode == GPU_SHARE_VMM_FD ? "vmm-fd" : gpu_pool_.mode == GPU_SHARE_CUDA_IPC ? "cuda-ipc" : "none",
         (unsigned long long)host_bytes, host_vmm_.ptr ? "pinned vmm-fd" : "shared memory");
    return 0;
}

int Daemon::setup_ipc() {
    ep_ = epoll_create1(EPOLL_CLOEXEC);
    if (ep_ < 0) return -errno;
    const NodeEntry &me = nf_.nodes[rank_];

    int r = box_.open(mq_daemon_name(rank_), sizeof(Message), 10, true);
    if (r) return r;
    add_fd(box_.fd(), FD_MQ);

    tcp_listen_ = tcp_listen("", me.ocm_port);
    if (tcp_listen_ < 0) {
        LOGE("cannot listen on port %d: %s", me.ocm_port, strerror(-tcp_listen_));
        return tcp_listen_;
    }
    set_nonblocking(tcp_listen_, true);
    add_fd(tcp_listen_, FD_TCP_LISTEN);

    // the unix socket was bound in init() (it doubles as the "this rank is taken" lock)
    set_nonblocking(uds_listen_, true);
    add_fd(uds_listen_, FD_UDS_LISTEN);

    timer_fd_ = timerfd_create(CLOCK_MONOTONIC, TFD_NONBLOCK | TFD_CLOEXEC);
    struct itimerspec its;
    memset(&its, 0, sizeof its);
    its.it_interval.tv_nsec = its.it_value.tv_nsec = 20 * 1000 * 1000; // 20 ms housekeeping
    timerfd_settime(timer_fd_, 0, &its, nullptr);
    add_fd(timer_fd_, FD_TIMER);

    sigset_t mask;
    sigemptyset(&mask);
    sigaddset(&mask, SIGINT);
    sigaddset(&mask, SIGTERM);
    pthread_sigmask(SIG_BLOCK, &mask, nullptr);
    signal(SIGPIPE, SIG_IGN);
    signal_fd_ = signalfd(-1, &mask, SFD_NONBLOCK | SFD_CLOEXEC);
    add_fd(signal_fd_, FD_SIGNAL);
    out_fd_.assign(n_ranks_, -1);
    return 0;
}

int Daemon::init() {
    std::string err;
    if (!parse_nodefile(opt_.nodefile, &nf_, &err, opt_.rank_override)) {
        LOGE("%s", err.c_str());
        return -EINVAL;
    }
    rank_ = nf_.my_rank;
    n_ranks_ = nf_.size();
    if (n_ranks_ > kMaxRanks) {
        LOGE("at most %d ranks are supported", kMaxRanks);
        return -EINVAL;
    }
    char tag[32];
    snprintf(tag, sizeof tag, "d%d", rank_);
    set_log_tag(tag);
    // block termination signals before any helper thread (CUDA, scrubber) exists
    sigset_t mask;
    sigemptyset(&mask);
    sigaddset(&mask, SIGINT);
    sigaddset(&mask, SIGTERM);
    pthread_sigmask(SIG_BLOCK, &mask, nullptr);

    // Claim the rank first: the abstract unix socket is exclusive and disappears with its process, so it cannot be
    // stale.  Everything after this replaces leftovers of a DEAD predecessor (tables, host pool, message queue) -
    // which must never happen to a live daemon that someone started a duplicate of.
    uds_listen_ = uds_listen(uds_name(rank_));
    if (uds_listen_ < 0) {
        LOGE("rank %d of namespace '%s' is already served by a running daemon (unix socket %s: %s)", rank_,
             ns_tag().c_str(), uds_name(rank_).c_str(), strerror(-uds_listen_));
        return uds_listen_;
    }

    int r = setup_pools();
    if (r) return r;
    r = setup_ipc();
    if (r) return r;

    if (rank_ == 0) {
        if (!shm_create(shm_cluster_name(), sizeof(ClusterTable), true, &cluster_map_, &err)) {
            LOGE("%s", err.c_str());
            return -EIO;
        }
        cluster_ = static_cast<ClusterTable *>(cluster_map_.base);
        cluster_format(cluster_, n_ranks_);
        if (mc_bytes_) {
            if (gpu::mc_create(n_ranks_, mc_bytes_, &mc_, &err)) {
                cluster_->mc_bytes = mc_bytes_;
                cluster_->mc_n_granules = (uint32_t)(mc_bytes_ / kMcGranule);
                // the signal words of the region (tables.hpp, kMcSignalGranules) are never handed out
                cluster_->mc_bitmap[0].fetch_or((1ull << kMcSignalGranules) - 1, std::memory_order_release);
                cluster_->mc_state.store(MC_CREATED, std::memory_order_release);
                LOGI("multicast object: %llu bytes over %d devices", (unsigned long long)mc_bytes_, n_ranks_);
            } else {
                LOGW("NVLS multicast unavailable: %s", err.c_str());
                cluster_->mc_state.store(MC_FAILED, std::memory_order_release);
                mc_bytes_ = 0;
            }
        }
    }
    scrubber_ = std::thread([this] { scrub_loop(); });
    join_deadline_ns_ = now_ns() + (uint64_t)opt_.join_timeout_ms * 1000000ull;
    join_cluster();
    return 0;
}

// MSG_ADD_NODE: tell rank 0 who we are (reference src/main.c:143-160, src/mem.c:121-130),
// with the node configuration actually filled in.
void Daemon::join_cluster() {
    Message m;
    m.init(MSG_ADD_NODE);
    m.rank = rank_;
    m.pid = (int)getpid();
    NodeConfig &c = m.u.node;
    snprintf(c.ip, sizeof c.ip, "%s", nf_.nodes[rank_].ip_eth.c_str());
    c.ram = tbl_->pool[TIER_HOST].bytes;
    c.num_gpu = gpu::device_count();
    c.gpu_index = gpu_;
    c.daemon_pid = (int)getpid();
    snprintf(c.pci_bdf, sizeof c.pci_bdf, "%s", topo_.pci_bdf);
    c.numa_node = (int16_t)host_numa_;
    c.peer_mask = (uint8_t)(topo_.peer_mask & 0xffu);
    c.caps = (uint8_t)((topo_.multicast ? NODE_CAP_MULTICAST : 0) | (topo_.vmm_fd_export ? NODE_CAP_VMM_FD : 0));
    for (int d = 0; d < c.num_gpu && d < kMaxGpus; ++d) {
        size_t fr = 0, tot = 0;
        if (d == gpu_ && gpu::mem_info(d, &fr, &tot)) c.gpu_mem[d] = tot;
    }
    uint64_t pools[TIER_COUNT] = {tbl_->pool[TIER_GPU].bytes, tbl_->pool[TIER_HOST].bytes};
    last_join_try_ns_ = now_ns();
    if (rank_ == 0) {
        governor_add_node(cluster_, 0, c, pools);
        cluster_->ready.store(1, std::memory_order_release);
        joined_ = true;
        tbl_->ready.store(1, std::memory_order_release);
        return;
    }
    // pool sizes travel in the gpu_mem/ram fields: gpu_mem[kMaxGpus-1] carries the GPU pool size
    c.gpu_mem[kMaxGpus - 1] = pools[TIER_GPU];
    m.seq = next_seq();
    if (send_peer(0, m)) {
        Pending p;
        p.orig = m;
        p.stage = ST_WAIT_JOIN;
        p.t0_ns = now_ns();
        pending_[m.seq] = p;
    } else {
        LOGD("rank 0 not reachable yet, will retry");
    }
}

// ------------------------------------------------------------------ peers --
int Daemon::peer_fd(int rank) {
    if (rank < 0 || rank >= n_ranks_) return -1;
    if (out_fd_[rank] >= 0) return out_fd_[rank];
    const NodeEntry &pe = nf_.nodes[rank];
    int fd = tcp_connect(pe.ip_eth, pe.ocm_port, 1000);
    if (fd < 0) {
        LOGD("connect to rank %d (%s:%d): %s", rank, pe.ip_eth.c_str(), pe.ocm_port, strerror(-fd));
        return -1;
    }
    set_nonblocking(fd, true);
    out_fd_[rank] = fd;
    conns_[fd] = PeerConn();
    conns_[fd].fd = fd;
    add_fd(fd, FD_PEER_OUT);
    return fd;
}

bool Daemon::send_peer(int rank, const Message &m) {
    int fd = peer_fd(rank);
    if (fd < 0) return false;
    int r = sock_put(fd, &m, sizeof m);
    if (r != 1) {
        close_peer(fd);
        return false;
    }
    msgs_out_.fetch_add(1, std::memory_order_relaxed);
    return true;
}

void Daemon::close_peer(int fd) {
    for (auto &o : out_fd_)
        if (o == fd) o = -1;
    del_fd(fd);
    conns_.erase(fd);
    sock_close(fd);
}

void Daemon::on_tcp_accept() {
    for (;;) {
        std::string peer;
        int fd = tcp_accept(tcp_listen_, &peer);
        if (fd < 0) break;
        set_nonblocking(fd, true);
        conns_[fd] = PeerConn();
        conns_[fd].fd = fd;
        add_fd(fd, FD_PEER_IN);
        LOGD("accepted daemon connection from %s", peer.c_str());
    }
}

void Daemon::on_peer_readable(int fd, bool outgoing) {
    auto it = conns_.find(fd);
    if (it == conns_.end()) return;
    PeerConn &c = it->second;
    uint8_t buf[sizeof(Message) * 8];
    for (;;) {
        ssize_t n = recv(fd, buf, sizeof buf, 0);
        if (n > 0) {
            c.inbuf.insert(c.inbuf.end(), buf, buf + n);
            continue;
        }
        if (n < 0 && (errno == EAGAIN || errno == EWOULDBLOCK)) break;
        if (n < 0 && errno == EINTR) continue;
        // closed or failed
        LOGD("peer connection fd %d closed", fd);
        close_peer(fd);
        return;
    }
    size_t off = 0;
    while (c.inbuf.size() - off >= sizeof(Message)) {
        Message m;
        memcpy(&m, c.inbuf.data() + off, sizeof m);
        off += sizeof m;
        msgs_in_.fetch_add(1, std::memory_order_relaxed);
        if (!m.valid()) {
            LOGW("dropping malformed message from fd %d", fd);
            continue;
        }
        if (outgoing && m.status == MSG_RESPONSE) on_peer_reply(m);
        else on_peer_request(fd, m);
        if (conns_.find(fd) == conns_.end()) return; // closed by a handler
    }
    auto it2 = conns_.find(fd);
    if (it2 != conns_.end() && off) it2->second.inbuf.erase(it2->second.inbuf.begin(), it2->second.inbuf.begin() + off);
}

// Requests arriving from other daemons (reference inbound_thread, src/mem.c:318-393).
void Daemon::on_peer_request(int fd, Message &m) {
    LOGD("peer request %s from rank %d", msg_type_str(m.type), m.rank);
    bool reply = true;
    m.err = 0;
    switch (m.type) {
    case MSG_ADD_NODE:
        if (rank_ != 0) { m.err = OCM_EPROTO; break; }
        {
            uint64_t pools[TIER_COUNT] = {m.u.node.gpu_mem[kMaxGpus - 1], m.u.node.ram};
            m.u.node.gpu_mem[kMaxGpus - 1] = 0;
            m.err = governor_add_node(cluster_, m.rank, m.u.node, pools);
        }
        break;
    case MSG_REQ_ALLOC:
        if (rank_ != 0) { m.err = OCM_EPROTO; break; }
        {
            AllocRequest req = m.u.req;
            Allocation a;
            m.err = governor_place(cluster_, req, &a);
            if (!m.err) m.u.alloc = a;
        }
        break;
    case MSG_DO_ALLOC:
        m.err = do_alloc_local(m);
        break;
    case MSG_DO_FREE:
        m.err = do_free_local(m.u.alloc);
        break;
    case MSG_NOTE_FREE:
        // rank and tier come off the wire: cluster_unreserve checks both before it indexes the table
        if (rank_ == 0 && m.u.alloc.tier < (uint32_t)TIER_COUNT)
            governor_release(cluster_, m.u.alloc.remote_rank, (int)m.u.alloc.tier, m.u.alloc.bytes);
        reply = false;
        break;
    case MSG_STATS:
        fill_stats(&m.u.stats);
        break;
    case MSG_PING:
        break;
    default:
        m.err = OCM_EPROTO;
        break;
    }
    if (!reply) return;
    m.status = MSG_RESPONSE;
    m.hops++;
    if (sock_put(fd, &m, sizeof m) != 1) close_peer(fd);
    else msgs_out_.fetch_add(1, std::memory_order_relaxed);
}

void Daemon::on_peer_reply(Message &m) {
    auto it = pending_.find(m.seq);
    if (it == pending_.end()) {
        LOGW("reply %s seq %u matches no pending request", msg_type_str(m.type), m.seq);
        return;
    }
    Pending p = it->second;
    pending_.erase(it);
    switch (p.stage) {
    case ST_WAIT_JOIN:
        if (m.err) {
            LOGE("rank 0 rejected join: %s", ocm_strerror(m.err));
            stop_.store(true);
        } else {
            joined_ = true;
            tbl_->ready.store(1, std::memory_order_release);
            LOGI("joined cluster as rank %d of %d", rank_, n_ranks_);
        }
        break;
    case ST_WAIT_PLACE: {
        Message r = p.orig;
        if (m.err) {
            finish_alloc(r, m.err);
        } else {
            r.u.alloc = m.u.alloc;
            placed(r);
        }
        break;
    }
    case ST_WAIT_DO_ALLOC: {
        Message r = p.orig;
        if (m.err) {
            // give the reservation back
            note_free(r.u.alloc.remote_rank, (int)r.u.alloc.tier, r.u.alloc.bytes);
            finish_alloc(r, m.err);
        } else {
            r.u.alloc = m.u.alloc;
            finish_alloc(r, 0);
        }
        break;
    }
    case ST_WAIT_DO_FREE: {
        Message r = p.orig;
        if (!m.err) note_free(r.u.alloc.remote_rank, (int)r.u.alloc.tier, r.u.alloc.bytes);
        r.type = MSG_RELEASE_APP;
        r.status = MSG_RESPONSE;
        r.err = m.err;
        reply_app(r);
        break;
    }
    case ST_WAIT_STATS: {
        Message r = p.orig;
        r.u.stats = m.u.stats;
        r.status = MSG_RESPONSE;
        r.err = m.err;
        reply_app(r);
        break;
    }
    }
}

// ------------------------------------------------------------------- apps --
void Daemon::reply_app(const Message &m) {
    auto it = apps_.find(m.pid);
    if (it == apps_.end()) {
        LOGW("reply for unknown app pid %d dropped", m.pid);
        return;
    }
    int r = box_.send(it->second.mq, &m, 2000);
    if (r) {
        LOGW("cannot reach app %d (%s): dropping it", m.pid, strerror(-r));
        drop_app(m.pid, true);
    } else {
        msgs_out_.fetch_add(1, std::memory_order_relaxed);
    }
}

void Daemon::drop_app(int pid, bool /*reclaim*/) {
    auto it = apps_.find(pid);
    if (it == apps_.end()) return;
    box_.detach(it->second.mq);
    apps_.erase(it);
}

void Daemon::on_mailbox() {
    Message m;
    for (int i = 0; i < 64; ++i) {
        int r = box_.recv(&m, 0);
        if (r != 1) break;
        msgs_in_.fetch_add(1, std::memory_order_relaxed);
        if (!m.valid()) {
            LOGW("malformed mailbox message dropped");
            continue;
        }
        on_app_msg(m);
    }
}

// Messages from applications (reference process_msg, src/main.c:57-103).
void Daemon::on_app_msg(Message &m) {
    LOGD("app %d: %s", m.pid, msg_type_str(m.type));
    switch (m.type) {
    case MSG_CONNECT: {
        std::string mq = mq_app_name(m.pid);
        int r = box_.attach(mq);
        if (r) {
            LOGW("cannot attach to %s: %s", mq.c_str(), strerror(-r));
            return;
        }
        apps_[m.pid] = App{m.pid, mq, now_ns()};
        Message c;
        c.init(MSG_CONNECT_CONFIRM, MSG_RESPONSE);
        c.pid = m.pid;
        c.rank = rank_;
        c.seq = m.seq;
        c.err = joined_ ? 0 : OCM_ENODAEMON;
        c.u.conn.rank = rank_;
        c.u.conn.n_ranks = n_ranks_;
        c.u.conn.gpu_index = gpu_;
        c.u.conn.daemon_pid = (int)getpid();
        snprintf(c.u.conn.ns, sizeof c.u.conn.ns, "%s", ns_tag().c_str());
        reply_app(c);
        break;
    }
    case MSG_DISCONNECT:
        // unknown pids are ignored (the reference walks off the list head)
        drop_app(m.pid, true);
        reclaim_dead();
        break;
    case MSG_REQ_ALLOC:
        start_alloc(m);
        break;
    case MSG_REQ_FREE:
        start_free(m);
        break;
    case MSG_STATS: {
        int target = m.rank;
        if (target < 0 || target == rank_ || target >= n_ranks_) {
            fill_stats(&m.u.stats);
            m.status = MSG_RESPONSE;
            m.err = 0;
            reply_app(m);
        } else {
            Message q = m;
            q.seq = next_seq();
            q.rank = rank_;
            Pending p;
            p.orig = m;
            p.stage = ST_WAIT_STATS;
            p.t0_ns = now_ns();
            if (send_peer(target, q)) pending_[q.seq] = p;
            else {
                m.status = MSG_RESPONSE;
                m.err = OCM_ENODAEMON;
                reply_app(m);
            }
        }
        break;
    }
    case MSG_PING:
        m.status = MSG_RESPONSE;
        reply_app(m);
        break;
    case MSG_SHUTDOWN:
        LOGI("shutdown requested by pid %d", m.pid);
        stop_.store(true);
        break;
    default:
        LOGW("unexpected %s from app %d", msg_type_str(m.type), m.pid);
        break;
    }
}

// ----------------------------------------------------------- alloc / free --
void Daemon::start_alloc(Message &m) {
    m.rank = rank_;
    m.u.req.orig_rank = rank_;
    if (!joined_) return finish_alloc(m, OCM_ENODAEMON);
    // rank 0 owns the cluster table; a daemon on rank 0's node sees the same table in shared memory and places the
    // request there itself (the reservation is one CAS loop, safe from any process)
    ClusterTable *ct = rank_ == 0 ? cluster_ : (node_local_ ? shared_cluster() : nullptr);
    if (ct) {
        AllocRequest req = m.u.req;
        Allocation a;
        int err = governor_place(ct, req, &a);
        if (err) return finish_alloc(m, err);
        m.u.alloc = a;
        return placed(m);
    }
    Message q = m;
    q.seq = next_seq();
    Pending p;
    p.orig = m;
    p.stage = ST_WAIT_PLACE;
    p.t0_ns = now_ns();
    if (send_peer(0, q)) pending_[q.seq] = p;
    else finish_alloc(m, OCM_ENODAEMON);
}

void Daemon::placed(Message &m) {
    Allocation &a = m.u.alloc;
    if (a.remote_rank == rank_) {
        int err = do_alloc_local(m);
        if (err) note_free(a.remote_rank, (int)a.tier, a.bytes);
        return finish_alloc(m, err);
    }
    if (TableHeader *pt = peer_table(a.remote_rank)) {
        BlockRef b;
        if (inject_alloc_fault()) {
            note_free(a.remote_rank, (int)a.tier, a.bytes);
            return finish_alloc(m, OCM_ENOMEM);
        }
        if (a.tier < TIER_COUNT && pool_alloc(pt, (int)a.tier, a.bytes, m.pid, m.rank, &b) == OCM_OK) {
            a.rem_alloc_id = b.id;
            a.offset = b.offset;
            a.block_bytes = b.block_bytes;
            a.region = b.region;
            a.generation = b.generation;
            return finish_alloc(m, 0);
        }
        // full (maybe only until its scrubber has run), or a refused request: let the owner decide and wait
    }
    Message q = m;
    q.type = MSG_DO_ALLOC;
    q.status = MSG_REQUEST;
    q.seq = next_seq();
    Pending p;
    p.orig = m;
    p.stage = ST_WAIT_DO_ALLOC;
    p.t0_ns = now_ns();
    if (send_peer(a.remote_rank, q)) pending_[q.seq] = p;
    else {
        note_free(a.remote_rank, (int)a.tier, a.bytes);
        finish_alloc(m, OCM_ENODAEMON);
    }
}

void Daemon::finish_alloc(Message &m, int err) {
    m.type = MSG_RELEASE_APP;
    m.status = MSG_RESPONSE;
    m.err = err;
    reply_app(m);
}

// fault injection for tests: OCM_FAULT_ALLOC_EVERY=N makes every N-th fulfilment fail
bool Daemon::inject_alloc_fault() {
    static const long fault_every = getenv("OCM_FAULT_ALLOC_EVERY") ? atol(getenv("OCM_FAULT_ALLOC_EVERY")) : 0;
    static long fault_count = 0;
    if (fault_every > 0 && ++fault_count % fault_every == 0) {
        LOGW("fault injection: failing allocation %ld", fault_count);
        return true;
    }
    return false;
}

// The tables of another rank of this node, or nullptr (other node, not up yet, restarted and not yet back).
TableHeader *Daemon::peer_table(int rank) {
    if (!node_local_ || rank < 0 || rank >= n_ranks_ || rank >= kMaxRanks || rank == rank_) return nullptr;
    ClusterTable *ct = shared_cluster();
    if (!ct || !ct->node[rank].joined.load(std::memory_order_acquire)) return nullptr;
    PeerTable &p = peer_tbl_[rank];
    if (p.hdr && p.hdr->daemon_pid != ct->node[rank].daemon_pid) {   // that daemon was restarted: new segment, same name
        shm_close(&p.map, false);
        p = PeerTable();
    }
    if (p.hdr) return p.hdr;
    const uint64_t now = now_ns();
    if (now < p.retry_ns) return nullptr;
    p.retry_ns = now + 200000000ull;
    std::string err;
    if (!shm_open_existing(shm_table_name(rank), &p.map, &err)) return nullptr;
    auto *h = static_cast<TableHeader *>(p.map.base);
    if (!table_validate(h, p.map.bytes) || h->rank != rank || h->daemon_pid != ct->node[rank].daemon_pid ||
        !h->ready.load(std::memory_order_acquire)) {
        shm_close(&p.map, false);
        p.map = ShmMapping();
        return nullptr;
    }
    p.hdr = h;
    LOGI("rank %d shares this node: its requests are served from the shared tables", rank);
    return h;
}

// Owner-side fulfilment (reference alloc_ate, src/alloc.c:150-222): pure table work.
int Daemon::do_alloc_local(Message &m) {
    Allocation &a = m.u.alloc;
    if (a.tier >= TIER_COUNT) return OCM_EINVAL;
    if (inject_alloc_fault()) return OCM_ENOMEM;
    BlockRef b;
    int err = pool_alloc(tbl_, (int)a.tier, a.bytes, m.pid, m.rank, &b);
    // blocks waiting to be zeroed still hold their space: give the scrubber a bounded moment
    for (uint64_t t0 = now_ns(); err == OCM_ENOMEM && pool_has_dirty(tbl_, (int)a.tier) && now_ns() - t0 < 50000000ull;) {
        if (a.tier == TIER_HOST) pool_scrub_host_inline(tbl_, host_map_.base);
        else usleep(100);
        err = pool_alloc(tbl_, (int)a.tier, a.bytes, m.pid, m.rank, &b);
    }
    if (err) return err;
    a.rem_alloc_id = b.id;
    a.offset = b.offset;
    a.block_bytes = b.block_bytes;
    a.region = b.region;
    a.generation = b.generation;
    return OCM_OK;
}

// Owner-side release (reference dealloc_ate, src/alloc.c:230-282).
int Daemon::do_free_local(const Allocation &a) {
    if (a.tier >= TIER_COUNT) return OCM_EINVAL;
    return pool_free(tbl_, (int)a.tier, a.region, a.rem_alloc_id);
}

void Daemon::note_free(int owner_rank, int tier, uint64_t bytes) {
    if (ClusterTable *ct = rank_ == 0 ? cluster_ : (node_local_ ? shared_cluster() : nullptr)) {
        governor_release(ct, owner_rank, tier, bytes);
        return;
    }
    Message n;
    n.init(MSG_NOTE_FREE, MSG_NO_STATUS);
    n.rank = rank_;
    n.u.alloc.remote_rank = owner_rank;
    n.u.alloc.tier = (uint32_t)tier;
    n.u.alloc.bytes = bytes;
    send_peer(0, n);
}

void Daemon::start_free(Message &m) {
    Allocation &a = m.u.alloc;
    m.rank = rank_;
    if (a.remote_rank == rank_) {
        int err = do_free_local(a);
        if (!err) note_free(rank_, (int)a.tier, a.bytes);
        m.type = MSG_RELEASE_APP;
        m.status = MSG_RESPONSE;
        m.err = err;
        return reply_app(m);
    }
    if (TableHeader *pt = peer_table(a.remote_rank)) {
        int err = a.tier < TIER_COUNT ? pool_free(pt, (int)a.tier, a.region, a.rem_alloc_id) : OCM_EINVAL;
        if (!err) note_free(a.remote_rank, (int)a.tier, a.bytes);
        m.type = MSG_RELEASE_APP;
        m.status = MSG_RESPONSE;
        m.err = err;
        return reply_app(m);
    }
    Message q = m;
    q.type = MSG_DO_FREE;
    q.status = MSG_REQUEST;
    q.seq = next_seq();
    Pending p;
    p.orig = m;
    p.stage = ST_WAIT_DO_FREE;
    p.t0_ns = now_ns();
    if (send_peer(a.remote_rank, q)) pending_[q.seq] = p;
    else {
        m.type = MSG_RELEASE_APP;
        m.status = MSG_RESPONSE;
        m.err = OCM_ENODAEMON;
        reply_app(m);
    }
}

void Daemon::fill_stats(StatsPayload *s) {
    memset(s, 0, sizeof *s);
    for (int t = 0; t < TIER_COUNT; ++t) {
        PoolDesc &p = tbl_->pool[t];
        s->pool_bytes[t] = p.bytes;
        s->pool_granule_bytes[t] = p.used_bytes.load(std::memory_order_relaxed);
        s->pool_used[t] = pool_payload_bytes(tbl_, t);   // what applications hold (empty cached slabs are free space)
        s->allocs_live += p.live.load(std::memory_order_relaxed);
        s->allocs_total += p.total_allocs.load(std::memory_order_relaxed);
        s->frees_total += p.total_frees.load(std::memory_order_relaxed);
    }
    s->apps = apps_.size();
    s->msgs_in = msgs_in_.load();
    s->msgs_out = msgs_out_.load();
    s->reclaimed_regions = reclaimed_.load();
    s->scrubbed_bytes = scrubbed_bytes_.load();
    s->table_faults = table_faults_.load();
}

// ------------------------------------------------------------ NVLS bring-up --
ClusterTable *Daemon::shared_cluster() {
    if (cluster_) return cluster_;
    if (!cluster_peer_map_.base) {
        std::string err;
        if (!shm_open_existing(shm_cluster_name(), &cluster_peer_map_, &err)) return nullptr;
    }
    auto *ct = static_cast<ClusterTable *>(cluster_peer_map_.base);
    if (cluster_peer_map_.bytes < sizeof(ClusterTable) || ct->magic != kTableMagic || ct->version != kTableVersion)
        return nullptr;
    return ct;
}

// Every device must be added to the multicast object before any rank binds memory, so the daemons
// advance in lock step through two counters in the cluster table:
//   add own device -> mc_added++ ... all added -> bind own pool tail -> mc_bound++ ... all bound -> READY
void Daemon::mc_step() {
    if (mc_done_ || !mc_bytes_ || !joined_) return;
    ClusterTable *ct = shared_cluster();
    if (!ct) return;
    const uint32_t st = ct->mc_state.load(std::memory_order_acquire);
    if (st == MC_FAILED || st == MC_NONE) {
        if (st == MC_FAILED || now_ns() > join_deadline_ns_) mc_done_ = true;   // rank 0 has no multicast object
        return;
    }
    std::string err;
    if (!mc_added_) {
        if (ct->mc_bytes != mc_bytes_) {
            LOGW("multicast size mismatch (rank 0: %llu, here: %llu): set the same OCM_MC_BYTES everywhere",
                 (unsigned long long)ct->mc_bytes, (unsigned long long)mc_bytes_);
            ct->mc_state.store(MC_FAILED, std::memory_order_release);
            mc_done_ = true;
            return;
        }
        if (rank_ != 0) {
            UdsReply rp;
            int fd = -1;
            if (!uds_fetch(uds_name(0), UDS_GET_MULTICAST, gpu_, &rp, &fd) || rp.err || fd < 0) {
                if (fd >= 0) close(fd);
                return;   // rank 0 not serving yet: next tick
            }
            bool ok = gpu::mc_import(fd, rp.bytes, &mc_, &err);
            close(fd);
            if (!ok) {
                LOGW("%s", err.c_str());
                ct->mc_state.store(MC_FAILED, std::memory_order_release);
                mc_done_ = true;
                return;
            }
        }
        if (!gpu::mc_add_device(&mc_, gpu_, &err)) {
            LOGW("%s", err.c_str());
            ct->mc_state.store(MC_FAILED, std::memory_order_release);
            mc_done_ = true;
            return;
        }
        mc_added_ = true;
        ct->mc_added.fetch_add(1, std::memory_order_acq_rel);
    }
    if (!mc_bound_ && ct->mc_added.load(std::memory_order_acquire) == (uint32_t)n_ranks_) {
        if (!gpu::mc_bind(&mc_, 0, &gpu_pool_, mc_tail_off_, mc_bytes_, &err)) {
            LOGW("%s", err.c_str());
            ct->mc_state.store(MC_FAILED, std::memory_order_release);
            mc_done_ = true;
            return;
        }
        mc_bound_ = true;
        ct->mc_bound.fetch_add(1, std::memory_order_acq_rel);
    }
    if (mc_bound_ && ct->mc_bound.load(std::memory_order_acquire) == (uint32_t)n_ranks_) {
        if (rank_ == 0) {
            ct->mc_state.store(MC_READY, std::memory_order_release);
            LOGI("NVLS multicast region ready: %llu bytes on %d GPUs", (unsigned long long)mc_bytes_, n_ranks_);
        }
        mc_done_ = true;
    }
}

// ------------------------------------------------------- fd server (UDS) ----
void Daemon::on_uds_accept() {
    for (;;) {
        int fd = accept4(uds_listen_, nullptr, nullptr, SOCK_CLOEXEC);
        if (fd < 0) break;
        UdsRequest rq;
        int r = sock_get(fd, &rq, sizeof rq, 1000);
        UdsReply rp;
        memset(&rp, 0, sizeof rp);
        rp.magic = kUdsMagic;
        int pass = -1;
        if (r != 1 || rq.magic != kUdsMagic) {
            rp.err = OCM_EPROTO;
        } else if (rq.what == UDS_GET_GPU_POOL) {
            rp.bytes = gpu_pool_.bytes;
            rp.owner_dev = gpu_;
            rp.mode = gpu_pool_.ptr ? gpu_pool_.mode : GPU_SHARE_NONE;
            if (!gpu_pool_.ptr) rp.err = OCM_ENOGPU;
            else if (gpu_pool_.mode == GPU_SHARE_VMM_FD) pass = gpu_pool_.share_fd;
        } else if (rq.what == UDS_GET_HOST_POOL) {
            rp.bytes = host_vmm_.bytes;
            rp.owner_dev = host_vmm_.owner_dev;       // NUMA node
            rp.mode = host_vmm_.ptr ? GPU_SHARE_VMM_FD : GPU_SHARE_NONE;
            if (host_vmm_.ptr) pass = host_vmm_.share_fd;
            else rp.err = OCM_EUNSUPPORTED;
        } else if (rq.what == UDS_GET_MULTICAST) {
            // rank 0 hands its multicast object to the other daemons and to applications
            if (rank_ == 0 && mc_.fd >= 0) {
                rp.bytes = mc_.bytes;
                rp.owner_dev = gpu_;
                rp.aux = cluster_ ? cluster_->mc_state.load() : 0;
                pass = mc_.fd;
            } else {
                rp.err = OCM_EUNSUPPORTED;
            }
        } else {
            rp.err = OCM_EINVAL;
        }
        uds_send_fd(fd, &rp, sizeof rp, pass);
        sock_close(fd);
    }
}

// ------------------------------------------------------------ housekeeping --
static bool pid_is_dead(int pid, int /*rank*/, void *) {
    if (pid <= 0) return false;
    return kill(pid, 0) != 0 && errno == ESRCH;
}

// Reclaim regions whose owner process is gone (README:68-69 of the reference
// lists this as future work) and forget dead applications.
void Daemon::reclaim_dead() {
    for (auto it = apps_.begin(); it != apps_.end();) {
        if (pid_is_dead(it->first, 0, nullptr)) {
            LOGI("app %d died, dropping", it->first);
            box_.detach(it->second.mq);
            mq_unlink(it->second.mq.c_str());
            it = apps_.erase(it);
        } else {
            ++it;
        }
    }
    for (int t = 0; t < TIER_COUNT; ++t) {
        std::vector<BlockRef> got;
        uint64_t bytes = 0;
        size_t n = pool_reclaim(tbl_, t, pid_is_dead, nullptr, &bytes, &got);
        if (!n) continue;
        reclaimed_.fetch_add(n);
        LOGI("reclaimed %zu block(s), %llu bytes in tier %d from dead owners", n, (unsigned long long)bytes, t);
        for (auto &b : got)
            if (const Region *rec = table_region_rec(tbl_, t, b.region)) note_free(rank_, t, rec->bytes);
    }
    // multicast ranges of this rank's dead applications: mark them; every daemon's scrubber zeroes its own slice
    if (mc_bytes_ && gpu_ >= 0) {
        ClusterTable *ct = shared_cluster();
        if (ct && ct->mc_state.load(std::memory_order_acquire) == MC_READY) {
            mc_cluster_.store(ct, std::memory_order_release);
            size_t n = mc_mark_dead(ct, rank_, [](int pid) { return pid_is_dead(pid, 0, nullptr); });
            if (n) LOGI("%zu multicast range(s) of dead owners marked for scrubbing", n);
        }
    }
}

void Daemon::on_timer() {
    uint64_t exp;
    while (read(timer_fd_, &exp, sizeof exp) > 0) {
    }
    uint64_t now = now_ns();
    tbl_->heartbeat_ns.store(now, std::memory_order_relaxed);
    if (joined_) {   // the placement code of every process reads this one (governor_rank_alive)
        if (ClusterTable *ct = shared_cluster()) {
            ct->node[rank_].heartbeat_ns.store(now, std::memory_order_relaxed);
            hb_cluster_.store(ct, std::memory_order_release);
        }
    }
    if (!joined_) {
        bool waiting = false;
        for (auto &kv : pending_) waiting |= kv.second.stage == ST_WAIT_JOIN;
        if (!waiting && now - last_join_try_ns_ > 100ull * 1000000ull) {
            if (now > join_deadline_ns_) {
                // reference: "start the master first" then raise(SIGINT) (src/mem.c:466-474)
                LOGE("could not reach rank 0 within %d ms - is the master daemon running?", opt_.join_timeout_ms);
                stop_.store(true);
                return;
            }
            join_cluster();
        }
    }
    // expire requests stuck on a dead peer
    for (auto it = pending_.begin(); it != pending_.end();) {
        if (now - it->second.t0_ns > 10ull * 1000000000ull) {
            Message r = it->second.orig;
            Stage st = it->second.stage;
            it = pending_.erase(it);
            if (st == ST_WAIT_JOIN) continue;
            r.type = st == ST_WAIT_STATS ? (uint16_t)MSG_STATS : (uint16_t)MSG_RELEASE_APP;
            r.status = MSG_RESPONSE;
            r.err = OCM_ENODAEMON;
            reply_app(r);
        } else {
            ++it;
        }
    }
    mc_step();
    if (now - last_audit_ns_ > 1000ull * 1000000ull) {
        last_audit_ns_ = now;
        audit_tables();
    }
    if (now - last_reclaim_ns_ > 250ull * 1000000ull) {
        last_reclaim_ns_ = now;
        reclaim_dead();
    }
}

// Every attached process can CAS this rank's tables (that is what makes ocm_alloc a handful of atomics), so a buggy
// client could leave them inconsistent.  Once a second the daemon checks that the bitmap, the byte counters and the
// records still tell the same story.  A disagreement only counts when it shows twice in a row with no allocation or
// release in between (other processes are allowed to be in the middle of an operation); it is logged and exported
// through MSG_STATS (table_faults) - the daemon does not try to repair a table somebody else scribbled on.
void Daemon::audit_tables() {
    for (int t = 0; t < TIER_COUNT; ++t) {
        PoolDesc &p = tbl_->pool[t];
        if (!p.n_granules) continue;
        const uint64_t ops_before = p.total_allocs.load() + p.total_frees.load();
        const uint64_t dirty_before = p.dirty_bytes.load();
        TableAudit a;
        const int problems = pool_audit(tbl_, t, &a);
        const bool quiet = ops_before == p.total_allocs.load() + p.total_frees.load() && dirty_before == p.dirty_bytes.load();
        if (!problems || !quiet) {
            audit_suspect_[t] = 0;
            continue;
        }
        if (audit_suspect_[t] == ops_before + 1) {   // the same picture as a second ago: nobody is in the middle of anything
            table_faults_.fetch_add(1);
            LOGE("tier %d tables are inconsistent: bitmap %llu B, used %llu B, dirty %llu B, live records %llu, live counter %llu",
                 t, (unsigned long long)a.bitmap_bytes, (unsigned long long)a.used_bytes, (unsigned long long)a.dirty_bytes,
                 (unsigned long long)a.live_records, (unsigned long long)a.live_counter);
            audit_suspect_[t] = 0;
        } else {
            audit_suspect_[t] = ops_before + 1;
        }
    }
}

// Zero released blocks off the allocation path and hand them back.
void Daemon::scrub_loop() {
    if (gpu_ >= 0) cudaSetDevice(gpu_);
    std::vector<DirtyBlock> blocks;
    std::vector<std::pair<uint32_t, uint32_t>> mc_dead;
    uint64_t last_mc_scan_ns = 0, last_beat_ns = 0;
    while (!stop_.load(std::memory_order_relaxed)) {
        bool worked = false;
        // second source of the liveness heartbeat (see hb_cluster_): "this process is alive", whatever the main
        // loop is doing right now
        const uint64_t t_now = now_ns();
        if (t_now - last_beat_ns > 10ull * 1000000ull) {
            last_beat_ns = t_now;
            if (tbl_) tbl_->heartbeat_ns.store(t_now, std::memory_order_relaxed);
            if (ClusterTable *hc = hb_cluster_.load(std::memory_order_acquire))
                hc->node[rank_].heartbeat_ns.store(t_now, std::memory_order_relaxed);
        }
        if (tbl_) {
            for (int t = 0; t < TIER_COUNT; ++t) {
                blocks.clear();
                if (!pool_take_dirty(tbl_, t, &blocks)) continue;
                worked = true;
                if (t == TIER_GPU && gpu_pool_.ptr) {
                    for (auto &b : blocks)
                        ocmk::launch_fill((char *)gpu_pool_.ptr + b.offset, b.bytes, 0, (cudaStream_t)scrub_stream_);
                    cudaStreamSynchronize((cudaStream_t)scrub_stream_);
                } else if (t == TIER_HOST && host_vmm_.ptr) {
                    for (auto &b : blocks) {
                        const Region *rec = table_region_rec(tbl_, t, b.region);
                        uint64_t n = std::min<uint64_t>(b.bytes, ((rec ? rec->bytes : b.bytes) + 4095) & ~4095ull);
                        if (n <= (64u << 10)) memset((char *)host_map_.base + b.offset, 0, n);
                        else ocmk::launch_fill((char *)host_map_.base + b.offset, n, 0, (cudaStream_t)scrub_stream_);
                    }
                    cudaStreamSynchronize((cudaStream_t)scrub_stream_);
                } else if (t == TIER_HOST && host_map_.base) {
                    for (auto &b : blocks) {
                        // only the part the owner could touch needs clearing
                        const Region *rec = table_region_rec(tbl_, t, b.region);
                        uint64_t n = std::min<uint64_t>(b.bytes, ((rec ? rec->bytes : b.bytes) + 4095) & ~4095ull);
                        memset((char *)host_map_.base + b.offset, 0, n);
                    }
                }
                for (auto &b : blocks) {
                    scrubbed_bytes_.fetch_add(b.bytes, std::memory_order_relaxed);
                    pool_release_scrubbed(tbl_, t, b.region);
                }
            }
        }
        // multicast ranges whose owner died: zero OUR GPU's slice; the daemon that finishes last frees the range
        ClusterTable *ct = mc_cluster_.load(std::memory_order_acquire);
        if (ct && gpu_pool_.ptr && tbl_ && now_ns() - last_mc_scan_ns > 50ull * 1000000ull) {
            last_mc_scan_ns = now_ns();
            mc_dead.clear();
            if (mc_dead_ranges(ct, rank_, &mc_dead)) {
                char *tail = (char *)gpu_pool_.ptr + tbl_->mc_tail_offset;
                for (auto &r : mc_dead) {
                    ocmk::launch_fill(tail + (uint64_t)r.first * kMcGranule, (uint64_t)r.second * kMcGranule, 0,
                                      (cudaStream_t)scrub_stream_);
                    ocmk::launch_fill(tail + (uint64_t)r.first * 8, 8, 0, (cudaStream_t)scrub_stream_);   // its signal word
                }
                cudaStreamSynchronize((cudaStream_t)scrub_stream_);
                for (auto &r : mc_dead) {
                    scrubbed_bytes_.fetch_add((uint64_t)r.second * kMcGranule, std::memory_order_relaxed);
                    if (mc_range_scrubbed(ct, rank_, r.first))
                        LOGI("multicast range at granule %u (%u granules) reclaimed from a dead owner", r.first, r.second);
                }
                worked = true;
            }
        }
        if (!worked) usleep(300);
    }
}

int Daemon::run() {
    LOGI("oncillamem rank %d/%d up (pid %d, ns '%s')", rank_, n_ranks_, (int)getpid(), ns_tag().c_str());
    struct epoll_event evs[32];
    // Requests come in bursts (an application's allocation loop, the two daemon hops of one request): after any event
    // the loop keeps looking without sleeping for a short while, so that the next message does not pay an epoll
    // wake-up (5-20 us each, four of them on one remote allocation).  OCM_DAEMON_SPIN_US=0 sleeps at once.
    const uint64_t spin_ns = (uint64_t)env_size("OCM_DAEMON_SPIN_US", 60) * 1000ull;
    uint64_t hot_until = 0;
    while (!stop_.load()) {
        int n = epoll_wait(ep_, evs, 32, hot_until ? 0 : 100);
        if (n == 0 && hot_until) {
            if (now_ns() >= hot_until) hot_until = 0;
#if defined(__x86_64__)
            else __builtin_ia32_pause();
#endif
            continue;
        }
        if (n > 0 && spin_ns) hot_until = now_ns() + spin_ns;
        if (n < 0 && errno == EINTR) continue;
        if (n < 0) {
            LOGE("epoll_wait: %s", strerror(errno));
            return -errno;
        }
        for (int i = 0; i < n; ++i) {
            int fd = evs[i].data.fd;
            auto it = fds_.find(fd);
            if (it == fds_.end()) continue;
            switch (it->second) {
            case FD_MQ: on_mailbox(); break;
            case FD_TCP_LISTEN: on_tcp_accept(); break;
            case FD_UDS_LISTEN: on_uds_accept(); break;
            case FD_PEER_OUT: on_peer_readable(fd, true); break;
            case FD_PEER_IN: on_peer_readable(fd, false); break;
            case FD_TIMER: on_timer(); break;
            case FD_SIGNAL: {
                struct signalfd_siginfo si;
                while (read(signal_fd_, &si, sizeof si) > 0) {
                }
                LOGI("signal received, shutting down");
                stop_.store(true);
                break;
            }
            }
        }
    }
    return 0;
}

} // namespace ocm
