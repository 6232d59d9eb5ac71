#include "pmsg.hpp"
#include <cstdlib>
#include "../common/log.hpp"
#include "../common/names.hpp"
#include <cerrno>
#include <cstring>
#include <dirent.h>
#include <fcntl.h>
#include <poll.h>
#include <sys/stat.h>
#include <time.h>
#include <vector>

namespace ocm {

Mailbox::~Mailbox() {
    for (auto &kv : peers_) mq_close(kv.second);
    peers_.clear();
    if (rq_ != (mqd_t)-1) close(true);
}

int Mailbox::open(const std::string &name, size_t msg_size, int depth, bool steal_stale) {
    if (rq_ != (mqd_t)-1) return -EBUSY;
    struct mq_attr attr;
    memset(&attr, 0, sizeof attr);
    attr.mq_maxmsg = depth;
    attr.mq_msgsize = (long)msg_size;
    for (int attempt = 0; attempt < 2; ++attempt) {
        mqd_t q = mq_open(name.c_str(), O_RDONLY | O_CREAT | O_EXCL | O_NONBLOCK, 0600, &attr);
        if (q != (mqd_t)-1) {
            rq_ = q;
            name_ = name;
            msg_size_ = msg_size;
            return 0;
        }
        if (errno == EEXIST && steal_stale && attempt == 0) {
            LOGW("stale mailbox %s: unlinking", name.c_str());
            mq_unlink(name.c_str());
            continue;
        }
        int e = errno;
        LOGE("mq_open(%s): %s", name.c_str(), strerror(e));
        return -e;
    }
    return -EEXIST;
}

int Mailbox::close(bool unlink_it) {
    if (rq_ == (mqd_t)-1) return -EBADF;
    mq_close(rq_);
    if (unlink_it) mq_unlink(name_.c_str());
    rq_ = (mqd_t)-1;
    return 0;
}

int Mailbox::recv(void *msg, int timeout_ms) {
    if (rq_ == (mqd_t)-1) return -EBADF;
    for (;;) {
        ssize_t n = mq_receive(rq_, static_cast<char *>(msg), msg_size_, nullptr);
        if (n >= 0) return 1;
        if (errno == EINTR) continue;
        if (errno != EAGAIN) return -errno;
        if (timeout_ms == 0) return 0;
        // A reply usually follows its request within tens of microseconds: look again for a short while before
        // sleeping in poll(), whose wake-up alone costs 5-20 us (OCM_MQ_SPIN_US, 0 = sleep at once).
        static const long spin_us = [] {
            const char *e = getenv("OCM_MQ_SPIN_US");
            return e && *e ? atol(e) : 40L;
        }();
        if (spin_us > 0) {
            struct timespec t0, t1;
            clock_gettime(CLOCK_MONOTONIC, &t0);
            bool got = false;
            for (;;) {
                n = mq_receive(rq_, static_cast<char *>(msg), msg_size_, nullptr);
                if (n >= 0) { got = true; break; }
                if (errno != EAGAIN && errno != EINTR) return -errno;
                clock_gettime(CLOCK_MONOTONIC, &t1);
                if ((t1.tv_sec - t0.tv_sec) * 1000000L + (t1.tv_nsec - t0.tv_nsec) / 1000 >= spin_us) break;
#if defined(__x86_64__)
                __builtin_ia32_pause();
#endif
            }
            if (got) return 1;
        }
        struct pollfd p = {(int)rq_, POLLIN, 0};
        int r = poll(&p, 1, timeout_ms);
        if (r == 0) return 0;
        if (r < 0 && errno != EINTR) return -errno;
    }
}

int Mailbox::pending() const {
    if (rq_ == (mqd_t)-1) return -EBADF;
    struct mq_attr a;
    if (mq_getattr(rq_, &a)) return -errno;
    return (int)a.mq_curmsgs;
}

int Mailbox::attach(const std::string &peer) {
    std::lock_guard<std::mutex> g(mu_);
    if (peers_.count(peer)) return 0;
    mqd_t q = mq_open(peer.c_str(), O_WRONLY | O_NONBLOCK);
    if (q == (mqd_t)-1) return -errno;
    peers_[peer] = q;
    return 0;
}

int Mailbox::detach(const std::string &peer) {
    std::lock_guard<std::mutex> g(mu_);
    auto it = peers_.find(peer);
    if (it == peers_.end()) return -ENOTCONN;
    mq_close(it->second);
    peers_.erase(it);
    return 0;
}

bool Mailbox::attached(const std::string &peer) const {
    std::lock_guard<std::mutex> g(mu_);
    return peers_.count(peer) != 0;
}

int Mailbox::send(const std::string &peer, const void *msg, int timeout_ms) {
    mqd_t q;
    {
        std::lock_guard<std::mutex> g(mu_);
        auto it = peers_.find(peer);
        if (it == peers_.end()) return -ENOTCONN;
        q = it->second;
    }
    struct timespec t0;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (;;) {
        if (mq_send(q, static_cast<const char *>(msg), msg_size_, 0) == 0) return 0;
        if (errno == EINTR) continue;
        if (errno != EAGAIN) return -errno;
        // receiver's queue is full: wait until it drains (writable)
        struct timespec now;
        clock_gettime(CLOCK_MONOTONIC, &now);
        long el = (now.tv_sec - t0.tv_sec) * 1000 + (now.tv_nsec - t0.tv_nsec) / 1000000;
        if (timeout_ms >= 0 && el >= timeout_ms) return -ETIMEDOUT;
        struct pollfd p = {(int)q, POLLOUT, 0};
        poll(&p, 1, timeout_ms < 0 ? 100 : (int)std::min<long>(100, timeout_ms - el));
    }
}

int Mailbox::cleanup(const std::string &prefix, const std::vector<std::string> &also) {
    int n = 0;
    for (auto &s : also)
        if (mq_unlink(s.c_str()) == 0) ++n;
    if (DIR *d = opendir("/dev/mqueue")) {
        std::string bare = prefix.size() && prefix[0] == '/' ? prefix.substr(1) : prefix;
        while (struct dirent *e = readdir(d)) {
            if (strncmp(e->d_name, bare.c_str(), bare.size()) == 0) {
                std::string nm = std::string("/") + e->d_name;
                if (mq_unlink(nm.c_str()) == 0) ++n;
            }
        }
        closedir(d);
    }
    return n;
}

} // namespace ocm

// ------------------------------------------------------------- C wrappers --
namespace {
ocm::Mailbox g_box;
size_t g_msg_size = 0;
std::string peer_name(pid_t pid) {
    if (pid == PMSG_DAEMON_PID) {
        const char *r = getenv("OCM_RANK");
        return ocm::mq_daemon_name(r ? atoi(r) : 0);
    }
    return ocm::mq_app_name((int)pid);
}
} // namespace

extern "C" {
int pmsg_init(size_t pmsg_size) {
    if (!pmsg_size) return -1;
    g_msg_size = pmsg_size;
    return 0;
}
int pmsg_open(pid_t self_pid) {
    if (!g_msg_size) return -1;
    return g_box.open(peer_name(self_pid), g_msg_size) ? -1 : 0;
}
int pmsg_close(void) { return g_box.close(true) ? -1 : 0; }
int pmsg_attach(pid_t to_pid) { return g_box.attach(peer_name(to_pid)) ? -1 : 0; }
int pmsg_detach(pid_t to_pid) { return g_box.detach(peer_name(to_pid)) ? -1 : 0; }
int pmsg_send(pid_t to_pid, void *msg) { return g_box.send(peer_name(to_pid), msg, -1) ? -1 : 0; }
int pmsg_recv(void *msg, bool block) {
    int r = g_box.recv(msg, block ? -1 : 0);
    return r == 1 ? 0 : -1;
}
int pmsg_cleanup(void) {
    return ocm::Mailbox::cleanup("/ocm_mq_" + ocm::ns_prefix(), {}) >= 0 ? 0 : -1;
}
int pmsg_pending(void) { return g_box.pending(); }
}
