// pmsg - process mailboxes over POSIX message queues.
// Same model as the reference (reference src/pmsg.c, inc/pmsg.h:31-51): every
// process owns ONE receive-only queue; to talk to a peer you attach to the
// peer's queue write-only.  Differences: blocking receive sleeps in poll(2) on
// the queue descriptor instead of spinning on EAGAIN (src/pmsg.c:134-151); the
// descriptor can be put into the daemon's epoll set; names carry an optional
// namespace and rank (names.hpp); stale queues are detected on open instead of
// sweeping mq_unlink over every pid up to pid_max (src/pmsg.c:495-548).
#pragma once
#include <cstddef>
#include <map>
#include <mutex>
#include <string>
#include <vector>
#include <mqueue.h>

namespace ocm {

class Mailbox {
  public:
    Mailbox() = default;
    ~Mailbox();
    Mailbox(const Mailbox &) = delete;
    Mailbox &operator=(const Mailbox &) = delete;

    // Create this process' receive queue.  `steal_stale`: if the name exists,
    // unlink it and retry (safe for pid-named queues: no other live process can
    // own our pid's name).  Returns 0 or -errno.
    int open(const std::string &name, size_t msg_size, int depth = 8, bool steal_stale = true);
    int close(bool unlink_it = true);
    bool is_open() const { return rq_ != (mqd_t)-1; }
    int fd() const { return (int)rq_; }
    const std::string &name() const { return name_; }
    size_t msg_size() const { return msg_size_; }

    // timeout_ms < 0: block; 0: poll once.  Returns 1 message received, 0 timeout, -errno error.
    int recv(void *msg, int timeout_ms);
    int pending() const;

    // Sender side.
    int attach(const std::string &peer);
    int detach(const std::string &peer);
    bool attached(const std::string &peer) const;
    // Returns 0, -ETIMEDOUT if the peer queue stayed full, -ENOTCONN if not attached, -errno.
    int send(const std::string &peer, const void *msg, int timeout_ms = 5000);

    // Remove queues whose name starts with `prefix` (needs /dev/mqueue mounted;
    // otherwise only `also` names are unlinked).  Returns number removed.
    static int cleanup(const std::string &prefix, const std::vector<std::string> &also);

  private:
    mqd_t rq_ = (mqd_t)-1;
    std::string name_;
    size_t msg_size_ = 0;
    mutable std::mutex mu_;
    std::map<std::string, mqd_t> peers_;
};

} // namespace ocm

// ---- C interface with the reference's signatures (inc/pmsg.h:31-51) -------
extern "C" {
#define PMSG_DAEMON_PID (-1)
int pmsg_init(size_t pmsg_size);
int pmsg_open(pid_t self_pid);
int pmsg_close(void);
int pmsg_attach(pid_t to_pid);
int pmsg_detach(pid_t to_pid);
int pmsg_send(pid_t to_pid, void *msg);
int pmsg_recv(void *msg, bool block);
int pmsg_cleanup(void);
int pmsg_pending(void);
}
