// Tiny request/reply exchanged on the daemon's abstract unix socket.  The reply
// may carry one file descriptor (the CUDA VMM export of the rank's HBM pool or of
// its pinned host pool, or rank 0's multicast object).
#pragma once
#include "../ipc/sock.hpp"
#include <cstdint>
#include <string>
#include <unistd.h>

namespace ocm {

constexpr uint32_t kUdsMagic = 0x4F434D55u; // "OCMU"
enum UdsWhat : uint32_t { UDS_GET_GPU_POOL = 1, UDS_GET_MULTICAST = 2, UDS_GET_HOST_POOL = 3 };

struct UdsRequest {
    uint32_t magic;
    uint32_t what;
    int32_t  pid;
    int32_t  device;
};
struct UdsReply {
    uint32_t magic;
    int32_t  err;
    uint64_t bytes;
    int32_t  owner_dev;
    uint32_t mode;       // GpuShareMode
    uint64_t aux;
};

// Client side: one exchange with the fd server called `name`.  true when a valid
// reply arrived; *fd_out is the received descriptor or -1 (caller closes it).
inline bool uds_fetch(const std::string &name, uint32_t what, int device, UdsReply *rp, int *fd_out, int timeout_ms = 2000) {
    *fd_out = -1;
    int s = uds_connect(name, timeout_ms);
    if (s < 0) return false;
    UdsRequest rq = {kUdsMagic, what, (int32_t)getpid(), device};
    bool ok = sock_put(s, &rq, sizeof rq) == 1 && uds_recv_fd(s, rp, sizeof *rp, fd_out, 5000) == 1 && rp->magic == kUdsMagic;
    sock_close(s);
    if (!ok && *fd_out >= 0) {
        close(*fd_out);
        *fd_out = -1;
    }
    return ok;
}

} // namespace ocm
