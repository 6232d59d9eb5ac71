// Tiny request/reply exchanged on the daemon's abstract unix socket.  The reply
// may carry one file descriptor (the CUDA VMM export of the rank's HBM pool).
#pragma once
#include <cstdint>

namespace ocm {

constexpr uint32_t kUdsMagic = 0x4F434D55u; // "OCMU"
enum UdsWhat : uint32_t { UDS_GET_GPU_POOL = 1, UDS_GET_MULTICAST = 2 };

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

} // namespace ocm
