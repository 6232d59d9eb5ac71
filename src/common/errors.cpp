// Error strings shared by libocm and the daemon.
#include "../../include/oncillamem.h"

extern "C" const char *ocm_strerror(int err) {
    if (err < 0) err = -err;
    switch (err) {
    case OCM_OK: return "ok";
    case OCM_EINVAL: return "invalid argument";
    case OCM_ENOMEM: return "out of pool memory";
    case OCM_ENODAEMON: return "daemon not reachable";
    case OCM_ENOGPU: return "no usable GPU";
    case OCM_ERANGE: return "offset/length out of range";
    case OCM_ECUDA: return "CUDA error";
    case OCM_EPROTO: return "protocol error";
    case OCM_ESTALE: return "stale handle (double free?)";
    case OCM_EUNSUPPORTED: return "unsupported";
    default: return "unknown error";
    }
}
