// Tracing hooks (the reference has no timers, counters or trace points at all).
// NVTX ranges around the public entry points, enabled with OCM_NVTX=1; header-only NVTX v3 resolves the
// profiler's injection library at run time, so there is no link dependency and no cost when disabled.
#pragma once
#include <cstdlib>
#include <nvtx3/nvToolsExt.h>

namespace ocm {

inline bool nvtx_enabled() {
    static const bool on = [] {
        const char *e = getenv("OCM_NVTX");
        return e && *e && *e != '0';
    }();
    return on;
}

struct TraceRange {
    bool on;
    explicit TraceRange(const char *name) : on(nvtx_enabled()) {
        if (on) nvtxRangePushA(name);
    }
    ~TraceRange() {
        if (on) nvtxRangePop();
    }
};

} // namespace ocm
