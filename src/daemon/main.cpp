// oncillamem - the per-rank daemon executable.
// Usage mirrors the reference (`bin/oncillamem <nodefile>`, reference
// src/main.c:162-166,187-224) with optional overrides.
#include "daemon.hpp"
#include "../common/log.hpp"
#include <csignal>
#include <cstdio>
#include <cstring>
#include <sys/prctl.h>
#include <unistd.h>

static void usage(const char *argv0) {
    fprintf(stderr,
            "Usage: %s <nodefile> [--rank N] [--gpu-bytes SIZE] [--host-bytes SIZE] [--no-gpu] [--join-timeout MS]\n"
            "  nodefile lines: rank dns ethernet_ip ocm_port [rdmacm_port] [gpu=<i>] [gpu_bytes=<n>] [host_bytes=<n>]\n"
            "  environment: OCM_RANK, OCM_NS, OCM_GPU_BYTES, OCM_HOST_BYTES, OCM_LOG|OCM_VERBOSE, OCM_NO_SCRUB, OCM_GPU_IPC\n",
            argv0);
}

int main(int argc, char **argv) {
    ocm::DaemonOptions o;
    for (int i = 1; i < argc; ++i) {
        const char *a = argv[i];
        auto need = [&](const char *name) -> const char * {
            if (i + 1 >= argc) {
                fprintf(stderr, "%s needs a value\n", name);
                exit(2);
            }
            return argv[++i];
        };
        if (!strcmp(a, "--rank")) o.rank_override = atoi(need(a));
        else if (!strcmp(a, "--gpu-bytes")) {
            if (!ocm::parse_size(need(a), &o.gpu_bytes)) { usage(argv[0]); return 2; }
        } else if (!strcmp(a, "--host-bytes")) {
            if (!ocm::parse_size(need(a), &o.host_bytes)) { usage(argv[0]); return 2; }
        } else if (!strcmp(a, "--no-gpu")) o.no_gpu = true;
        else if (!strcmp(a, "--join-timeout")) o.join_timeout_ms = atoi(need(a));
        else if (!strcmp(a, "-h") || !strcmp(a, "--help")) { usage(argv[0]); return 0; }
        else if (a[0] == '-') { usage(argv[0]); return 2; }
        else o.nodefile = a;
    }
    if (o.nodefile.empty()) {
        usage(argv[0]);
        return 2;
    }
    // launched by a harness (LocalCluster): do not outlive it, so a killed benchmark cannot leak pools
    if (getenv("OCM_DIE_WITH_PARENT")) {
        prctl(PR_SET_PDEATHSIG, SIGTERM);
        if (getppid() == 1) return 1;   // the parent is already gone
    }
    ocm::Daemon d(o);
    int r = d.init();
    if (r) {
        LOGE("daemon initialisation failed (%d)", r);
        return 1;
    }
    // readiness line for launch scripts
    printf("oncillamem ready\n");
    fflush(stdout);
    r = d.run();
    return r ? 1 : 0;
}
