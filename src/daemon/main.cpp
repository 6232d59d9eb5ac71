// oncillamem - the per-rank daemon executable.
// Usage mirrors the reference (`bin/oncillamem <nodefile>`, reference
// src/main.c:162-166,187-224) with optional overrides.
#include "daemon.hpp"
#include "../common/gpu_rt.hpp"
#include "../common/log.hpp"
#include "../common/topo.hpp"
#include <csignal>
#include <cstdio>
#include <cstring>
#include <sys/prctl.h>
#include <unistd.h>

static void usage(const char *argv0) {
    fprintf(stderr,
            "Usage: %s <nodefile> [--rank N] [--gpu-bytes SIZE] [--host-bytes SIZE] [--no-gpu] [--join-timeout MS] [--check]\n"
            "  --check: parse the nodefile, print the topology and this process's rank, then exit\n"
            "  nodefile lines: rank dns ethernet_ip ocm_port [rdmacm_port] [gpu=<i>] [gpu_bytes=<n>] [host_bytes=<n>] [numa=<node>]\n"
            "  environment: OCM_RANK, OCM_NS, OCM_GPU_BYTES, OCM_HOST_BYTES, OCM_LOG|OCM_VERBOSE, OCM_NO_SCRUB, OCM_GPU_IPC\n",
            argv0);
}

// --check: parse the nodefile, show the topology and the rank this process would serve, start nothing.
static int check_nodefile(const ocm::DaemonOptions &o) {
    ocm::Nodefile nf;
    std::string err;
    if (!ocm::parse_nodefile(o.nodefile, &nf, &err, o.rank_override)) {
        fprintf(stderr, "%s: %s\n", o.nodefile.c_str(), err.c_str());
        return 1;
    }
    printf("%d rank(s); this process would be rank %d%s\n", nf.size(), nf.my_rank,
           nf.my_rank < 0 ? " (no OCM_RANK and no line for this hostname)" : nf.my_rank == 0 ? " (the master)" : "");
    printf("%4s  %-20s %-15s %6s  %-5s %12s %12s\n", "rank", "dns", "ethernet_ip", "port", "gpu", "gpu_bytes", "host_bytes");
    for (const auto &n : nf.nodes) {
        char gpu[16];
        if (n.gpu == -2) snprintf(gpu, sizeof gpu, "auto");
        else if (n.gpu == -1) snprintf(gpu, sizeof gpu, "none");
        else snprintf(gpu, sizeof gpu, "%d", n.gpu);
        printf("%4d  %-20s %-15s %6d  %-5s %12llu %12llu\n", n.rank, n.dns.c_str(), n.ip_eth.c_str(), n.ocm_port, gpu,
               (unsigned long long)n.gpu_bytes, (unsigned long long)n.host_bytes);
    }
    // what this machine offers: host NUMA nodes, and per GPU its PCI address, NUMA node and peer-access row
    auto nodes = ocm::topo::numa_nodes();
    printf("host NUMA nodes:");
    for (int n : nodes) printf(" %d(%zu cpus)", n, ocm::topo::node_cpus(n).size());
    printf("\n");
    const int ndev = o.no_gpu ? 0 : ocm::gpu::device_count();
    if (ndev > 0) {
        printf("%4s  %-13s %5s %5s %6s  peer access (P = direct load/store, . = none)\n", "gpu", "pci", "numa", "nvls", "vmmfd");
        for (int d = 0; d < ndev; ++d) {
            ocm::topo::GpuTopo t;
            if (!ocm::topo::discover_gpu(d, &t)) continue;
            printf("%4d  %-13s %5d %5s %6s  ", d, t.pci_bdf, t.numa_node, t.multicast ? "yes" : "no", t.vmm_fd_export ? "yes" : "no");
            for (int p = 0; p < ndev; ++p) printf("%c", p == d ? 'X' : (t.peer_mask >> p) & 1u ? 'P' : '.');
            printf("\n");
        }
    } else {
        printf("no CUDA device visible: host tier only\n");
    }
    return nf.my_rank < 0 ? 1 : 0;
}

int main(int argc, char **argv) {
    ocm::DaemonOptions o;
    bool check_only = false;
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
        else if (!strcmp(a, "--check")) check_only = true;
        else if (!strcmp(a, "-h") || !strcmp(a, "--help")) { usage(argv[0]); return 0; }
        else if (a[0] == '-') { usage(argv[0]); return 2; }
        else o.nodefile = a;
    }
    if (o.nodefile.empty()) {
        usage(argv[0]);
        return 2;
    }
    if (check_only) return check_nodefile(o);
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
