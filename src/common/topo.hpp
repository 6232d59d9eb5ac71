// Topology discovery for placement and for the pinned-host tier.
//
// The reference's node configuration (`struct alloc_node_config {ib_ip, ram, gpu_mem[8], num_gpu}`,
// reference inc/alloc.h:57-64) is declared, sent at join (src/main.c:143-160) and then never used
// for any decision.  Here every daemon records, for the GPU it serves,
//   * which other GPUs of the box that GPU can reach directly (cudaDeviceCanAccessPeer row),
//   * whether it supports NVLS multicast,
//   * its PCI address and the host NUMA node it hangs off (sysfs),
// and rank 0 keeps that in the cluster table so that `governor_place` never puts an HBM far side
// on a GPU the requester cannot load from / store to, and so that the rank's pinned host pool is
// first-touched and bound on the GPU's own NUMA node (PCIe reads then stay on one socket).
// Everything degrades to "unknown" (-1 / empty) on virtual machines that expose no NUMA layout:
// the B200 VMs this was measured on report numa_node = -1 for every GPU and a single node.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace ocm {
namespace topo {

// "0-3,8,10-11" -> {0,1,2,3,8,10,11}; garbage yields what could be parsed.
std::vector<int> parse_cpulist(const std::string &s);

// Host NUMA nodes that have CPUs, ascending.  {0} on machines without a NUMA directory.
std::vector<int> numa_nodes();
std::vector<int> node_cpus(int node);

// NUMA node of a PCI device ("0000:1b:00.0", any case), -1 when the kernel does not know.
// root: sysfs mount point (tests pass a fake tree).
int pci_numa_node(const std::string &bdf, const std::string &root = "/sys");

// Restrict the calling thread to the CPUs of `node` (no-op and false when unknown / not permitted).
bool bind_thread_to_node(int node);
// Prefer `node` for the pages of [p, p+bytes) (mbind MPOL_PREFERRED; pages already touched stay where
// they are).  False when the kernel refuses (containers without CAP_SYS_NICE) - callers then rely on
// first touch from a thread bound with bind_thread_to_node.
bool prefer_node_for_range(void *p, size_t bytes, int node);
// First-touch [p, p+bytes) from `threads` helper threads bound to `node` (node < 0: unbound).
void touch_range(void *p, size_t bytes, int node, int threads = 4);

struct GpuTopo {
    int      device = -1;
    char     pci_bdf[16] = {0};
    int      numa_node = -1;
    uint32_t peer_mask = 0;        // bit d: this GPU can access device d directly (own bit always set)
    bool     multicast = false;
    bool     vmm_fd_export = false;
};
// Fills what the CUDA runtime and sysfs know about `device`.  False without a usable GPU.
bool discover_gpu(int device, GpuTopo *out);

} // namespace topo
} // namespace ocm
