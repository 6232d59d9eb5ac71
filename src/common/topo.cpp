#include "topo.hpp"
#include "gpu_rt.hpp"
#include "log.hpp"
#include <algorithm>
#include <cctype>
#include <cstring>
#include <cuda_runtime.h>
#include <dirent.h>
#include <fstream>
#include <sched.h>
#include <sys/syscall.h>
#include <thread>
#include <unistd.h>

namespace ocm {
namespace topo {

std::vector<int> parse_cpulist(const std::string &s) {
    std::vector<int> out;
    size_t i = 0;
    while (i < s.size()) {
        while (i < s.size() && !isdigit((unsigned char)s[i])) ++i;
        if (i >= s.size()) break;
        long a = 0;
        while (i < s.size() && isdigit((unsigned char)s[i])) a = a * 10 + (s[i++] - '0');
        long b = a;
        if (i < s.size() && s[i] == '-') {
            ++i;
            b = 0;
            bool any = false;
            while (i < s.size() && isdigit((unsigned char)s[i])) {
                b = b * 10 + (s[i++] - '0');
                any = true;
            }
            if (!any) b = a;
        }
        for (long c = a; c <= b && c < 8192; ++c) out.push_back((int)c);
    }
    return out;
}

static std::string slurp(const std::string &path) {
    std::ifstream f(path);
    std::string s;
    if (f) std::getline(f, s);
    return s;
}

std::vector<int> numa_nodes() {
    std::vector<int> out;
    if (DIR *d = opendir("/sys/devices/system/node")) {
        while (struct dirent *e = readdir(d)) {
            if (strncmp(e->d_name, "node", 4) || !isdigit((unsigned char)e->d_name[4])) continue;
            const int n = atoi(e->d_name + 4);
            if (!node_cpus(n).empty()) out.push_back(n);
        }
        closedir(d);
    }
    if (out.empty()) out.push_back(0);
    std::sort(out.begin(), out.end());
    return out;
}

std::vector<int> node_cpus(int node) {
    if (node < 0) return {};
    return parse_cpulist(slurp("/sys/devices/system/node/node" + std::to_string(node) + "/cpulist"));
}

int pci_numa_node(const std::string &bdf, const std::string &root) {
    std::string b = bdf;
    for (auto &ch : b) ch = (char)tolower((unsigned char)ch);
    // cudaDeviceGetPCIBusId prints an 8-digit domain ("00000000:1B:00.0"); sysfs uses 4
    if (b.size() > 12 && b.compare(0, 4, "0000") == 0 && b[8] == ':') b = b.substr(4);
    std::string s = slurp(root + "/bus/pci/devices/" + b + "/numa_node");
    if (s.empty()) return -1;
    char *end = nullptr;
    long v = strtol(s.c_str(), &end, 10);
    return end == s.c_str() ? -1 : (int)v;
}

bool bind_thread_to_node(int node) {
    std::vector<int> cpus = node_cpus(node);
    if (cpus.empty()) return false;
    cpu_set_t allowed, want;
    CPU_ZERO(&allowed);
    CPU_ZERO(&want);
    if (sched_getaffinity(0, sizeof allowed, &allowed) != 0) return false;
    int n = 0;
    for (int c : cpus)
        if (c < CPU_SETSIZE && CPU_ISSET(c, &allowed)) {
            CPU_SET(c, &want);
            ++n;
        }
    return n > 0 && sched_setaffinity(0, sizeof want, &want) == 0;
}

bool prefer_node_for_range(void *p, size_t bytes, int node) {
    if (node < 0 || node >= 1024 || !p || !bytes) return false;
#ifdef SYS_mbind
    unsigned long mask[16] = {0};
    mask[node / (8 * sizeof(unsigned long))] |= 1ul << (node % (8 * sizeof(unsigned long)));
    const int kMpolPreferred = 1;
    return syscall(SYS_mbind, p, bytes, kMpolPreferred, mask, (unsigned long)(sizeof mask * 8), 0u) == 0;
#else
    return false;
#endif
}

void touch_range(void *p, size_t bytes, int node, int threads) {
    if (!p || !bytes) return;
    threads = std::max(1, std::min(threads, 64));
    const size_t page = 4096;
    const size_t per = ((bytes + threads - 1) / threads + page - 1) / page * page;
    std::vector<std::thread> ts;
    for (int t = 0; t < threads; ++t) {
        const size_t lo = (size_t)t * per;
        if (lo >= bytes) break;
        const size_t hi = std::min(bytes, lo + per);
        ts.emplace_back([=] {
            if (node >= 0) bind_thread_to_node(node);
            volatile char *c = static_cast<volatile char *>(p);
            for (size_t o = lo; o < hi; o += page) c[o] = c[o];   // a write fault places the page, the content stays
        });
    }
    for (auto &t : ts) t.join();
}

bool discover_gpu(int device, GpuTopo *out) {
    *out = GpuTopo();
    const int n = gpu::device_count();
    if (device < 0 || device >= n) return false;
    out->device = device;
    char bdf[32] = {0};
    if (cudaDeviceGetPCIBusId(bdf, sizeof bdf, device) == cudaSuccess) {
        std::string b = bdf;
        for (auto &ch : b) ch = (char)tolower((unsigned char)ch);
        if (b.size() > 12 && b.compare(0, 4, "0000") == 0 && b[8] == ':') b = b.substr(4);
        snprintf(out->pci_bdf, sizeof out->pci_bdf, "%s", b.c_str());
        out->numa_node = pci_numa_node(b);
    } else {
        cudaGetLastError();
    }
    out->peer_mask = 1u << device;
    for (int d = 0; d < n && d < 32; ++d)
        if (d != device && gpu::can_access_peer(device, d)) out->peer_mask |= 1u << d;
    out->multicast = gpu::mc_supported(device);
    out->vmm_fd_export = gpu::vmm_export_supported(device);
    return true;
}

} // namespace topo
} // namespace ocm
