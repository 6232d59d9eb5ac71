// GPU memory plumbing shared by the daemon (owner side) and libocm (importer
// side).  Replaces memory registration + key exchange of the reference
// transports (ibv_reg_mr + {va,rkey,len} in RDMA-CM private data, reference
// src/rdma_server.c:109-158; rma2_register + {node,vpid,NLA}, src/extoll_server.c:74-111):
// the owner creates its pool once with the CUDA virtual-memory API and exports
// a POSIX file descriptor; importers map it and grant their own device access,
// after which peer HBM is a plain device pointer reachable over NVLink.  A
// legacy cudaIpc path (64-byte handle that fits in the table header) is kept
// for systems without fd export.  No driver call happens per allocation.
//
// libcuda is never linked: driver entry points are resolved through the
// runtime (cudaGetDriverEntryPoint), so the library loads on CPU-only hosts.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>

namespace ocm {
namespace gpu {

// Number of usable CUDA devices; 0 on CPU-only hosts (never fails).
int device_count();
bool set_device(int dev, std::string *err);
int current_device();
bool get_uuid(int dev, uint8_t out[16]);
bool mem_info(int dev, size_t *free_b, size_t *total_b);

struct PoolMapping {
    void    *ptr = nullptr;       // device pointer usable from `access_dev`
    size_t   bytes = 0;
    int      owner_dev = -1;
    int      access_dev = -1;
    int      share_fd = -1;       // owner: exported descriptor (VMM mode)
    uint64_t vmm_handle = 0;      // CUmemGenericAllocationHandle
    uint32_t mode = 0;            // GpuShareMode
    bool     imported = false;
};

bool vmm_export_supported(int dev);
size_t vmm_granularity(int dev);

// Owner: create a pool of `bytes` (rounded up to the VMM granularity) on `dev`,
// map it read/write for `dev` and export it.  mode_pref: 0 auto, 1 VMM, 2 cudaIpc.
bool pool_create(int dev, size_t bytes, int mode_pref, PoolMapping *out, uint8_t ipc_handle_out[64], std::string *err);
// Importer: map a pool exported by another process for use from `access_dev`.
bool pool_import_fd(int fd, size_t bytes, int owner_dev, int access_dev, PoolMapping *out, std::string *err);
bool pool_import_ipc(const uint8_t ipc_handle[64], size_t bytes, int owner_dev, int access_dev, PoolMapping *out,
                     std::string *err);
void pool_release(PoolMapping *m);
// Let one more device of this process access an existing VMM mapping (same virtual address).
// Pinned host memory as a VMM allocation (cuMemCreate, location = host NUMA node `numa`): created and pinned once by
// the owner, shared by file descriptor like an HBM pool.  An importing process maps it in milliseconds per GiB,
// where cudaHostRegister of a shared-memory segment pins every page again in every process (measured on the B200
// box, 4 GiB: 0.12 s against 0.88 s; bench/hostvmm_probe.cu).  The mapping is readable and writable by `access_dev`
// and by the CPU.  PoolMapping.owner_dev carries the NUMA node.
bool host_pool_create(int numa, size_t bytes, int access_dev, PoolMapping *out, std::string *err);
bool host_pool_import_fd(int fd, size_t bytes, int numa, int access_dev, PoolMapping *out, std::string *err);
bool pool_grant_access(const PoolMapping *m, int dev, std::string *err);

// ---- NVLS multicast objects (cuMulticast*) ------------------------------------------------------
// One object spans the same number of bytes on every GPU of the cluster; a store to its virtual address
// (multimem.st) is replicated by the NVSwitch into the memory each device bound to it.
struct McObject {
    uint64_t handle = 0;      // CUmemGenericAllocationHandle of the multicast object
    size_t   bytes = 0;
    int      fd = -1;         // creator: exported descriptor
    void    *va = nullptr;    // mapped multicast address (mc_map)
    int      va_dev = -1;
};
bool   mc_supported(int dev);
size_t mc_granularity(int n_devices, size_t bytes, bool recommended);
bool   mc_create(int n_devices, size_t bytes, McObject *out, std::string *err);
bool   mc_import(int fd, size_t bytes, McObject *out, std::string *err);
bool   mc_add_device(McObject *mc, int dev, std::string *err);
// Bind [mem_off, mem_off+bytes) of a VMM pool created by pool_create to [mc_off, mc_off+bytes) of the object.
bool   mc_bind(McObject *mc, size_t mc_off, const PoolMapping *pool, size_t mem_off, size_t bytes, std::string *err);
bool   mc_map(McObject *mc, int dev, std::string *err);
void   mc_release(McObject *mc);

// Pin + map an existing host range (shared-memory pools, tables) so kernels and
// async copies can use it.  Idempotent per range.
bool host_register(void *p, size_t bytes, std::string *err);
void host_unregister(void *p);

bool can_access_peer(int dev, int peer);
// Legacy peer enablement (needed for cudaIpc mappings and cudaMalloc'ed memory).
bool enable_peer_access(int dev, int peer);

const char *last_error_string();

} // namespace gpu
} // namespace ocm
