#include "gpu_rt.hpp"
#include "log.hpp"
#include "tables.hpp"
#include <cuda.h>
#include <cuda_runtime.h>
#include <map>
#include <mutex>
#include <unistd.h>

namespace ocm {
namespace gpu {

namespace {
thread_local char g_err[256] = "";
void set_err(std::string *err, const std::string &s) {
    snprintf(g_err, sizeof g_err, "%s", s.c_str());
    if (err) *err = s;
}
std::string cu_str(CUresult r) { return "CUresult " + std::to_string((int)r); }

struct Driver {
    bool ok = false;
    decltype(&cuMemCreate) MemCreate = nullptr;
    decltype(&cuMemRelease) MemRelease = nullptr;
    decltype(&cuMemAddressReserve) MemAddressReserve = nullptr;
    decltype(&cuMemAddressFree) MemAddressFree = nullptr;
    decltype(&cuMemMap) MemMap = nullptr;
    decltype(&cuMemUnmap) MemUnmap = nullptr;
    decltype(&cuMemSetAccess) MemSetAccess = nullptr;
    decltype(&cuMemExportToShareableHandle) MemExport = nullptr;
    decltype(&cuMemImportFromShareableHandle) MemImport = nullptr;
    decltype(&cuMemGetAllocationGranularity) MemGranularity = nullptr;
    decltype(&cuDeviceGetAttribute) DeviceGetAttribute = nullptr;
    // multicast entry points are optional (older drivers): mc_ok tells
    bool mc_ok = false;
    decltype(&cuMulticastCreate) MulticastCreate = nullptr;
    decltype(&cuMulticastAddDevice) MulticastAddDevice = nullptr;
    decltype(&cuMulticastBindMem) MulticastBindMem = nullptr;
    decltype(&cuMulticastGetGranularity) MulticastGetGranularity = nullptr;
};

template <typename F> bool load(const char *name, F *fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult st;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &st) != cudaSuccess || !p ||
        st != cudaDriverEntryPointSuccess) {
        cudaGetLastError();
        return false;
    }
    *fn = reinterpret_cast<F>(p);
    return true;
}

Driver &drv() {
    static Driver d = [] {
        Driver x;
        if (device_count() <= 0) return x;
        x.ok = load("cuMemCreate", &x.MemCreate) && load("cuMemRelease", &x.MemRelease) &&
               load("cuMemAddressReserve", &x.MemAddressReserve) && load("cuMemAddressFree", &x.MemAddressFree) &&
               load("cuMemMap", &x.MemMap) && load("cuMemUnmap", &x.MemUnmap) &&
               load("cuMemSetAccess", &x.MemSetAccess) && load("cuMemExportToShareableHandle", &x.MemExport) &&
               load("cuMemImportFromShareableHandle", &x.MemImport) &&
               load("cuMemGetAllocationGranularity", &x.MemGranularity) &&
               load("cuDeviceGetAttribute", &x.DeviceGetAttribute);
        x.mc_ok = x.ok && load("cuMulticastCreate", &x.MulticastCreate) &&
                  load("cuMulticastAddDevice", &x.MulticastAddDevice) && load("cuMulticastBindMem", &x.MulticastBindMem) &&
                  load("cuMulticastGetGranularity", &x.MulticastGetGranularity);
        return x;
    }();
    return d;
}

CUmemAllocationProp pool_prop(int dev) {
    CUmemAllocationProp p = {};
    p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    p.location.id = dev;
    p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return p;
}
} // namespace

const char *last_error_string() { return g_err; }

int device_count() {
    static int n = [] {
        if (getenv("OCM_NO_GPU")) return 0;
        int c = 0;
        cudaError_t e = cudaGetDeviceCount(&c);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return 0;
        }
        return c;
    }();
    return n;
}

bool set_device(int dev, std::string *err) {
    if (dev < 0 || dev >= device_count()) {
        set_err(err, "no such CUDA device " + std::to_string(dev));
        return false;
    }
    cudaError_t e = cudaSetDevice(dev);
    if (e == cudaSuccess) e = cudaFree(0);
    if (e != cudaSuccess) {
        set_err(err, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
        return false;
    }
    return true;
}

int current_device() {
    int d = -1;
    if (device_count() <= 0 || cudaGetDevice(&d) != cudaSuccess) return -1;
    return d;
}

bool get_uuid(int dev, uint8_t out[16]) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) return false;
    memcpy(out, p.uuid.bytes, 16);
    return true;
}

bool mem_info(int dev, size_t *free_b, size_t *total_b) {
    int cur = current_device();
    if (cur != dev && cudaSetDevice(dev) != cudaSuccess) return false;
    bool ok = cudaMemGetInfo(free_b, total_b) == cudaSuccess;
    if (cur >= 0 && cur != dev) cudaSetDevice(cur);
    return ok;
}

bool vmm_export_supported(int dev) {
    if (!drv().ok) return false;
    int v = 0;
    if (drv().DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev) != CUDA_SUCCESS || !v)
        return false;
    v = 0;
    if (drv().DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev) != CUDA_SUCCESS)
        return false;
    return v != 0;
}

size_t vmm_granularity(int dev) {
    if (!drv().ok) return 2 << 20;
    CUmemAllocationProp p = pool_prop(dev);
    size_t g = 0;
    if (drv().MemGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || !g) g = 2 << 20;
    return g;
}

static bool map_handle(CUmemGenericAllocationHandle h, size_t bytes, int access_dev, void **ptr, std::string *err) {
    CUdeviceptr va = 0;
    size_t gran = 2 << 20;
    CUresult r = drv().MemAddressReserve(&va, bytes, gran, 0, 0);
    if (r != CUDA_SUCCESS) {
        set_err(err, "cuMemAddressReserve: " + cu_str(r));
        return false;
    }
    r = drv().MemMap(va, bytes, 0, h, 0);
    if (r != CUDA_SUCCESS) {
        drv().MemAddressFree(va, bytes);
        set_err(err, "cuMemMap: " + cu_str(r));
        return false;
    }
    CUmemAccessDesc ad = {};
    ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ad.location.id = access_dev;
    ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = drv().MemSetAccess(va, bytes, &ad, 1);
    if (r != CUDA_SUCCESS) {
        drv().MemUnmap(va, bytes);
        drv().MemAddressFree(va, bytes);
        set_err(err, "cuMemSetAccess(dev " + std::to_string(access_dev) + "): " + cu_str(r));
        return false;
    }
    *ptr = reinterpret_cast<void *>(va);
    return true;
}

bool pool_create(int dev, size_t bytes, int mode_pref, PoolMapping *out, uint8_t ipc_handle_out[64], std::string *err) {
    if (!set_device(dev, err)) return false;
    if (const char *e = getenv("OCM_GPU_IPC")) {
        if (!strcmp(e, "legacy") || !strcmp(e, "cudaipc")) mode_pref = GPU_SHARE_CUDA_IPC;
        else if (!strcmp(e, "vmm")) mode_pref = GPU_SHARE_VMM_FD;
    }
    bool vmm = mode_pref == GPU_SHARE_VMM_FD || (mode_pref == 0 && vmm_export_supported(dev));
    if (vmm) {
        size_t g = vmm_granularity(dev);
        size_t sz = (bytes + g - 1) / g * g;
        CUmemAllocationProp p = pool_prop(dev);
        CUmemGenericAllocationHandle h = 0;
        CUresult r = drv().ok ? drv().MemCreate(&h, sz, &p, 0) : CUDA_ERROR_NOT_SUPPORTED;
        if (r == CUDA_SUCCESS) {
            void *ptr = nullptr;
            int fd = -1;
            if (map_handle(h, sz, dev, &ptr, err)) {
                r = drv().MemExport(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
                if (r == CUDA_SUCCESS) {
                    out->ptr = ptr; out->bytes = sz; out->owner_dev = dev; out->access_dev = dev;
                    out->share_fd = fd; out->vmm_handle = (uint64_t)h; out->mode = GPU_SHARE_VMM_FD;
                    out->imported = false;
                    if (ipc_handle_out) memset(ipc_handle_out, 0, 64);
                    return true;
                }
                set_err(err, "cuMemExportToShareableHandle: " + cu_str(r));
                drv().MemUnmap((CUdeviceptr)ptr, sz);
                drv().MemAddressFree((CUdeviceptr)ptr, sz);
            }
            drv().MemRelease(h);
        } else {
            set_err(err, "cuMemCreate(" + std::to_string(sz) + "): " + cu_str(r));
        }
        if (mode_pref == GPU_SHARE_VMM_FD) return false;
        LOGW("VMM pool creation failed (%s); falling back to cudaIpc", g_err);
    }
    void *ptr = nullptr;
    cudaError_t e = cudaMalloc(&ptr, bytes);
    if (e != cudaSuccess) {
        set_err(err, std::string("cudaMalloc pool: ") + cudaGetErrorString(e));
        return false;
    }
    cudaIpcMemHandle_t ih;
    e = cudaIpcGetMemHandle(&ih, ptr);
    if (e != cudaSuccess) {
        set_err(err, std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e));
        cudaFree(ptr);
        return false;
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle must fit the table header");
    if (ipc_handle_out) memcpy(ipc_handle_out, &ih, 64);
    out->ptr = ptr; out->bytes = bytes; out->owner_dev = dev; out->access_dev = dev;
    out->share_fd = -1; out->vmm_handle = 0; out->mode = GPU_SHARE_CUDA_IPC; out->imported = false;
    return true;
}

bool pool_import_fd(int fd, size_t bytes, int owner_dev, int access_dev, PoolMapping *out, std::string *err) {
    if (!drv().ok) {
        set_err(err, "CUDA driver VMM entry points unavailable");
        return false;
    }
    CUmemGenericAllocationHandle h = 0;
    CUresult r = drv().MemImport(&h, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (r != CUDA_SUCCESS) {
        set_err(err, "cuMemImportFromShareableHandle: " + cu_str(r));
        return false;
    }
    void *ptr = nullptr;
    if (!map_handle(h, bytes, access_dev, &ptr, err)) {
        drv().MemRelease(h);
        return false;
    }
    out->ptr = ptr; out->bytes = bytes; out->owner_dev = owner_dev; out->access_dev = access_dev;
    out->share_fd = -1; out->vmm_handle = (uint64_t)h; out->mode = GPU_SHARE_VMM_FD; out->imported = true;
    return true;
}

bool pool_import_ipc(const uint8_t ipc_handle[64], size_t bytes, int owner_dev, int access_dev, PoolMapping *out,
                     std::string *err) {
    cudaIpcMemHandle_t ih;
    memcpy(&ih, ipc_handle, 64);
    void *ptr = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&ptr, ih, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_err(err, std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e));
        return false;
    }
    out->ptr = ptr; out->bytes = bytes; out->owner_dev = owner_dev; out->access_dev = access_dev;
    out->share_fd = -1; out->vmm_handle = 0; out->mode = GPU_SHARE_CUDA_IPC; out->imported = true;
    return true;
}

bool pool_grant_access(const PoolMapping *m, int dev, std::string *err) {
    if (!m->ptr || m->mode != GPU_SHARE_VMM_FD || !drv().ok) {
        set_err(err, "not a VMM mapping");
        return false;
    }
    CUmemAccessDesc ad = {};
    ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ad.location.id = dev;
    ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CUresult r = drv().MemSetAccess((CUdeviceptr)m->ptr, m->bytes, &ad, 1);
    if (r != CUDA_SUCCESS) {
        set_err(err, "cuMemSetAccess(dev " + std::to_string(dev) + "): " + cu_str(r));
        return false;
    }
    return true;
}

// ------------------------------------------------------------ host pool (VMM) --
static bool map_host_handle(CUmemGenericAllocationHandle h, size_t bytes, int numa, int access_dev, void **ptr, std::string *err) {
    CUdeviceptr va = 0;
    CUresult r = drv().MemAddressReserve(&va, bytes, 2 << 20, 0, 0);
    if (r != CUDA_SUCCESS) {
        set_err(err, "cuMemAddressReserve(host pool): " + cu_str(r));
        return false;
    }
    r = drv().MemMap(va, bytes, 0, h, 0);
    if (r != CUDA_SUCCESS) {
        drv().MemAddressFree(va, bytes);
        set_err(err, "cuMemMap(host pool): " + cu_str(r));
        return false;
    }
    CUmemAccessDesc ad[2] = {};
    ad[0].location.type = CU_MEM_LOCATION_TYPE_HOST_NUMA;
    ad[0].location.id = numa;
    ad[0].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    ad[1].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ad[1].location.id = access_dev;
    ad[1].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = drv().MemSetAccess(va, bytes, ad, access_dev >= 0 ? 2 : 1);
    if (r != CUDA_SUCCESS) {
        drv().MemUnmap(va, bytes);
        drv().MemAddressFree(va, bytes);
        set_err(err, "cuMemSetAccess(host pool, dev " + std::to_string(access_dev) + "): " + cu_str(r));
        return false;
    }
    *ptr = reinterpret_cast<void *>(va);
    return true;
}

bool host_pool_create(int numa, size_t bytes, int access_dev, PoolMapping *out, std::string *err) {
    if (!drv().ok || access_dev < 0) {
        set_err(err, "CUDA driver VMM entry points unavailable");
        return false;
    }
    if (!set_device(access_dev, err)) return false;
    if (numa < 0) numa = 0;
    CUmemAllocationProp p = {};
    p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    p.location.type = CU_MEM_LOCATION_TYPE_HOST_NUMA;
    p.location.id = numa;
    p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t g = 0;
    if (drv().MemGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || !g) g = 2 << 20;
    const size_t sz = (bytes + g - 1) / g * g;
    CUmemGenericAllocationHandle h = 0;
    CUresult r = drv().MemCreate(&h, sz, &p, 0);
    if (r != CUDA_SUCCESS) {
        set_err(err, "cuMemCreate(host NUMA " + std::to_string(numa) + ", " + std::to_string(sz) + "): " + cu_str(r));
        return false;
    }
    void *ptr = nullptr;
    if (!map_host_handle(h, sz, numa, access_dev, &ptr, err)) {
        drv().MemRelease(h);
        return false;
    }
    int fd = -1;
    r = drv().MemExport(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        set_err(err, "cuMemExportToShareableHandle(host pool): " + cu_str(r));
        drv().MemUnmap((CUdeviceptr)ptr, sz);
        drv().MemAddressFree((CUdeviceptr)ptr, sz);
        drv().MemRelease(h);
        return false;
    }
    out->ptr = ptr; out->bytes = sz; out->owner_dev = numa; out->access_dev = access_dev;
    out->share_fd = fd; out->vmm_handle = (uint64_t)h; out->mode = GPU_SHARE_VMM_FD; out->imported = false;
    return true;
}

bool host_pool_import_fd(int fd, size_t bytes, int numa, int access_dev, PoolMapping *out, std::string *err) {
    if (!drv().ok) {
        set_err(err, "CUDA driver VMM entry points unavailable");
        return false;
    }
    CUmemGenericAllocationHandle h = 0;
    CUresult r = drv().MemImport(&h, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (r != CUDA_SUCCESS) {
        set_err(err, "cuMemImportFromShareableHandle(host pool): " + cu_str(r));
        return false;
    }
    void *ptr = nullptr;
    if (!map_host_handle(h, bytes, numa < 0 ? 0 : numa, access_dev, &ptr, err)) {
        drv().MemRelease(h);
        return false;
    }
    out->ptr = ptr; out->bytes = bytes; out->owner_dev = numa; out->access_dev = access_dev;
    out->share_fd = -1; out->vmm_handle = (uint64_t)h; out->mode = GPU_SHARE_VMM_FD; out->imported = true;
    return true;
}

void pool_release(PoolMapping *m) {
    if (!m->ptr) return;
    if (m->mode == GPU_SHARE_VMM_FD && drv().ok) {
        drv().MemUnmap((CUdeviceptr)m->ptr, m->bytes);
        drv().MemAddressFree((CUdeviceptr)m->ptr, m->bytes);
        drv().MemRelease((CUmemGenericAllocationHandle)m->vmm_handle);
        if (m->share_fd >= 0) close(m->share_fd);
    } else if (m->mode == GPU_SHARE_CUDA_IPC) {
        if (m->imported) cudaIpcCloseMemHandle(m->ptr);
        else cudaFree(m->ptr);
    }
    m->ptr = nullptr;
    m->share_fd = -1;
}

// ---------------------------------------------------------------- multicast --
bool mc_supported(int dev) {
    if (!drv().mc_ok || dev < 0) return false;
    int v = 0;
    if (drv().DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return false;
    return v != 0;
}

size_t mc_granularity(int n_devices, size_t bytes, bool recommended) {
    if (!drv().mc_ok) return 0;
    CUmulticastObjectProp p = {};
    p.numDevices = (unsigned)n_devices;
    p.size = bytes;
    p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t g = 0;
    if (drv().MulticastGetGranularity(&g, &p, recommended ? CU_MULTICAST_GRANULARITY_RECOMMENDED
                                                           : CU_MULTICAST_GRANULARITY_MINIMUM) != CUDA_SUCCESS)
        return 0;
    return g;
}

bool mc_create(int n_devices, size_t bytes, McObject *out, std::string *err) {
    if (!drv().mc_ok) {
        set_err(err, "multicast entry points unavailable");
        return false;
    }
    CUmulticastObjectProp p = {};
    p.numDevices = (unsigned)n_devices;
    p.size = bytes;
    p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUmemGenericAllocationHandle h = 0;
    CUresult r = drv().MulticastCreate(&h, &p);
    if (r != CUDA_SUCCESS) {
        set_err(err, "cuMulticastCreate(" + std::to_string(n_devices) + " devices, " + std::to_string(bytes) + " B): " + cu_str(r));
        return false;
    }
    int fd = -1;
    r = drv().MemExport(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        drv().MemRelease(h);
        set_err(err, "export of the multicast object: " + cu_str(r));
        return false;
    }
    out->handle = (uint64_t)h;
    out->bytes = bytes;
    out->fd = fd;
    return true;
}

bool mc_import(int fd, size_t bytes, McObject *out, std::string *err) {
    if (!drv().mc_ok) {
        set_err(err, "multicast entry points unavailable");
        return false;
    }
    CUmemGenericAllocationHandle h = 0;
    CUresult r = drv().MemImport(&h, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (r != CUDA_SUCCESS) {
        set_err(err, "import of the multicast object: " + cu_str(r));
        return false;
    }
    out->handle = (uint64_t)h;
    out->bytes = bytes;
    out->fd = -1;
    return true;
}

bool mc_add_device(McObject *mc, int dev, std::string *err) {
    CUresult r = drv().MulticastAddDevice((CUmemGenericAllocationHandle)mc->handle, dev);
    if (r != CUDA_SUCCESS) {
        set_err(err, "cuMulticastAddDevice(" + std::to_string(dev) + "): " + cu_str(r));
        return false;
    }
    return true;
}

bool mc_bind(McObject *mc, size_t mc_off, const PoolMapping *pool, size_t mem_off, size_t bytes, std::string *err) {
    if (pool->mode != GPU_SHARE_VMM_FD) {
        set_err(err, "multicast needs a VMM pool");
        return false;
    }
    CUresult r = drv().MulticastBindMem((CUmemGenericAllocationHandle)mc->handle, mc_off,
                                        (CUmemGenericAllocationHandle)pool->vmm_handle, mem_off, bytes, 0);
    if (r != CUDA_SUCCESS) {
        set_err(err, "cuMulticastBindMem(off " + std::to_string(mem_off) + ", " + std::to_string(bytes) + " B): " + cu_str(r));
        return false;
    }
    return true;
}

bool mc_map(McObject *mc, int dev, std::string *err) {
    void *ptr = nullptr;
    if (!map_handle((CUmemGenericAllocationHandle)mc->handle, mc->bytes, dev, &ptr, err)) return false;
    mc->va = ptr;
    mc->va_dev = dev;
    return true;
}

void mc_release(McObject *mc) {
    if (!drv().ok) return;
    if (mc->va) {
        drv().MemUnmap((CUdeviceptr)mc->va, mc->bytes);
        drv().MemAddressFree((CUdeviceptr)mc->va, mc->bytes);
        mc->va = nullptr;
    }
    if (mc->handle) drv().MemRelease((CUmemGenericAllocationHandle)mc->handle);
    if (mc->fd >= 0) close(mc->fd);
    mc->handle = 0;
    mc->fd = -1;
}

namespace {
std::mutex g_reg_mu;
std::map<void *, size_t> g_registered;
} // namespace

bool host_register(void *p, size_t bytes, std::string *err) {
    if (device_count() <= 0) return true; // nothing to do on CPU-only hosts
    std::lock_guard<std::mutex> g(g_reg_mu);
    if (g_registered.count(p)) return true;
    cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
    if (e == cudaErrorHostMemoryAlreadyRegistered) {
        cudaGetLastError();
        e = cudaSuccess;
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_err(err, std::string("cudaHostRegister: ") + cudaGetErrorString(e));
        return false;
    }
    g_registered[p] = bytes;
    return true;
}

void host_unregister(void *p) {
    if (device_count() <= 0) return;
    std::lock_guard<std::mutex> g(g_reg_mu);
    auto it = g_registered.find(p);
    if (it == g_registered.end()) return;
    cudaHostUnregister(p);
    g_registered.erase(it);
}

bool can_access_peer(int dev, int peer) {
    if (dev == peer) return true;
    int ok = 0;
    if (cudaDeviceCanAccessPeer(&ok, dev, peer) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return ok != 0;
}

bool enable_peer_access(int dev, int peer) {
    if (dev == peer) return true;
    int cur = current_device();
    if (cur != dev) cudaSetDevice(dev);
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        e = cudaSuccess;
    }
    if (cur >= 0 && cur != dev) cudaSetDevice(cur);
    return e == cudaSuccess;
}

} // namespace gpu
} // namespace ocm
