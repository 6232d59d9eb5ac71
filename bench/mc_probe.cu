// mc_probe - does NVLS multicast (cuMulticast* + multimem.st) work on this box?
// Single process, N GPUs: create a multicast object over all devices, bind 2 MiB+ of each device's memory,
// map the multicast VA on device 0, store a pattern with multimem.st (kernel K6b) and check every replica.
// Prints one JSON line; exit code 0 even when multicast is unsupported (that is a result, not a failure).
#include "kernels/kernels.h"
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

template <typename F> static bool load(const char *name, F *fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult st;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &st) != cudaSuccess || !p) return false;
    *fn = reinterpret_cast<F>(p);
    return true;
}

#define DRV(call)                                                                 \
    do {                                                                          \
        CUresult r_ = (call);                                                     \
        if (r_ != CUDA_SUCCESS) {                                                 \
            printf("{\"kind\": \"mc_probe\", \"ok\": false, \"where\": \"%s\", \"curesult\": %d}\n", #call, (int)r_); \
            return 0;                                                             \
        }                                                                         \
    } while (0)

__global__ void check_kernel(const uint32_t *p, size_t n, uint32_t want, unsigned long long *bad) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    unsigned long long b = 0;
    for (; i < n; i += stride) b += p[i] != want;
    if (b) atomicAdd(bad, b);
}

int main() {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
        printf("{\"kind\": \"mc_probe\", \"ok\": false, \"where\": \"no device\"}\n");
        return 0;
    }
    for (int d = 0; d < ndev; ++d) {
        cudaSetDevice(d);
        cudaFree(0);
    }
    cudaSetDevice(0);
    decltype(&cuMulticastCreate) MulticastCreate;
    decltype(&cuMulticastAddDevice) MulticastAddDevice;
    decltype(&cuMulticastBindMem) MulticastBindMem;
    decltype(&cuMulticastGetGranularity) MulticastGetGranularity;
    decltype(&cuMemCreate) MemCreate;
    decltype(&cuMemAddressReserve) MemAddressReserve;
    decltype(&cuMemMap) MemMap;
    decltype(&cuMemSetAccess) MemSetAccess;
    decltype(&cuDeviceGetAttribute) DeviceGetAttribute;
    decltype(&cuMemExportToShareableHandle) MemExport;
    decltype(&cuMemImportFromShareableHandle) MemImport;
    if (!load("cuMulticastCreate", &MulticastCreate) || !load("cuMulticastAddDevice", &MulticastAddDevice) ||
        !load("cuMulticastBindMem", &MulticastBindMem) || !load("cuMulticastGetGranularity", &MulticastGetGranularity) ||
        !load("cuMemCreate", &MemCreate) || !load("cuMemAddressReserve", &MemAddressReserve) || !load("cuMemMap", &MemMap) ||
        !load("cuMemSetAccess", &MemSetAccess) || !load("cuDeviceGetAttribute", &DeviceGetAttribute) ||
        !load("cuMemExportToShareableHandle", &MemExport) || !load("cuMemImportFromShareableHandle", &MemImport)) {
        printf("{\"kind\": \"mc_probe\", \"ok\": false, \"where\": \"driver entry points\"}\n");
        return 0;
    }
    int sup = 0;
    DRV(DeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, 0));
    if (!sup) {
        printf("{\"kind\": \"mc_probe\", \"ok\": false, \"where\": \"CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED=0\", \"n_gpus\": %d}\n", ndev);
        return 0;
    }
    CUmulticastObjectProp mp = {};
    mp.numDevices = (unsigned)ndev;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    mp.size = 64ull << 20;
    size_t gran = 0, gran_min = 0;
    DRV(MulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    DRV(MulticastGetGranularity(&gran_min, &mp, CU_MULTICAST_GRANULARITY_MINIMUM));
    mp.size = (mp.size + gran - 1) / gran * gran;
    CUmemGenericAllocationHandle mc;
    DRV(MulticastCreate(&mc, &mp));
    // the fd round trip is what the multi-process runtime would do
    int fd = -1;
    DRV(MemExport(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    CUmemGenericAllocationHandle mc2;
    DRV(MemImport(&mc2, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    for (int d = 0; d < ndev; ++d) DRV(MulticastAddDevice(mc, d));
    std::vector<CUdeviceptr> uni(ndev);
    for (int d = 0; d < ndev; ++d) {
        CUmemAllocationProp ap = {};
        ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ap.location.id = d;
        ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        CUmemGenericAllocationHandle h;
        DRV(MemCreate(&h, mp.size, &ap, 0));
        DRV(MulticastBindMem(mc, 0, h, 0, mp.size, 0));
        DRV(MemAddressReserve(&uni[d], mp.size, gran, 0, 0));
        DRV(MemMap(uni[d], mp.size, 0, h, 0));
        CUmemAccessDesc ad = {};
        ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ad.location.id = 0;   // device 0 reads every replica back (peer access through the VMM mapping)
        ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        DRV(MemSetAccess(uni[d], mp.size, &ad, 1));
        if (d != 0) {
            ad.location.id = d;
            DRV(MemSetAccess(uni[d], mp.size, &ad, 1));
        }
    }
    CUdeviceptr mcva;
    DRV(MemAddressReserve(&mcva, mp.size, gran, 0, 0));
    DRV(MemMap(mcva, mp.size, 0, mc2, 0));
    CUmemAccessDesc ad = {};
    ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ad.location.id = 0;
    ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    DRV(MemSetAccess(mcva, mp.size, &ad, 1));

    uint8_t *src;
    unsigned long long *bad;
    unsigned *ticket;
    cudaMalloc(&src, mp.size);
    cudaMalloc(&bad, 8);
    cudaMalloc(&ticket, 4);
    cudaMemset(src, 0x5A, mp.size);
    cudaMemset(bad, 0, 8);
    cudaMemset(ticket, 0, 4);
    for (int d = 0; d < ndev; ++d) cudaMemset((void *)uni[d], 0, mp.size);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaError_t ce = ocmk::launch_put_multimem((void *)mcva, src, mp.size, nullptr, ticket, 0);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int i = 0; i < 10; ++i) ce = ocmk::launch_put_multimem((void *)mcva, src, mp.size, nullptr, ticket, 0);
    cudaEventRecord(e1);
    cudaError_t se = cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    for (int d = 0; d < ndev; ++d) check_kernel<<<256, 256>>>((const uint32_t *)uni[d], mp.size / 4, 0x5A5A5A5Au, bad);
    cudaDeviceSynchronize();
    unsigned long long hb = 0;
    cudaMemcpy(&hb, bad, 8, cudaMemcpyDeviceToHost);
    // gather-reduce through the switch: every replica holds 1.5f, the multicast load returns the sum over the GPUs
    for (int d = 0; d < ndev; ++d) ocmk::launch_fill((void *)uni[d], mp.size, 0x3FC000003FC00000ull, 0);
    cudaDeviceSynchronize();
    float *red;
    cudaMalloc(&red, mp.size);
    ocmk::launch_get_reduce_f32(red, (const void *)mcva, mp.size, 0);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int i = 0; i < 10; ++i) ocmk::launch_get_reduce_f32(red, (const void *)mcva, mp.size, 0);
    cudaEventRecord(e1);
    cudaError_t re = cudaDeviceSynchronize();
    float rms = 0;
    cudaEventElapsedTime(&rms, e0, e1);
    float probe[2] = {0, 0};
    cudaMemcpy(&probe[0], red, 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(&probe[1], red + mp.size / 4 - 1, 4, cudaMemcpyDeviceToHost);
    const bool red_ok = re == cudaSuccess && probe[0] == 1.5f * ndev && probe[1] == 1.5f * ndev;
    printf("{\"kind\": \"mc_probe_reduce\", \"ok\": %s, \"n_gpus\": %d, \"bytes\": %zu, \"us_per_reduce_get\": %.1f, "
           "\"result_GBps\": %.1f, \"gathered_GBps\": %.1f}\n",
           red_ok ? "true" : "false", ndev, (size_t)mp.size, rms * 100.0, mp.size / (rms / 10 * 1e-3) / 1e9,
           ndev * mp.size / (rms / 10 * 1e-3) / 1e9);
    printf("{\"kind\": \"mc_probe\", \"ok\": %s, \"n_gpus\": %d, \"bytes\": %zu, \"gran_rec\": %zu, \"gran_min\": %zu, "
           "\"launch\": \"%s\", \"sync\": \"%s\", \"bad_words\": %llu, \"us_per_put\": %.1f, \"src_GBps\": %.1f, "
           "\"delivered_GBps\": %.1f}\n",
           (ce == cudaSuccess && se == cudaSuccess && hb == 0) ? "true" : "false", ndev, (size_t)mp.size, gran, gran_min,
           cudaGetErrorString(ce), cudaGetErrorString(se), hb, ms * 100.0, mp.size / (ms / 10 * 1e-3) / 1e9,
           ndev * mp.size / (ms / 10 * 1e-3) / 1e9);
    return 0;
}
