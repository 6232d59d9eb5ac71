// hostvmm_probe - can the pinned host tier be a VMM allocation (cuMemCreate, location HOST_NUMA) shared by file
// descriptor, as the HBM pools are?  Parent creates and exports, child (forked before any CUDA call) imports, maps,
// lets a kernel write, reads with the CPU.  Prints how long each step takes, next to cudaHostRegister of shared memory.
//   hostvmm_probe [GiB]
#include <cuda.h>
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sys/mman.h>
#include <sys/socket.h>
#include <sys/wait.h>
#include <unistd.h>

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
#define CU(x)                                                              \
    do {                                                                   \
        CUresult r_ = (x);                                                 \
        if (r_ != CUDA_SUCCESS) {                                          \
            const char *s_ = nullptr;                                      \
            cuGetErrorString(r_, &s_);                                     \
            printf("{\"step\": \"%s\", \"error\": \"%s\"}\n", #x, s_ ? s_ : "?"); \
            fflush(stdout);                                                \
            _exit(3);                                                      \
        }                                                                  \
    } while (0)

__global__ void stamp(uint64_t *p, size_t n, uint64_t v) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v + i;
}

static int send_fd(int sock, int fd) {
    char b = 'x', ctl[CMSG_SPACE(sizeof(int))];
    iovec io = {&b, 1};
    msghdr m = {};
    m.msg_iov = &io; m.msg_iovlen = 1; m.msg_control = ctl; m.msg_controllen = sizeof ctl;
    cmsghdr *c = CMSG_FIRSTHDR(&m);
    c->cmsg_level = SOL_SOCKET; c->cmsg_type = SCM_RIGHTS; c->cmsg_len = CMSG_LEN(sizeof(int));
    memcpy(CMSG_DATA(c), &fd, sizeof fd);
    return sendmsg(sock, &m, 0) == 1 ? 0 : -1;
}
static int recv_fd(int sock) {
    char b, ctl[CMSG_SPACE(sizeof(int))];
    iovec io = {&b, 1};
    msghdr m = {};
    m.msg_iov = &io; m.msg_iovlen = 1; m.msg_control = ctl; m.msg_controllen = sizeof ctl;
    if (recvmsg(sock, &m, 0) != 1) return -1;
    int fd = -1;
    memcpy(&fd, CMSG_DATA(CMSG_FIRSTHDR(&m)), sizeof fd);
    return fd;
}

static CUmemAllocationProp host_prop() {
    CUmemAllocationProp p = {};
    p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    p.location.type = CU_MEM_LOCATION_TYPE_HOST_NUMA;
    p.location.id = 0;
    p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return p;
}

static void *map_it(CUmemGenericAllocationHandle h, size_t sz, size_t gran, int dev) {
    CUdeviceptr va = 0;
    CU(cuMemAddressReserve(&va, sz, gran, 0, 0));
    CU(cuMemMap(va, sz, 0, h, 0));
    CUmemAccessDesc d[2] = {};
    d[0].location.type = CU_MEM_LOCATION_TYPE_DEVICE; d[0].location.id = dev; d[0].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    d[1].location.type = CU_MEM_LOCATION_TYPE_HOST_NUMA; d[1].location.id = 0; d[1].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CU(cuMemSetAccess(va, sz, d, 2));
    return (void *)va;
}

int main(int argc, char **argv) {
    const size_t sz = (size_t)(argc > 1 ? atof(argv[1]) : 4.0) * (1ull << 30);
    int sp[2];
    if (socketpair(AF_UNIX, SOCK_STREAM, 0, sp)) return 1;
    pid_t child = fork();
    if (child == 0) {   // importer
        int fd = recv_fd(sp[1]);
        CU(cuInit(0));
        cudaSetDevice(0);
        cudaFree(0);
        double t0 = now();
        CUmemGenericAllocationHandle h;
        CU(cuMemImportFromShareableHandle(&h, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
        CUmemAllocationProp p = host_prop();
        size_t gran = 0;
        CU(cuMemGetAllocationGranularity(&gran, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
        uint64_t *va = (uint64_t *)map_it(h, sz, gran, 0);
        double t1 = now();
        stamp<<<296, 256>>>(va, sz / 8, 7);
        cudaError_t e = cudaDeviceSynchronize();
        double t2 = now();
        bool ok = e == cudaSuccess && va[0] == 7 && va[sz / 8 - 1] == 7 + sz / 8 - 1 && va[12345] == 7 + 12345;
        va[1] = 99;      // CPU store
        uint64_t *dbuf;
        cudaMalloc(&dbuf, 1 << 20);
        double t3 = now();
        cudaError_t e2 = cudaMemcpy(dbuf, va, 1 << 20, cudaMemcpyHostToDevice);
        cudaPointerAttributes at = {};
        cudaPointerGetAttributes(&at, va);
        printf("{\"step\": \"import+map in a second process\", \"GiB\": %.1f, \"seconds\": %.4f, \"kernel_write_GBps\": %.1f, "
               "\"cpu_reads_what_gpu_wrote\": %s, \"memcpy_from_it\": \"%s\", \"ptr_type\": %d, \"memcpy_s\": %.5f}\n",
               sz / 1073741824.0, t1 - t0, sz / (t2 - t1) / 1e9, ok ? "true" : "false", cudaGetErrorString(e2), (int)at.type, now() - t3);
        fflush(stdout);
        _exit(ok ? 0 : 4);
    }
    CU(cuInit(0));
    cudaSetDevice(0);
    cudaFree(0);
    CUmemAllocationProp p = host_prop();
    size_t gran = 0;
    CU(cuMemGetAllocationGranularity(&gran, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    double t0 = now();
    CUmemGenericAllocationHandle h;
    CU(cuMemCreate(&h, sz, &p, 0));
    double t1 = now();
    uint64_t *va = (uint64_t *)map_it(h, sz, gran, 0);
    double t2 = now();
    memset(va, 0, sz);
    double t3 = now();
    int fd = -1;
    CU(cuMemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    printf("{\"step\": \"create\", \"GiB\": %.1f, \"granularity\": %zu, \"create_s\": %.4f, \"map_s\": %.4f, \"cpu_memset_GBps\": %.1f}\n",
           sz / 1073741824.0, gran, t1 - t0, t2 - t1, sz / (t3 - t2) / 1e9);
    fflush(stdout);
    send_fd(sp[0], fd);
    int st = 0;
    waitpid(child, &st, 0);
    // the alternative in use today: shared memory, pinned by each process with cudaHostRegister
    void *shm = mmap(nullptr, sz, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    memset(shm, 0, sz);
    double r0 = now();
    cudaError_t e = cudaHostRegister(shm, sz, cudaHostRegisterPortable | cudaHostRegisterMapped);
    double r1 = now();
    printf("{\"step\": \"cudaHostRegister of touched shared memory\", \"GiB\": %.1f, \"seconds\": %.4f, \"result\": \"%s\"}\n",
           sz / 1073741824.0, r1 - r0, cudaGetErrorString(e));
    return WIFEXITED(st) ? WEXITSTATUS(st) : 5;
}
