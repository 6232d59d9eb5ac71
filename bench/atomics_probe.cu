// atomics_probe - are system-scope reductions from SEVERAL GPUs to one word atomic with respect to each other?
//   host word : pinned, portable host memory (where the completion signals of the allocation tables live)
//   hbm word  : device memory of GPU 0, reached by the peers over NVLink
// Every GPU (and, in a third experiment, a CPU thread as well) adds 1 to the same word N times concurrently;
// lost updates show up as a final count below the expected one.  One JSON line per experiment.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>
#include <cuda_runtime.h>

#define CK(x)                                                                                  \
    do {                                                                                       \
        cudaError_t e_ = (x);                                                                  \
        if (e_ != cudaSuccess) {                                                               \
            fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); \
            exit(1);                                                                           \
        }                                                                                      \
    } while (0)

__global__ void hammer(unsigned long long *w, int iters) {
    for (int i = 0; i < iters; ++i) asm volatile("red.release.sys.global.add.u64 [%0], 1;" ::"l"(w) : "memory");
}

static unsigned long long run(unsigned long long *w, int ndev, int iters, bool with_cpu, unsigned long long *cpu_adds) {
    const int blocks = 8, threads = 128;
    for (int d = 0; d < ndev; ++d) {
        CK(cudaSetDevice(d));
        hammer<<<blocks, threads>>>(w, iters);
    }
    unsigned long long n_cpu = 0;
    if (with_cpu) {
        auto *aw = reinterpret_cast<std::atomic<unsigned long long> *>(w);
        for (int i = 0; i < 200000; ++i) {
            aw->fetch_add(1, std::memory_order_relaxed);
            ++n_cpu;
        }
    }
    for (int d = 0; d < ndev; ++d) {
        CK(cudaSetDevice(d));
        CK(cudaDeviceSynchronize());
    }
    if (cpu_adds) *cpu_adds = n_cpu;
    return (unsigned long long)ndev * blocks * threads * iters + n_cpu;
}

int main(int argc, char **argv) {
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (argc > 1) ndev = std::min(ndev, atoi(argv[1]));
    const int iters = 2000;
    std::vector<int> native(ndev);
    for (int d = 0; d < ndev; ++d) CK(cudaDeviceGetAttribute(&native[d], cudaDevAttrHostNativeAtomicSupported, d));
    unsigned long long *hw;
    CK(cudaHostAlloc(&hw, 4096, cudaHostAllocPortable | cudaHostAllocMapped));
    for (int use = 1; use <= ndev; use = use == ndev ? ndev + 1 : std::min(use * 2, ndev)) {
        *hw = 0;
        const unsigned long long want = run(hw, use, iters, false, nullptr);
        printf("{\"word\": \"pinned host\", \"gpus\": %d, \"cpu_thread\": false, \"expected\": %llu, \"got\": %llu, \"lost\": %lld, "
               "\"hostNativeAtomicSupported\": %d}\n", use, want, *hw, (long long)want - (long long)*hw, native[0]);
    }
    {
        *hw = 0;
        unsigned long long n_cpu = 0;
        const unsigned long long want = run(hw, ndev, iters, true, &n_cpu);
        printf("{\"word\": \"pinned host\", \"gpus\": %d, \"cpu_thread\": true, \"expected\": %llu, \"got\": %llu, \"lost\": %lld}\n",
               ndev, want, *hw, (long long)want - (long long)*hw);
    }
    // a word in GPU 0's HBM, peers reach it over NVLink
    CK(cudaSetDevice(0));
    unsigned long long *dw;
    CK(cudaMalloc(&dw, 4096));
    CK(cudaMemset(dw, 0, 4096));
    for (int d = 1; d < ndev; ++d) {
        CK(cudaSetDevice(d));
        cudaError_t e = cudaDeviceEnablePeerAccess(0, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CK(e);
        cudaGetLastError();
    }
    const unsigned long long want = run(dw, ndev, iters, false, nullptr);
    unsigned long long got = 0;
    CK(cudaSetDevice(0));
    CK(cudaMemcpy(&got, dw, 8, cudaMemcpyDeviceToHost));
    printf("{\"word\": \"HBM of GPU 0\", \"gpus\": %d, \"cpu_thread\": false, \"expected\": %llu, \"got\": %llu, \"lost\": %lld}\n", ndev,
           want, got, (long long)want - (long long)got);
    return 0;
}
