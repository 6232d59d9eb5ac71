// peer_kernel - "remote memory is just a pointer": an application's OWN kernel reads and writes the far
// side of an OCM_REMOTE_GPU allocation (HBM of the neighbour rank's GPU, mapped into this process over
// NVLink) through ocm_remote_ptr(), and uses the runtime's kernels through include/ocm_kernels.h to check
// the result.  The reference has no way to do this: its remote memory is only reachable through
// ib_write/ib_read on a bounce buffer (reference src/lib.c:526-547), and OCM_REMOTE_GPU is rejected
// (reference src/lib.c:207-211).
//
//   build/bin/peer_kernel            (needs a running cluster with a GPU per rank)
#include <ocm_kernels.h>
#include <oncillamem.h>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__global__ void saxpy_into_far(float *far, const float *x, float a, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        far[i] = a * x[i] + far[i];   // read-modify-write of the peer GPU's HBM, straight from this kernel
}

int main() {
    const size_t n = 1 << 22;
    if (ocm_init() != 0) {
        fprintf(stderr, "no oncillamem daemon: %s\n", ocm_strerror(ocm_last_error()));
        return 1;
    }
    if (ocm_device() < 0) {
        fprintf(stderr, "this example needs a GPU\n");
        return 1;
    }
    cudaSetDevice(ocm_device());
    struct ocm_alloc_params ap = {};
    ap.kind = OCM_REMOTE_GPU;
    ap.local_alloc_bytes = n * sizeof(float);
    ap.rem_alloc_bytes = n * sizeof(float);
    ocm_alloc_t a = ocm_alloc(&ap);
    if (!a || ocm_alloc_kind(a) != OCM_REMOTE_GPU) {
        fprintf(stderr, "no remote-GPU allocation: %s\n", ocm_strerror(ocm_last_error()));
        return 1;
    }
    void *win = nullptr;
    size_t len = 0;
    ocm_localbuf(a, &win, &len);                       // window: HBM of our GPU
    float *far = static_cast<float *>(ocm_remote_ptr(a));   // far side: HBM of rank ocm_remote_rank(a)
    printf("window %p on GPU %d, far side %p on rank %d\n", win, ocm_device(), (void *)far, ocm_remote_rank(a));

    // x = 0x3f800000 pattern = 1.0f everywhere (runtime fill kernel), far side starts zeroed by the allocator
    ocmk_fill(win, n * sizeof(float), 0x3f8000003f800000ull, nullptr);
    cudaDeviceSynchronize();
    saxpy_into_far<<<592, 256>>>(far, static_cast<const float *>(win), 2.0f, n);   // far = 2 * 1 + 0
    saxpy_into_far<<<592, 256>>>(far, static_cast<const float *>(win), 0.5f, n);   // far = 0.5 * 1 + 2
    cudaDeviceSynchronize();

    // bring it back with a one-sided get and verify with the runtime's verify kernel: 2.5f = 0x40200000
    struct ocm_params cp = {};
    cp.bytes = n * sizeof(float);
    cp.op_flag = 0;
    if (ocm_copy_onesided(a, &cp) != 0) return 2;
    uint64_t *bad;
    cudaMallocManaged(&bad, 8);
    *bad = 0;
    ocmk_verify(win, n * sizeof(float), 0x4020000040200000ull, bad, nullptr);
    cudaDeviceSynchronize();
    const unsigned long long mism = *bad;
    cudaFree(bad);
    ocm_free(a);
    ocm_tini();
    printf(mism ? "FAIL: %llu mismatching words\n" : "pass: peer_kernel\n", mism);
    return mism ? 3 : 0;
}
