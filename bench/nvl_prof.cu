// nvl_prof - one launch of every kernel that moves bytes over NVLink, for Nsight Compute (one process, 2 GPUs: kernels
// run on GPU 0, the peer memory lives on GPU 1).  Run it plainly first (it checks the results), then under
//   ncu --set full --metrics nvltx__bytes.sum,nvlrx__bytes.sum,nvltx__bytes_data_user.sum,nvlrx__bytes_data_user.sum ...
// Each kernel runs twice: the first launch warms it up, ncu's -k/-s/-c filters pick the second.
#include "kernels/kernels.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x)                                                                                  \
    do {                                                                                       \
        cudaError_t e_ = (x);                                                                  \
        if (e_ != cudaSuccess) {                                                               \
            fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); \
            exit(1);                                                                           \
        }                                                                                      \
    } while (0)

int main() {
    using namespace ocmk;
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (ndev < 2) {
        fprintf(stderr, "needs 2 GPUs\n");
        return 2;
    }
    const size_t n = 256ull << 20;
    uint8_t *loc, *loc2, *peer;
    float *scales;
    CK(cudaSetDevice(1));
    CK(cudaMalloc(&peer, n));
    CK(cudaMalloc(&scales, 8 << 20));
    CK(cudaMemset(peer, 0, n));
    CK(cudaDeviceEnablePeerAccess(0, 0));
    CK(cudaSetDevice(0));
    CK(cudaDeviceEnablePeerAccess(1, 0));
    CK(cudaMalloc(&loc, n));
    CK(cudaMalloc(&loc2, n));
    CK(cudaMemset(loc, 0x3c, n));
    unsigned *ticket;
    unsigned long long *bad;
    CK(cudaMalloc(&ticket, 4096));
    CK(cudaMemset(ticket, 0, 4096));
    CK(cudaMallocManaged(&bad, 8));
    cudaStream_t st;
    CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    auto check = [&](const char *what, const void *p, size_t bytes, uint64_t pattern) {
        *bad = 0;
        CK(launch_verify(p, bytes, pattern, bad, st));
        CK(cudaStreamSynchronize(st));
        printf("%-28s %s\n", what, *bad ? "MISMATCH" : "ok");
        if (*bad) exit(1);
    };
    CopyTuning bulk;
    bulk.mode = COPY_BULK;
    bulk.bulk_min = 0;
    for (int rep = 0; rep < 2; ++rep) {
        CK(launch_copy(peer, nullptr, loc, n, nullptr, ticket, st));             // put, 256-bit vector engine
        CK(launch_copy(loc2, nullptr, peer, n, nullptr, ticket, st));            // get
        CK(launch_copy(peer, nullptr, loc, n, nullptr, ticket, st, &bulk));      // put, TMA bulk engine
        CK(cudaStreamSynchronize(st));
    }
    check("put/get 256 MiB", loc2, n, 0x3c3c3c3c3c3c3c3cull);
    // strided pack to the peer and fp32 -> fp8 quantising put
    StridedDesc d;
    memset(&d, 0, sizeof d);
    d.src = loc;
    d.dst = peer;
    d.row_bytes = 4096;
    d.n[0] = 16384; d.n[1] = d.n[2] = 1;
    d.src_stride[0] = 8192;
    d.dst_stride[0] = 4096;
    d.src_dtype = d.dst_dtype = DT_U8;
    StridedDesc q = d;
    q.row_bytes = 4096 * 4;
    q.n[0] = 8192;
    q.src_stride[0] = 4096 * 4;
    q.dst_stride[0] = 4096;
    q.src_dtype = DT_F32;
    q.dst_dtype = DT_F8E4M3;
    q.scale_block = 32;
    q.scales = scales;
    for (int rep = 0; rep < 2; ++rep) {
        CK(launch_strided(d, nullptr, ticket, st));
        CK(launch_strided(q, nullptr, ticket, st));
        CK(cudaStreamSynchronize(st));
    }
    // GUPS over a table striped across both GPUs (half of the updates cross NVLink)
    GupsTable t;
    memset(&t, 0, sizeof t);
    CK(cudaMemsetAsync(loc2, 0, n, st));
    CK(cudaMemsetAsync(peer, 0, n, st));
    t.base[0] = reinterpret_cast<uint64_t *>(loc2);
    t.base[1] = reinterpret_cast<uint64_t *>(peer);
    t.n_ranks = 2;
    t.words_per_stripe = 1 << 16;
    t.total_words = 2 * n / 8;
    for (int rep = 0; rep < 2; ++rep) CK(launch_gups(t, 1ull << 26, 7, st));   // twice: the xor stream undoes itself
    CK(cudaStreamSynchronize(st));
    check("gups involution (local)", loc2, n, 0);
    check("gups involution (peer)", peer, n, 0);
    printf("nvl_prof done\n");
    return 0;
}
