// sbench - device-timed strided pack / cast kernels (K4/K5), local and peer destination.
//   [OCM_STRIDED_MODE=tma|wide|bulk|vec] sbench [peer]
// Every case rotates over 4 buffer sets so that no repetition finds its data in the 126 MB L2.
// Roofline per case: bytes read + bytes written over the measured HBM copy bandwidth (6544 GB/s, MEASURED_PEAKS.json)
// for a local destination; the slower of that and bytes written / 900 GB/s (NVLink 5 per direction) for a peer one.
#include "kernels/kernels.h"
#include <algorithm>
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

static cudaStream_t g_st;
static const double kHbm = 6544.0, kNvl = 900.0;

struct Case {
    const char *name;
    uint64_t row_bytes, n[3];
    int64_t sstride[3], dstride[3];
    int sdt, ddt, blk;
};

int main(int argc, char **argv) {
    const bool peer = argc > 1 && !strcmp(argv[1], "peer");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (peer && ndev < 2) return 2;
    const char *mode = getenv("OCM_STRIDED_MODE") ? getenv("OCM_STRIDED_MODE") : "auto";
    constexpr int kSets = 4;
    const size_t src_span = 272ull << 20, dst_span = 272ull << 20;
    uint8_t *src[kSets], *dst[kSets];
    float *scales[kSets];
    CK(cudaSetDevice(0));
    if (peer) CK(cudaDeviceEnablePeerAccess(1, 0));
    for (int i = 0; i < kSets; ++i) {
        CK(cudaMalloc(&src[i], src_span));
        CK(cudaMemset(src[i], 0x3c, src_span));
        if (peer) CK(cudaSetDevice(1));
        CK(cudaMalloc(&dst[i], dst_span));
        CK(cudaMalloc(&scales[i], 8 << 20));
        CK(cudaSetDevice(0));
    }
    unsigned *ticket;
    CK(cudaMalloc(&ticket, 4096));
    CK(cudaMemset(ticket, 0, 4096));
    CK(cudaStreamCreateWithFlags(&g_st, cudaStreamNonBlocking));
    using namespace ocmk;
    std::vector<Case> cases = {
        {"pack 16384 x 4KiB pitch 8KiB", 4096, {16384, 1, 1}, {8192, 0, 0}, {4096, 0, 0}, DT_U8, DT_U8, 0},
        {"pack 65536 x 4KiB pitch 4KiB+64", 4096, {65536, 1, 1}, {4160, 0, 0}, {4096, 0, 0}, DT_U8, DT_U8, 0},
        {"pack 262144 x 256B pitch 512B", 256, {262144, 1, 1}, {512, 0, 0}, {256, 0, 0}, DT_U8, DT_U8, 0},
        {"pack 1048576 x 64B pitch 128B", 64, {1048576, 1, 1}, {128, 0, 0}, {64, 0, 0}, DT_U8, DT_U8, 0},
        {"pack 1024 x 64KiB pitch 128KiB", 65536, {1024, 1, 1}, {131072, 0, 0}, {65536, 0, 0}, DT_U8, DT_U8, 0},
        {"unpack 16384 x 4KiB -> pitch 8KiB", 4096, {16384, 1, 1}, {4096, 0, 0}, {8192, 0, 0}, DT_U8, DT_U8, 0},
        {"box 3-D 64 x 64 x 16 rows of 1KiB", 1024, {64, 64, 16}, {2048, 160 << 10, 12 << 20}, {1024, 64 << 10, 4 << 20}, DT_U8, DT_U8, 0},
        {"cast f32->fp8 blk32 8192 x 4096", 4096 * 4, {8192, 1, 1}, {4096 * 4, 0, 0}, {4096, 0, 0}, DT_F32, DT_F8E4M3, 32},
        {"cast f32->fp8 blk128 8192 x 4096", 4096 * 4, {8192, 1, 1}, {4096 * 4, 0, 0}, {4096, 0, 0}, DT_F32, DT_F8E4M3, 128},
        {"cast f32->bf16 8192 x 4096", 4096 * 4, {8192, 1, 1}, {4096 * 4, 0, 0}, {4096 * 2, 0, 0}, DT_F32, DT_BF16, 0},
        {"cast f32->fp8 blk32 strided 8192 x 4096 pitch 8192 el", 4096 * 4, {8192, 1, 1}, {8192 * 4, 0, 0}, {4096, 0, 0}, DT_F32, DT_F8E4M3, 32},
        {"cast fp8->f32 blk32 8192 x 4096", 4096, {8192, 1, 1}, {4096, 0, 0}, {4096 * 4, 0, 0}, DT_F8E4M3, DT_F32, 32},
    };
    for (auto &c : cases) {
        auto desc = [&](int set) {
            StridedDesc d;
            d.src = src[set];
            d.dst = dst[set];
            d.row_bytes = c.row_bytes;
            for (int i = 0; i < 3; ++i) {
                d.n[i] = c.n[i];
                d.src_stride[i] = c.sstride[i];
                d.dst_stride[i] = c.dstride[i];
            }
            d.src_dtype = c.sdt;
            d.dst_dtype = c.ddt;
            d.scale_block = c.blk;
            d.scales = scales[set];
            return d;
        };
        for (int i = 0; i < kSets; ++i) CK(launch_strided(desc(i), nullptr, ticket, g_st));
        CK(cudaStreamSynchronize(g_st));
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0));
        CK(cudaEventCreate(&e1));
        const int reps = 12;
        CK(cudaEventRecord(e0, g_st));
        for (int i = 0; i < reps; ++i) CK(launch_strided(desc(i % kSets), nullptr, ticket, g_st));
        CK(cudaEventRecord(e1, g_st));
        CK(cudaStreamSynchronize(g_st));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        const double us = ms * 1e3 / reps;
        const uint64_t rows = c.n[0] * c.n[1] * c.n[2];
        const uint64_t elems = c.row_bytes / dtype_size(c.sdt);
        const double rd = (double)rows * c.row_bytes + (c.sdt == DT_F8E4M3 ? (double)rows * elems / c.blk * 4 : 0);
        const double wr = (double)rows * elems * dtype_size(c.ddt) + (c.ddt == DT_F8E4M3 ? (double)rows * elems / c.blk * 4 : 0);
        const double roof_us = std::max((rd + wr) / (kHbm * 1e3), peer ? wr / (kNvl * 1e3) : 0.0);
        printf("{\"kind\": \"sbench\", \"mode\": \"%s\", \"dst\": \"%s\", \"case\": \"%s\", \"us\": %.1f, \"read_MB\": %.1f, \"write_MB\": %.1f, "
               "\"GBps_read\": %.0f, \"roofline_us\": %.1f, \"frac_of_roofline\": %.2f}\n",
               mode, peer ? "peer" : "local", c.name, us, rd / 1e6, wr / 1e6, rd / us / 1e3, roof_us, roof_us / us);
        fflush(stdout);
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
    }
    return 0;
}
