// kbench - native micro-benchmark of the copy kernels (no Python in the loop).
//   kbench [peer|local] [quick]
// "peer": source on GPU 0, destination on GPU 1 (put) and the reverse (get), kernels always run on GPU 0.
// For every size it times R back-to-back launches with CUDA events and prints JSON lines, for
//   memcpy      cudaMemcpyAsync (the library baseline: copy engines)
//   bulk:<c>x<a>x<s>x<t>   our TMA kernel with an explicit tuning
//   vec16       our 16-byte vector kernel
//   vec32       experimental 32-byte (256-bit LDG/STG) vector kernel
#include "kernels/kernels.h"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>

#define CK(x)                                                                          \
    do {                                                                               \
        cudaError_t e_ = (x);                                                          \
        if (e_ != cudaSuccess) {                                                       \
            fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); \
            exit(1);                                                                   \
        }                                                                              \
    } while (0)

__global__ void __launch_bounds__(256) vec32_kernel(const uint8_t *s, uint8_t *d, uint64_t bytes) {
    const uint64_t n = bytes / 32;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + stride < n; i += 2 * stride) {
        uint32_t a[8], b[8];
        asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]), "=r"(a[4]), "=r"(a[5]), "=r"(a[6]), "=r"(a[7])
                     : "l"(s + i * 32));
        asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3]), "=r"(b[4]), "=r"(b[5]), "=r"(b[6]), "=r"(b[7])
                     : "l"(s + (i + stride) * 32));
        asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(d + i * 32), "r"(a[0]),
                     "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7])
                     : "memory");
        asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(d + (i + stride) * 32),
                     "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7])
                     : "memory");
    }
    for (; i < n; i += stride) {
        uint32_t a[8];
        asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]), "=r"(a[4]), "=r"(a[5]), "=r"(a[6]), "=r"(a[7])
                     : "l"(s + i * 32));
        asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(d + i * 32), "r"(a[0]),
                     "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7])
                     : "memory");
    }
}

// experimental dual-destination variants (window + far side inside one HBM): unroll depth and store policy
template <int U, int POLICY>   // POLICY 0: L1::no_allocate, 1: .cs (streaming), 2: plain
__global__ void __launch_bounds__(256) dual_kernel(const uint8_t *s, uint8_t *d, uint8_t *d2, uint64_t bytes) {
    const uint64_t n = bytes / 32;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    auto ld = [&](uint64_t k, uint32_t *a) {
        asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]), "=r"(a[4]), "=r"(a[5]), "=r"(a[6]), "=r"(a[7])
                     : "l"(s + k * 32));
    };
    auto st = [&](uint8_t *base, uint64_t k, const uint32_t *a) {
        if (POLICY == 0)
            asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(base + k * 32), "r"(a[0]),
                         "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]) : "memory");
        else if (POLICY == 1)
            asm volatile("st.global.cs.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(base + k * 32), "r"(a[0]), "r"(a[1]),
                         "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]) : "memory");
        else
            asm volatile("st.global.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(base + k * 32), "r"(a[0]), "r"(a[1]),
                         "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]) : "memory");
    };
    for (; i + (U - 1) * stride < n; i += U * stride) {
        uint32_t v[U][8];
#pragma unroll
        for (int u = 0; u < U; ++u) ld(i + u * stride, v[u]);
#pragma unroll
        for (int u = 0; u < U; ++u) st(d, i + u * stride, v[u]);
#pragma unroll
        for (int u = 0; u < U; ++u) st(d2, i + u * stride, v[u]);
    }
    for (; i < n; i += stride) {
        uint32_t v[8];
        ld(i, v);
        st(d, i, v);
        st(d2, i, v);
    }
}

struct Variant {
    std::string name;
    int mode;   // 0 memcpy, 1 bulk, 2 vec16, 3 vec32
    int ctas, agents, stages, tile;
};

static double run(const Variant &v, uint8_t *dst, const uint8_t *src, size_t n, int reps, cudaStream_t st, unsigned *ticket) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    auto once = [&]() {
        if (v.mode == 0) {
            CK(cudaMemcpyAsync(dst, src, n, cudaMemcpyDefault, st));
        } else if (v.mode == 4) {   // the product's default dispatch
            CK(ocmk::launch_copy(dst, nullptr, src, n, nullptr, ticket, st, nullptr));
        } else if (v.mode == 3) {
            int ctas = v.ctas ? v.ctas : 296;
            vec32_kernel<<<ctas, 256, 0, st>>>(src, dst, n);
        } else {
            ocmk::CopyTuning t;
            t.mode = v.mode == 1 ? ocmk::COPY_BULK : ocmk::COPY_VEC;
            t.bulk_min = 0;
            t.ctas = v.ctas;
            if (v.agents) t.agents = v.agents;
            if (v.stages) t.stages = v.stages;
            if (v.tile) t.tile_bytes = v.tile;
            CK(ocmk::launch_copy(dst, nullptr, src, n, nullptr, ticket, st, &t));
        }
    };
    for (int i = 0; i < 3; ++i) once();
    CK(cudaStreamSynchronize(st));
    CK(cudaEventRecord(e0, st));
    for (int i = 0; i < reps; ++i) once();
    CK(cudaEventRecord(e1, st));
    CK(cudaStreamSynchronize(st));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return ms * 1e-3 / reps;
}

int main(int argc, char **argv) {
    bool peer = argc > 1 && !strcmp(argv[1], "peer");
    bool quick = argc > 2 && !strcmp(argv[2], "quick");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (peer && ndev < 2) {
        fprintf(stderr, "peer mode needs 2 GPUs\n");
        return 2;
    }
    const size_t maxn = 1ull << 30;
    uint8_t *a0, *b0, *r1 = nullptr;
    CK(cudaSetDevice(0));
    CK(cudaMalloc(&a0, maxn));
    CK(cudaMalloc(&b0, maxn));
    CK(cudaMemset(a0, 1, maxn));
    if (peer) {
        CK(cudaDeviceEnablePeerAccess(1, 0));
        CK(cudaSetDevice(1));
        CK(cudaMalloc(&r1, maxn));
        CK(cudaMemset(r1, 2, maxn));
        CK(cudaDeviceEnablePeerAccess(0, 0));
        CK(cudaSetDevice(0));
    }
    unsigned *ticket;
    CK(cudaMalloc(&ticket, 4096));
    CK(cudaMemset(ticket, 0, 4096));
    cudaStream_t st;
    CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));

    {
        // library baseline for ocm_alloc: what the reference's GPU tier calls per allocation (ref src/lib.c:240,365)
        for (size_t bytes : {(size_t)(64 << 10), (size_t)(1 << 20), (size_t)(64 << 20), (size_t)(1 << 30)}) {
            std::vector<double> am, af;
            for (int i = 0; i < 200; ++i) {
                void *p = nullptr;
                struct timespec t0, t1, t2;
                clock_gettime(CLOCK_MONOTONIC, &t0);
                CK(cudaMalloc(&p, bytes));
                clock_gettime(CLOCK_MONOTONIC, &t1);
                CK(cudaFree(p));
                clock_gettime(CLOCK_MONOTONIC, &t2);
                am.push_back((t1.tv_sec - t0.tv_sec) * 1e6 + (t1.tv_nsec - t0.tv_nsec) * 1e-3);
                af.push_back((t2.tv_sec - t1.tv_sec) * 1e6 + (t2.tv_nsec - t1.tv_nsec) * 1e-3);
            }
            std::sort(am.begin(), am.end());
            std::sort(af.begin(), af.end());
            printf("{\"kind\": \"cudamalloc_baseline\", \"bytes\": %zu, \"malloc_p50_us\": %.2f, \"free_p50_us\": %.2f}\n", bytes,
                   am[am.size() / 2], af[af.size() / 2]);
        }
    }
    if (argc > 1 && !strcmp(argv[1], "dual")) {
        // window + far side inside one HBM (bench.py at N=1): product kernel vs experimental variants
        uint8_t *c0;
        CK(cudaMalloc(&c0, maxn));
        const size_t n = 256ull << 20;
        auto time_it = [&](const char *name, auto &&launch) {
            for (int i = 0; i < 3; ++i) launch();
            CK(cudaStreamSynchronize(st));
            cudaEvent_t e0, e1;
            CK(cudaEventCreate(&e0));
            CK(cudaEventCreate(&e1));
            CK(cudaEventRecord(e0, st));
            for (int i = 0; i < 20; ++i) launch();
            CK(cudaEventRecord(e1, st));
            CK(cudaStreamSynchronize(st));
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            printf("{\"kind\": \"dual\", \"bytes\": %zu, \"variant\": \"%s\", \"us\": %.1f, \"GBps\": %.1f, \"hbm_traffic_GBps\": %.1f}\n", n,
                   name, ms * 1e3 / 20, n / (ms * 1e-3 / 20) / 1e9, 3.0 * n / (ms * 1e-3 / 20) / 1e9);
            fflush(stdout);
        };
        time_it("product", [&] { CK(ocmk::launch_copy(b0, c0, a0, n, nullptr, ticket, st, nullptr)); });
        time_it("two_memcpy", [&] {
            CK(cudaMemcpyAsync(b0, a0, n, cudaMemcpyDefault, st));
            CK(cudaMemcpyAsync(c0, a0, n, cudaMemcpyDefault, st));
        });
        for (int ctas : {296, 592, 1184, 2368}) {
            char nm[64];
            snprintf(nm, sizeof nm, "u2_na_%d", ctas);
            time_it(nm, [&] { dual_kernel<2, 0><<<ctas, 256, 0, st>>>(a0, b0, c0, n); });
            snprintf(nm, sizeof nm, "u4_na_%d", ctas);
            time_it(nm, [&] { dual_kernel<4, 0><<<ctas, 256, 0, st>>>(a0, b0, c0, n); });
            snprintf(nm, sizeof nm, "u2_cs_%d", ctas);
            time_it(nm, [&] { dual_kernel<2, 1><<<ctas, 256, 0, st>>>(a0, b0, c0, n); });
            snprintf(nm, sizeof nm, "u1_plain_%d", ctas);
            time_it(nm, [&] { dual_kernel<1, 2><<<ctas, 256, 0, st>>>(a0, b0, c0, n); });
        }
        return 0;
    }
    std::vector<Variant> vars = {
        {"memcpy", 0, 0, 0, 0, 0},
        {"auto", 4, 0, 0, 0, 0},
        {"bulk:148x4x3x16K", 1, 148, 4, 3, 16384},
        {"bulk:148x2x3x32K", 1, 148, 2, 3, 32768},
        {"bulk:148x1x3x64K", 1, 148, 1, 3, 65536},
        {"bulk:64x1x4x32K", 1, 64, 1, 4, 32768},
        {"bulk:32x2x3x32K", 1, 32, 2, 3, 32768},
        {"bulk:148x8x3x8K", 1, 148, 8, 3, 8192},
        {"bulk:148x4x6x8K", 1, 148, 4, 6, 8192},
        {"bulk:296x2x3x16K", 1, 296, 2, 3, 16384},
        {"vec16:296", 2, 296, 0, 0, 0},
        {"vec16:592", 2, 592, 0, 0, 0},
        {"vec32:296", 3, 296, 0, 0, 0},
        {"vec32:592", 3, 592, 0, 0, 0},
        {"vec32:1184", 3, 1184, 0, 0, 0},
    };
    if (quick) vars.resize(4);
    std::vector<size_t> sizes = {4096, 65536, 262144, 1 << 20, 4 << 20, 16 << 20, 64 << 20, 256 << 20, 1 << 30};
    struct Dir { const char *name; uint8_t *dst; const uint8_t *src; };
    std::vector<Dir> dirs;
    if (peer) {
        dirs.push_back({"put", r1, a0});
        dirs.push_back({"get", b0, r1});
    } else {
        dirs.push_back({"local", b0, a0});
    }
    // synchronous small-message latency (host clock around enqueue + stream sync): the shape of the
    // reference's own BW test, which calls a blocking ocm_copy_onesided per message
    for (auto &d : dirs)
        for (size_t n : {(size_t)64, (size_t)4096, (size_t)65536, (size_t)(1 << 20)})
            for (int which = 0; which < 2; ++which) {
                std::vector<double> us;
                for (int i = 0; i < 220; ++i) {
                    struct timespec t0, t1;
                    clock_gettime(CLOCK_MONOTONIC, &t0);
                    if (which == 0) CK(cudaMemcpyAsync(d.dst, d.src, n, cudaMemcpyDefault, st));
                    else CK(ocmk::launch_copy(d.dst, nullptr, d.src, n, nullptr, ticket, st, nullptr));
                    CK(cudaStreamSynchronize(st));
                    clock_gettime(CLOCK_MONOTONIC, &t1);
                    if (i >= 20) us.push_back((t1.tv_sec - t0.tv_sec) * 1e6 + (t1.tv_nsec - t0.tv_nsec) * 1e-3);
                }
                std::sort(us.begin(), us.end());
                printf("{\"kind\": \"sync_latency\", \"dir\": \"%s\", \"bytes\": %zu, \"engine\": \"%s\", \"p50_us\": %.2f, \"p99_us\": %.2f}\n",
                       d.name, n, which ? "ours" : "memcpy", us[us.size() / 2], us[us.size() * 99 / 100]);
            }
    for (auto &d : dirs)
        for (size_t n : sizes)
            for (auto &v : vars) {
                if (v.mode == 3 && n < 65536) continue;
                int reps = (int)std::max<size_t>(4, std::min<size_t>(200, (8ull << 30) / n / 4));
                double s = run(v, d.dst, d.src, n, reps, st, ticket);
                printf("{\"kind\": \"kbench\", \"dir\": \"%s\", \"bytes\": %zu, \"variant\": \"%s\", \"us\": %.2f, \"GBps\": %.1f}\n",
                       d.name, n, v.name.c_str(), s * 1e6, n / s / 1e9);
                fflush(stdout);
            }
    return 0;
}
