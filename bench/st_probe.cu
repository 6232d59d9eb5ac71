// st_probe - which SM store flavour moves the most bytes over NVLink?  (2 GPUs: source on GPU 0, destination on GPU 1,
// kernels on GPU 0.)  profiles/README.md section 2 found every SM-issued write stream saturating near 714 GB/s while the
// copy engines reach 780 and SM-issued reads 775-782; this sweeps the remaining knobs: cache operator, vector width,
// unroll depth, address layout per warp, grid size - plus the read direction and cudaMemcpyAsync as controls.
//   st_probe [bytes=1073741824]
#include "kernels/kernels.h"
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x)                                                                                      \
    do {                                                                                           \
        cudaError_t e_ = (x);                                                                      \
        if (e_ != cudaSuccess) {                                                                   \
            fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_));     \
            exit(1);                                                                               \
        }                                                                                          \
    } while (0)

struct V32 {
    uint32_t w[8];
};
__device__ __forceinline__ V32 ld32(const uint8_t *p) {
    V32 v;
    asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3]), "=r"(v.w[4]), "=r"(v.w[5]), "=r"(v.w[6]), "=r"(v.w[7])
                 : "l"(p));
    return v;
}
__device__ __forceinline__ V32 ld32_256(const uint8_t *p) {
    V32 v;
    asm volatile("ld.global.L1::no_allocate.L2::256B.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3]), "=r"(v.w[4]), "=r"(v.w[5]), "=r"(v.w[6]), "=r"(v.w[7])
                 : "l"(p));
    return v;
}
// get: LD 0 plain v8, 1 v8 with the L2::256B prefetch-size hint
template <int LD, int U>
__global__ void __launch_bounds__(256) get_kernel(const uint8_t *__restrict__ s, uint8_t *__restrict__ d, uint64_t bytes) {
    const uint64_t n = bytes / 32;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < n; i += U * stride) {
        V32 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = LD ? ld32_256(s + (i + u * stride) * 32) : ld32(s + (i + u * stride) * 32);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            uint8_t *p = d + (i + u * stride) * 32;
            asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v[u].w[0]), "r"(v[u].w[1]),
                         "r"(v[u].w[2]), "r"(v[u].w[3]), "r"(v[u].w[4]), "r"(v[u].w[5]), "r"(v[u].w[6]), "r"(v[u].w[7])
                         : "memory");
        }
    }
}

#define ST8(OP)                                                                                                         \
    asm volatile(OP " [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v.w[0]), "r"(v.w[1]), "r"(v.w[2]), "r"(v.w[3]),  \
                 "r"(v.w[4]), "r"(v.w[5]), "r"(v.w[6]), "r"(v.w[7])                                                     \
                 : "memory")
#define ST4(OP, q, a, b, c, d) asm volatile(OP " [%0], {%1,%2,%3,%4};" ::"l"(q), "r"(a), "r"(b), "r"(c), "r"(d) : "memory")

template <int POLICY> __device__ __forceinline__ void st32(uint8_t *p, const V32 &v, uint64_t pol) {
    if (POLICY == 0) ST8("st.global.L1::no_allocate.v8.u32");
    else if (POLICY == 1) ST8("st.global.v8.u32");
    else if (POLICY == 2) ST8("st.global.cs.v8.u32");
    else if (POLICY == 3) ST8("st.global.cg.v8.u32");
    else if (POLICY == 4) ST8("st.global.wt.v8.u32");
    else if (POLICY == 5) {
        ST4("st.global.L1::no_allocate.v4.u32", p, v.w[0], v.w[1], v.w[2], v.w[3]);
        ST4("st.global.L1::no_allocate.v4.u32", p + 16, v.w[4], v.w[5], v.w[6], v.w[7]);
    } else if (POLICY == 6) {
        asm volatile("st.global.L2::cache_hint.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8}, %9;" ::"l"(p), "r"(v.w[0]), "r"(v.w[1]),
                     "r"(v.w[2]), "r"(v.w[3]), "r"(v.w[4]), "r"(v.w[5]), "r"(v.w[6]), "r"(v.w[7]), "l"(pol)
                     : "memory");
    } else if (POLICY == 7) {
        ST4("st.relaxed.sys.global.v4.u32", p, v.w[0], v.w[1], v.w[2], v.w[3]);
        ST4("st.relaxed.sys.global.v4.u32", p + 16, v.w[4], v.w[5], v.w[6], v.w[7]);
    } else if (POLICY == 8) {
        ST4("st.volatile.global.v4.u32", p, v.w[0], v.w[1], v.w[2], v.w[3]);
        ST4("st.volatile.global.v4.u32", p + 16, v.w[4], v.w[5], v.w[6], v.w[7]);
    }
}

// LAYOUT 0: grid-stride, consecutive threads 32 B apart (a warp instruction covers 1 KiB), U independent chunks in flight
// LAYOUT 1: every warp owns a contiguous run of U KiB per iteration (its U instructions hit consecutive KiB)
template <int POLICY, int U, int LAYOUT>
__global__ void __launch_bounds__(256) put_kernel(const uint8_t *__restrict__ s, uint8_t *__restrict__ d, uint64_t bytes) {
    uint64_t pol = 0;
    if (POLICY == 6) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    const uint64_t n = bytes / 32;
    if (LAYOUT == 0) {
        const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
        uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + (U - 1) * stride < n; i += U * stride) {
            V32 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) v[u] = ld32(s + (i + u * stride) * 32);
#pragma unroll
            for (int u = 0; u < U; ++u) st32<POLICY>(d + (i + u * stride) * 32, v[u], pol);
        }
        for (; i < n; i += stride) {
            V32 v = ld32(s + i * 32);
            st32<POLICY>(d + i * 32, v, pol);
        }
    } else {
        const uint64_t warps = (uint64_t)gridDim.x * (blockDim.x / 32);
        const uint64_t w = (uint64_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
        const uint64_t lane = threadIdx.x & 31;
        for (uint64_t base = w * 32 * U; base + 32 * U <= n; base += warps * 32 * U) {
            V32 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) v[u] = ld32(s + (base + u * 32 + lane) * 32);
#pragma unroll
            for (int u = 0; u < U; ++u) st32<POLICY>(d + (base + u * 32 + lane) * 32, v[u], pol);
        }
    }
}

static cudaStream_t g_st;
template <typename F> static void time_it(const char *dir, const char *name, size_t n, F &&launch) {
    for (int i = 0; i < 2; ++i) launch();
    CK(cudaStreamSynchronize(g_st));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int reps = 6;
    CK(cudaEventRecord(e0, g_st));
    for (int i = 0; i < reps; ++i) launch();
    CK(cudaEventRecord(e1, g_st));
    CK(cudaStreamSynchronize(g_st));
    CK(cudaGetLastError());
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("{\"kind\": \"st_probe\", \"dir\": \"%s\", \"bytes\": %zu, \"variant\": \"%s\", \"us\": %.1f, \"GBps\": %.1f}\n", dir, n, name,
           ms * 1e3 / reps, n / (ms * 1e-3 / reps) / 1e9);
    fflush(stdout);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
}

int main(int argc, char **argv) {
    const size_t n = argc > 1 ? strtoull(argv[1], nullptr, 0) : (1ull << 30);
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (ndev < 2) {
        fprintf(stderr, "needs 2 GPUs\n");
        return 2;
    }
    uint8_t *a0, *b0, *r1;
    CK(cudaSetDevice(1));
    CK(cudaMalloc(&r1, n));
    CK(cudaMemset(r1, 2, n));
    CK(cudaDeviceEnablePeerAccess(0, 0));
    CK(cudaSetDevice(0));
    CK(cudaDeviceEnablePeerAccess(1, 0));
    CK(cudaMalloc(&a0, n));
    CK(cudaMalloc(&b0, n));
    CK(cudaMemset(a0, 1, n));
    CK(cudaStreamCreateWithFlags(&g_st, cudaStreamNonBlocking));

    time_it("put", "memcpy", n, [&] { CK(cudaMemcpyAsync(r1, a0, n, cudaMemcpyDefault, g_st)); });
    time_it("get", "memcpy", n, [&] { CK(cudaMemcpyAsync(b0, r1, n, cudaMemcpyDefault, g_st)); });
    time_it("get", "ld_v8_u2_592", n, [&] { put_kernel<0, 2, 0><<<592, 256, 0, g_st>>>(r1, b0, n); });
    time_it("get", "ld_v8_u4_592", n, [&] { put_kernel<0, 4, 0><<<592, 256, 0, g_st>>>(r1, b0, n); });
#define P(POL, U, L, CT, NAME) time_it("put", NAME, n, [&] { put_kernel<POL, U, L><<<CT, 256, 0, g_st>>>(a0, r1, n); })
    P(0, 2, 0, 592, "na_v8_u2_592");
    P(0, 4, 0, 592, "na_v8_u4_592");
    P(0, 8, 0, 296, "na_v8_u8_296");
    P(0, 1, 0, 1184, "na_v8_u1_1184");
    P(1, 2, 0, 592, "wb_v8_u2_592");
    P(2, 2, 0, 592, "cs_v8_u2_592");
    P(3, 2, 0, 592, "cg_v8_u2_592");
    P(4, 2, 0, 592, "wt_v8_u2_592");
    P(5, 2, 0, 592, "na_v4x2_u2_592");
    P(6, 2, 0, 592, "l2evictfirst_v8_u2_592");
    P(7, 2, 0, 592, "relaxed_sys_v4x2_u2_592");
    P(8, 2, 0, 592, "volatile_v4x2_u2_592");
    P(0, 4, 1, 592, "na_v8_warp4K_592");
    P(0, 8, 1, 296, "na_v8_warp8K_296");
    P(0, 4, 1, 148, "na_v8_warp4K_148");
    P(0, 2, 0, 36, "na_v8_u2_36");
    P(0, 2, 0, 72, "na_v8_u2_72");
    P(0, 2, 0, 144, "na_v8_u2_144");
    P(0, 4, 0, 18, "na_v8_u4_18");

    // ---- both directions at once (what a ring of puts/gets does to every link): GPU 0 and GPU 1 run the same thing ----
    uint8_t *a1 = r1, *b1;
    cudaStream_t st1;
    CK(cudaSetDevice(1));
    CK(cudaMalloc(&b1, n));
    CK(cudaStreamCreateWithFlags(&st1, cudaStreamNonBlocking));
    unsigned *tk0, *tk1;
    CK(cudaMalloc(&tk1, 4096));
    CK(cudaMemset(tk1, 0, 4096));
    CK(cudaSetDevice(0));
    CK(cudaMalloc(&tk0, 4096));
    CK(cudaMemset(tk0, 0, 4096));
    auto bidir = [&](const char *name, auto &&launch0, auto &&launch1) {
        cudaEvent_t e[4];
        for (int rep = 0; rep < 2; ++rep) {
            const int reps = rep ? 6 : 2;
            CK(cudaSetDevice(0));
            if (rep) { CK(cudaEventCreate(&e[0])); CK(cudaEventCreate(&e[1])); CK(cudaEventRecord(e[0], g_st)); }
            CK(cudaSetDevice(1));
            if (rep) { CK(cudaEventCreate(&e[2])); CK(cudaEventCreate(&e[3])); CK(cudaEventRecord(e[2], st1)); }
            for (int i = 0; i < reps; ++i) {
                CK(cudaSetDevice(0));
                launch0();
                CK(cudaSetDevice(1));
                launch1();
            }
            CK(cudaSetDevice(0));
            if (rep) CK(cudaEventRecord(e[1], g_st));
            CK(cudaSetDevice(1));
            if (rep) CK(cudaEventRecord(e[3], st1));
            CK(cudaStreamSynchronize(st1));
            CK(cudaSetDevice(0));
            CK(cudaStreamSynchronize(g_st));
        }
        float ms0 = 0, ms1 = 0;
        CK(cudaEventElapsedTime(&ms0, e[0], e[1]));
        CK(cudaEventElapsedTime(&ms1, e[2], e[3]));
        printf("{\"kind\": \"st_probe_bidir\", \"bytes\": %zu, \"variant\": \"%s\", \"gpu0_GBps\": %.1f, \"gpu1_GBps\": %.1f}\n", n, name,
               n / (ms0 * 1e-3 / 6) / 1e9, n / (ms1 * 1e-3 / 6) / 1e9);
        fflush(stdout);
    };
    bidir("ce_put", [&] { CK(cudaMemcpyAsync(a1, a0, n, cudaMemcpyDefault, g_st)); },
          [&] { CK(cudaMemcpyAsync(b0, b1, n, cudaMemcpyDefault, st1)); });
    bidir("ce_get", [&] { CK(cudaMemcpyAsync(b0, a1, n, cudaMemcpyDefault, g_st)); },
          [&] { CK(cudaMemcpyAsync(b1, a0, n, cudaMemcpyDefault, st1)); });
    bidir("put_v8_u2_592", [&] { put_kernel<0, 2, 0><<<592, 256, 0, g_st>>>(a0, a1, n); },
          [&] { put_kernel<0, 2, 0><<<592, 256, 0, st1>>>(b1, b0, n); });
    bidir("get_v8_u2_592", [&] { get_kernel<0, 2><<<592, 256, 0, g_st>>>(a1, b0, n); },
          [&] { get_kernel<0, 2><<<592, 256, 0, st1>>>(a0, b1, n); });
    bidir("get_v8_u4_592", [&] { get_kernel<0, 4><<<592, 256, 0, g_st>>>(a1, b0, n); },
          [&] { get_kernel<0, 4><<<592, 256, 0, st1>>>(a0, b1, n); });
    bidir("get_v8_u8_296", [&] { get_kernel<0, 8><<<296, 256, 0, g_st>>>(a1, b0, n); },
          [&] { get_kernel<0, 8><<<296, 256, 0, st1>>>(a0, b1, n); });
    bidir("get_v8_L2_256B_u2_592", [&] { get_kernel<1, 2><<<592, 256, 0, g_st>>>(a1, b0, n); },
          [&] { get_kernel<1, 2><<<592, 256, 0, st1>>>(a0, b1, n); });
    bidir("get_v8_L2_256B_u4_592", [&] { get_kernel<1, 4><<<592, 256, 0, g_st>>>(a1, b0, n); },
          [&] { get_kernel<1, 4><<<592, 256, 0, st1>>>(a0, b1, n); });
    for (int tile : {16384, 32768, 65536}) {
        ocmk::CopyTuning t;
        t.mode = ocmk::COPY_BULK;
        t.bulk_min = 0;
        t.ctas = 148;
        t.agents = tile == 65536 ? 1 : tile == 32768 ? 2 : 4;
        t.stages = 3;
        t.tile_bytes = tile;
        char nm[64];
        snprintf(nm, sizeof nm, "get_tma_bulk_%dK", tile >> 10);
        bidir(nm, [&] { CK(ocmk::launch_copy(b0, nullptr, a1, n, nullptr, tk0, g_st, &t)); },
              [&] { CK(ocmk::launch_copy(b1, nullptr, a0, n, nullptr, tk1, st1, &t)); });
        snprintf(nm, sizeof nm, "put_tma_bulk_%dK", tile >> 10);
        bidir(nm, [&] { CK(ocmk::launch_copy(a1, nullptr, a0, n, nullptr, tk0, g_st, &t)); },
              [&] { CK(ocmk::launch_copy(b0, nullptr, b1, n, nullptr, tk1, st1, &t)); });
    }

    // ---- mid-size messages, one direction: grid shape x unroll, pipelined launches (us per message) ----
    for (size_t m : {(size_t)(256 << 10), (size_t)(1 << 20), (size_t)(4 << 20), (size_t)(16 << 20)}) {
        auto mid = [&](const char *dir, const char *name, int ctas, auto &&launch) {
            for (int i = 0; i < 3; ++i) launch();
            CK(cudaStreamSynchronize(g_st));
            cudaEvent_t e0, e1;
            CK(cudaEventCreate(&e0));
            CK(cudaEventCreate(&e1));
            const int reps = 40;
            CK(cudaEventRecord(e0, g_st));
            for (int i = 0; i < reps; ++i) launch();
            CK(cudaEventRecord(e1, g_st));
            CK(cudaStreamSynchronize(g_st));
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            printf("{\"kind\": \"st_probe_mid\", \"dir\": \"%s\", \"bytes\": %zu, \"variant\": \"%s\", \"ctas\": %d, \"us\": %.2f, \"GBps\": %.1f}\n",
                   dir, m, name, ctas, ms * 1e3 / reps, m / (ms * 1e-3 / reps) / 1e9);
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
        };
        mid("put", "memcpy", 0, [&] { CK(cudaMemcpyAsync(a1, a0, m, cudaMemcpyDefault, g_st)); });
        mid("get", "memcpy", 0, [&] { CK(cudaMemcpyAsync(b0, a1, m, cudaMemcpyDefault, g_st)); });
        mid("put", "product", 0, [&] { CK(ocmk::launch_copy(a1, nullptr, a0, m, nullptr, tk0, g_st, nullptr)); });
        mid("get", "product", 0, [&] { CK(ocmk::launch_copy(b0, nullptr, a1, m, nullptr, tk0, g_st, nullptr)); });
        for (int ctas : {8, 16, 32, 64, 128, 148, 296, 592}) {
            if ((size_t)ctas * 256 * 32 > m) continue;
            mid("put", "u1", ctas, [&] { put_kernel<0, 1, 0><<<ctas, 256, 0, g_st>>>(a0, a1, m); });
            mid("put", "u2", ctas, [&] { put_kernel<0, 2, 0><<<ctas, 256, 0, g_st>>>(a0, a1, m); });
            mid("put", "u4", ctas, [&] { put_kernel<0, 4, 0><<<ctas, 256, 0, g_st>>>(a0, a1, m); });
            mid("put", "u8", ctas, [&] { put_kernel<0, 8, 0><<<ctas, 256, 0, g_st>>>(a0, a1, m); });
            mid("get", "u1", ctas, [&] { get_kernel<0, 1><<<ctas, 256, 0, g_st>>>(a1, b0, m); });
            mid("get", "u2", ctas, [&] { get_kernel<0, 2><<<ctas, 256, 0, g_st>>>(a1, b0, m); });
            mid("get", "u4", ctas, [&] { get_kernel<0, 4><<<ctas, 256, 0, g_st>>>(a1, b0, m); });
            mid("get", "u8", ctas, [&] { get_kernel<0, 8><<<ctas, 256, 0, g_st>>>(a1, b0, m); });
        }
    }
    return 0;
}
