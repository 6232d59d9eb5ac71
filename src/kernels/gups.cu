// K7 - GUPS-style random read-modify-write over a table striped across the
// HBM pools of every rank (BASELINE.json config 3: pooled-HBM random access).
// The reference only plans allocations that span several nodes (reference
// src/lib.c:42-45, src/mem.c:252); here the striped allocation exists and the
// update loop issues 8-byte xor reductions straight to the owning GPU over
// NVLink (RED.XOR on peer addresses), no staging and no per-peer connection.
// CPU restatement: ocm_oracle_gups (src/oracle/oracle.cpp).
#include "kernels.h"
#include "ptx.cuh"

namespace ocmk {

// Grid: four CTAs of 256 threads per SM.  The update stream is a function of (thread, step) with the thread count
// as a parameter, so the CPU oracle (ocm_oracle_gups) is told the same count (gups_threads()).
constexpr int kGupsCtasPerSm = 4;
uint64_t gups_threads() { return (uint64_t)num_sms() * kGupsCtasPerSm * 256; }

__device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// word index in the logical table -> address in the owning rank's slab
// (shifts and 32-bit division: a 64-bit divide per update would cost more than the update itself)
__device__ __forceinline__ uint64_t *gups_addr(const GupsTable &t, uint64_t idx, unsigned shift) {
    const uint32_t stripe = (uint32_t)(idx >> shift);
    const uint64_t within = idx & (t.words_per_stripe - 1);
    const uint32_t rank = stripe % (uint32_t)t.n_ranks;
    const uint32_t local_stripe = stripe / (uint32_t)t.n_ranks;
    return t.base[rank] + ((uint64_t)local_stripe << shift) + within;
}

__global__ void __launch_bounds__(256) ocm_gups_kernel(GupsTable t, uint64_t n_updates, uint64_t seed, unsigned shift) {
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t T = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t mask = t.total_words - 1;
    uint64_t step = 0;
    // four updates per iteration: all four values and addresses are computed before the first reduction is issued,
    // so the hashing of update k+1..k+3 overlaps the issue of update k (a RED returns nothing: nothing waits on it)
    uint64_t u = tid;
    for (; u + 3 * T < n_updates; u += 4 * T, step += 4) {
        uint64_t v[4], *p[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            v[k] = mix64(seed + tid * 0x9E3779B97F4A7C15ull + (step + k) * 0xD1B54A32D192ED03ull);
            p[k] = gups_addr(t, v[k] & mask, shift);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) asm volatile("red.global.xor.b64 [%0], %1;" ::"l"(p[k]), "l"(v[k]) : "memory");
    }
    for (; u < n_updates; u += T, ++step) {
        uint64_t v = mix64(seed + tid * 0x9E3779B97F4A7C15ull + step * 0xD1B54A32D192ED03ull);
        uint64_t *p = gups_addr(t, v & mask, shift);
        asm volatile("red.global.xor.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
    }
}

__global__ void __launch_bounds__(256) ocm_gups_init_kernel(GupsTable t, int my_rank) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t stripes = t.total_words / t.words_per_stripe;
    const uint64_t my_stripes = (stripes + t.n_ranks - 1 - my_rank) / t.n_ranks;
    const uint64_t my_words = my_stripes * t.words_per_stripe;
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < my_words; j += stride) {
        uint64_t local_stripe = j / t.words_per_stripe, within = j & (t.words_per_stripe - 1);
        uint64_t idx = (local_stripe * t.n_ranks + my_rank) * t.words_per_stripe + within;
        t.base[my_rank][j] = idx;
    }
}

cudaError_t launch_gups(const GupsTable &t, uint64_t n_updates, uint64_t seed, cudaStream_t stream) {
    if (t.n_ranks < 1 || t.n_ranks > 64 || (t.total_words & (t.total_words - 1)) ||
        (t.words_per_stripe & (t.words_per_stripe - 1)) || t.words_per_stripe > t.total_words)
        return cudaErrorInvalidValue;
    if (t.total_words / t.words_per_stripe > 0xFFFFFFFFull) return cudaErrorInvalidValue;
    const unsigned shift = (unsigned)__builtin_ctzll(t.words_per_stripe);
    ocm_gups_kernel<<<(unsigned)(gups_threads() / 256), 256, 0, stream>>>(t, n_updates, seed, shift);
    return cudaGetLastError();
}

cudaError_t launch_gups_init(const GupsTable &t, int my_rank, cudaStream_t stream) {
    if (my_rank < 0 || my_rank >= t.n_ranks) return cudaErrorInvalidValue;
    ocm_gups_init_kernel<<<num_sms() * 4, 256, 0, stream>>>(t, my_rank);
    return cudaGetLastError();
}

} // namespace ocmk
