// K9 - pattern fill, verify and checksum.
// Zero-on-release of pool blocks (the reference zero-fills remote buffers with
// calloc/memset on the CPU, reference src/alloc.c:171, src/extoll_client.c:46)
// and the 0xdeadbeef-style pattern checks of its transport tests
// (test/ib_client.c:144-188) as bandwidth-bound device kernels that work on
// local or peer addresses.
#include "kernels.h"
#include "ptx.cuh"
#include <algorithm>

namespace ocmk {

__global__ void __launch_bounds__(256) ocm_fill_kernel(uint8_t *dst, uint64_t bytes, uint64_t pattern) {
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    // head up to 16-byte alignment, pattern phase follows the absolute offset
    uint64_t head = (16 - ((uintptr_t)dst & 15)) & 15;
    if (head > bytes) head = bytes;
    if (tid < head) dst[tid] = (uint8_t)(pattern >> (8 * (tid & 7)));
    const uint64_t body = (bytes - head) & ~15ull;
    // rotate so that byte k of the buffer holds pattern byte (k & 7)
    const unsigned rot = (unsigned)((head & 7) * 8);
    const uint64_t p = rot ? ((pattern >> rot) | (pattern << (64 - rot))) : pattern;
    uint4 v = make_uint4((uint32_t)p, (uint32_t)(p >> 32), (uint32_t)p, (uint32_t)(p >> 32));
    uint4 *d = reinterpret_cast<uint4 *>(dst + head);
    for (uint64_t i = tid; i < body / 16; i += stride) d[i] = v;
    const uint64_t tail0 = head + body;
    if (tid < bytes - tail0) dst[tail0 + tid] = (uint8_t)(pattern >> (8 * ((tail0 + tid) & 7)));
}

__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// mode 0: count bytes that differ from the pattern; mode 1: wrapping sum of
// little-endian 8-byte words (the tail is zero-extended).
template <int MODE>
__global__ void __launch_bounds__(256) ocm_scan_kernel(const uint8_t *p, uint64_t bytes, uint64_t pattern,
                                                       unsigned long long *out) {
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    unsigned long long acc = 0;
    const bool aligned = ((uintptr_t)p & 7) == 0;
    const uint64_t nwords = aligned ? bytes / 8 : 0;
    const uint64_t *w = reinterpret_cast<const uint64_t *>(p);
    for (uint64_t i = tid; i < nwords; i += stride) {
        uint64_t v = w[i];
        if (MODE == 0) {
            uint64_t x = v ^ pattern;
            if (x) {
#pragma unroll
                for (int b = 0; b < 8; ++b) acc += ((x >> (8 * b)) & 0xff) != 0;
            }
        } else {
            acc += v;
        }
    }
    // unaligned buffers and the tail go byte by byte
    for (uint64_t i = nwords * 8 + tid; i < bytes; i += stride) {
        uint8_t v = p[i];
        if (MODE == 0) acc += v != (uint8_t)(pattern >> (8 * (i & 7)));
        else acc += (unsigned long long)v << (8 * (i & 7));
    }
    acc = warp_sum(acc);
    __shared__ unsigned long long part[8];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long s = 0;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += part[i];
        if (s) atomicAdd(out, s);
    }
}

static int grid_for(uint64_t bytes) {
    uint64_t want = (bytes / 16 + 255) / 256;
    uint64_t cap = (uint64_t)num_sms() * 8;
    return (int)std::max<uint64_t>(1, std::min(want, cap));
}

cudaError_t launch_fill(void *dst, size_t bytes, uint64_t pattern, cudaStream_t stream) {
    if (!bytes) return cudaSuccess;
    ocm_fill_kernel<<<grid_for(bytes), 256, 0, stream>>>(static_cast<uint8_t *>(dst), bytes, pattern);
    return cudaGetLastError();
}

cudaError_t launch_verify(const void *p, size_t bytes, uint64_t pattern, unsigned long long *mismatches,
                          cudaStream_t stream) {
    if (!bytes) return cudaSuccess;
    ocm_scan_kernel<0><<<grid_for(bytes), 256, 0, stream>>>(static_cast<const uint8_t *>(p), bytes, pattern, mismatches);
    return cudaGetLastError();
}

cudaError_t launch_checksum(const void *p, size_t bytes, unsigned long long *sum, cudaStream_t stream) {
    if (!bytes) return cudaSuccess;
    ocm_scan_kernel<1><<<grid_for(bytes), 256, 0, stream>>>(static_cast<const uint8_t *>(p), bytes, 0, sum);
    return cudaGetLastError();
}

} // namespace ocmk
