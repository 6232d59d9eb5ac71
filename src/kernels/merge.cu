// K8 (second half) - merge of sorted runs of 64-bit keys.
//
// The out-of-core sort (oncilla_b200/models/sort.py, BASELINE.json config 4) forms sorted runs, exchanges their
// slices between ranks and is then left with k sorted slices per partition.  Re-sorting them would throw the order
// away (eight radix passes); merging keeps it: ceil(log2 k) passes of a pairwise merge.  The reference has no
// counterpart (it moves bytes).
//
// Pairwise merge, merge-path partitioned: a tile is 2048 consecutive OUTPUT keys.  Two threads of the CTA find where
// the tile's first and last diagonal cut the two inputs (binary search in global memory), the CTA stages exactly
// those input pieces in shared memory, every thread finds its own 8-key diagonal in the staged pieces (binary
// search in shared memory), merges 8 keys in registers, and the tile leaves through shared memory again so that
// the global stores are whole 2 KiB rows.  Stable: on equal keys the left input goes first.
#include "kernels.h"
#include <algorithm>

namespace ocmk {

constexpr int kMergeThreads = 256;
constexpr int kMergeVT = 8;
constexpr int kMergeTile = kMergeThreads * kMergeVT;

// number of keys taken from `a` among the first `diag` outputs of merge(a, b)
template <typename T>
__device__ __forceinline__ uint32_t merge_path(const T *a, uint32_t na, const T *b, uint32_t nb, uint32_t diag) {
    uint32_t lo = diag > nb ? diag - nb : 0, hi = diag < na ? diag : na;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (a[mid] <= b[diag - 1 - mid]) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}
__device__ __forceinline__ uint64_t merge_path64(const uint64_t *a, uint64_t na, const uint64_t *b, uint64_t nb, uint64_t diag) {
    uint64_t lo = diag > nb ? diag - nb : 0, hi = diag < na ? diag : na;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (a[mid] <= b[diag - 1 - mid]) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}

struct MergePair {
    const uint64_t *a, *b;
    uint64_t *out;
    uint64_t na, nb;
    uint64_t first_tile;       // tiles [first_tile, first_tile of the next pair) belong to this pair
};
constexpr int kMergeMaxPairs = 32;
struct MergeArgs {
    MergePair pair[kMergeMaxPairs];
    uint32_t n_pairs;
    uint64_t n_tiles;
};

__global__ void __launch_bounds__(kMergeThreads) ocm_merge_kernel(const __grid_constant__ MergeArgs g) {
    __shared__ uint64_t in[kMergeTile + 2];
    __shared__ uint64_t outb[kMergeTile];
    __shared__ uint64_t cut[2];
    for (uint64_t tile = blockIdx.x; tile < g.n_tiles; tile += gridDim.x) {
        uint32_t pi = 0;
        while (pi + 1 < g.n_pairs && tile >= g.pair[pi + 1].first_tile) ++pi;
        const MergePair &p = g.pair[pi];
        const uint64_t total = p.na + p.nb;
        const uint64_t d0 = (tile - p.first_tile) * kMergeTile;
        const uint64_t d1 = d0 + kMergeTile < total ? d0 + kMergeTile : total;
        if (threadIdx.x < 2) cut[threadIdx.x] = merge_path64(p.a, p.na, p.b, p.nb, threadIdx.x ? d1 : d0);
        __syncthreads();
        const uint64_t a0 = cut[0], a1 = cut[1], b0 = d0 - a0, b1 = d1 - a1;
        const uint32_t ca = (uint32_t)(a1 - a0), cb = (uint32_t)(b1 - b0), n = ca + cb;
        for (uint32_t i = threadIdx.x; i < ca; i += kMergeThreads) in[i] = p.a[a0 + i];
        for (uint32_t i = threadIdx.x; i < cb; i += kMergeThreads) in[ca + i] = p.b[b0 + i];
        __syncthreads();
        const uint64_t *sa = in, *sb = in + ca;
        const uint32_t my = min((uint32_t)threadIdx.x * kMergeVT, n);
        uint32_t ia = merge_path(sa, ca, sb, cb, my), ib = my - ia;
        uint64_t r[kMergeVT];
#pragma unroll
        for (int k = 0; k < kMergeVT; ++k) {
            const bool has_a = ia < ca, has_b = ib < cb;
            const uint64_t va = has_a ? sa[ia] : ~0ull, vb = has_b ? sb[ib] : ~0ull;
            const bool take_a = has_a && (!has_b || va <= vb);
            r[k] = take_a ? va : vb;
            ia += take_a;
            ib += !take_a && has_b;
        }
#pragma unroll
        for (int k = 0; k < kMergeVT; ++k)
            if (my + k < n) outb[my + k] = r[k];
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < n; i += kMergeThreads) p.out[d0 + i] = outb[i];
        __syncthreads();
    }
}

// One launch merges up to kMergeMaxPairs independent pairs (one level of a merge tree).
cudaError_t launch_merge_pairs(const MergePairDesc *pairs, int n_pairs, cudaStream_t stream, unsigned long long *launches) {
    for (int base = 0; base < n_pairs; base += kMergeMaxPairs) {
        const int m = std::min(kMergeMaxPairs, n_pairs - base);
        MergeArgs g;
        uint64_t tiles = 0;
        for (int i = 0; i < m; ++i) {
            const MergePairDesc &d = pairs[base + i];
            g.pair[i].a = d.a; g.pair[i].b = d.b; g.pair[i].out = d.out; g.pair[i].na = d.na; g.pair[i].nb = d.nb;
            g.pair[i].first_tile = tiles;
            tiles += (d.na + d.nb + kMergeTile - 1) / kMergeTile;
        }
        g.n_pairs = (uint32_t)m;
        g.n_tiles = tiles;
        if (!tiles) continue;
        const int ctas = (int)std::min<uint64_t>(tiles, (uint64_t)num_sms() * 6);
        ocm_merge_kernel<<<ctas, kMergeThreads, 0, stream>>>(g);
        if (launches) ++*launches;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

} // namespace ocmk
