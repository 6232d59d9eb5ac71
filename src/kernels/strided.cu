// K4/K5 - strided scatter/gather fused with pack/unpack and optional cast.
//
// Generalises the reference's two-hop staging (memcpy into a bounce buffer at
// one offset, then a transport write from another: reference src/lib.c:526-547)
// to an N-D box (up to 3 outer dimensions of contiguous rows) moved in ONE pass:
//   * byte rows, 16-byte aligned: each warp gathers a batch of rows with one
//     bulk (TMA) load per row into a shared-memory stage - the pack step - and
//     scatters them with bulk stores to the (peer) destination;
//   * any alignment: 16/8/4/1-byte vector path;
//   * cast rows: fp32/bf16 <-> bf16/fp32, and block-scaled fp8-e4m3
//     quantise/dequantise (per-block amax by warp shuffles, one fp32 scale per
//     block) fused into the same pass, so a strided fp8 put reads the source
//     once and sends a quarter of the bytes over NVLink.
// The CPU restatement is ocm_oracle_strided_copy (src/oracle/oracle.cpp); the
// two must agree bit for bit.
#include "kernels.h"
#include "ptx.cuh"
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp8.h>
#include <mutex>

namespace ocmk {

struct RowMap {
    uint64_t n0, n1, n2;          // extents of the outer dimensions
    int64_t  ss0, ss1, ss2;       // source strides (bytes)
    int64_t  ds0, ds1, ds2;       // destination strides (bytes)
};

// 64-bit integer division is ~100 instructions on the SM: take the 32-bit path whenever the operands fit
// (the cast kernel was issue-bound on its index arithmetic, profiles/README.md §5).
__device__ __forceinline__ void divmod(uint64_t a, uint64_t b, uint64_t *q, uint64_t *r) {
    if (((a | b) >> 32) == 0) {
        const uint32_t a32 = (uint32_t)a, b32 = (uint32_t)b;
        *q = a32 / b32;
        *r = a32 % b32;
    } else {
        *q = a / b;
        *r = a % b;
    }
}

__device__ __forceinline__ void row_offsets(const RowMap &m, uint64_t row, int64_t *so, int64_t *dof) {
    if (m.n1 == 1 && m.n2 == 1) {   // plain 2-D case: no division at all
        *so = (int64_t)row * m.ss0;
        *dof = (int64_t)row * m.ds0;
        return;
    }
    uint64_t i1, t, i2, i3;
    divmod(row, m.n0, &t, &i1);
    divmod(t, m.n1, &i3, &i2);
    *so = (int64_t)i1 * m.ss0 + (int64_t)i2 * m.ss1 + (int64_t)i3 * m.ss2;
    *dof = (int64_t)i1 * m.ds0 + (int64_t)i2 * m.ds1 + (int64_t)i3 * m.ds2;
}

struct StridedArgs {
    const uint8_t *src;
    uint8_t       *dst;
    RowMap         map;
    uint64_t       rows;
    uint64_t       row_bytes;       // source side
    uint32_t       tile, stages;    // bulk path
    uint32_t       rows_per_batch, chunks_per_row, chunk_bytes;
    uint64_t      *signal;
    unsigned      *ticket;
};

constexpr int kSAgents = 8;
constexpr int kSStages = 8;

// ---- bulk rows -------------------------------------------------------------------
// A "batch" is either up to 32 whole rows that fit one stage, or one chunk of a
// long row.  Every lane of the warp owns one row of the batch.
__global__ void __launch_bounds__(kSAgents * 32, 1) ocm_strided_bulk_kernel(StridedArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[kSAgents][kSStages];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int agents = blockDim.x >> 5;
    if (lane == 0) {
        for (uint32_t s = 0; s < a.stages; ++s) mbar_init(&full[warp][s], 1);
        mbar_fence_init();
    }
    __syncthreads();

    const bool long_rows = a.chunks_per_row > 1;
    const uint64_t n_batches =
        long_rows ? a.rows * a.chunks_per_row : (a.rows + a.rows_per_batch - 1) / a.rows_per_batch;
    const uint64_t ag = (uint64_t)blockIdx.x * agents + warp;
    const uint64_t total = (uint64_t)gridDim.x * agents;
    uint8_t *my = smem + (size_t)warp * a.stages * a.tile;

    if (ag < n_batches) {
        const uint64_t n_my = (n_batches - ag + total - 1) / total;
        // describe this lane's piece of batch j: global offsets, smem slot, bytes (0 = lane idle)
        auto piece = [&](uint64_t j, int64_t *so, int64_t *dof, uint32_t *soff, uint32_t *len) {
            uint64_t b = ag + j * total;
            *len = 0;
            if (long_rows) {
                if (lane != 0) return;
                uint64_t row = b / a.chunks_per_row, ch = b % a.chunks_per_row;
                row_offsets(a.map, row, so, dof);
                uint64_t o = ch * (uint64_t)a.chunk_bytes;
                *so += (int64_t)o;
                *dof += (int64_t)o;
                *soff = 0;
                *len = (uint32_t)(a.row_bytes - o < a.chunk_bytes ? a.row_bytes - o : a.chunk_bytes);
            } else {
                uint64_t row = b * a.rows_per_batch + lane;
                if ((uint32_t)lane >= a.rows_per_batch || row >= a.rows) return;
                row_offsets(a.map, row, so, dof);
                *soff = (uint32_t)lane * (uint32_t)a.row_bytes;
                *len = (uint32_t)a.row_bytes;
            }
        };
        auto issue_load = [&](uint64_t j) {
            uint32_t s = (uint32_t)(j % a.stages);
            int64_t so = 0, dof = 0;
            uint32_t soff = 0, len = 0;
            piece(j, &so, &dof, &soff, &len);
            uint32_t tot = len;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
            if (lane == 0) mbar_arrive_expect_tx(&full[warp][s], tot);
            __syncwarp();
            if (len) bulk_g2s(my + (size_t)s * a.tile + soff, a.src + so, len, &full[warp][s]);
        };
        const uint64_t npro = n_my < a.stages ? n_my : a.stages;
        for (uint64_t j = 0; j < npro; ++j) issue_load(j);
        for (uint64_t i = 0; i < n_my; ++i) {
            uint32_t s = (uint32_t)(i % a.stages);
            mbar_wait(&full[warp][s], (uint32_t)((i / a.stages) & 1));
            int64_t so = 0, dof = 0;
            uint32_t soff = 0, len = 0;
            piece(i, &so, &dof, &soff, &len);
            if (len) bulk_s2g(a.dst + dof, my + (size_t)s * a.tile + soff, len);
            bulk_commit();   // every lane commits (possibly empty) so group counts stay aligned
            if (i >= 1 && i - 1 + a.stages < n_my) {
                bulk_wait_read<1>();
                __syncwarp();
                issue_load(i - 1 + a.stages);
            }
        }
        bulk_wait_all<0>();
        fence_proxy_async_all();
        __threadfence_system();
    }
    __syncthreads();
    if (threadIdx.x == 0) finish_and_signal(a.signal, a.ticket);
}

// ---- tensor-map rows (TMA tiled mode) ------------------------------------------------
// Source and destination boxes are described to the TMA unit as rank-4 tensors of 8-byte elements
// (cuTensorMapEncodeTiled: extents {row_bytes/8, n0, n1, n2}, byte strides {ss0, ss1, ss2}); one elected lane per
// warp pulls a {bx x by} tile into shared memory with cp.async.bulk.tensor (SASS UTMALDG, completion on an mbarrier)
// and pushes it to the destination box with the store form (UTMASTG).  The address arithmetic of the gather and of
// the scatter - the pack/unpack of reference src/lib.c:526-547 - is done by the copy engine, not by SM instructions;
// edge tiles are clipped by the hardware.  Needs 16-byte aligned bases, row lengths and strides, positive strides.
struct TmaStridedArgs {
    uint32_t tx, ty, n1, n2;       // tiles along the row / along dimension n0, outer extents
    uint32_t bx, by;               // tile: bx 8-byte elements x by rows
    uint32_t tile_bytes, stages;
    uint32_t pitch;                // shared-memory distance between stages (tile_bytes rounded up to 128)
    uint64_t n_tiles;
    uint64_t *signal;
    unsigned *ticket;
};

__device__ __forceinline__ void tma_load_4d(void *smem_dst, const CUtensorMap *map, uint32_t c0, uint32_t c1, uint32_t c2,
                                            uint32_t c3, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap *map, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                             const void *smem_src) {
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%1, %2, %3, %4}], [%5];" ::"l"(
                     reinterpret_cast<uint64_t>(map)),
                 "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(smem_u32(smem_src))
                 : "memory");
}

__global__ void __launch_bounds__(kSAgents * 32, 1)
    ocm_strided_tma_kernel(const __grid_constant__ CUtensorMap src_map, const __grid_constant__ CUtensorMap dst_map, TmaStridedArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[kSAgents][kSStages];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int agents = blockDim.x >> 5;
    if (lane == 0) {
        for (uint32_t s = 0; s < a.stages; ++s) mbar_init(&full[warp][s], 1);
        mbar_fence_init();
    }
    __syncthreads();
    const uint64_t ag = (uint64_t)blockIdx.x * agents + warp;
    const uint64_t total = (uint64_t)gridDim.x * agents;
    if (lane == 0 && ag < a.n_tiles) {
        // 128-byte alignment of every stage: tile_bytes is a multiple of 128
        uint8_t *my = smem + (size_t)warp * a.stages * a.pitch;
        const uint64_t n_my = (a.n_tiles - ag + total - 1) / total;
        auto coords = [&](uint64_t j, uint32_t *c) {
            uint64_t t = ag + j * total;
            const uint32_t ix = (uint32_t)(t % a.tx);
            t /= a.tx;
            const uint32_t iy = (uint32_t)(t % a.ty);
            t /= a.ty;
            c[0] = ix * a.bx;
            c[1] = iy * a.by;
            c[2] = (uint32_t)(t % a.n1);
            c[3] = (uint32_t)(t / a.n1);
        };
        auto issue_load = [&](uint64_t j) {
            const uint32_t s = (uint32_t)(j % a.stages);
            uint32_t c[4];
            coords(j, c);
            mbar_arrive_expect_tx(&full[warp][s], a.tile_bytes);   // the unit always delivers the whole box (zero-filled past the edge)
            tma_load_4d(my + (size_t)s * a.pitch, &src_map, c[0], c[1], c[2], c[3], &full[warp][s]);
        };
        const uint64_t npro = n_my < a.stages ? n_my : a.stages;
        for (uint64_t j = 0; j < npro; ++j) issue_load(j);
        for (uint64_t i = 0; i < n_my; ++i) {
            const uint32_t s = (uint32_t)(i % a.stages);
            mbar_wait(&full[warp][s], (uint32_t)((i / a.stages) & 1));
            uint32_t c[4];
            coords(i, c);
            tma_store_4d(&dst_map, c[0], c[1], c[2], c[3], my + (size_t)s * a.pitch);
            bulk_commit();
            if (i >= 1 && i - 1 + a.stages < n_my) {
                bulk_wait_read<1>();          // the store of tile i-1 has read its stage: refill it
                issue_load(i - 1 + a.stages);
            }
        }
        bulk_wait_all<0>();
        fence_proxy_async_all();
        __threadfence_system();
    }
    __syncthreads();
    if (threadIdx.x == 0) finish_and_signal(a.signal, a.ticket);
}

// ---- wide vector rows: 32 (or 16) bytes per thread per access, U accesses in flight -------------------
// For rows that are whole 32-byte (16-byte) units on both sides: consecutive threads take consecutive units of the
// flattened (row, unit) space, so a warp covers up to 1 KiB of one row - or several short rows - per instruction,
// and every thread has U independent loads in flight before the first store (the 1-D bulk kernel above keeps one
// lane per row busy and stalls on its barrier ring for L2-resident boxes: profiles/README.md section 5).
struct W32 {
    uint32_t w[8];
};
template <int VB> struct WideVec;
template <> struct WideVec<32> {
    using T = W32;
    static __device__ __forceinline__ T ld(const uint8_t *p) {
        T v;
        asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3]), "=r"(v.w[4]), "=r"(v.w[5]), "=r"(v.w[6]), "=r"(v.w[7])
                     : "l"(p));
        return v;
    }
    static __device__ __forceinline__ void st(uint8_t *p, const T &v) {
        asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v.w[0]), "r"(v.w[1]),
                     "r"(v.w[2]), "r"(v.w[3]), "r"(v.w[4]), "r"(v.w[5]), "r"(v.w[6]), "r"(v.w[7])
                     : "memory");
    }
};
template <> struct WideVec<16> {
    using T = uint4;
    static __device__ __forceinline__ T ld(const uint8_t *p) { return ldg_stream(p); }
    static __device__ __forceinline__ void st(uint8_t *p, const T &v) { stg_stream(p, v); }
};

template <int VB, int U>
__global__ void __launch_bounds__(256) ocm_strided_wide_kernel(StridedArgs a) {
    using V = WideVec<VB>;
    const uint64_t upr = a.row_bytes / VB;          // units per row
    const uint64_t total = a.rows * upr;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    auto addr = [&](uint64_t u, const uint8_t **s, uint8_t **d) {
        uint64_t row, k;
        divmod(u, upr, &row, &k);
        int64_t so, dof;
        row_offsets(a.map, row, &so, &dof);
        *s = a.src + so + (int64_t)(k * VB);
        *d = a.dst + dof + (int64_t)(k * VB);
    };
    uint64_t u = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; u + (U - 1) * stride < total; u += U * stride) {
        const uint8_t *s[U];
        uint8_t *d[U];
        typename V::T v[U];
#pragma unroll
        for (int i = 0; i < U; ++i) addr(u + i * stride, &s[i], &d[i]);
#pragma unroll
        for (int i = 0; i < U; ++i) v[i] = V::ld(s[i]);
#pragma unroll
        for (int i = 0; i < U; ++i) V::st(d[i], v[i]);
    }
    for (; u < total; u += stride) {
        const uint8_t *s;
        uint8_t *d;
        addr(u, &s, &d);
        V::st(d, V::ld(s));
    }
    __syncthreads();
    if (threadIdx.x == 0) finish_and_signal(a.signal, a.ticket);
}

// ---- vector rows (any alignment) ------------------------------------------------
template <typename VT>
__global__ void __launch_bounds__(256) ocm_strided_vec_kernel(StridedArgs a) {
    constexpr uint64_t V = sizeof(VT);
    const uint64_t vpr = a.row_bytes / V;        // host guarantees row_bytes % V == 0
    const uint64_t total = a.rows * vpr;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        uint64_t row, v;
        divmod(i, vpr, &row, &v);
        int64_t so, dof;
        row_offsets(a.map, row, &so, &dof);
        *reinterpret_cast<VT *>(a.dst + dof + (int64_t)(v * V)) =
            *reinterpret_cast<const VT *>(a.src + so + (int64_t)(v * V));
    }
    __syncthreads();
    if (threadIdx.x == 0) finish_and_signal(a.signal, a.ticket);
}

// ---- cast rows ----------------------------------------------------------------------
struct CastArgs {
    const uint8_t *src;
    uint8_t       *dst;
    RowMap         map;
    uint64_t       rows, elems;
    int            src_dtype, dst_dtype;
    uint32_t       blk;             // scale block (elements); 0 = plain cast
    uint32_t       bpr;             // blocks per row
    float         *scales;
    uint64_t      *signal;
    unsigned      *ticket;
};

__device__ __forceinline__ void load8(const uint8_t *p, int dt, float v[8]) {
    if (dt == DT_F32) {
        uint4 a = *reinterpret_cast<const uint4 *>(p), b = *reinterpret_cast<const uint4 *>(p + 16);
        v[0] = __uint_as_float(a.x); v[1] = __uint_as_float(a.y); v[2] = __uint_as_float(a.z); v[3] = __uint_as_float(a.w);
        v[4] = __uint_as_float(b.x); v[5] = __uint_as_float(b.y); v[6] = __uint_as_float(b.z); v[7] = __uint_as_float(b.w);
    } else if (dt == DT_BF16) {
        uint4 a = *reinterpret_cast<const uint4 *>(p);
        const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[2 * i] = __uint_as_float(w[i] << 16);
            v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    } else { // e4m3
        uint2 a = *reinterpret_cast<const uint2 *>(p);
        const uint32_t w[2] = {a.x, a.y};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __nv_fp8x2_storage_t pr = (__nv_fp8x2_storage_t)((w[i >> 1] >> (16 * (i & 1))) & 0xffffu);
            __half2_raw h = __nv_cvt_fp8x2_to_halfraw2(pr, __NV_E4M3);
            float2 f = __half22float2(*reinterpret_cast<__half2 *>(&h));
            v[2 * i] = f.x;
            v[2 * i + 1] = f.y;
        }
    }
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    __nv_bfloat162 b = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t *>(&b);
}

__device__ __forceinline__ void store8(uint8_t *p, int dt, const float v[8]) {
    if (dt == DT_F32) {
        *reinterpret_cast<uint4 *>(p) = make_uint4(__float_as_uint(v[0]), __float_as_uint(v[1]), __float_as_uint(v[2]), __float_as_uint(v[3]));
        *reinterpret_cast<uint4 *>(p + 16) = make_uint4(__float_as_uint(v[4]), __float_as_uint(v[5]), __float_as_uint(v[6]), __float_as_uint(v[7]));
    } else if (dt == DT_BF16) {
        *reinterpret_cast<uint4 *>(p) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
    } else {
        uint32_t w[2] = {0, 0};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __nv_fp8x2_storage_t pr = __nv_cvt_float2_to_fp8x2(make_float2(v[2 * i], v[2 * i + 1]), __NV_SATFINITE, __NV_E4M3);
            w[i >> 1] |= (uint32_t)pr << (16 * (i & 1));
        }
        *reinterpret_cast<uint2 *>(p) = make_uint2(w[0], w[1]);
    }
}

// One thread = 8 consecutive elements per chunk, U chunks per loop iteration: all U loads are issued before the
// first conversion, so every thread keeps U x 32 bytes (fp32 source) in flight.  With quantisation the G = blk/8
// lanes of an aligned lane group share one scale block.  The chunk -> (row, column) split is a shift when the row
// holds a power-of-two number of chunks (POW2), a 32-bit division otherwise.
template <int U, bool POW2>
__global__ void __launch_bounds__(256) ocm_cast_kernel(CastArgs a) {
    const bool quant = a.dst_dtype == DT_F8E4M3 && a.src_dtype != DT_F8E4M3;
    const bool dequant = a.src_dtype == DT_F8E4M3 && a.dst_dtype != DT_F8E4M3;
    const uint32_t G = a.blk ? a.blk / 8 : 1;                 // 1, 4 or 16: always a power of two
    const uint32_t gshift = G == 16 ? 4 : G == 4 ? 2 : 0;
    const uint64_t cpr = a.blk ? (uint64_t)a.bpr * G : (a.elems + 7) / 8;   // chunks per row, padded to whole blocks
    const uint32_t cshift = POW2 ? (uint32_t)(63 - __clzll((long long)cpr)) : 0;
    const uint64_t total = a.rows * cpr;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const size_t ss = dtype_size(a.src_dtype), ds = dtype_size(a.dst_dtype);
    // all lanes of a warp iterate together (shuffles below): the loop bound is the warp's first chunk
    const uint64_t lane = threadIdx.x & 31;
    for (uint64_t c0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; c0 - lane < total; c0 += (uint64_t)U * stride) {
        float v[U][8];
        uint64_t row[U], ch[U];
        int64_t dof[U];
        bool valid[U];
#pragma unroll
        for (int i = 0; i < U; ++i) {
            const uint64_t c = c0 + (uint64_t)i * stride;
            if (POW2) {
                row[i] = c >> cshift;
                ch[i] = c & (cpr - 1);
            } else {
                divmod(c, cpr, &row[i], &ch[i]);
            }
            valid[i] = c < total && ch[i] * 8 < a.elems;
            int64_t so = 0;
            dof[i] = 0;
            if (valid[i]) {
                row_offsets(a.map, row[i], &so, &dof[i]);
                load8(a.src + so + (int64_t)(ch[i] * 8 * ss), a.src_dtype, v[i]);
            } else {
#pragma unroll
                for (int k = 0; k < 8; ++k) v[i][k] = 0.f;
            }
        }
#pragma unroll
        for (int i = 0; i < U; ++i) {
            if (c0 - lane + (uint64_t)i * stride >= total) break;      // warp-uniform: nothing left for this warp
            if (quant) {
                float amax = 0.f;
#pragma unroll
                for (int k = 0; k < 8; ++k) amax = fmaxf(amax, fabsf(v[i][k]));
                for (uint32_t o = G >> 1; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
                // scale = amax/448 and its reciprocal: two IEEE divisions per thread, then multiplies
                // (the oracle performs the same operations in the same order, so results match bit for bit)
                const float scale = amax > 1e-30f ? __fdiv_rn(amax, 448.0f) : 1.0f;   // (near-)zero block: scale 1
                const float inv = __fdiv_rn(1.0f, scale);
                if (valid[i]) {
                    if ((ch[i] & (G - 1)) == 0) a.scales[row[i] * a.bpr + (ch[i] >> gshift)] = scale;
#pragma unroll
                    for (int k = 0; k < 8; ++k) v[i][k] = __fmul_rn(v[i][k], inv);
                }
            } else if (dequant && valid[i]) {
                const float scale = a.scales[row[i] * a.bpr + (ch[i] >> gshift)];
#pragma unroll
                for (int k = 0; k < 8; ++k) v[i][k] = __fmul_rn(v[i][k], scale);
            }
            if (valid[i]) store8(a.dst + dof[i] + (int64_t)(ch[i] * 8 * ds), a.dst_dtype, v[i]);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) finish_and_signal(a.signal, a.ticket);
}

// ---- cast rows, 32 elements per thread ------------------------------------------------------------
// The throughput path for rows that are whole 32-element chunks with 32-byte aligned rows on both sides.  A thread
// owns 32 consecutive elements: 128 / 64 / 32 bytes of fp32 / bf16 / fp8, moved with 256-bit accesses, U chunks in
// flight.  With block-scaled fp8 a scale block is one thread (block 32: no shuffle at all) or four neighbouring
// lanes (block 128: two shuffles), so the two IEEE divisions that define the scale (the oracle's formula, bit for
// bit) are paid once per 32 elements instead of once per 8, and a dequantising thread loads its scale once per 128
// output bytes.  Source/destination types are template parameters: no per-element type dispatch in the loop.
template <int DT> struct ElemIO;
template <> struct ElemIO<DT_F32> {
    static constexpr int kBytes = 128;
    static __device__ __forceinline__ void load(const uint8_t *p, float (&v)[32]) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const W32 w = WideVec<32>::ld(p + 32 * q);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[8 * q + k] = __uint_as_float(w.w[k]);
        }
    }
    static __device__ __forceinline__ void store(uint8_t *p, const float (&v)[32]) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            W32 w;
#pragma unroll
            for (int k = 0; k < 8; ++k) w.w[k] = __float_as_uint(v[8 * q + k]);
            WideVec<32>::st(p + 32 * q, w);
        }
    }
};
template <> struct ElemIO<DT_BF16> {
    static constexpr int kBytes = 64;
    static __device__ __forceinline__ void load(const uint8_t *p, float (&v)[32]) {
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const W32 w = WideVec<32>::ld(p + 32 * q);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                v[16 * q + 2 * k] = __uint_as_float(w.w[k] << 16);
                v[16 * q + 2 * k + 1] = __uint_as_float(w.w[k] & 0xffff0000u);
            }
        }
    }
    static __device__ __forceinline__ void store(uint8_t *p, const float (&v)[32]) {
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            W32 w;
#pragma unroll
            for (int k = 0; k < 8; ++k) w.w[k] = pack_bf16x2(v[16 * q + 2 * k], v[16 * q + 2 * k + 1]);
            WideVec<32>::st(p + 32 * q, w);
        }
    }
};
template <> struct ElemIO<DT_F8E4M3> {
    static constexpr int kBytes = 32;
    static __device__ __forceinline__ void load(const uint8_t *p, float (&v)[32]) {
        const W32 w = WideVec<32>::ld(p);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const __nv_fp8x2_storage_t pr = (__nv_fp8x2_storage_t)((w.w[k] >> (16 * h)) & 0xffffu);
                __half2_raw hr = __nv_cvt_fp8x2_to_halfraw2(pr, __NV_E4M3);
                const float2 f = __half22float2(*reinterpret_cast<__half2 *>(&hr));
                v[4 * k + 2 * h] = f.x;
                v[4 * k + 2 * h + 1] = f.y;
            }
        }
    }
    static __device__ __forceinline__ void store(uint8_t *p, const float (&v)[32]) {
        W32 w;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const uint32_t lo = __nv_cvt_float2_to_fp8x2(make_float2(v[4 * k], v[4 * k + 1]), __NV_SATFINITE, __NV_E4M3);
            const uint32_t hi = __nv_cvt_float2_to_fp8x2(make_float2(v[4 * k + 2], v[4 * k + 3]), __NV_SATFINITE, __NV_E4M3);
            w.w[k] = lo | (hi << 16);
        }
        WideVec<32>::st(p, w);
    }
};

// Store the 32 converted elements of every lane so that each store instruction of the warp writes consecutive
// 32-byte pieces.  A lane's chunk is KB = 128 (fp32) or 64 (bf16) bytes of destination: storing it directly would
// put the lanes of one instruction KB bytes apart, i.e. a quarter (half) of every 128-byte line per instruction -
// harmless inside one HBM, where the L2 merges the sectors, but peer writes bypass the L2 and go out over NVLink as
// partial-line packets (measured 371 GB/s for an fp32 -> bf16 put against 660+ for full lines).  So the warp
// transposes through shared memory: rows of KB bytes with a 16-byte pad (conflict-free 128-bit accesses), read back
// as the warp-contiguous 32-byte piece q * 1 KiB + lane * 32 whose owner lane supplies the destination address.
template <int DDT> struct WideStore {
    static constexpr int KB = ElemIO<DDT>::kBytes;
    static constexpr int kPitch = KB + 16;
    static constexpr int kWarpBytes = 32 * kPitch;
    static __device__ __forceinline__ void run(uint8_t *w, uint32_t lane, const float (&v)[32], uint8_t *my_dst) {
        uint32_t words[KB / 4];
        if (DDT == DT_F32) {
#pragma unroll
            for (int k = 0; k < 32; ++k) words[k % (KB / 4)] = __float_as_uint(v[k]);
        } else {
#pragma unroll
            for (int k = 0; k < 16; ++k) words[k % (KB / 4)] = pack_bf16x2(v[2 * k], v[2 * k + 1]);
        }
        uint8_t *row = w + lane * kPitch;
#pragma unroll
        for (int j = 0; j < KB / 16; ++j)
            *reinterpret_cast<uint4 *>(row + 16 * j) = make_uint4(words[4 * j], words[4 * j + 1], words[4 * j + 2], words[4 * j + 3]);
        __syncwarp();
        const uint64_t base = reinterpret_cast<uint64_t>(my_dst);
#pragma unroll
        for (int q = 0; q < KB / 32; ++q) {
            const uint32_t x = (uint32_t)q * 1024u + lane * 32u;
            const uint32_t j = x / KB, off = x % KB;
            const uint64_t bj = __shfl_sync(0xffffffffu, base, (int)j);
            const uint4 lo = *reinterpret_cast<const uint4 *>(w + j * kPitch + off);
            const uint4 hi = *reinterpret_cast<const uint4 *>(w + j * kPitch + off + 16);
            if (bj) {
                W32 o;
                o.w[0] = lo.x; o.w[1] = lo.y; o.w[2] = lo.z; o.w[3] = lo.w;
                o.w[4] = hi.x; o.w[5] = hi.y; o.w[6] = hi.z; o.w[7] = hi.w;
                WideVec<32>::st(reinterpret_cast<uint8_t *>(bj) + off, o);
            }
        }
        __syncwarp();
    }
};

template <int SDT, int DDT, int U>
__global__ void __launch_bounds__(256, 2) ocm_cast32_kernel(CastArgs a) {
    constexpr bool wide_dst = DDT != DT_F8E4M3;
    __shared__ __align__(16) uint8_t xpose[wide_dst ? 8 * (32 * (ElemIO<wide_dst ? DDT : DT_F32>::kBytes + 16)) : 16];
    constexpr bool quant = DDT == DT_F8E4M3, dequant = SDT == DT_F8E4M3;
    const uint32_t G = a.blk == 128 ? 4u : 1u;                  // threads per scale block
    const uint64_t cpr = a.elems / 32;                          // chunks per row (host guarantees whole chunks)
    const uint64_t total = a.rows * cpr;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t lane = threadIdx.x & 31;
    for (uint64_t c0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; c0 - lane < total; c0 += (uint64_t)U * stride) {
        float v[U][32];
        uint64_t row[U], ch[U];
        int64_t dof[U];
        bool valid[U];
#pragma unroll
        for (int i = 0; i < U; ++i) {
            const uint64_t c = c0 + (uint64_t)i * stride;
            divmod(c, cpr, &row[i], &ch[i]);
            valid[i] = c < total;
            int64_t so = 0;
            dof[i] = 0;
            if (valid[i]) {
                row_offsets(a.map, row[i], &so, &dof[i]);
                ElemIO<SDT>::load(a.src + so + (int64_t)(ch[i] * ElemIO<SDT>::kBytes), v[i]);
            } else {
#pragma unroll
                for (int k = 0; k < 32; ++k) v[i][k] = 0.f;
            }
        }
#pragma unroll
        for (int i = 0; i < U; ++i) {
            if (c0 - lane + (uint64_t)i * stride >= total) break;       // warp-uniform
            if (quant) {
                float amax = 0.f;
#pragma unroll
                for (int k = 0; k < 32; ++k) amax = fmaxf(amax, fabsf(v[i][k]));
                for (uint32_t o = G >> 1; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
                const float scale = amax > 1e-30f ? __fdiv_rn(amax, 448.0f) : 1.0f;   // (near-)zero block: scale 1
                const float inv = __fdiv_rn(1.0f, scale);
                if (valid[i]) {
                    if ((ch[i] & (G - 1)) == 0) a.scales[row[i] * a.bpr + ch[i] / G] = scale;
#pragma unroll
                    for (int k = 0; k < 32; ++k) v[i][k] = __fmul_rn(v[i][k], inv);
                }
            } else if (dequant && valid[i]) {
                const float scale = a.scales[row[i] * a.bpr + ch[i] / G];
#pragma unroll
                for (int k = 0; k < 32; ++k) v[i][k] = __fmul_rn(v[i][k], scale);
            }
            uint8_t *my_dst = valid[i] ? a.dst + dof[i] + (int64_t)(ch[i] * ElemIO<DDT>::kBytes) : nullptr;
            if constexpr (wide_dst) {
                WideStore<DDT>::run(xpose + (threadIdx.x >> 5) * WideStore<DDT>::kWarpBytes, (uint32_t)lane, v[i], my_dst);
            } else {
                if (my_dst) ElemIO<DDT>::store(my_dst, v[i]);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) finish_and_signal(a.signal, a.ticket);
}

template <int U> static bool launch_cast32(const CastArgs &c, int ctas, cudaStream_t stream) {
#define OCM_CAST32(S, D)                                                     \
    if (c.src_dtype == S && c.dst_dtype == D) {                              \
        ocm_cast32_kernel<S, D, U><<<ctas, 256, 0, stream>>>(c);             \
        return true;                                                         \
    }
    OCM_CAST32(DT_F32, DT_BF16)
    OCM_CAST32(DT_BF16, DT_F32)
    OCM_CAST32(DT_F32, DT_F8E4M3)
    OCM_CAST32(DT_BF16, DT_F8E4M3)
    OCM_CAST32(DT_F8E4M3, DT_F32)
    OCM_CAST32(DT_F8E4M3, DT_BF16)
#undef OCM_CAST32
    return false;
}

// Scalar fallback for rows whose length or alignment does not fit 8-element
// chunks: one thread per scale block (or per 8 elements for plain casts).
__device__ __forceinline__ float load1(const uint8_t *p, int dt) {
    if (dt == DT_F32) return *reinterpret_cast<const float *>(p);
    if (dt == DT_BF16) return __uint_as_float((uint32_t)*reinterpret_cast<const uint16_t *>(p) << 16);
    __half_raw h = __nv_cvt_fp8_to_halfraw(*p, __NV_E4M3);
    return __half2float(*reinterpret_cast<__half *>(&h));
}
__device__ __forceinline__ void store1(uint8_t *p, int dt, float f) {
    if (dt == DT_F32) *reinterpret_cast<float *>(p) = f;
    else if (dt == DT_BF16) {
        __nv_bfloat16 b = __float2bfloat16_rn(f);
        *reinterpret_cast<uint16_t *>(p) = *reinterpret_cast<uint16_t *>(&b);
    } else *p = (uint8_t)__nv_cvt_float_to_fp8(f, __NV_SATFINITE, __NV_E4M3);
}

__global__ void __launch_bounds__(256) ocm_cast_scalar_kernel(CastArgs a) {
    const bool quant = a.dst_dtype == DT_F8E4M3 && a.src_dtype != DT_F8E4M3;
    const bool dequant = a.src_dtype == DT_F8E4M3 && a.dst_dtype != DT_F8E4M3;
    const uint32_t span = a.blk ? a.blk : 8;
    const uint64_t spr = (a.elems + span - 1) / span;
    const uint64_t total = a.rows * spr;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const size_t ss = dtype_size(a.src_dtype), ds = dtype_size(a.dst_dtype);
    for (uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; c < total; c += stride) {
        uint64_t row, b;
        divmod(c, spr, &row, &b);
        const uint64_t lo = b * span, hi = lo + span < a.elems ? lo + span : a.elems;
        int64_t so, dof;
        row_offsets(a.map, row, &so, &dof);
        const uint8_t *s = a.src + so;
        uint8_t *d = a.dst + dof;
        float scale = 1.0f;
        if (quant) {
            float amax = 0.f;
            for (uint64_t e = lo; e < hi; ++e) amax = fmaxf(amax, fabsf(load1(s + e * ss, a.src_dtype)));
            scale = amax > 1e-30f ? __fdiv_rn(amax, 448.0f) : 1.0f;
            a.scales[row * a.bpr + b] = scale;
        } else if (dequant) {
            scale = a.scales[row * a.bpr + b];
        }
        const float inv = quant ? __fdiv_rn(1.0f, scale) : 1.0f;
        for (uint64_t e = lo; e < hi; ++e) {
            float f = load1(s + e * ss, a.src_dtype);
            if (quant) f = __fmul_rn(f, inv);
            else if (dequant) f = __fmul_rn(f, scale);
            store1(d + e * ds, a.dst_dtype, f);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) finish_and_signal(a.signal, a.ticket);
}

// ------------------------------------------------------------------ host ----
// cuTensorMapEncodeTiled through the runtime's driver-entry-point lookup (libcuda is never linked).
using EncodeTiledFn = CUresult (*)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                   const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult st;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &st) != cudaSuccess || !p ||
            st != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            p = nullptr;
        }
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// Tensor maps are cached per (base, extents, strides, tile): an application that moves the same box again and
// again (halo exchange, paged KV blocks) pays the ~1 us encoding once.
struct TmapKey {
    const void *base;
    uint64_t row_bytes, n[3];
    int64_t  stride[3];
    uint32_t bx, by;
    bool operator==(const TmapKey &o) const { return memcmp(this, &o, sizeof *this) == 0; }
};
struct TmapEntry {
    TmapKey key;
    CUtensorMap map;
};
static std::mutex g_tmap_mu;
static TmapEntry g_tmap_cache[32];
static int g_tmap_n = 0;

static bool tmap_get(const void *base, uint64_t row_bytes, const uint64_t n[3], const int64_t stride[3], uint32_t bx, uint32_t by,
                     CUtensorMap *out) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return false;
    TmapKey k;
    memset(&k, 0, sizeof k);
    k.base = base; k.row_bytes = row_bytes; k.bx = bx; k.by = by;
    for (int i = 0; i < 3; ++i) { k.n[i] = n[i]; k.stride[i] = n[i] > 1 ? stride[i] : 0; }
    {
        std::lock_guard<std::mutex> g(g_tmap_mu);
        for (int i = 0; i < g_tmap_n; ++i)
            if (g_tmap_cache[i].key == k) {
                *out = g_tmap_cache[i].map;
                if (i) std::swap(g_tmap_cache[i], g_tmap_cache[i - 1]);   // drift towards the front
                return true;
            }
    }
    cuuint64_t dims[4] = {row_bytes / 8, n[0], n[1], n[2]};
    cuuint64_t strides[3];
    uint64_t prev = (row_bytes + 15) & ~15ull;
    for (int i = 0; i < 3; ++i) {
        // a dimension of extent 1 never uses its stride: give it any legal one
        strides[i] = n[i] > 1 ? (cuuint64_t)stride[i] : prev;
        prev = strides[i] * std::max<uint64_t>(n[i], 1);
        if (prev >= (1ull << 40)) prev = 16;
    }
    cuuint32_t box[4] = {bx, by, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUtensorMap m;
    CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT64, 4, const_cast<void *>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    *out = m;
    std::lock_guard<std::mutex> g(g_tmap_mu);
    const int slot = g_tmap_n < 32 ? g_tmap_n++ : 31;
    g_tmap_cache[slot].key = k;
    g_tmap_cache[slot].map = m;
    return true;
}

static bool tma_side_ok(const void *base, uint64_t row_bytes, const uint64_t n[3], const int64_t stride[3]) {
    if (((uintptr_t)base & 15) || (row_bytes & 15) || row_bytes / 8 >= (1ull << 32)) return false;
    for (int i = 0; i < 3; ++i) {
        if (n[i] >= (1ull << 32)) return false;
        if (n[i] > 1 && (stride[i] <= 0 || (stride[i] & 15) || (uint64_t)stride[i] >= (1ull << 40))) return false;
    }
    return n[0] > 1 ? (uint64_t)stride[0] >= row_bytes : true;   // rows must not overlap (a box of by rows is one transfer)
}

enum StridedMode { SM_AUTO = 0, SM_TMA, SM_WIDE, SM_BULK, SM_VEC };
static StridedMode strided_mode() {
    static const StridedMode m = [] {
        const char *e = getenv("OCM_STRIDED_MODE");
        if (!e) return SM_AUTO;
        if (!strcmp(e, "tma")) return SM_TMA;
        if (!strcmp(e, "wide")) return SM_WIDE;
        if (!strcmp(e, "bulk")) return SM_BULK;
        if (!strcmp(e, "vec")) return SM_VEC;
        return SM_AUTO;
    }();
    return m;
}
static int env_int(const char *k, int dflt) {
    const char *e = getenv(k);
    return (e && *e) ? atoi(e) : dflt;
}

static cudaError_t launch_strided_tma(const StridedDesc &d, const RowMap &m, uint64_t rows, uint64_t *signal, unsigned *ticket,
                                      cudaStream_t stream, bool *done) {
    *done = false;
    const uint64_t n[3] = {m.n0, m.n1, m.n2};
    if (!tma_side_ok(d.src, d.row_bytes, n, d.src_stride) || !tma_side_ok(d.dst, d.row_bytes, n, d.dst_stride)) return cudaSuccess;
    static const int tile_target = env_int("OCM_TMA_TILE", 16384);
    static const int agents = std::min(std::max(env_int("OCM_TMA_AGENTS", 4), 1), kSAgents);
    static const int stages_env = std::min(std::max(env_int("OCM_TMA_STAGES", 3), 2), kSStages);
    TmaStridedArgs a;
    a.bx = (uint32_t)std::min<uint64_t>(d.row_bytes / 8, 256);
    a.by = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(std::min<uint64_t>(256, m.n0), (uint64_t)tile_target / (a.bx * 8ull)));
    a.tx = (uint32_t)((d.row_bytes / 8 + a.bx - 1) / a.bx);
    a.ty = (uint32_t)((m.n0 + a.by - 1) / a.by);
    a.n1 = (uint32_t)m.n1;
    a.n2 = (uint32_t)m.n2;
    a.tile_bytes = a.bx * 8u * a.by;
    a.pitch = (a.tile_bytes + 127u) & ~127u;
    a.n_tiles = (uint64_t)a.tx * a.ty * m.n1 * m.n2;
    int stages = stages_env;
    while ((size_t)agents * stages * a.pitch > (size_t)(200 << 10) && stages > 2) --stages;
    if ((size_t)agents * stages * a.pitch > (size_t)(200 << 10)) return cudaSuccess;
    a.stages = (uint32_t)stages;
    a.signal = signal;
    a.ticket = ticket;
    CUtensorMap sm, dm;
    if (!tmap_get(d.src, d.row_bytes, n, d.src_stride, a.bx, a.by, &sm) ||
        !tmap_get(d.dst, d.row_bytes, n, d.dst_stride, a.bx, a.by, &dm))
        return cudaSuccess;
    const size_t smem = (size_t)agents * stages * a.pitch;
    const int sms = num_sms();
    const int per_sm = smem <= (size_t)(100 << 10) ? 2 : 1;
    int ctas = (int)std::min<uint64_t>((a.n_tiles + agents - 1) / agents, (uint64_t)sms * per_sm);
    if (ctas > 1 && signal && !ticket) return cudaErrorInvalidValue;
    static bool attr[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 64 && !attr[dev]) {
        cudaError_t e = cudaFuncSetAttribute(ocm_strided_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 << 10);
        if (e != cudaSuccess) return e;
        attr[dev] = true;
    }
    (void)rows;
    ocm_strided_tma_kernel<<<ctas, agents * 32, smem, stream>>>(sm, dm, a);
    *done = true;
    return cudaGetLastError();
}

static bool all_aligned(const StridedDesc &d, uint64_t a_src, uint64_t a_dst) {
    uintptr_t s = (uintptr_t)d.src, t = (uintptr_t)d.dst;
    for (int i = 0; i < 3; ++i) {
        if (d.n[i] > 1) {
            s |= (uintptr_t)d.src_stride[i];
            t |= (uintptr_t)d.dst_stride[i];
        }
    }
    return (s & (a_src - 1)) == 0 && (t & (a_dst - 1)) == 0;
}

cudaError_t launch_strided(const StridedDesc &d, uint64_t *signal, unsigned *ticket, cudaStream_t stream,
                           unsigned long long *launches) {
    const uint64_t rows = std::max<uint64_t>(d.n[0], 1) * std::max<uint64_t>(d.n[1], 1) * std::max<uint64_t>(d.n[2], 1);
    if (!rows || !d.row_bytes) return cudaSuccess;
    RowMap m;
    m.n0 = std::max<uint64_t>(d.n[0], 1); m.n1 = std::max<uint64_t>(d.n[1], 1); m.n2 = std::max<uint64_t>(d.n[2], 1);
    m.ss0 = d.src_stride[0]; m.ss1 = d.src_stride[1]; m.ss2 = d.src_stride[2];
    m.ds0 = d.dst_stride[0]; m.ds1 = d.dst_stride[1]; m.ds2 = d.dst_stride[2];
    const int sms = num_sms();
    if (launches) ++*launches;

    if (d.src_dtype == d.dst_dtype) {
        StridedArgs a;
        a.src = static_cast<const uint8_t *>(d.src);
        a.dst = static_cast<uint8_t *>(d.dst);
        a.map = m; a.rows = rows; a.row_bytes = d.row_bytes; a.signal = signal; a.ticket = ticket;
        a.tile = a.stages = a.rows_per_batch = a.chunks_per_row = a.chunk_bytes = 0;
        const StridedMode mode = strided_mode();
        // wide vector rows: whole 32-byte (or 16-byte) units on both sides
        const int wide = (d.row_bytes % 32 == 0 && all_aligned(d, 32, 32)) ? 32 : (d.row_bytes % 16 == 0 && all_aligned(d, 16, 16)) ? 16 : 0;
        const bool tma_first = mode == SM_TMA || (mode == SM_AUTO && !wide);
        if (tma_first && d.row_bytes >= 64 && rows * d.row_bytes >= (64 << 10)) {
            bool done = false;
            cudaError_t e = launch_strided_tma(d, m, rows, signal, ticket, stream, &done);
            if (e != cudaSuccess || done) return e;
        }
        if (wide && (mode == SM_AUTO || mode == SM_WIDE || mode == SM_TMA)) {
            const uint64_t units = rows * (d.row_bytes / wide);
            // 4 units per thread per iteration; up to 8 CTAs per SM keep ~1 MiB per SM in flight
            int ctas = (int)std::min<uint64_t>(std::max<uint64_t>((units + 1023) / 1024, 1), (uint64_t)sms * 8);
            if (ctas > 1 && signal && !ticket) return cudaErrorInvalidValue;
            if (wide == 32) ocm_strided_wide_kernel<32, 4><<<ctas, 256, 0, stream>>>(a);
            else ocm_strided_wide_kernel<16, 4><<<ctas, 256, 0, stream>>>(a);
            return cudaGetLastError();
        }
        const bool bulk_ok = d.row_bytes % 16 == 0 && d.row_bytes >= 256 && all_aligned(d, 16, 16) &&
                             rows * d.row_bytes >= (64 << 10) && mode != SM_VEC;
        if (bulk_ok) {
            const uint32_t tile = 16384;
            int agents = 4, stages = 3;
            a.tile = tile; a.stages = stages;
            if (d.row_bytes <= tile) {
                a.chunks_per_row = 1;
                a.chunk_bytes = (uint32_t)d.row_bytes;
                a.rows_per_batch = (uint32_t)std::min<uint64_t>(32, tile / d.row_bytes);
            } else {
                a.chunk_bytes = tile;
                a.chunks_per_row = (uint32_t)((d.row_bytes + tile - 1) / tile);
                a.rows_per_batch = 1;
            }
            uint64_t n_batches = a.chunks_per_row > 1 ? rows * a.chunks_per_row
                                                      : (rows + a.rows_per_batch - 1) / a.rows_per_batch;
            int ctas = (int)std::min<uint64_t>((n_batches + agents - 1) / agents, (uint64_t)sms);
            if (ctas > 1 && signal && !ticket) return cudaErrorInvalidValue;
            static bool attr[64] = {false};
            int dev = 0;
            cudaGetDevice(&dev);
            if (dev < 64 && !attr[dev]) {
                cudaError_t e = cudaFuncSetAttribute(ocm_strided_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 << 10);
                if (e != cudaSuccess) return e;
                attr[dev] = true;
            }
            ocm_strided_bulk_kernel<<<ctas, agents * 32, (size_t)agents * stages * tile, stream>>>(a);
            return cudaGetLastError();
        }
        int width = 16;
        while (width > 1 && (!all_aligned(d, width, width) || d.row_bytes % width)) width >>= 1;
        if (width == 2) width = 1;
        uint64_t total = rows * (d.row_bytes / width);
        int ctas = (int)std::min<uint64_t>(std::max<uint64_t>((total + 1023) / 1024, 1), (uint64_t)sms * 8);
        if (ctas > 1 && signal && !ticket) return cudaErrorInvalidValue;
        switch (width) {
        case 16: ocm_strided_vec_kernel<uint4><<<ctas, 256, 0, stream>>>(a); break;
        case 8: ocm_strided_vec_kernel<uint2><<<ctas, 256, 0, stream>>>(a); break;
        case 4: ocm_strided_vec_kernel<uint32_t><<<ctas, 256, 0, stream>>>(a); break;
        default: ocm_strided_vec_kernel<uint8_t><<<ctas, 256, 0, stream>>>(a); break;
        }
        return cudaGetLastError();
    }

    // casts
    const size_t ss = dtype_size(d.src_dtype), ds = dtype_size(d.dst_dtype);
    if (d.src_dtype == DT_U8 || d.dst_dtype == DT_U8 || d.row_bytes % ss) return cudaErrorInvalidValue;
    const bool fp8 = d.src_dtype == DT_F8E4M3 || d.dst_dtype == DT_F8E4M3;
    if (fp8 && (!d.scales || (d.scale_block != 32 && d.scale_block != 128))) return cudaErrorInvalidValue;
    CastArgs c;
    c.src = static_cast<const uint8_t *>(d.src);
    c.dst = static_cast<uint8_t *>(d.dst);
    c.map = m; c.rows = rows; c.elems = d.row_bytes / ss;
    c.src_dtype = d.src_dtype; c.dst_dtype = d.dst_dtype;
    c.blk = fp8 ? (uint32_t)d.scale_block : 0;
    c.bpr = fp8 ? (uint32_t)((c.elems + c.blk - 1) / c.blk) : 0;
    c.scales = d.scales; c.signal = signal; c.ticket = ticket;
    // throughput path: whole 32-element chunks (whole 128-element blocks for block 128), 32-byte aligned rows
    if (c.elems % 32 == 0 && (c.blk != 128 || c.elems % 128 == 0) && all_aligned(d, 32, 32) && !getenv("OCM_CAST_NARROW")) {
        const uint64_t chunks = rows * (c.elems / 32);
        // 2 chunks per thread per iteration; 2 CTAs per SM are resident (register budget), a few waves deep
        int ctas = (int)std::min<uint64_t>(std::max<uint64_t>((chunks + 511) / 512, 1), (uint64_t)sms * 8);
        if (ctas > 1 && signal && !ticket) return cudaErrorInvalidValue;
        if (launch_cast32<2>(c, ctas, stream)) return cudaGetLastError();
    }
    const bool fast = c.elems % 8 == 0 && all_aligned(d, 8 * ss >= 16 ? 16 : 8 * ss, 8 * ds >= 16 ? 16 : 8 * ds);
    uint64_t work = fast ? rows * (fp8 ? (uint64_t)c.bpr * (c.blk / 8) : c.elems / 8)
                         : rows * ((c.elems + (c.blk ? c.blk : 8) - 1) / (c.blk ? c.blk : 8));
    // the fast kernel moves 4 chunks per thread per iteration: size the grid for that, up to 8 CTAs per SM
    int ctas = (int)std::min<uint64_t>(std::max<uint64_t>((work + (fast ? 1023 : 255)) / (fast ? 1024 : 256), 1), (uint64_t)sms * 8);
    if (ctas > 1 && signal && !ticket) return cudaErrorInvalidValue;
    if (fast) {
        const uint64_t cpr = fp8 ? (uint64_t)c.bpr * (c.blk / 8) : c.elems / 8;
        if ((cpr & (cpr - 1)) == 0) ocm_cast_kernel<4, true><<<ctas, 256, 0, stream>>>(c);
        else ocm_cast_kernel<4, false><<<ctas, 256, 0, stream>>>(c);
    } else {
        ocm_cast_scalar_kernel<<<ctas, 256, 0, stream>>>(c);
    }
    return cudaGetLastError();
}

} // namespace ocmk
