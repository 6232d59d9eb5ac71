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
#include <cuda_bf16.h>
#include <cuda_fp8.h>

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

// One thread = 8 consecutive elements.  With quantisation the G = blk/8 lanes of
// an aligned lane group share one scale block.
__global__ void __launch_bounds__(256) ocm_cast_kernel(CastArgs a) {
    const bool quant = a.dst_dtype == DT_F8E4M3 && a.src_dtype != DT_F8E4M3;
    const bool dequant = a.src_dtype == DT_F8E4M3 && a.dst_dtype != DT_F8E4M3;
    const uint32_t G = a.blk ? a.blk / 8 : 1;                 // 1, 4 or 16: always a power of two
    const uint32_t gshift = G == 16 ? 4 : G == 4 ? 2 : 0;
    const uint64_t cpr = a.blk ? (uint64_t)a.bpr * G : (a.elems + 7) / 8;   // chunks per row, padded to whole blocks
    const uint64_t total = a.rows * cpr;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const size_t ss = dtype_size(a.src_dtype), ds = dtype_size(a.dst_dtype);
    // all lanes of a warp iterate together (shuffles below)
    const uint64_t total_pad = (total + 31) & ~31ull;
    for (uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; c < total_pad; c += stride) {
        uint64_t row, ch;
        divmod(c, cpr, &row, &ch);
        const uint64_t e0 = ch * 8;
        const bool valid = c < total && e0 < a.elems;
        float v[8];
        int64_t so = 0, dof = 0;
        if (valid) {
            row_offsets(a.map, row, &so, &dof);
            load8(a.src + so + (int64_t)(e0 * ss), a.src_dtype, v);
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = 0.f;
        }
        if (quant) {
            float amax = 0.f;
#pragma unroll
            for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(v[i]));
            for (uint32_t o = G >> 1; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
            // scale = amax/448 and its reciprocal: two IEEE divisions per thread, then multiplies
            // (the oracle performs the same operations in the same order, so results match bit for bit)
            const float scale = amax > 1e-30f ? __fdiv_rn(amax, 448.0f) : 1.0f;   // (near-)zero block: scale 1
            const float inv = __fdiv_rn(1.0f, scale);
            if (valid) {
                if ((ch & (G - 1)) == 0) a.scales[row * a.bpr + (ch >> gshift)] = scale;
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = __fmul_rn(v[i], inv);
            }
        } else if (dequant && valid) {
            const float scale = a.scales[row * a.bpr + (ch >> gshift)];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = __fmul_rn(v[i], scale);
        }
        if (valid) store8(a.dst + dof + (int64_t)(e0 * ds), a.dst_dtype, v);
    }
    __syncthreads();
    if (threadIdx.x == 0) finish_and_signal(a.signal, a.ticket);
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
        const bool bulk_ok = d.row_bytes % 16 == 0 && d.row_bytes >= 256 && all_aligned(d, 16, 16) &&
                             rows * d.row_bytes >= (64 << 10);
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
    const bool fast = c.elems % 8 == 0 && all_aligned(d, 8 * ss >= 16 ? 16 : 8 * ss, 8 * ds >= 16 ? 16 : 8 * ds);
    uint64_t work = fast ? rows * (fp8 ? (uint64_t)c.bpr * (c.blk / 8) : c.elems / 8)
                         : rows * ((c.elems + (c.blk ? c.blk : 8) - 1) / (c.blk ? c.blk : 8));
    int ctas = (int)std::min<uint64_t>(std::max<uint64_t>((work + 255) / 256, 1), (uint64_t)sms * 8);
    if (ctas > 1 && signal && !ticket) return cudaErrorInvalidValue;
    if (fast) ocm_cast_kernel<<<ctas, 256, 0, stream>>>(c);
    else ocm_cast_scalar_kernel<<<ctas, 256, 0, stream>>>(c);
    return cudaGetLastError();
}

} // namespace ocmk
