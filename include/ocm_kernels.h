/*
 * ocm_kernels.h -- C ABI of the sm_100a data-movement kernels in libocm.so.
 *
 * The ocm_* calls in oncillamem.h use these kernels internally on OCM
 * allocations; this header exposes them for raw device-usable pointers
 * (local HBM, peer HBM mapped into the process, or pinned+registered host
 * memory), e.g. for torch tensors through the Python binding.
 * All calls enqueue on `stream` (a cudaStream_t; NULL = default stream) of the
 * CURRENT device and return 0 or a negative enum ocm_err.  The reference has
 * no device code at all; citations name the host-side operation
 * each kernel replaces.
 */
#ifndef OCM_KERNELS_H_
#define OCM_KERNELS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* K1/K2/K3 contiguous put/get/copy (replaces ib_write/ib_read/ib_poll,
 * reference src/rdma.c:241-302, and extoll_write/read, src/extoll.c:286-310).
 * dst2 may be NULL.  signal may be NULL; otherwise a device-visible uint64
 * that receives +1 (release, system scope) after the last byte is visible. */
int ocmk_copy(void *dst, void *dst2, const void *src, size_t bytes, uint64_t *signal, void *stream);
/* Same with explicit tuning: mode 0 auto / 1 bulk(TMA) / 2 vector; 0 = default for the others. */
int ocmk_copy_tuned(void *dst, void *dst2, const void *src, size_t bytes, uint64_t *signal, void *stream, int mode,
                    int ctas, int agents, int stages, int tile_bytes);

/* K9 fill / verify / checksum (replaces calloc/memset zero-fill, reference
 * src/alloc.c:171, and the pattern checks of test/ib_client.c:144-188).
 * out_* point to device-visible uint64 accumulators that the caller zeroes. */
int ocmk_fill(void *dst, size_t bytes, uint64_t pattern, void *stream);
int ocmk_verify(const void *p, size_t bytes, uint64_t pattern, uint64_t *out_mismatches, void *stream);
int ocmk_checksum(const void *p, size_t bytes, uint64_t *out_sum, void *stream);
int ocmk_signal(uint64_t *signal, void *stream);

/* K4/K5 strided pack/unpack with optional cast (generalises the *_offset_2
 * two-hop staging of reference src/lib.c:526-547).  dtypes: 0 u8, 1 f32,
 * 2 bf16, 3 f8e4m3 (block-scaled, scales = one fp32 per scale_block elements
 * per row, row-major). */
struct ocmk_strided {
    const void *src;
    void       *dst;
    uint64_t    row_bytes;      /* contiguous bytes per row, source side */
    uint64_t    n[3];           /* outer extents (>= 1) */
    int64_t     src_stride[3];  /* bytes */
    int64_t     dst_stride[3];  /* bytes */
    int         src_dtype, dst_dtype;
    int         scale_block;    /* 32 or 128 */
    float      *scales;
};
int ocmk_strided_copy(const struct ocmk_strided *d, uint64_t *signal, void *stream);

/* K6 replicated put: one source to n_dst destinations (one CTA of a cluster per destination, TMA bulk tiles). */
int ocmk_put_replicated(void *const *dsts, int n_dst, const void *src, size_t bytes, void *stream);

/* K7 GUPS over a table striped across n_ranks slabs. */
int ocmk_gups(uint64_t *const *bases, int n_ranks, uint64_t words_per_stripe, uint64_t total_words,
              uint64_t n_updates, uint64_t seed, void *stream);
uint64_t ocmk_gups_threads(uint64_t n_updates);

/* K8 device radix sort of 64-bit keys (in place; tmp has room for n keys). */
int ocmk_sort_u64(uint64_t *keys, uint64_t *tmp, uint64_t n, void *stream);
/* Same sort for callers that must not allocate device memory: applications whose cluster daemons
 * have claimed nearly all HBM for their pools (a cudaMalloc for scratch could fail), and callers
 * inside latency-critical sections where cudaMalloc/cudaFree could stall the device.  The scratch is
 * a fixed 388 KiB array inside the library, shared by every call on a GPU: at most
 * OCMK_SORT_NOALLOC_MAX_KEYS keys per call (more is refused with -OCM_ERANGE), calls that overlap on
 * one GPU race on the scratch, and the first call on a device (which looks up the array's device
 * address) should happen before any stream capture begins.  It has no counterpart in the reference. */
#define OCMK_SORT_NOALLOC_MAX_KEYS (1ull << 24)
int ocmk_sort_u64_noalloc(uint64_t *keys, uint64_t *tmp, uint64_t n, void *stream);

/* K8 merge: out = stable merge of two ascending unsigned 64-bit sequences (merge-path tiles).  ocmk_merge_runs_u64
 * merges k runs laid end to end (run i = keys[offsets[i] .. offsets[i+1]), offsets on the host) by ceil(log2 k)
 * levels of pairwise merges between keys and tmp; it returns 0 if the result ends in keys, 1 if in tmp. */
int ocmk_merge_u64(const uint64_t *a, uint64_t na, const uint64_t *b, uint64_t nb, uint64_t *out, void *stream);
int ocmk_merge_runs_u64(uint64_t *keys, uint64_t *tmp, const uint64_t *offsets, int k, void *stream);

/* number of kernels launched through this ABI and the ocm_* calls so far */
uint64_t ocmk_launch_count(void);

/* CPU oracles (plain C++ restatements, src/oracle/) */
void     ocm_oracle_copy(void *dst, void *dst2, const void *src, size_t bytes);
void     ocm_oracle_fill(void *dst, size_t bytes, uint64_t pattern);
uint64_t ocm_oracle_verify(const void *p, size_t bytes, uint64_t pattern);
uint64_t ocm_oracle_checksum(const void *p, size_t bytes);
int      ocm_oracle_strided_copy(const struct ocmk_strided *d);
void     ocm_oracle_gups(uint64_t *table, uint64_t total_words, uint64_t n_updates, uint64_t seed, uint64_t n_threads);
void     ocm_oracle_sort_u64(uint64_t *keys, size_t n);

#ifdef __cplusplus
}
#endif
#endif /* OCM_KERNELS_H_ */
