// Plain C++ restatements of every device kernel: the test oracle
// and the timed single-thread CPU baseline.  For K1/K2 the oracle is the
// reference's own semantics - memcpy with the same offsets (reference
// src/lib.c:523); the reference has no CPU or GPU code for the other kernels.
#pragma once
#include <cstddef>
#include <cstdint>

extern "C" {

// K1/K2/K3
void ocm_oracle_copy(void *dst, void *dst2, const void *src, size_t bytes);
// K9: byte k of the buffer holds byte (k & 7) of the little-endian pattern
void ocm_oracle_fill(void *dst, size_t bytes, uint64_t pattern);
uint64_t ocm_oracle_verify(const void *p, size_t bytes, uint64_t pattern);   // mismatching bytes
uint64_t ocm_oracle_checksum(const void *p, size_t bytes);                   // wrapping sum of LE u64 words

// K4/K5: descriptor mirrors ocmk::StridedDesc (kernels.h)
struct ocm_oracle_strided {
    const void *src;
    void       *dst;
    uint64_t    row_bytes;
    uint64_t    n[3];
    int64_t     src_stride[3];
    int64_t     dst_stride[3];
    int         src_dtype, dst_dtype;   // 0 u8, 1 f32, 2 bf16, 3 f8e4m3
    int         scale_block;
    float      *scales;
};
int ocm_oracle_strided_copy(const struct ocm_oracle_strided *d);

// scalar conversions used by the cast kernels (round-to-nearest-even, saturating)
uint16_t ocm_oracle_f32_to_bf16(float f);
float    ocm_oracle_bf16_to_f32(uint16_t h);
uint8_t  ocm_oracle_f32_to_e4m3(float f);
float    ocm_oracle_e4m3_to_f32(uint8_t v);

// K7: apply n_updates xor updates of the GUPS stream to a flat table
void ocm_oracle_gups(uint64_t *table, uint64_t total_words, uint64_t n_updates, uint64_t seed, uint64_t n_threads);
uint64_t ocm_oracle_gups_value(uint64_t seed, uint64_t thread, uint64_t step);

// K8: reference ordering for the sort pipeline
void ocm_oracle_sort_u64(uint64_t *keys, size_t n);
}
