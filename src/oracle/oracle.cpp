#include "oracle.hpp"
#include <algorithm>
#include <cmath>
#include <cstring>

extern "C" {

void ocm_oracle_copy(void *dst, void *dst2, const void *src, size_t bytes) {
    memcpy(dst, src, bytes);
    if (dst2) memcpy(dst2, src, bytes);
}

void ocm_oracle_fill(void *dst, size_t bytes, uint64_t pattern) {
    uint8_t *d = static_cast<uint8_t *>(dst);
    for (size_t k = 0; k < bytes; ++k) d[k] = (uint8_t)(pattern >> (8 * (k & 7)));
}

uint64_t ocm_oracle_verify(const void *p, size_t bytes, uint64_t pattern) {
    const uint8_t *d = static_cast<const uint8_t *>(p);
    uint64_t bad = 0;
    for (size_t k = 0; k < bytes; ++k) bad += d[k] != (uint8_t)(pattern >> (8 * (k & 7)));
    return bad;
}

uint64_t ocm_oracle_checksum(const void *p, size_t bytes) {
    const uint8_t *d = static_cast<const uint8_t *>(p);
    uint64_t sum = 0;
    size_t k = 0;
    for (; k + 8 <= bytes; k += 8) {
        uint64_t w;
        memcpy(&w, d + k, 8);
        sum += w;
    }
    for (; k < bytes; ++k) sum += (uint64_t)d[k] << (8 * (k & 7));
    return sum;
}

// ---- scalar conversions ------------------------------------------------------
uint16_t ocm_oracle_f32_to_bf16(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)0x7fff;          // NaN (cvt.rn.bf16.f32 canonical NaN)
    uint32_t lsb = (u >> 16) & 1u;
    u += 0x7fffu + lsb;                                                     // round to nearest even
    return (uint16_t)(u >> 16);
}

float ocm_oracle_bf16_to_f32(uint16_t h) {
    uint32_t u = (uint32_t)h << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

// e4m3 (fn): 1-4-3, bias 7, no infinities, max 448, 0x7f/0xff = NaN.
float ocm_oracle_e4m3_to_f32(uint8_t v) {
    int sign = v >> 7, exp = (v >> 3) & 0xf, man = v & 7;
    float r;
    if (exp == 0xf && man == 7) r = NAN;
    else if (exp == 0) r = ldexpf((float)man, -9);                          // subnormal: man * 2^-3 * 2^-6
    else r = ldexpf(1.0f + (float)man / 8.0f, exp - 7);
    return sign ? -r : r;
}

uint8_t ocm_oracle_f32_to_e4m3(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    uint8_t sign = (uint8_t)((u >> 24) & 0x80);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint8_t)(sign | 0x7f);    // NaN
    float a = fabsf(f);
    if (a >= 448.0f) return (uint8_t)(sign | 0x7e);                         // satfinite (also +-inf)
    if (a == 0.0f) return sign;
    // quantum of the target binade: subnormals share 2^-9
    int e;
    frexpf(a, &e);                       // a = m * 2^e, m in [0.5,1) -> unbiased exponent e-1
    int ue = e - 1;
    if (ue < -6) ue = -6;
    float q = ldexpf(1.0f, ue - 3);      // spacing of representable values
    float n = a / q;                     // exact (power-of-two scaling)
    float r = nearbyintf(n);             // default rounding mode: nearest even
    float val = r * q;
    if (val >= 448.0f) return (uint8_t)(sign | 0x7e);
    if (val == 0.0f) return sign;
    int e2;
    float m2 = frexpf(val, &e2);         // val = m2 * 2^e2
    int ue2 = e2 - 1;
    uint8_t out;
    if (ue2 < -6) {
        out = (uint8_t)lrintf(val * 512.0f);                                // subnormal mantissa = val / 2^-9
    } else {
        int man = (int)lrintf((m2 * 2.0f - 1.0f) * 8.0f);
        out = (uint8_t)(((ue2 + 7) << 3) | (man & 7));
    }
    return (uint8_t)(sign | out);
}

// ---- K4/K5 ---------------------------------------------------------------------
static size_t dsize(int dt) { return dt == 1 ? 4 : dt == 2 ? 2 : 1; }

static float load_elem(const uint8_t *p, int dt) {
    if (dt == 1) {
        float f;
        memcpy(&f, p, 4);
        return f;
    }
    if (dt == 2) {
        uint16_t h;
        memcpy(&h, p, 2);
        return ocm_oracle_bf16_to_f32(h);
    }
    return 0.f;
}
static void store_elem(uint8_t *p, int dt, float f) {
    if (dt == 1) memcpy(p, &f, 4);
    else if (dt == 2) {
        uint16_t h = ocm_oracle_f32_to_bf16(f);
        memcpy(p, &h, 2);
    }
}

int ocm_oracle_strided_copy(const struct ocm_oracle_strided *d) {
    const int sd = d->src_dtype, dd = d->dst_dtype;
    if (d->row_bytes % dsize(sd)) return -1;
    const uint64_t elems = d->row_bytes / dsize(sd);
    const bool quant = dd == 3 && sd != 3, dequant = sd == 3 && dd != 3;
    const uint64_t blk = (quant || dequant) ? (uint64_t)d->scale_block : 0;
    if ((quant || dequant) && (blk == 0 || !d->scales)) return -1;
    if ((quant && sd == 0) || (dequant && dd == 0)) return -1;
    if (!quant && !dequant && sd != dd && (sd == 0 || dd == 0 || sd == 3 || dd == 3)) return -1;
    const uint64_t bpr = blk ? (elems + blk - 1) / blk : 0;
    uint64_t row = 0;
    for (uint64_t i3 = 0; i3 < d->n[2]; ++i3)
        for (uint64_t i2 = 0; i2 < d->n[1]; ++i2)
            for (uint64_t i1 = 0; i1 < d->n[0]; ++i1, ++row) {
                const uint8_t *s = (const uint8_t *)d->src + (int64_t)i1 * d->src_stride[0] +
                                   (int64_t)i2 * d->src_stride[1] + (int64_t)i3 * d->src_stride[2];
                uint8_t *t = (uint8_t *)d->dst + (int64_t)i1 * d->dst_stride[0] + (int64_t)i2 * d->dst_stride[1] +
                             (int64_t)i3 * d->dst_stride[2];
                if (sd == dd) {
                    memcpy(t, s, d->row_bytes);
                } else if (quant) {
                    for (uint64_t b = 0; b < bpr; ++b) {
                        uint64_t lo = b * blk, hi = std::min(elems, lo + blk);
                        float amax = 0.f;
                        for (uint64_t e = lo; e < hi; ++e) amax = fmaxf(amax, fabsf(load_elem(s + e * dsize(sd), sd)));
                        float scale = amax > 1e-30f ? amax / 448.0f : 1.0f;   // (near-)zero block: scale 1
                        d->scales[row * bpr + b] = scale;
                        const float inv = 1.0f / scale;   // quantise by multiplying with the reciprocal (as the kernel)
                        for (uint64_t e = lo; e < hi; ++e)
                            t[e] = ocm_oracle_f32_to_e4m3(load_elem(s + e * dsize(sd), sd) * inv);
                    }
                } else if (dequant) {
                    for (uint64_t e = 0; e < elems; ++e) {
                        float scale = d->scales[row * bpr + e / blk];
                        store_elem(t + e * dsize(dd), dd, ocm_oracle_e4m3_to_f32(s[e]) * scale);
                    }
                } else {
                    for (uint64_t e = 0; e < elems; ++e) store_elem(t + e * dsize(dd), dd, load_elem(s + e * dsize(sd), sd));
                }
            }
    return 0;
}

// ---- K7 ----------------------------------------------------------------------------
static inline uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
uint64_t ocm_oracle_gups_value(uint64_t seed, uint64_t thread, uint64_t step) {
    return mix64(seed + thread * 0x9E3779B97F4A7C15ull + step * 0xD1B54A32D192ED03ull);
}
void ocm_oracle_gups(uint64_t *table, uint64_t total_words, uint64_t n_updates, uint64_t seed, uint64_t n_threads) {
    for (uint64_t u = 0; u < n_updates; ++u) {
        uint64_t v = ocm_oracle_gups_value(seed, u % n_threads, u / n_threads);
        table[v & (total_words - 1)] ^= v;
    }
}

void ocm_oracle_sort_u64(uint64_t *keys, size_t n) { std::sort(keys, keys + n); }

} // extern "C"
