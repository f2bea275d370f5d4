/* oracle_quants.c — TEST INFRASTRUCTURE ONLY. CPU restatement of the reference block
 * quantizers/dequantizers (reference ggml/src/ggml-quants.c, ggml-common.h, ggml-impl.h).
 * Pinned bit-for-bit against oracle/_ref/libggml-base.so by tests/test_oracle_pin.py. */
#include "oracle.h"

#include <math.h>
#include <string.h>

/* ---- fp16 (IEEE binary16, round-to-nearest-even), ggml-impl.h GGML_FP32_TO_FP16 semantics ---- */
float orc_fp16_to_fp32(uint16_t h) {
    const uint32_t sign = (uint32_t) (h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1f, man = h & 0x3ffu, bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else {
            int e = -1;
            do { man <<= 1; ++e; } while (!(man & 0x400u));
            bits = sign | ((uint32_t) (127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
        }
    } else if (exp == 31) bits = sign | 0x7f800000u | (man << 13);
    else bits = sign | ((exp + 112) << 23) | (man << 13);
    float f; memcpy(&f, &bits, 4); return f;
}

uint16_t orc_fp32_to_fp16(float f) {
    uint32_t x; memcpy(&x, &f, 4);
    const uint16_t sign = (uint16_t) ((x >> 16) & 0x8000u);
    const uint32_t absx = x & 0x7fffffffu;
    if (absx >= 0x7f800000u) return (uint16_t) (sign | 0x7c00u | (absx > 0x7f800000u ? 0x200u : 0));
    if (absx >= 0x477ff000u) return (uint16_t) (sign | 0x7c00u);           /* rounds to inf */
    if (absx < 0x33000001u) return sign;                                    /* rounds to zero (< 2^-25 or == 2^-25 tie-to-even) */
    int32_t e = (int32_t) (absx >> 23) - 127;
    uint32_t m = (absx & 0x7fffffu) | 0x800000u;
    int shift; uint32_t base;
    if (e < -14) { shift = 13 + (-14 - e); base = 0; }                      /* subnormal half */
    else         { shift = 13;             base = (uint32_t) (e + 15) << 10; m &= 0x7fffffu; }
    uint32_t r = m >> shift;
    const uint32_t rem = m & ((1u << shift) - 1), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (r & 1))) ++r;                        /* RNE; carry into exponent is correct */
    return (uint16_t) (sign | (base + r));
}

float orc_e8m0_to_fp32_half(uint8_t e) {   /* ggml-impl.h:467-488 */
    uint32_t bits = e < 2 ? (0x00200000u << e) : ((uint32_t) (e - 1) << 23);
    float f; memcpy(&f, &bits, 4); return f;
}

size_t orc_row_size(int type, int64_t k) {
    switch (type) {
        case ORC_F32: return (size_t) k * 4;
        case ORC_F16: return (size_t) k * 2;
        case ORC_Q4_0: return (size_t) (k / 32) * 18;
        case ORC_Q8_0: return (size_t) (k / 32) * 34;
        case ORC_Q8_1: return (size_t) (k / 32) * 36;
        case ORC_MXFP4: return (size_t) (k / 32) * 17;
        case ORC_Q4_K: return (size_t) (k / 256) * 144;
        case ORC_Q6_K: return (size_t) (k / 256) * 210;
        default: return 0;
    }
}

static void put16(uint8_t * p, uint16_t v) { p[0] = (uint8_t) v; p[1] = (uint8_t) (v >> 8); }
static uint16_t get16(const uint8_t * p) { return (uint16_t) (p[0] | (p[1] << 8)); }

/* ---- quantizers ---- */
void orc_quantize_row_q4_0(const float * x, void * vy, int64_t k) {   /* ggml-quants.c:35-64 */
    uint8_t * y = (uint8_t *) vy;
    for (int64_t i = 0; i < k / 32; ++i, y += 18) {
        float amax = 0.0f, max = 0.0f;
        for (int j = 0; j < 32; ++j) { const float v = x[i * 32 + j]; if (amax < fabsf(v)) { amax = fabsf(v); max = v; } }
        const float d = max / -8, id = d ? 1.0f / d : 0.0f;
        put16(y, orc_fp32_to_fp16(d));
        for (int j = 0; j < 16; ++j) {
            const float x0 = x[i * 32 + j] * id, x1 = x[i * 32 + 16 + j] * id;
            int a = (int8_t) (x0 + 8.5f), b = (int8_t) (x1 + 8.5f);
            const uint8_t xi0 = (uint8_t) (a < 15 ? a : 15), xi1 = (uint8_t) (b < 15 ? b : 15);
            y[2 + j] = (uint8_t) (xi0 | (xi1 << 4));
        }
    }
}

void orc_quantize_row_q8_0(const float * x, void * vy, int64_t k) {   /* ggml-quants.c:197-223 */
    uint8_t * y = (uint8_t *) vy;
    for (int64_t i = 0; i < k / 32; ++i, y += 34) {
        float amax = 0.0f;
        for (int j = 0; j < 32; ++j) { const float v = fabsf(x[i * 32 + j]); amax = amax > v ? amax : v; }
        const float d = amax / 127, id = d ? 1.0f / d : 0.0f;
        put16(y, orc_fp32_to_fp16(d));
        for (int j = 0; j < 32; ++j) y[2 + j] = (uint8_t) (int8_t) roundf(x[i * 32 + j] * id);
    }
}

void orc_quantize_row_q8_1(const float * x, void * vy, int64_t k) {   /* ggml-quants.c:225-258 */
    uint8_t * y = (uint8_t *) vy;
    for (int64_t i = 0; i < k / 32; ++i, y += 36) {
        float amax = 0.0f;
        for (int j = 0; j < 32; ++j) { const float v = fabsf(x[i * 32 + j]); amax = amax > v ? amax : v; }
        const float d = amax / 127, id = d ? 1.0f / d : 0.0f;
        put16(y, orc_fp32_to_fp16(d));
        int sum = 0;
        for (int j = 0; j < 32; ++j) { const int8_t q = (int8_t) roundf(x[i * 32 + j] * id); y[4 + j] = (uint8_t) q; sum += q; }
        put16(y + 2, orc_fp32_to_fp16(sum * d));
    }
}

static const int8_t kvalues_mxfp4[16] = {0, 1, 2, 3, 4, 6, 8, 12, 0, -1, -2, -3, -4, -6, -8, -12};   /* ggml-common.h:1094 */

static int best_index_mxfp4(float x, float e) {   /* ggml-quants.c:260-271 */
    int best = 0; float best_err = fabsf(kvalues_mxfp4[0] * e - x);
    for (int i = 1; i < 16; ++i) { const float err = fabsf(kvalues_mxfp4[i] * e - x); if (err < best_err) { best = i; best_err = err; } }
    return best;
}

void orc_quantize_row_mxfp4(const float * x, void * vy, int64_t k) {   /* ggml-quants.c:273-305 */
    uint8_t * y = (uint8_t *) vy;
    for (int64_t i = 0; i < k / 32; ++i, y += 17) {
        float amax = 0.0f;
        for (int j = 0; j < 32; ++j) { const float v = fabsf(x[i * 32 + j]); if (amax < v) amax = v; }
        const uint8_t e = amax > 0.0f ? (uint8_t) (floorf(log2f(amax)) - 2 + 127) : 0;
        const float d = orc_e8m0_to_fp32_half(e);
        y[0] = e;
        for (int j = 0; j < 16; ++j)
            y[1 + j] = (uint8_t) (best_index_mxfp4(x[i * 32 + j], d) | (best_index_mxfp4(x[i * 32 + 16 + j], d) << 4));
    }
}

/* ---- dequantizers ---- */
static void scale_min_k4(int j, const uint8_t * q, uint8_t * d, uint8_t * m) {   /* ggml-quants.c:703-711 */
    if (j < 4) { *d = q[j] & 63; *m = q[j + 4] & 63; }
    else { *d = (uint8_t) ((q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4)); *m = (uint8_t) ((q[j + 4] >> 4) | ((q[j] >> 6) << 4)); }
}

int orc_dequantize_row(int type, const void * vx, float * y, int64_t k) {
    const uint8_t * x = (const uint8_t *) vx;
    switch (type) {
        case ORC_F32: memcpy(y, x, (size_t) k * 4); return 0;
        case ORC_F16: for (int64_t i = 0; i < k; ++i) y[i] = orc_fp16_to_fp32(get16(x + 2 * i)); return 0;
        case ORC_Q4_0:
            for (int64_t i = 0; i < k / 32; ++i, x += 18) {
                const float d = orc_fp16_to_fp32(get16(x));
                for (int j = 0; j < 16; ++j) {
                    y[i * 32 + j] = ((x[2 + j] & 0x0F) - 8) * d;
                    y[i * 32 + j + 16] = ((x[2 + j] >> 4) - 8) * d;
                }
            }
            return 0;
        case ORC_Q8_0:
            for (int64_t i = 0; i < k / 32; ++i, x += 34) {
                const float d = orc_fp16_to_fp32(get16(x));
                for (int j = 0; j < 32; ++j) y[i * 32 + j] = (int8_t) x[2 + j] * d;
            }
            return 0;
        case ORC_MXFP4:
            for (int64_t i = 0; i < k / 32; ++i, x += 17) {
                const float d = orc_e8m0_to_fp32_half(x[0]);
                for (int j = 0; j < 16; ++j) {
                    y[i * 32 + j] = kvalues_mxfp4[x[1 + j] & 0x0F] * d;
                    y[i * 32 + j + 16] = kvalues_mxfp4[x[1 + j] >> 4] * d;
                }
            }
            return 0;
        case ORC_Q4_K:
            for (int64_t i = 0; i < k / 256; ++i, x += 144) {
                const float d = orc_fp16_to_fp32(get16(x)), min = orc_fp16_to_fp32(get16(x + 2));
                const uint8_t * sc = x + 4, * q = x + 16;
                int is = 0; uint8_t s, m;
                for (int j = 0; j < 256; j += 64) {
                    scale_min_k4(is + 0, sc, &s, &m); const float d1 = d * s, m1 = min * m;
                    scale_min_k4(is + 1, sc, &s, &m); const float d2 = d * s, m2 = min * m;
                    for (int l = 0; l < 32; ++l) *y++ = d1 * (q[l] & 0xF) - m1;
                    for (int l = 0; l < 32; ++l) *y++ = d2 * (q[l] >> 4) - m2;
                    q += 32; is += 2;
                }
            }
            return 0;
        case ORC_Q6_K:
            for (int64_t i = 0; i < k / 256; ++i, x += 210) {
                const float d = orc_fp16_to_fp32(get16(x + 208));
                const uint8_t * ql = x, * qh = x + 128; const int8_t * sc = (const int8_t *) (x + 192);
                for (int n = 0; n < 256; n += 128) {
                    for (int l = 0; l < 32; ++l) {
                        const int is = l / 16;
                        const int8_t q1 = (int8_t) ((ql[l] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                        const int8_t q2 = (int8_t) ((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                        const int8_t q3 = (int8_t) ((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                        const int8_t q4 = (int8_t) ((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                        y[l] = d * sc[is + 0] * q1; y[l + 32] = d * sc[is + 2] * q2;
                        y[l + 64] = d * sc[is + 4] * q3; y[l + 96] = d * sc[is + 6] * q4;
                    }
                    y += 128; ql += 64; qh += 32; sc += 8;
                }
            }
            return 0;
        default: return -1;
    }
}
