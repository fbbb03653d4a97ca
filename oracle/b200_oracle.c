/*
 * b200_oracle.c — CPU ORACLE, TEST INFRASTRUCTURE ONLY (see b200_oracle.h).
 *
 * Plain-C restatement of the reference's CPU definitions of truth for the hot path.  Written from
 * the block-format definitions; each function cites the reference lines it follows.  Compile with
 * -ffp-contract=off: the reference's dequantizers live in libggml-base, which is built without
 * -march flags, so `d1 * q - m1` is two roundings (mul then sub), never an FMA.
 *
 * Pinned bit-for-bit against oracle/_ref (the unmodified reference compiled here) by
 * tests/test_oracle_pinning.py and against tests/golden/ fixtures generated from the reference.
 */
#include "b200_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ---------------------------------------------------------------- block layouts
 * Byte-level views of src/ggml-common.h:170-175 (q4_0), 219-224 (q8_0), 226-237 (q8_1),
 * 295-306 (q4_K), 312-324 (q5_K), 330-336 (q6_K).  We address fields by byte offset so that no
 * struct padding assumptions are involved. */
enum {
    Q4_0_BYTES = 18,  /* f16 d | u8 qs[16]                                   */
    Q8_0_BYTES = 34,  /* f16 d | i8 qs[32]                                   */
    Q8_1_BYTES = 36,  /* f16 d | f16 s | i8 qs[32]                           */
    Q4_K_BYTES = 144, /* f16 d | f16 dmin | u8 scales[12] | u8 qs[128]       */
    Q5_K_BYTES = 176, /* f16 d | f16 dmin | u8 scales[12] | u8 qh[32] | u8 qs[128] */
    Q6_K_BYTES = 210, /* u8 ql[128] | u8 qh[64] | i8 scales[16] | f16 d      */
    /* the formats only the general kernels serve (src/ggml-common.h:178-187, 198-202, 206-216, 416-419) */
    Q4_1_BYTES = 20,  /* f16 d | f16 m | u8 qs[16]                           */
    Q5_0_BYTES = 22,  /* f16 d | u8 qh[4] | u8 qs[16]                        */
    Q5_1_BYTES = 24,  /* f16 d | f16 m | u8 qh[4] | u8 qs[16]                */
    IQ4_NL_BYTES = 18, /* f16 d | u8 qs[16]                                  */
};

int b200o_block_elems(int type) {
    switch (type) {
        case B200O_F32: case B200O_F16: return 1;
        case B200O_Q4_0: case B200O_Q8_0: case B200O_Q8_1: return 32;
        case B200O_Q4_1: case B200O_Q5_0: case B200O_Q5_1: case B200O_IQ4_NL: return 32;
        case B200O_Q4_K: case B200O_Q5_K: case B200O_Q6_K: return 256;
        default: return 0;
    }
}

size_t b200o_block_bytes(int type) {
    switch (type) {
        case B200O_F32:  return 4;
        case B200O_F16:  return 2;
        case B200O_Q4_0: return Q4_0_BYTES;
        case B200O_Q8_0: return Q8_0_BYTES;
        case B200O_Q8_1: return Q8_1_BYTES;
        case B200O_Q4_K: return Q4_K_BYTES;
        case B200O_Q5_K: return Q5_K_BYTES;
        case B200O_Q6_K: return Q6_K_BYTES;
        case B200O_Q4_1: return Q4_1_BYTES;
        case B200O_Q5_0: return Q5_0_BYTES;
        case B200O_Q5_1: return Q5_1_BYTES;
        case B200O_IQ4_NL: return IQ4_NL_BYTES;
        default: return 0;
    }
}

size_t b200o_row_bytes(int type, int64_t k) {
    return (size_t)(k / b200o_block_elems(type)) * b200o_block_bytes(type);
}

/* ---------------------------------------------------------------- fp16
 * IEEE-754 binary16 <-> binary32.  The reference (src/ggml-impl.h:358-405) uses a branch-free
 * bit trick; the values are the standard ones: exact widening, round-to-nearest-even narrowing,
 * overflow -> inf, NaN -> sign | 0x7E00. */
static uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static float    u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

float b200o_fp16_to_fp32(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    const uint32_t exp  = (h >> 10) & 0x1Fu;
    const uint32_t man  = h & 0x3FFu;
    if (exp == 0) {
        /* zero / subnormal: value = man * 2^-24, exactly representable in f32 */
        const float v = (float) man * 0x1.0p-24f;
        return u2f(sign | f2u(v));
    }
    if (exp == 31) {
        return u2f(sign | 0x7F800000u | (man << 13)); /* inf / nan */
    }
    return u2f(sign | ((exp + 112u) << 23) | (man << 13));
}

uint16_t b200o_fp32_to_fp16(float f) {
    const uint32_t w    = f2u(f);
    const uint16_t sign = (uint16_t)((w >> 16) & 0x8000u);
    const uint32_t absw = w & 0x7FFFFFFFu;
    if (absw > 0x7F800000u) {
        return (uint16_t)(sign | 0x7E00u);            /* NaN (src/ggml-impl.h:404) */
    }
    if (absw >= 0x47800000u) {                        /* >= 65536 (incl. inf) -> inf; 65520..65536 handled below */
        return (uint16_t)(sign | 0x7C00u);
    }
    if (absw >= 0x38800000u) {                        /* normal half range: 2^-14 .. */
        /* drop 13 mantissa bits with round-to-nearest-even; a carry may bump the exponent,
         * and from the top binade rounds up to inf (0x7C00) correctly */
        const uint32_t rebias = absw - 0x38000000u;   /* exponent - 112 */
        const uint32_t lsb    = (rebias >> 13) & 1u;
        const uint32_t r      = rebias + 0x0FFFu + lsb;
        return (uint16_t)(sign | (r >> 13));
    }
    if (absw < 0x33000000u) {                         /* < 2^-25 -> 0 (2^-25 itself ties to even = 0) */
        return sign;
    }
    {
        /* subnormal half: result = RNE(|f| * 2^24) as an integer 0..1024 */
        const uint32_t e    = absw >> 23;                      /* 102 .. 112 */
        const uint32_t man  = (absw & 0x7FFFFFu) | 0x800000u;  /* 24-bit significand */
        const uint32_t sh   = 126u - e;                        /* 14 .. 24 */
        const uint32_t q    = man >> sh;
        const uint32_t rem  = man & ((1u << sh) - 1u);
        const uint32_t half = 1u << (sh - 1u);
        uint32_t r = q;
        if (rem > half || (rem == half && (q & 1u))) {
            r++;
        }
        return (uint16_t)(sign | r);
    }
}

static uint16_t rd16(const uint8_t * p) { return (uint16_t)(p[0] | (p[1] << 8)); }
static void     wr16(uint8_t * p, uint16_t v) { p[0] = (uint8_t)(v & 0xFF); p[1] = (uint8_t)(v >> 8); }

/* ---------------------------------------------------------------- dequantization */

/* src/ggml-quants.c:307-325 — y[j] = ((qs[j] & 15) - 8) * d ; y[j+16] = ((qs[j] >> 4) - 8) * d */
static void dq_q4_0(const uint8_t * x, float * y, int64_t k) {
    const int64_t nb = k / 32;
    for (int64_t i = 0; i < nb; i++) {
        const uint8_t * b = x + i * Q4_0_BYTES;
        const float d = b200o_fp16_to_fp32(rd16(b));
        for (int j = 0; j < 16; j++) {
            const int x0 = (b[2 + j] & 0x0F) - 8;
            const int x1 = (b[2 + j] >> 4) - 8;
            y[i * 32 + j]      = (float) x0 * d;
            y[i * 32 + j + 16] = (float) x1 * d;
        }
    }
}

/* src/ggml-quants.c:401-415 — y[j] = qs[j] * d */
static void dq_q8_0(const uint8_t * x, float * y, int64_t k) {
    const int64_t nb = k / 32;
    for (int64_t i = 0; i < nb; i++) {
        const uint8_t * b = x + i * Q8_0_BYTES;
        const float d = b200o_fp16_to_fp32(rd16(b));
        for (int j = 0; j < 32; j++) {
            y[i * 32 + j] = (float)(int8_t) b[2 + j] * d;
        }
    }
}

/* src/ggml-quants.c:327-346 — y[j] = (qs[j] & 15) * d + m ; y[j+16] = (qs[j] >> 4) * d + m */
static void dq_q4_1(const uint8_t * x, float * y, int64_t k) {
    for (int64_t i = 0; i < k / 32; i++) {
        const uint8_t * b = x + i * Q4_1_BYTES;
        const float d = b200o_fp16_to_fp32(rd16(b)), m = b200o_fp16_to_fp32(rd16(b + 2));
        for (int j = 0; j < 16; j++) {
            const int x0 = b[4 + j] & 0x0F, x1 = b[4 + j] >> 4;
            y[i * 32 + j]      = (float) x0 * d + m;
            y[i * 32 + j + 16] = (float) x1 * d + m;
        }
    }
}

/* src/ggml-quants.c:348-372 — fifth bit j (elements 0..15) / j + 16 (elements 16..31) of the 32-bit qh, value - 16 */
static void dq_q5_0(const uint8_t * x, float * y, int64_t k) {
    for (int64_t i = 0; i < k / 32; i++) {
        const uint8_t * b = x + i * Q5_0_BYTES;
        const float d = b200o_fp16_to_fp32(rd16(b));
        const uint32_t qh = (uint32_t) b[2] | ((uint32_t) b[3] << 8) | ((uint32_t) b[4] << 16) | ((uint32_t) b[5] << 24);
        for (int j = 0; j < 16; j++) {
            const uint8_t xh_0 = (uint8_t) (((qh >> (j + 0)) << 4) & 0x10);
            const uint8_t xh_1 = (uint8_t) ((qh >> (j + 12)) & 0x10);
            const int32_t x0 = ((b[6 + j] & 0x0F) | xh_0) - 16;
            const int32_t x1 = ((b[6 + j] >> 4) | xh_1) - 16;
            y[i * 32 + j]      = (float) x0 * d;
            y[i * 32 + j + 16] = (float) x1 * d;
        }
    }
}

/* src/ggml-quants.c:374-399 */
static void dq_q5_1(const uint8_t * x, float * y, int64_t k) {
    for (int64_t i = 0; i < k / 32; i++) {
        const uint8_t * b = x + i * Q5_1_BYTES;
        const float d = b200o_fp16_to_fp32(rd16(b)), m = b200o_fp16_to_fp32(rd16(b + 2));
        const uint32_t qh = (uint32_t) b[4] | ((uint32_t) b[5] << 8) | ((uint32_t) b[6] << 16) | ((uint32_t) b[7] << 24);
        for (int j = 0; j < 16; j++) {
            const uint8_t xh_0 = (uint8_t) (((qh >> (j + 0)) << 4) & 0x10);
            const uint8_t xh_1 = (uint8_t) ((qh >> (j + 12)) & 0x10);
            const int x0 = (b[8 + j] & 0x0F) | xh_0;
            const int x1 = (b[8 + j] >> 4) | xh_1;
            y[i * 32 + j]      = (float) x0 * d + m;
            y[i * 32 + j + 16] = (float) x1 * d + m;
        }
    }
}

/* src/ggml-quants.c:2512-2528 with kvalues_iq4nl (src/ggml-common.h:1088-1090) */
static const int8_t iq4nl_values[16] = { -127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113 };
static void dq_iq4_nl(const uint8_t * x, float * y, int64_t k) {
    for (int64_t i = 0; i < k / 32; i++) {
        const uint8_t * b = x + i * IQ4_NL_BYTES;
        const float d = b200o_fp16_to_fp32(rd16(b));
        for (int j = 0; j < 16; j++) {
            y[i * 32 + j]      = d * (float) iq4nl_values[b[2 + j] & 0xf];
            y[i * 32 + j + 16] = d * (float) iq4nl_values[b[2 + j] >> 4];
        }
    }
}

/* 6-bit scale/min unpack, src/ggml-quants.c:703-710 */
static void scale_min_k4(int j, const uint8_t * q, uint8_t * sc, uint8_t * mn) {
    if (j < 4) {
        *sc = q[j] & 63;
        *mn = q[j + 4] & 63;
    } else {
        *sc = (uint8_t)((q[j + 4] & 0x0F) | ((q[j - 4] >> 6) << 4));
        *mn = (uint8_t)((q[j + 4] >> 4)   | ((q[j]     >> 6) << 4));
    }
}

/* src/ggml-quants.c:1352-1374 — per 64 elements: d1 = d*sc, m1 = dmin*m (each rounded to f32),
 * y = d1 * nibble - m1 (mul rounded, then sub rounded) */
static void dq_q4_K(const uint8_t * x, float * y, int64_t k) {
    const int64_t nb = k / 256;
    for (int64_t i = 0; i < nb; i++) {
        const uint8_t * b  = x + i * Q4_K_BYTES;
        const float d      = b200o_fp16_to_fp32(rd16(b));
        const float dmin   = b200o_fp16_to_fp32(rd16(b + 2));
        const uint8_t * sc = b + 4;
        const uint8_t * q  = b + 16;
        for (int g = 0; g < 4; g++) {
            uint8_t s1, m1, s2, m2;
            scale_min_k4(2 * g,     sc, &s1, &m1);
            scale_min_k4(2 * g + 1, sc, &s2, &m2);
            const float d1 = d * (float) s1, mm1 = dmin * (float) m1;
            const float d2 = d * (float) s2, mm2 = dmin * (float) m2;
            for (int l = 0; l < 32; l++) {
                const float p = d1 * (float)(q[32 * g + l] & 0x0F);
                y[i * 256 + 64 * g + l] = p - mm1;
            }
            for (int l = 0; l < 32; l++) {
                const float p = d2 * (float)(q[32 * g + l] >> 4);
                y[i * 256 + 64 * g + 32 + l] = p - mm2;
            }
        }
    }
}

/* src/ggml-quants.c:1554-1579 — as q4_K with a 5th bit from qh: bit 2g for the low-nibble group,
 * bit 2g+1 for the high-nibble group of 64-element chunk g */
static void dq_q5_K(const uint8_t * x, float * y, int64_t k) {
    const int64_t nb = k / 256;
    for (int64_t i = 0; i < nb; i++) {
        const uint8_t * b  = x + i * Q5_K_BYTES;
        const float d      = b200o_fp16_to_fp32(rd16(b));
        const float dmin   = b200o_fp16_to_fp32(rd16(b + 2));
        const uint8_t * sc = b + 4;
        const uint8_t * qh = b + 16;
        const uint8_t * ql = b + 48;
        for (int g = 0; g < 4; g++) {
            uint8_t s1, m1, s2, m2;
            scale_min_k4(2 * g,     sc, &s1, &m1);
            scale_min_k4(2 * g + 1, sc, &s2, &m2);
            const float d1 = d * (float) s1, mm1 = dmin * (float) m1;
            const float d2 = d * (float) s2, mm2 = dmin * (float) m2;
            for (int l = 0; l < 32; l++) {
                const int q = (ql[32 * g + l] & 0x0F) + (((qh[l] >> (2 * g)) & 1) ? 16 : 0);
                const float p = d1 * (float) q;
                y[i * 256 + 64 * g + l] = p - mm1;
            }
            for (int l = 0; l < 32; l++) {
                const int q = (ql[32 * g + l] >> 4) + (((qh[l] >> (2 * g + 1)) & 1) ? 16 : 0);
                const float p = d2 * (float) q;
                y[i * 256 + 64 * g + 32 + l] = p - mm2;
            }
        }
    }
}

/* src/ggml-quants.c:1762-1791 — per 128 elements: 4 interleaved 32-element streams; 6-bit value
 * = low nibble | 2 high bits, minus 32; y = (d * sc) * q, both products rounded to f32 */
static void dq_q6_K(const uint8_t * x, float * y, int64_t k) {
    const int64_t nb = k / 256;
    for (int64_t i = 0; i < nb; i++) {
        const uint8_t * b  = x + i * Q6_K_BYTES;
        const uint8_t * ql = b;
        const uint8_t * qh = b + 128;
        const int8_t  * sc = (const int8_t *)(b + 192);
        const float d      = b200o_fp16_to_fp32(rd16(b + 208));
        for (int h = 0; h < 2; h++) {
            float * yy = y + i * 256 + 128 * h;
            for (int l = 0; l < 32; l++) {
                const int is = l / 16;
                const int8_t q1 = (int8_t)((ql[64 * h + l]      & 0x0F) | (((qh[32 * h + l] >> 0) & 3) << 4)) - 32;
                const int8_t q2 = (int8_t)((ql[64 * h + l + 32] & 0x0F) | (((qh[32 * h + l] >> 2) & 3) << 4)) - 32;
                const int8_t q3 = (int8_t)((ql[64 * h + l]      >> 4)   | (((qh[32 * h + l] >> 4) & 3) << 4)) - 32;
                const int8_t q4 = (int8_t)((ql[64 * h + l + 32] >> 4)   | (((qh[32 * h + l] >> 6) & 3) << 4)) - 32;
                const float a1 = d * (float) sc[8 * h + is + 0];
                const float a2 = d * (float) sc[8 * h + is + 2];
                const float a3 = d * (float) sc[8 * h + is + 4];
                const float a4 = d * (float) sc[8 * h + is + 6];
                yy[l +  0] = a1 * (float) q1;
                yy[l + 32] = a2 * (float) q2;
                yy[l + 64] = a3 * (float) q3;
                yy[l + 96] = a4 * (float) q4;
            }
        }
    }
}

void b200o_dequantize_row(int type, const void * x, float * y, int64_t k) {
    switch (type) {
        case B200O_Q4_0: dq_q4_0((const uint8_t *) x, y, k); break;
        case B200O_Q8_0: dq_q8_0((const uint8_t *) x, y, k); break;
        case B200O_Q4_K: dq_q4_K((const uint8_t *) x, y, k); break;
        case B200O_Q5_K: dq_q5_K((const uint8_t *) x, y, k); break;
        case B200O_Q6_K: dq_q6_K((const uint8_t *) x, y, k); break;
        case B200O_Q4_1: dq_q4_1((const uint8_t *) x, y, k); break;
        case B200O_Q5_0: dq_q5_0((const uint8_t *) x, y, k); break;
        case B200O_Q5_1: dq_q5_1((const uint8_t *) x, y, k); break;
        case B200O_IQ4_NL: dq_iq4_nl((const uint8_t *) x, y, k); break;
        case B200O_Q8_1: {
            /* d * q (used for the mode-1 MUL_MAT oracle) */
            const uint8_t * p = (const uint8_t *) x;
            for (int64_t i = 0; i < k / 32; i++) {
                const float d = b200o_fp16_to_fp32(rd16(p + i * Q8_1_BYTES));
                for (int j = 0; j < 32; j++) {
                    y[i * 32 + j] = (float)(int8_t) p[i * Q8_1_BYTES + 4 + j] * d;
                }
            }
        } break;
        default: abort();
    }
}

void b200o_dequantize_row_f16(int type, const void * x, uint16_t * y, int64_t k) {
    float * tmp = (float *) malloc(sizeof(float) * (size_t) k);
    b200o_dequantize_row(type, x, tmp, k);
    for (int64_t i = 0; i < k; i++) {
        y[i] = b200o_fp32_to_fp16(tmp[i]);
    }
    free(tmp);
}

/* ---------------------------------------------------------------- activation quantization */

/* src/ggml-quants.c:225-258: amax over 32, d = amax/127, id = d ? 1/d : 0, q = roundf(x*id)
 * (half away from zero), stored d = fp16(d), s = fp16(sum(q) * d) with the UNROUNDED f32 d */
void b200o_quantize_row_q8_1(const float * x, void * vy, int64_t k) {
    uint8_t * y = (uint8_t *) vy;
    const int64_t nb = k / 32;
    for (int64_t i = 0; i < nb; i++) {
        float amax = 0.0f;
        for (int j = 0; j < 32; j++) {
            const float a = fabsf(x[i * 32 + j]);
            amax = amax > a ? amax : a;      /* MAX(amax, fabsf(v)): NaN in v is dropped */
        }
        const float d  = amax / 127.0f;
        const float id = d ? 1.0f / d : 0.0f;
        uint8_t * b = y + i * Q8_1_BYTES;
        wr16(b, b200o_fp32_to_fp16(d));
        int sum = 0;
        for (int j = 0; j < 32; j++) {
            const float v = x[i * 32 + j] * id;
            const int8_t q = (int8_t) roundf(v);
            b[4 + j] = (uint8_t) q;
            sum += q;
        }
        wr16(b + 2, b200o_fp32_to_fp16((float) sum * d));
    }
}

/* quantize_row_q8_K_ref, src/ggml-quants.c:2555-2592 (block_q8_K: f32 d | i8 qs[256] | i16 bsums[16], src/ggml-common.h:338-343;
 * nearest_int :444-449 = round to nearest even through the 1.5 * 2^23 magic number): per 256 values the FIRST element of
 * largest magnitude maps to -127 (iscale = -127 / max, max signed), q = min(127, nearest_int(iscale * x)), d = 1 / iscale.
 * This is what ggml-cpu quantizes the activations of a K-quant MUL_MAT to (vec_dot_type Q8_K, src/ggml-cpu/ggml-cpu.c:270-300).
 * Planar output: q[k] int8, d[k/256] f32, bsums[k/16] int16 (sums of 16 quants). */
static int nearest_int_q(float fval) {
    float val = fval + 12582912.f;
    int i; memcpy(&i, &val, sizeof(int));
    return (i & 0x007fffff) - 0x00400000;
}
void b200o_quantize_row_q8_K(const float * x, int8_t * q, float * d, int16_t * bsums, int64_t k) {
    const int64_t nb = k / 256;
    for (int64_t i = 0; i < nb; i++, x += 256, q += 256, bsums += 16) {
        float max = 0, amax = 0;
        for (int j = 0; j < 256; j++) {
            const float ax = fabsf(x[j]);
            if (ax > amax) { amax = ax; max = x[j]; }
        }
        if (!amax) {
            d[i] = 0;
            memset(q, 0, 256);
            memset(bsums, 0, 16 * sizeof(int16_t));   /* the reference leaves bsums unwritten here; its consumers multiply them by d = 0 */
            continue;
        }
        const float iscale = -127.f / max;
        for (int j = 0; j < 256; j++) {
            const int v = nearest_int_q(iscale * x[j]);
            q[j] = (int8_t) (v < 127 ? v : 127);
        }
        for (int j = 0; j < 16; j++) {
            int sum = 0;
            for (int ii = 0; ii < 16; ii++) sum += q[j * 16 + ii];
            bsums[j] = (int16_t) sum;
        }
        d[i] = 1 / iscale;
    }
}

/* src/ggml-quants.c:199-222 */
void b200o_quantize_row_q8_0(const float * x, void * vy, int64_t k) {
    uint8_t * y = (uint8_t *) vy;
    const int64_t nb = k / 32;
    for (int64_t i = 0; i < nb; i++) {
        float amax = 0.0f;
        for (int j = 0; j < 32; j++) {
            const float a = fabsf(x[i * 32 + j]);
            amax = amax > a ? amax : a;
        }
        const float d  = amax / 127.0f;
        const float id = d ? 1.0f / d : 0.0f;
        uint8_t * b = y + i * Q8_0_BYTES;
        wr16(b, b200o_fp32_to_fp16(d));
        for (int j = 0; j < 32; j++) {
            b[2 + j] = (uint8_t)(int8_t) roundf(x[i * 32 + j] * id);
        }
    }
}

/* src/ggml-quants.c:36-72: d = (signed value of largest magnitude) / -8; q = min(15, (int8)(x/d + 8.5)) */
void b200o_quantize_row_q4_0(const float * x, void * vy, int64_t k) {
    uint8_t * y = (uint8_t *) vy;
    const int64_t nb = k / 32;
    for (int64_t i = 0; i < nb; i++) {
        float amax = 0.0f, max = 0.0f;
        for (int j = 0; j < 32; j++) {
            const float v = x[i * 32 + j];
            if (amax < fabsf(v)) { amax = fabsf(v); max = v; }
        }
        const float d  = max / -8.0f;
        const float id = d ? 1.0f / d : 0.0f;
        uint8_t * b = y + i * Q4_0_BYTES;
        wr16(b, b200o_fp32_to_fp16(d));
        for (int j = 0; j < 16; j++) {
            const float x0 = x[i * 32 + j] * id;
            const float x1 = x[i * 32 + 16 + j] * id;
            int8_t a0 = (int8_t)(x0 + 8.5f), a1 = (int8_t)(x1 + 8.5f);
            const uint8_t q0 = (uint8_t)(a0 < 15 ? a0 : 15);
            const uint8_t q1 = (uint8_t)(a1 < 15 ? a1 : 15);
            b[2 + j] = (uint8_t)(q0 | (q1 << 4));
        }
    }
}

/* ---------------------------------------------------------------- GET_ROWS / MUL_MAT */

void b200o_get_rows(int type, const void * table, size_t row_stride, int64_t k,
                    const int32_t * rows, int64_t nrows_out, float * dst) {
    for (int64_t i = 0; i < nrows_out; i++) {
        b200o_dequantize_row(type, (const uint8_t *) table + (size_t) rows[i] * row_stride, dst + i * k, k);
    }
}

void b200o_mul_mat(int type, const void * w, size_t row_stride, int64_t m, int64_t k,
                   const float * x, int64_t n, float * dst, int mode) {
    float * wrow = (float *) malloc(sizeof(float) * (size_t) k);
    float * xq   = (float *) malloc(sizeof(float) * (size_t)(k * n));
    if (mode == 1) {
        uint8_t * q = (uint8_t *) malloc((size_t)(k / 32) * Q8_1_BYTES);
        for (int64_t j = 0; j < n; j++) {
            b200o_quantize_row_q8_1(x + j * k, q, k);
            b200o_dequantize_row(B200O_Q8_1, q, xq + j * k, k);
        }
        free(q);
    } else if (mode == 2) {
        int8_t  * q  = (int8_t *)  malloc((size_t) k);
        float   * d  = (float *)   malloc(sizeof(float) * (size_t)(k / 256));
        int16_t * bs = (int16_t *) malloc(sizeof(int16_t) * (size_t)(k / 16));
        for (int64_t j = 0; j < n; j++) {
            b200o_quantize_row_q8_K(x + j * k, q, d, bs, k);
            for (int64_t t = 0; t < k; t++) xq[j * k + t] = d[t / 256] * (float) q[t];   /* dequantize_row_q8_K, :2594-2604 */
        }
        free(q); free(d); free(bs);
    } else {
        memcpy(xq, x, sizeof(float) * (size_t)(k * n));
    }
    for (int64_t i = 0; i < m; i++) {
        b200o_dequantize_row(type, (const uint8_t *) w + (size_t) i * row_stride, wrow, k);
        for (int64_t j = 0; j < n; j++) {
            double acc = 0.0;
            const float * xr = xq + j * k;
            for (int64_t t = 0; t < k; t++) {
                acc += (double) wrow[t] * (double) xr[t];
            }
            dst[j * m + i] = (float) acc;
        }
    }
    free(wrow);
    free(xq);
}

double b200o_nmse(const float * a, const float * b, size_t n) {
    double mse_a_b = 0.0, mse_a_0 = 0.0;
    for (size_t i = 0; i < n; i++) {
        const float ai = a[i], bi = b[i];
        mse_a_b += (double)((ai - bi) * (ai - bi));
        mse_a_0 += (double)(ai * ai);
    }
    return mse_a_b / mse_a_0;
}
