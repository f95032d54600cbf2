/* oracle/ggml_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see ggml_oracle.h).
 *
 * Plain-C restatement of the reference's CPU algorithm for the hot path.  Organisation is our
 * own (one "expand a block to integers + per-group scales" helper per weight type, shared by
 * dequantisation, the q8_0/q8_K dot products and the q8_1 dot products); the floating-point
 * operation ORDER follows the reference exactly where bit-exactness is claimed
 * (dequantize_row_*, quantize_row_q8_*_ref) -- compile with -ffp-contract=off, as the reference
 * is (it builds with -std=c11, which disables contraction in gcc).
 *
 * Parity status: PINNED against oracle/_ref (tests/test_oracle_vs_ref.py, tests/golden/).
 */
#include "ggml_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

#define QK   32   /* QK4_0 = QK8_0 = QK8_1, ggml-common.h:144,186,194 */
#define QK_K 256  /* ggml-common.h:69 */

/* ------------------------------------------------------------------ fp16 */
/* IEEE binary16 <-> binary32, round-to-nearest-even; same results as the reference's
 * ggml_compute_fp16_to_fp32 / ggml_compute_fp32_to_fp16 (ggml-impl.h) and as F16C. */
float orc_fp16_to_fp32(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    const uint32_t exp  = (h >> 10) & 0x1Fu;
    uint32_t man        = h & 0x3FFu;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) {
            bits = sign;
        } else { /* subnormal: normalise */
            int e = -1;
            do { man <<= 1; e++; } while (!(man & 0x400u));
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3FFu) << 13);
        }
    } else if (exp == 31) {
        bits = sign | 0x7F800000u | (man << 13);
    } else {
        bits = sign | ((exp + 127 - 15) << 23) | (man << 13);
    }
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

uint16_t orc_fp32_to_fp16(float f) {
    uint32_t x;
    memcpy(&x, &f, 4);
    const uint16_t sign = (uint16_t)((x >> 16) & 0x8000u);
    const uint32_t ax   = x & 0x7FFFFFFFu;
    if (ax >= 0x7F800000u) { /* inf / nan */
        return (uint16_t)(sign | 0x7C00u | (ax > 0x7F800000u ? 0x200u : 0));
    }
    if (ax >= 0x477FF000u) { /* rounds to >= 65520 -> inf */
        return (uint16_t)(sign | 0x7C00u);
    }
    if (ax < 0x33000001u) { /* < 2^-25 (or == 2^-25 exactly, ties to even 0) */
        return sign;
    }
    int32_t  e = (int32_t)(ax >> 23) - 127;
    uint32_t m = (ax & 0x7FFFFFu) | 0x800000u; /* 24-bit significand */
    int shift;
    uint32_t base;
    if (e < -14) { /* subnormal half */
        shift = 13 + (-14 - e);
        base  = 0;
    } else {
        shift = 13;
        base  = (uint32_t)(e + 15) << 10;
        m &= 0x7FFFFFu;
    }
    uint32_t q   = m >> shift;
    uint32_t rem = m & ((1u << shift) - 1);
    uint32_t half = 1u << (shift - 1);
    if (rem > half || (rem == half && (q & 1u))) q++;
    return (uint16_t)(sign | (base + q)); /* carry into exponent is the correct rounding */
}

/* ------------------------------------------------------------------ block layouts */
/* byte offsets inside each block (ggml-common.h:144-335); we address bytes directly so this
 * file needs no packed-struct pragmas. */
enum {
    Q4_0_BYTES = 18,  /* half d; u8 qs[16]                                  :144-149 */
    Q8_0_BYTES = 34,  /* half d; i8 qs[32]                                  :186-191 */
    Q4_1_BYTES = 20,  /* half d; half m; u8 qs[16]                          :155-161 */
    Q5_0_BYTES = 22,  /* half d; u8 qh[4]; u8 qs[16]                        :164-169 */
    Q5_1_BYTES = 24,  /* half d; half m; u8 qh[4]; u8 qs[16]                :172-183 */
    Q8_1_BYTES = 36,  /* half d; half s; i8 qs[32]                          :194-204 */
    Q2_K_BYTES = 84,  /* u8 scales[16]; u8 qs[64]; half d; half dmin         :257-268 */
    Q3_K_BYTES = 110, /* u8 hmask[32]; u8 qs[64]; u8 scales[12]; half d      :274-280 */
    Q4_K_BYTES = 144, /* half d; half dmin; u8 scales[12]; u8 qs[128]       :286-298 */
    Q5_K_BYTES = 176, /* half d; half dmin; u8 scales[12]; u8 qh[32]; u8 qs[128] :303-316 */
    Q6_K_BYTES = 210, /* u8 ql[128]; u8 qh[64]; i8 scales[16]; half d       :321-327 */
    Q8_K_BYTES = 292, /* float d; i8 qs[256]; i16 bsums[16]                 :330-335 */
    IQ4_NL_BYTES = 18,  /* half d; u8 qs[16]                                :406-410 */
    IQ4_XS_BYTES = 136, /* half d; u16 scales_h; u8 scales_l[4]; u8 qs[128] :413-418 */
};

int orc_blck_size(int type) {
    switch (type) {
        case ORC_TYPE_F32: case ORC_TYPE_F16: return 1;
        case ORC_TYPE_Q4_0: case ORC_TYPE_Q8_0: case ORC_TYPE_Q8_1: return QK;
        case ORC_TYPE_Q4_1: case ORC_TYPE_Q5_0: case ORC_TYPE_Q5_1: case ORC_TYPE_IQ4_NL: return QK;
        case ORC_TYPE_IQ4_XS:
        case ORC_TYPE_Q2_K: case ORC_TYPE_Q3_K:
        case ORC_TYPE_Q4_K: case ORC_TYPE_Q5_K: case ORC_TYPE_Q6_K: case ORC_TYPE_Q8_K: return QK_K;
    }
    return 0;
}

size_t orc_type_size(int type) {
    switch (type) {
        case ORC_TYPE_F32:  return 4;
        case ORC_TYPE_F16:  return 2;
        case ORC_TYPE_Q4_0: return Q4_0_BYTES;
        case ORC_TYPE_Q4_1: return Q4_1_BYTES;
        case ORC_TYPE_Q5_0: return Q5_0_BYTES;
        case ORC_TYPE_Q5_1: return Q5_1_BYTES;
        case ORC_TYPE_Q8_0: return Q8_0_BYTES;
        case ORC_TYPE_Q8_1: return Q8_1_BYTES;
        case ORC_TYPE_Q2_K: return Q2_K_BYTES;
        case ORC_TYPE_Q3_K: return Q3_K_BYTES;
        case ORC_TYPE_Q4_K: return Q4_K_BYTES;
        case ORC_TYPE_Q5_K: return Q5_K_BYTES;
        case ORC_TYPE_Q6_K: return Q6_K_BYTES;
        case ORC_TYPE_Q8_K: return Q8_K_BYTES;
        case ORC_TYPE_IQ4_NL: return IQ4_NL_BYTES;
        case ORC_TYPE_IQ4_XS: return IQ4_XS_BYTES;
    }
    return 0;
}

size_t orc_row_size(int type, int64_t ne) {
    const int bs = orc_blck_size(type);
    return bs ? orc_type_size(type) * (size_t)(ne / bs) : 0;
}

static inline uint16_t rd16(const uint8_t *p) { uint16_t v; memcpy(&v, p, 2); return v; }
static inline float    rdf32(const uint8_t *p) { float v; memcpy(&v, p, 4); return v; }
static inline float    rdh(const uint8_t *p) { return orc_fp16_to_fp32(rd16(p)); }
static inline void     wr16(uint8_t *p, uint16_t v) { memcpy(p, &v, 2); }

/* 6-bit (scale, min) pair j of a q4_K / q5_K super-block: get_scale_min_k4, ggml-quants.c:1899-1907 */
static inline void k4_scale_min(int j, const uint8_t *s, int *sc, int *mn) {
    if (j < 4) {
        *sc = s[j] & 63;
        *mn = s[j + 4] & 63;
    } else {
        *sc = (s[j + 4] & 0x0F) | ((s[j - 4] >> 6) << 4);
        *mn = (s[j + 4] >> 4)   | ((s[j]     >> 6) << 4);
    }
}

/* Expand one weight block into integer quants.  Conventions per type:
 *   Q4_0: q[32] in [-8,7]            (dequantize_row_q4_0: low nibbles are values 0..15, high 16..31)
 *   Q8_0: q[32] in [-128,127]
 *   Q4_K: q[256] in [0,15],  sc[8], mn[8]   (value order: 64-value chunks, low nibbles then high)
 *   Q5_K: q[256] in [0,31],  sc[8], mn[8]
 *   Q6_K: q[256] in [-32,31], sc[16] signed (value order per 128: ql lo | ql+32 lo | ql hi | ql+32 hi)
 */
static void expand_q4_0(const uint8_t *b, int8_t *q) {
    const uint8_t *qs = b + 2;
    for (int j = 0; j < 16; j++) {
        q[j]      = (int8_t)((qs[j] & 0x0F) - 8);
        q[j + 16] = (int8_t)((qs[j] >> 4) - 8);
    }
}
static void expand_q8_0(const uint8_t *b, int8_t *q) { memcpy(q, b + 2, 32); }
/* IQ4_NL / IQ4_XS: a nibble indexes 16 non-uniform int8 levels (kvalues_iq4nl, ggml-common.h); values 0..15 of a 32-value block
 * are the low nibbles, 16..31 the high nibbles (dequantize_row_iq4_nl, ggml-quants.c:3743-3760) */
static const int8_t iq4_levels[16] = {-127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113};
static void expand_iq4(const uint8_t *qs, int8_t *q) {
    for (int j = 0; j < 16; j++) {
        q[j]      = iq4_levels[qs[j] & 0x0F];
        q[j + 16] = iq4_levels[qs[j] >> 4];
    }
}
/* IQ4_XS: q[256] levels, sc[8] = 6-bit sub-block scale - 32 (low 4 bits in scales_l, high 2 in scales_h; :3762-3785) */
static void expand_iq4_xs(const uint8_t *b, int8_t *q, int *sc) {
    const unsigned h = rd16(b + 2);
    for (int ib = 0; ib < 8; ib++) {
        sc[ib] = (int)(((b[4 + ib / 2] >> (4 * (ib % 2))) & 0xF) | (((h >> (2 * ib)) & 3) << 4)) - 32;
        expand_iq4(b + 8 + 16 * ib, q + 32 * ib);
    }
}
/* Q4_1: q[32] in [0,15] (values 0..15 = low nibbles, 16..31 = high nibbles), ggml-quants.c:1543-1562 */
static void expand_q4_1(const uint8_t *b, int8_t *q) {
    const uint8_t *qs = b + 4;
    for (int j = 0; j < 16; j++) {
        q[j]      = (int8_t)(qs[j] & 0x0F);
        q[j + 16] = (int8_t)(qs[j] >> 4);
    }
}
/* Q5_0 / Q5_1: the fifth bit of value j is bit j of qh (ggml-quants.c:1564-1588, 1590-1615); off = 16 (Q5_0) or 0 (Q5_1) */
static void expand_q5(const uint8_t *qh4, const uint8_t *qs, int off, int8_t *q) {
    uint32_t qh;
    memcpy(&qh, qh4, 4);
    for (int j = 0; j < 16; j++) {
        const int xh_0 = (int)(((qh >> (j + 0)) << 4) & 0x10);
        const int xh_1 = (int)((qh >> (j + 12)) & 0x10);
        q[j]      = (int8_t)(((qs[j] & 0x0F) | xh_0) - off);
        q[j + 16] = (int8_t)(((qs[j] >> 4) | xh_1) - off);
    }
}

static void expand_q4_K(const uint8_t *b, int8_t *q, int *sc, int *mn) {
    const uint8_t *scales = b + 4, *qs = b + 16;
    for (int j = 0; j < 8; j++) k4_scale_min(j, scales, &sc[j], &mn[j]);
    for (int c = 0; c < 4; c++) {
        for (int l = 0; l < 32; l++) {
            q[64 * c + l]      = (int8_t)(qs[32 * c + l] & 0x0F);
            q[64 * c + 32 + l] = (int8_t)(qs[32 * c + l] >> 4);
        }
    }
}
static void expand_q5_K(const uint8_t *b, int8_t *q, int *sc, int *mn) {
    const uint8_t *scales = b + 4, *qh = b + 16, *qs = b + 48;
    for (int j = 0; j < 8; j++) k4_scale_min(j, scales, &sc[j], &mn[j]);
    for (int c = 0; c < 4; c++) {
        for (int l = 0; l < 32; l++) {
            const int h0 = (qh[l] >> (2 * c)) & 1, h1 = (qh[l] >> (2 * c + 1)) & 1;
            q[64 * c + l]      = (int8_t)((qs[32 * c + l] & 0x0F) + 16 * h0);
            q[64 * c + 32 + l] = (int8_t)((qs[32 * c + l] >> 4) + 16 * h1);
        }
    }
}
/* Q2_K (dequantize_row_q2_K, ggml-quants.c:1980-2010): value v of the block lives in qs[32 (v / 128) + v % 32], bits
 * 2 ((v % 128) / 32); scale byte v / 16: low nibble = scale, high nibble = min.  q[256] in [0,3], sc[16], mn[16]. */
static void expand_q2_K(const uint8_t *b, int8_t *q, int *sc, int *mn) {
    const uint8_t *scales = b, *qs = b + 16;
    for (int j = 0; j < 16; j++) { sc[j] = scales[j] & 0xF; mn[j] = scales[j] >> 4; }
    for (int v = 0; v < QK_K; v++) q[v] = (int8_t)((qs[32 * (v / 128) + v % 32] >> (2 * ((v % 128) / 32))) & 3);
}
/* Q3_K (dequantize_row_q3_K, :2328-2373): low 2 bits like Q2_K, high bit = bit v / 32 of hmask[v % 32] (set: +0, clear: -4);
 * 16 six-bit scales packed in 12 bytes (the aux[] shuffle of :2345-2350), stored value - 32.  q[256] in [-4,3], sc[16] signed. */
static void expand_q3_K(const uint8_t *b, int8_t *q, int *sc) {
    const uint8_t *hm = b, *qs = b + 32;
    uint32_t aux[4];
    memcpy(aux, b + 96, 12);
    const uint32_t kmask1 = 0x03030303, kmask2 = 0x0f0f0f0f, tmp = aux[2];
    aux[2] = ((aux[0] >> 4) & kmask2) | (((tmp >> 4) & kmask1) << 4);
    aux[3] = ((aux[1] >> 4) & kmask2) | (((tmp >> 6) & kmask1) << 4);
    aux[0] = (aux[0] & kmask2) | (((tmp >> 0) & kmask1) << 4);
    aux[1] = (aux[1] & kmask2) | (((tmp >> 2) & kmask1) << 4);
    const int8_t *scales = (const int8_t *)aux;
    for (int j = 0; j < 16; j++) sc[j] = scales[j] - 32;
    for (int v = 0; v < QK_K; v++) {
        const int lo = (qs[32 * (v / 128) + v % 32] >> (2 * ((v % 128) / 32))) & 3;
        q[v] = (int8_t)(lo - ((hm[v % 32] >> (v / 32)) & 1 ? 0 : 4));
    }
}
static void expand_q6_K(const uint8_t *b, int8_t *q, int *sc) {
    const uint8_t *ql = b, *qh = b + 128;
    const int8_t *scales = (const int8_t *)(b + 192);
    for (int j = 0; j < 16; j++) sc[j] = scales[j];
    for (int h = 0; h < 2; h++) {
        for (int l = 0; l < 32; l++) {
            const uint8_t hb = qh[32 * h + l];
            q[128 * h + l]      = (int8_t)(((ql[64 * h + l]      & 0x0F) | (((hb >> 0) & 3) << 4)) - 32);
            q[128 * h + 32 + l] = (int8_t)(((ql[64 * h + 32 + l] & 0x0F) | (((hb >> 2) & 3) << 4)) - 32);
            q[128 * h + 64 + l] = (int8_t)(((ql[64 * h + l]      >> 4)   | (((hb >> 4) & 3) << 4)) - 32);
            q[128 * h + 96 + l] = (int8_t)(((ql[64 * h + 32 + l] >> 4)   | (((hb >> 6) & 3) << 4)) - 32);
        }
    }
}

/* ------------------------------------------------------------------ dequantisation */
int orc_dequantize_row(int type, const void *src, float *dst, int64_t k) {
    const uint8_t *b = (const uint8_t *)src;
    int8_t q[QK_K];
    int sc[16], mn[8];
    switch (type) {
        case ORC_TYPE_F32: memcpy(dst, src, (size_t)k * 4); return 0;
        case ORC_TYPE_F16:
            for (int64_t i = 0; i < k; i++) dst[i] = rdh(b + 2 * i);
            return 0;
        case ORC_TYPE_Q4_0: /* ggml-quants.c:1523-1541: y = x0*d */
            if (k % QK) return -1;
            for (int64_t i = 0; i < k / QK; i++, b += Q4_0_BYTES, dst += QK) {
                const float d = rdh(b);
                expand_q4_0(b, q);
                for (int j = 0; j < QK; j++) dst[j] = (float)q[j] * d;
            }
            return 0;
        case ORC_TYPE_IQ4_NL: /* ggml-quants.c:3743-3760: y = d * level */
            if (k % QK) return -1;
            for (int64_t i = 0; i < k / QK; i++, b += IQ4_NL_BYTES, dst += QK) {
                const float d = rdh(b);
                expand_iq4(b + 2, q);
                for (int j = 0; j < QK; j++) dst[j] = d * (float)q[j];
            }
            return 0;
        case ORC_TYPE_IQ4_XS: /* :3762-3785: y = (d * (ls - 32)) * level */
            if (k % QK_K) return -1;
            for (int64_t i = 0; i < k / QK_K; i++, b += IQ4_XS_BYTES, dst += QK_K) {
                const float d = rdh(b);
                expand_iq4_xs(b, q, sc);
                for (int ib = 0; ib < 8; ib++) {
                    const float dl = d * (float)sc[ib];
                    for (int j = 0; j < 32; j++) dst[32 * ib + j] = dl * (float)q[32 * ib + j];
                }
            }
            return 0;
        case ORC_TYPE_Q4_1: /* :1543-1562: y = x0*d + m (a product, then a sum) */
        case ORC_TYPE_Q5_1: /* :1590-1615 */
            if (k % QK) return -1;
            for (int64_t i = 0; i < k / QK; i++, b += orc_type_size(type), dst += QK) {
                const float d = rdh(b), m = rdh(b + 2);
                if (type == ORC_TYPE_Q4_1) expand_q4_1(b, q); else expand_q5(b + 4, b + 8, 0, q);
                for (int j = 0; j < QK; j++) {
                    const float prod = (float)q[j] * d;
                    dst[j] = prod + m;
                }
            }
            return 0;
        case ORC_TYPE_Q5_0: /* :1564-1588: y = x0*d */
            if (k % QK) return -1;
            for (int64_t i = 0; i < k / QK; i++, b += Q5_0_BYTES, dst += QK) {
                const float d = rdh(b);
                expand_q5(b + 2, b + 6, 16, q);
                for (int j = 0; j < QK; j++) dst[j] = (float)q[j] * d;
            }
            return 0;
        case ORC_TYPE_Q8_0: /* :1617-1632 */
            if (k % QK) return -1;
            for (int64_t i = 0; i < k / QK; i++, b += Q8_0_BYTES, dst += QK) {
                const float d = rdh(b);
                expand_q8_0(b, q);
                for (int j = 0; j < QK; j++) dst[j] = (float)q[j] * d;
            }
            return 0;
        case ORC_TYPE_Q4_K: /* :2556-2578: y = (d*sc)*q - (dmin*m), two roundings then the sub */
        case ORC_TYPE_Q5_K: /* :2764-2789 */
            if (k % QK_K) return -1;
            for (int64_t i = 0; i < k / QK_K; i++, dst += QK_K) {
                const float d = rdh(b), dmin = rdh(b + 2);
                if (type == ORC_TYPE_Q4_K) { expand_q4_K(b, q, sc, mn); b += Q4_K_BYTES; }
                else                       { expand_q5_K(b, q, sc, mn); b += Q5_K_BYTES; }
                for (int j = 0; j < 8; j++) {
                    const float dj = d * (float)sc[j];
                    const float mj = dmin * (float)mn[j];
                    for (int l = 0; l < 32; l++) {
                        const float prod = dj * (float)q[32 * j + l];
                        dst[32 * j + l] = prod - mj;
                    }
                }
            }
            return 0;
        case ORC_TYPE_Q2_K: /* :1980-2010: y = (d*sc)*q - (dmin*m) */
            if (k % QK_K) return -1;
            for (int64_t i = 0; i < k / QK_K; i++, b += Q2_K_BYTES, dst += QK_K) {
                int mn16[16];
                const float d = rdh(b + 80), dmin = rdh(b + 82);
                expand_q2_K(b, q, sc, mn16);
                for (int j = 0; j < 16; j++) {
                    const float dl = d * (float)sc[j], ml = dmin * (float)mn16[j];
                    for (int l = 0; l < 16; l++) {
                        const float prod = dl * (float)q[16 * j + l];
                        dst[16 * j + l] = prod - ml;
                    }
                }
            }
            return 0;
        case ORC_TYPE_Q3_K: /* :2328-2373: y = (d*(sc-32))*q */
            if (k % QK_K) return -1;
            for (int64_t i = 0; i < k / QK_K; i++, b += Q3_K_BYTES, dst += QK_K) {
                const float d = rdh(b + 108);
                expand_q3_K(b, q, sc);
                for (int j = 0; j < 16; j++) {
                    const float dl = d * (float)sc[j];
                    for (int l = 0; l < 16; l++) dst[16 * j + l] = dl * (float)q[16 * j + l];
                }
            }
            return 0;
        case ORC_TYPE_Q6_K: /* :2978-3007: y = (d*sc)*q */
            if (k % QK_K) return -1;
            for (int64_t i = 0; i < k / QK_K; i++, b += Q6_K_BYTES, dst += QK_K) {
                const float d = rdh(b + 208);
                expand_q6_K(b, q, sc);
                for (int j = 0; j < 16; j++) {
                    const float dj = d * (float)sc[j];
                    for (int l = 0; l < 16; l++) dst[16 * j + l] = dj * (float)q[16 * j + l];
                }
            }
            return 0;
        case ORC_TYPE_Q8_K: /* :3825-3835 */
            if (k % QK_K) return -1;
            for (int64_t i = 0; i < k / QK_K; i++, b += Q8_K_BYTES, dst += QK_K) {
                const float d = rdf32(b);
                for (int j = 0; j < QK_K; j++) dst[j] = d * (float)((const int8_t *)(b + 4))[j];
            }
            return 0;
    }
    return -2;
}

/* ------------------------------------------------------------------ quantisation */
/* nearest_int, ggml-quants.c:1640-1645 (round-half-even via the 1.5*2^23 magic constant) */
static inline int nearest_int(float fval) {
    float val = fval + 12582912.f;
    int i;
    memcpy(&i, &val, sizeof(int));
    return (i & 0x007fffff) - 0x00400000;
}

int orc_quantize_row(int type, const float *x, void *dst, int64_t k) {
    uint8_t *b = (uint8_t *)dst;
    switch (type) {
        case ORC_TYPE_Q4_0: /* quantize_row_q4_0_ref, ggml-quants.c:669-704 */
            if (k % QK) return -1;
            for (int64_t i = 0; i < k / QK; i++, x += QK, b += Q4_0_BYTES) {
                float amax = 0.0f, max = 0.0f;
                for (int j = 0; j < QK; j++) {
                    if (amax < fabsf(x[j])) { amax = fabsf(x[j]); max = x[j]; }
                }
                const float d  = max / -8;
                const float id = d ? 1.0f / d : 0.0f;
                wr16(b, orc_fp32_to_fp16(d));
                for (int j = 0; j < QK / 2; j++) {
                    const float x0 = x[j] * id, x1 = x[QK / 2 + j] * id;
                    int q0 = (int8_t)(x0 + 8.5f), q1 = (int8_t)(x1 + 8.5f);
                    if (q0 > 15) q0 = 15;
                    if (q1 > 15) q1 = 15;
                    b[2 + j] = (uint8_t)((uint8_t)q0 | ((uint8_t)q1 << 4));
                }
            }
            return 0;
        case ORC_TYPE_Q8_0: /* quantize_row_q8_0_ref, :849-872 */
        case ORC_TYPE_Q8_1: /* quantize_row_q8_1_ref, :1160-1193 */
            if (k % QK) return -1;
            for (int64_t i = 0; i < k / QK; i++, x += QK) {
                float amax = 0.0f;
                for (int j = 0; j < QK; j++) { const float a = fabsf(x[j]); if (a > amax) amax = a; }
                const float d  = amax / ((1 << 7) - 1);
                const float id = d ? 1.0f / d : 0.0f;
                wr16(b, orc_fp32_to_fp16(d));
                int8_t *qs = (int8_t *)(b + (type == ORC_TYPE_Q8_1 ? 4 : 2));
                int sum = 0;
                for (int j = 0; j < QK; j++) {
                    qs[j] = (int8_t)roundf(x[j] * id);
                    sum += qs[j];
                }
                if (type == ORC_TYPE_Q8_1) {
                    wr16(b + 2, orc_fp32_to_fp16((float)sum * d)); /* s = d * sum(q), :1190 */
                    b += Q8_1_BYTES;
                } else {
                    b += Q8_0_BYTES;
                }
            }
            return 0;
        case ORC_TYPE_Q8_K: /* quantize_row_q8_K_ref, :3786-3823 */
            if (k % QK_K) return -1;
            for (int64_t i = 0; i < k / QK_K; i++, x += QK_K, b += Q8_K_BYTES) {
                float max = 0, amax = 0;
                for (int j = 0; j < QK_K; j++) {
                    const float ax = fabsf(x[j]);
                    if (ax > amax) { amax = ax; max = x[j]; }
                }
                int8_t *qs = (int8_t *)(b + 4);
                int16_t bsums[16];
                if (!amax) {
                    const float z = 0;
                    memcpy(b, &z, 4);
                    memset(qs, 0, QK_K);
                    /* the reference leaves bsums untouched here (:3800-3805); the work buffer is
                     * reused, so their content is unspecified -- we zero them (they multiply d==0) */
                    memset(b + 4 + QK_K, 0, 32);
                    continue;
                }
                const float iscale = -127.f / max;
                for (int j = 0; j < QK_K; j++) {
                    int v = nearest_int(iscale * x[j]);
                    qs[j] = (int8_t)(v < 127 ? v : 127);
                }
                for (int j = 0; j < 16; j++) {
                    int s = 0;
                    for (int l = 0; l < 16; l++) s += qs[16 * j + l];
                    bsums[j] = (int16_t)s;
                }
                memcpy(b + 4 + QK_K, bsums, 32);
                const float d = 1 / iscale;
                memcpy(b, &d, 4);
            }
            return 0;
    }
    return -2;
}

/* ------------------------------------------------------------------ dot products (CPU semantics) */
static int vec_dot_act_type(int wtype) {
    switch (wtype) {
        case ORC_TYPE_Q4_0: case ORC_TYPE_Q8_0: case ORC_TYPE_Q5_0: return ORC_TYPE_Q8_0; /* ggml.c:793-808, 873-889 */
        case ORC_TYPE_IQ4_NL: return ORC_TYPE_Q8_0; /* ggml.c type_traits[GGML_TYPE_IQ4_NL].vec_dot_type */
        case ORC_TYPE_IQ4_XS: return ORC_TYPE_Q8_K;
        case ORC_TYPE_Q4_1: case ORC_TYPE_Q5_1: return ORC_TYPE_Q8_1; /* vec_dot_type of q4_1 / q5_1, ggml.c type_traits */
        case ORC_TYPE_Q2_K: case ORC_TYPE_Q3_K:
        case ORC_TYPE_Q4_K: case ORC_TYPE_Q5_K: case ORC_TYPE_Q6_K: return ORC_TYPE_Q8_K; /* :924-959 */
    }
    return -1;
}

float orc_vec_dot(int wtype, int64_t n, const void *wv, const void *av) {
    const uint8_t *w = (const uint8_t *)wv, *a = (const uint8_t *)av;
    int8_t q[QK_K];
    int sc[16], mn[8];
    if (wtype == ORC_TYPE_Q4_0 || wtype == ORC_TYPE_Q8_0) {
        const size_t wb = orc_type_size(wtype);
        float sumf = 0;
        for (int64_t i = 0; i < n / QK; i++, w += wb, a += Q8_0_BYTES) {
            if (wtype == ORC_TYPE_Q4_0) expand_q4_0(w, q); else expand_q8_0(w, q);
            const int8_t *y = (const int8_t *)(a + 2);
            int sumi = 0;
            for (int j = 0; j < QK; j++) sumi += q[j] * y[j];
            if (wtype == ORC_TYPE_Q4_0) {
                /* ggml-quants.c:4496: sumf += sumi*d_x*d_y evaluated left to right */
                sumf += (float)sumi * rdh(w) * rdh(a);
            } else {
                /* :5855: sumf += sumi*(d_x*d_y) */
                sumf += (float)sumi * (rdh(w) * rdh(a));
            }
        }
        return sumf;
    }
    if (wtype == ORC_TYPE_IQ4_NL) { /* ggml_vec_dot_iq4_nl_q8_0 scalar tail, ggml-quants.c:12660-12669 */
        float sumf = 0;
        for (int64_t i = 0; i < n / QK; i++, w += IQ4_NL_BYTES, a += Q8_0_BYTES) {
            expand_iq4(w + 2, q);
            const int8_t *y = (const int8_t *)(a + 2);
            const float d = rdh(a) * rdh(w);
            int sumi1 = 0, sumi2 = 0;
            for (int j = 0; j < 16; j++) { sumi1 += y[j] * q[j]; sumi2 += y[j + 16] * q[j + 16]; }
            sumf += d * (float)(sumi1 + sumi2);
        }
        return sumf;
    }
    if (wtype == ORC_TYPE_IQ4_XS) { /* ggml_vec_dot_iq4_xs_q8_K scalar branch, ggml-quants.c:12975-13004 */
        float sumf = 0;
        for (int64_t i = 0; i < n / QK_K; i++, w += IQ4_XS_BYTES, a += Q8_K_BYTES) {
            expand_iq4_xs(w, q, sc);
            const float d4d8 = rdh(w) * rdf32(a);
            const int8_t *y = (const int8_t *)(a + 4);
            for (int ib = 0; ib < 8; ib++) {
                const float dl = d4d8 * (float)sc[ib];
                int sumi1 = 0, sumi2 = 0;
                for (int j = 0; j < 16; j++) { sumi1 += y[32 * ib + j] * q[32 * ib + j]; sumi2 += y[32 * ib + 16 + j] * q[32 * ib + 16 + j]; }
                sumf += dl * (float)(sumi1 + sumi2);
            }
        }
        return sumf;
    }
    if (wtype == ORC_TYPE_Q5_0) { /* ggml_vec_dot_q5_0_q8_0 scalar tail, ggml-quants.c:5120-5143 */
        float sumf = 0;
        for (int64_t i = 0; i < n / QK; i++, w += Q5_0_BYTES, a += Q8_0_BYTES) {
            expand_q5(w + 2, w + 6, 16, q);
            const int8_t *y = (const int8_t *)(a + 2);
            int sumi = 0;
            for (int j = 0; j < QK; j++) sumi += q[j] * y[j];
            sumf += (rdh(w) * rdh(a)) * (float)sumi;
        }
        return sumf;
    }
    if (wtype == ORC_TYPE_Q4_1 || wtype == ORC_TYPE_Q5_1) { /* ggml_vec_dot_q4_1_q8_1 :4771-4785, q5_1_q8_1 scalar tail */
        const size_t wb = orc_type_size(wtype);
        float sumf = 0;
        for (int64_t i = 0; i < n / QK; i++, w += wb, a += Q8_1_BYTES) {
            if (wtype == ORC_TYPE_Q4_1) expand_q4_1(w, q); else expand_q5(w + 4, w + 8, 0, q);
            const int8_t *y = (const int8_t *)(a + 4);
            int sumi = 0;
            for (int j = 0; j < QK; j++) sumi += q[j] * y[j];
            const float t1 = (rdh(w) * rdh(a)) * (float)sumi;
            const float t2 = rdh(w + 2) * rdh(a + 2);
            sumf += t1 + t2;
        }
        return sumf;
    }
    /* K-quants vs q8_K.  The scalar reference keeps 8 interleaved int32 lanes (aux32[l], l = value
     * index mod 8) per super-block and 8 float lanes across super-blocks, then adds the lanes
     * (:8233-8277, :8865-8913, :9526-9565); we keep that association so the float result matches. */
    float sums[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float sumf = 0;
    const size_t wb = orc_type_size(wtype);
    for (int64_t i = 0; i < n / QK_K; i++, w += wb, a += Q8_K_BYTES) {
        const float yd = rdf32(a);
        const int8_t *y = (const int8_t *)(a + 4);
        int16_t bsums[16];
        memcpy(bsums, a + 4 + QK_K, 32);
        int32_t lane[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        float d;
        if (wtype == ORC_TYPE_Q2_K) { /* ggml_vec_dot_q2_K_q8_K scalar branch, ggml-quants.c:6892-6930: no lanes */
            int mn16[16];
            expand_q2_K(w, q, sc, mn16);
            int summs = 0, isum = 0;
            for (int j = 0; j < 16; j++) summs += bsums[j] * mn16[j];
            for (int j = 0; j < 16; j++) {
                int isuml = 0;
                for (int l = 0; l < 16; l++) isuml += y[16 * j + l] * q[16 * j + l];
                isum += sc[j] * isuml;
            }
            const float dall = yd * rdh(w + 80), dmn = yd * rdh(w + 82);
            const float t1 = dall * (float)isum, t2 = dmn * (float)summs;
            sumf += t1 - t2;
            continue;
        }
        if (wtype == ORC_TYPE_Q3_K) { /* :7651-7707 */
            expand_q3_K(w, q, sc);
            for (int j = 0; j < QK_K; j++) lane[j & 7] += sc[j / 16] * (int32_t)(int16_t)(y[j] * q[j]);
            d = rdh(w + 108) * yd;
        } else if (wtype == ORC_TYPE_Q6_K) {
            expand_q6_K(w, q, sc);
            for (int j = 0; j < QK_K; j++) lane[j & 7] += sc[j / 16] * (int32_t)(int16_t)(y[j] * q[j]);
            d = rdh(w + 208) * yd;
        } else {
            if (wtype == ORC_TYPE_Q4_K) expand_q4_K(w, q, sc, mn); else expand_q5_K(w, q, sc, mn);
            int sumi = 0;
            for (int j = 0; j < 16; j++) sumi += bsums[j] * mn[j / 2];
            for (int j = 0; j < QK_K; j++) lane[j & 7] += sc[j / 32] * (int32_t)(int16_t)(y[j] * q[j]);
            d = rdh(w) * yd;
            const float dmin = rdh(w + 2) * yd;
            sumf -= dmin * (float)sumi;
        }
        for (int l = 0; l < 8; l++) sums[l] += d * (float)lane[l];
    }
    for (int l = 0; l < 8; l++) sumf += sums[l];
    return sumf;
}

/* GPU-path arithmetic: per 32-wide q8_1 block (d8, s8 = d8*sum(q8), q8[32]):
 *   Q4_0: d4 * (d8*sum((v-8+8)*u) - 8*s8)  -> written as d4*(d8*sumi_unsigned - 8*s8)   vecdotq.cuh:23-42
 *   Q8_0: d*d8*sum(v*u)                                                                  :151-163
 *   Q4_K/Q5_K: d*sum_j sc_j*d8_j*sum(v*u) - dmin*sum_j m_j*s8_j                          :357-379, 413-442
 *   Q6_K: d*sum_g sc_g*d8*sum((v-32)*u) over 16-wide groups                              :476-496
 * (the reference kernel uses sum(u) via dp4a for the min term; with s8 = d8*sum(u) from
 * quantize_row_q8_1_ref the two agree up to the fp16 rounding of s8 -- we follow the s8 form,
 * which is what our kernel does).  Accumulated in double: this is a value oracle, not an
 * order-of-summation oracle. */
double orc_vec_dot_q8_1(int wtype, int64_t n, const void *wv, const void *av) {
    const uint8_t *w = (const uint8_t *)wv, *a = (const uint8_t *)av;
    int8_t q[QK_K];
    int sc[16], mn[8];
    double acc = 0;
    if (wtype == ORC_TYPE_Q4_0 || wtype == ORC_TYPE_Q8_0) {
        const size_t wb = orc_type_size(wtype);
        for (int64_t i = 0; i < n / QK; i++, w += wb, a += Q8_1_BYTES) {
            const int8_t *y = (const int8_t *)(a + 4);
            const double d8 = rdh(a), s8 = rdh(a + 2), dw = rdh(w);
            int sumi = 0;
            if (wtype == ORC_TYPE_Q4_0) {
                const uint8_t *qs = w + 2;
                for (int j = 0; j < 16; j++) sumi += (qs[j] & 0x0F) * y[j] + (qs[j] >> 4) * y[j + 16];
                acc += dw * (d8 * sumi - 8.0 * s8);
            } else {
                expand_q8_0(w, q);
                for (int j = 0; j < QK; j++) sumi += q[j] * y[j];
                acc += dw * d8 * sumi;
            }
        }
        return acc;
    }
    if (wtype == ORC_TYPE_IQ4_NL) { /* d * d8 * sum(level * u), vecdotq.cuh:1086-1105 */
        for (int64_t i = 0; i < n / QK; i++, w += IQ4_NL_BYTES, a += Q8_1_BYTES) {
            const int8_t *y = (const int8_t *)(a + 4);
            int sumi = 0;
            expand_iq4(w + 2, q);
            for (int j = 0; j < QK; j++) sumi += q[j] * y[j];
            acc += (double)rdh(w) * (double)rdh(a) * sumi;
        }
        return acc;
    }
    if (wtype == ORC_TYPE_IQ4_XS) { /* d * sum_ib (ls_ib - 32) * d8_ib * sum(level * u), vecdotq.cuh:1110-1140 */
        for (int64_t i = 0; i < n / QK_K; i++, w += IQ4_XS_BYTES, a += 8 * Q8_1_BYTES) {
            expand_iq4_xs(w, q, sc);
            double blk = 0;
            for (int ib = 0; ib < 8; ib++) {
                const int8_t *y = (const int8_t *)(a + ib * Q8_1_BYTES + 4);
                int sumi = 0;
                for (int j = 0; j < 32; j++) sumi += q[32 * ib + j] * y[j];
                blk += (double)sc[ib] * (double)rdh(a + ib * Q8_1_BYTES) * sumi;
            }
            acc += (double)rdh(w) * blk;
        }
        return acc;
    }
    if (wtype == ORC_TYPE_Q4_1 || wtype == ORC_TYPE_Q5_0 || wtype == ORC_TYPE_Q5_1) {
        /* Q4_1 / Q5_1: d*d8*sum(v*u) + m*s8 (vecdotq.cuh:44-76, 110-149); Q5_0: d*(d8*sum(v5*u) - 16*s8) (:78-108) */
        const size_t wb = orc_type_size(wtype);
        for (int64_t i = 0; i < n / QK; i++, w += wb, a += Q8_1_BYTES) {
            const int8_t *y = (const int8_t *)(a + 4);
            const double d8 = rdh(a), s8 = rdh(a + 2), dw = rdh(w);
            if (wtype == ORC_TYPE_Q4_1) expand_q4_1(w, q);
            else if (wtype == ORC_TYPE_Q5_1) expand_q5(w + 4, w + 8, 0, q);
            else expand_q5(w + 2, w + 6, 0, q);
            int sumi = 0;
            for (int j = 0; j < QK; j++) sumi += q[j] * y[j];
            if (wtype == ORC_TYPE_Q5_0) acc += dw * (d8 * sumi - 16.0 * s8);
            else acc += dw * d8 * sumi + (double)rdh(w + 2) * s8;
        }
        return acc;
    }
    const size_t wb = orc_type_size(wtype);
    for (int64_t i = 0; i < n / QK_K; i++, w += wb, a += 8 * Q8_1_BYTES) {
        if (wtype == ORC_TYPE_Q2_K || wtype == ORC_TYPE_Q3_K) {
            /* per 16-wide group g (q8_1 block g / 2, half g & 1):  Q2_K: d sc d8 sum(v u) - dmin m d8 sum(u);
             * Q3_K: d sc d8 sum(v u), v in [-4, 3]  (vecdotq.cuh:230-355 use the same per-16 integer sums) */
            int mn16[16];
            const double d = rdh(w + (wtype == ORC_TYPE_Q2_K ? 80 : 108)), dmin = wtype == ORC_TYPE_Q2_K ? rdh(w + 82) : 0.0;
            if (wtype == ORC_TYPE_Q2_K) expand_q2_K(w, q, sc, mn16); else { expand_q3_K(w, q, sc); memset(mn16, 0, sizeof(mn16)); }
            double sd = 0, sm = 0;
            for (int g = 0; g < 16; g++) {
                const uint8_t *blk = a + (g / 2) * Q8_1_BYTES;
                const int8_t *y = (const int8_t *)(blk + 4) + 16 * (g & 1);
                int sumi = 0, sumu = 0;
                for (int l = 0; l < 16; l++) { sumi += q[16 * g + l] * y[l]; sumu += y[l]; }
                sd += (double)rdh(blk) * sc[g] * sumi;
                sm += (double)rdh(blk) * mn16[g] * sumu;
            }
            acc += d * sd - dmin * sm;
            continue;
        }
        if (wtype == ORC_TYPE_Q6_K) {
            expand_q6_K(w, q, sc);
            const double d = rdh(w + 208);
            double s = 0;
            for (int g = 0; g < 16; g++) {
                const uint8_t *blk = a + (g / 2) * Q8_1_BYTES;
                const int8_t *y = (const int8_t *)(blk + 4) + 16 * (g & 1);
                int sumi = 0;
                for (int l = 0; l < 16; l++) sumi += q[16 * g + l] * y[l];
                s += (double)rdh(blk) * sc[g] * sumi;
            }
            acc += d * s;
        } else {
            if (wtype == ORC_TYPE_Q4_K) expand_q4_K(w, q, sc, mn); else expand_q5_K(w, q, sc, mn);
            const double d = rdh(w), dmin = rdh(w + 2);
            double sd = 0, sm = 0;
            for (int j = 0; j < 8; j++) {
                const uint8_t *blk = a + j * Q8_1_BYTES;
                const int8_t *y = (const int8_t *)(blk + 4);
                int sumi = 0;
                for (int l = 0; l < 32; l++) sumi += q[32 * j + l] * y[l];
                sd += (double)rdh(blk) * sc[j] * sumi;
                sm += (double)rdh(blk + 2) * mn[j];
            }
            acc += d * sd - dmin * sm;
        }
    }
    return acc;
}

/* ------------------------------------------------------------------ tiny parallel-for */
typedef void (*orc_job_fn)(void *ctx, int64_t begin, int64_t end);
typedef struct { orc_job_fn fn; void *ctx; int64_t begin, end; } orc_job;
static void *orc_job_entry(void *p) { orc_job *j = (orc_job *)p; j->fn(j->ctx, j->begin, j->end); return NULL; }

static int orc_resolve_threads(int nthreads, int64_t n) {
    if (nthreads <= 0) {
        long c = sysconf(_SC_NPROCESSORS_ONLN);
        nthreads = c > 0 ? (int)c : 1;
    }
    if (nthreads > 256) nthreads = 256;
    if ((int64_t)nthreads > n) nthreads = n > 0 ? (int)n : 1;
    return nthreads;
}

static void orc_parallel_for(int64_t n, int nthreads, orc_job_fn fn, void *ctx) {
    nthreads = orc_resolve_threads(nthreads, n);
    if (nthreads == 1) { fn(ctx, 0, n); return; }
    pthread_t th[256];
    orc_job jobs[256];
    const int64_t per = (n + nthreads - 1) / nthreads;
    int started = 0;
    for (int t = 0; t < nthreads; t++) {
        jobs[t].fn = fn; jobs[t].ctx = ctx;
        jobs[t].begin = t * per;
        jobs[t].end = jobs[t].begin + per < n ? jobs[t].begin + per : n;
        if (jobs[t].begin >= jobs[t].end) break;
        if (pthread_create(&th[t], NULL, orc_job_entry, &jobs[t]) != 0) { fn(ctx, jobs[t].begin, jobs[t].end); th[t] = 0; }
        started = t + 1;
    }
    for (int t = 0; t < started; t++) if (th[t]) pthread_join(th[t], NULL);
}

/* ------------------------------------------------------------------ MUL_MAT */
typedef struct {
    int wtype, q81;
    const uint8_t *w; size_t w_row_bytes;
    const uint8_t *act; size_t act_row_bytes;
    float *dst; double *dst64;
    int64_t K, M, N;
} mm_ctx;

static void mm_rows(void *vp, int64_t r0, int64_t r1) {
    mm_ctx *c = (mm_ctx *)vp;
    for (int64_t m = r0; m < r1; m++) {
        const uint8_t *wr = c->w + (size_t)m * c->w_row_bytes;
        for (int64_t n = 0; n < c->N; n++) {
            const uint8_t *ar = c->act + (size_t)n * c->act_row_bytes;
            if (c->q81) c->dst64[n * c->M + m] = orc_vec_dot_q8_1(c->wtype, c->K, wr, ar);
            else        c->dst[n * c->M + m]   = orc_vec_dot(c->wtype, c->K, wr, ar);
        }
    }
}

static int mm_common(int wtype, const void *w, size_t w_row_bytes, const float *x, float *dst, double *dst64,
                     int64_t K, int64_t M, int64_t N, int nthreads, int q81) {
    const int at = q81 ? ORC_TYPE_Q8_1 : vec_dot_act_type(wtype);
    if (at < 0 || K % orc_blck_size(wtype) || K % orc_blck_size(at)) return -1;
    const size_t arow = orc_row_size(at, K);
    uint8_t *act = (uint8_t *)malloc(arow * (size_t)(N > 0 ? N : 1));
    if (!act) return -3;
    for (int64_t n = 0; n < N; n++) orc_quantize_row(at, x + n * K, act + n * arow, K); /* ggml.c:12562-12590 */
    mm_ctx c = { wtype, q81, (const uint8_t *)w, w_row_bytes, act, arow, dst, dst64, K, M, N };
    orc_parallel_for(M, nthreads, mm_rows, &c);
    free(act);
    return 0;
}

int orc_mul_mat(int wtype, const void *w, size_t w_row_bytes, const float *x, float *dst,
                int64_t K, int64_t M, int64_t N, int nthreads) {
    return mm_common(wtype, w, w_row_bytes, x, dst, NULL, K, M, N, nthreads, 0);
}

int orc_mul_mat_q8_1(int wtype, const void *w, size_t w_row_bytes, const float *x, double *dst,
                     int64_t K, int64_t M, int64_t N, int nthreads) {
    return mm_common(wtype, w, w_row_bytes, x, NULL, dst, K, M, N, nthreads, 1);
}

/* ------------------------------------------------------------------ MUL_MAT_ID */
int orc_mul_mat_id(int wtype, const void *as, size_t w_row_bytes, size_t w_expert_bytes,
                   const int32_t *ids, const float *b, float *dst,
                   int64_t K, int64_t M, int64_t n_expert, int64_t n_used, int64_t n_b_rows,
                   int64_t T, int nthreads) {
    /* ggml.c:12812-12912: rows are grouped per expert; per (token, slot) the arithmetic is the
     * MUL_MAT one with src1 row (slot % ne11, token).  We evaluate it pair by pair. */
    for (int64_t t = 0; t < T; t++) {
        for (int64_t e = 0; e < n_used; e++) {
            const int32_t ex = ids[t * n_used + e];
            if (ex < 0 || ex >= n_expert) return -4; /* GGML_ASSERT(i02 >= 0 && i02 < n_as), :12830 */
            const uint8_t *w = (const uint8_t *)as + (size_t)ex * w_expert_bytes;
            const float *x = b + (t * n_b_rows + (e % n_b_rows)) * K;
            int rc = orc_mul_mat(wtype, w, w_row_bytes, x, dst + (t * n_used + e) * M, K, M, 1, nthreads);
            if (rc) return rc;
        }
    }
    return 0;
}

/* ------------------------------------------------------------------ FLASH_ATTN_EXT */
typedef struct {
    const orc_fattn_params *p;
    const uint8_t *q, *k, *v, *mask;
    float *dst; double *dst64;
} fa_ctx;

static float alibi_slope(const orc_fattn_params *p, int64_t h) {
    /* ggml.c:15743-15747, 15766 */
    if (!(p->max_bias > 0.0f)) return 1.0f;
    const uint32_t n_head_log2 = 1u << (uint32_t)floor(log2((double)p->n_head));
    const float m0 = powf(2.0f, -(p->max_bias) / n_head_log2);
    const float m1 = powf(2.0f, -(p->max_bias / 2.0f) / n_head_log2);
    return (uint32_t)h < n_head_log2 ? powf(m0, (float)(h + 1)) : powf(m1, (float)(2 * (h - n_head_log2) + 1));
}

static void fa_rows(void *vp, int64_t r0, int64_t r1) {
    fa_ctx *c = (fa_ctx *)vp;
    const orc_fattn_params *p = c->p;
    const int64_t D = p->D;
    float scale = p->scale;
    if (p->logit_softcap != 0) scale /= p->logit_softcap; /* :15738-15740 */
    const int64_t gqa = p->n_head / p->n_head_kv;

    float    *VKQ32 = (float *)malloc(sizeof(float) * D);
    float    *V32   = (float *)malloc(sizeof(float) * D);
    uint16_t *VKQ16 = (uint16_t *)malloc(sizeof(uint16_t) * D);
    uint8_t  *Qq    = (uint8_t *)malloc(sizeof(float) * D + 64);

    for (int64_t ir = r0; ir < r1; ir++) {
        const int64_t h = ir / p->N, iq1 = ir % p->N; /* :15761-15763 with neq3 == 1 */
        const float slope = alibi_slope(p, h);
        float S = 0.0f, M = -INFINITY;
        const int v_f16 = p->type_v == ORC_TYPE_F16;
        if (v_f16) memset(VKQ16, 0, sizeof(uint16_t) * D); else memset(VKQ32, 0, sizeof(float) * D);
        const uint8_t *mp = c->mask ? c->mask + iq1 * p->mask_nb1 : NULL;
        const int64_t hk = h / gqa;
        const float *pq = (const float *)(c->q + iq1 * p->q_nb1 + h * p->q_nb2);
        /* Q -> K's vec_dot_type (:15788): f16 for f16 K, q8_0 for a q8_0 or q4_0 K */
        if (p->type_k == ORC_TYPE_F16) {
            for (int64_t d = 0; d < D; d++) wr16(Qq + 2 * d, orc_fp32_to_fp16(pq[d]));
        } else {
            orc_quantize_row(ORC_TYPE_Q8_0, pq, Qq, D);
        }
        for (int64_t ic = 0; ic < p->n_kv; ic++) {
            const float mv = mp ? slope * rdh(mp + 2 * ic) : 0.0f;
            if (mv == -INFINITY) continue; /* :15795 */
            const uint8_t *kr = c->k + ic * p->k_nb1 + hk * p->k_nb2;
            float s;
            if (p->type_k == ORC_TYPE_F16) {
                /* ggml_vec_dot_f16 scalar branch: double accumulator of float products (ggml.c:2265,2300) */
                double acc = 0.0;
                for (int64_t d = 0; d < D; d++) acc += (double)(rdh(kr + 2 * d) * rdh(Qq + 2 * d));
                s = (float)acc;
            } else {
                s = orc_vec_dot(p->type_k, D, kr, Qq); /* q8_0 . q8_0 or q4_0 . q8_0: both have vec_dot_type q8_0 */
            }
            s = s * scale;
            if (p->logit_softcap != 0.0f) s = p->logit_softcap * tanhf(s);
            s += mv;
            const float Mold = M;
            float ms = 1.0f, vs = 1.0f;
            const uint8_t *vr = c->v + ic * p->v_nb1 + hk * p->v_nb2;
            const int newmax = s > M; /* :15819, :15836 */
            if (newmax) { M = s; ms = expf(Mold - M); } else { vs = expf(s - M); }
            if (v_f16) {
                if (newmax) { /* ggml_vec_scale_f16, ggml.c:2501-2528 */
                    for (int64_t d = 0; d < D; d++) VKQ16[d] = orc_fp32_to_fp16(orc_fp16_to_fp32(VKQ16[d]) * ms);
                }
                for (int64_t d = 0; d < D; d++) { /* ggml_vec_mad_f16, ggml.c:2386-2415 */
                    VKQ16[d] = orc_fp32_to_fp16(orc_fp16_to_fp32(VKQ16[d]) + rdh(vr + 2 * d) * vs);
                }
            } else {
                if (newmax) {
                    for (int64_t d = 0; d < D; d++) VKQ32[d] *= ms;
                }
                orc_dequantize_row(p->type_v, vr, V32, D);
                for (int64_t d = 0; d < D; d++) VKQ32[d] += V32[d] * vs;
            }
            S = S * ms + vs;
        }
        if (v_f16) for (int64_t d = 0; d < D; d++) VKQ32[d] = orc_fp16_to_fp32(VKQ16[d]);
        const float S_inv = 1.0f / S;
        float *out = c->dst + (iq1 * p->n_head + h) * D; /* permute(0,2,1,3), :15875 */
        for (int64_t d = 0; d < D; d++) out[d] = VKQ32[d] * S_inv;
    }
    free(VKQ32); free(V32); free(VKQ16); free(Qq);
}

int orc_flash_attn_ext(const orc_fattn_params *p, const void *q, const void *k, const void *v,
                       const void *mask, float *dst, int nthreads) {
    if (p->type_k != ORC_TYPE_F16 && p->type_k != ORC_TYPE_Q8_0 && p->type_k != ORC_TYPE_Q4_0) return -1;
    if (p->type_v != ORC_TYPE_F16 && p->type_v != ORC_TYPE_Q8_0 && p->type_v != ORC_TYPE_Q4_0) return -1;
    if (p->n_head % p->n_head_kv) return -1;
    /* a q8_0 row is whole blocks of 32; an f16 cache has no block constraint (head sizes 80 and 112 exist) */
    if ((p->type_k != ORC_TYPE_F16 || p->type_v != ORC_TYPE_F16) && p->D % 32) return -1;
    fa_ctx c = { p, (const uint8_t *)q, (const uint8_t *)k, (const uint8_t *)v, (const uint8_t *)mask, dst, NULL };
    orc_parallel_for(p->n_head * p->N, nthreads, fa_rows, &c);
    return 0;
}

static void fa_rows_f64(void *vp, int64_t r0, int64_t r1) {
    fa_ctx *c = (fa_ctx *)vp;
    const orc_fattn_params *p = c->p;
    const int64_t D = p->D, gqa = p->n_head / p->n_head_kv;
    double scale = p->scale;
    if (p->logit_softcap != 0) scale /= p->logit_softcap;
    float  *row = (float *)malloc(sizeof(float) * D);
    double *sc  = (double *)malloc(sizeof(double) * p->n_kv);
    double *acc = (double *)malloc(sizeof(double) * D);
    for (int64_t ir = r0; ir < r1; ir++) {
        const int64_t h = ir / p->N, iq1 = ir % p->N, hk = h / gqa;
        const double slope = alibi_slope(p, h);
        const uint8_t *mp = c->mask ? c->mask + iq1 * p->mask_nb1 : NULL;
        const float *pq = (const float *)(c->q + iq1 * p->q_nb1 + h * p->q_nb2);
        double mx = -INFINITY;
        for (int64_t ic = 0; ic < p->n_kv; ic++) {
            const double mv = mp ? slope * (double)rdh(mp + 2 * ic) : 0.0;
            if (mv == -INFINITY) { sc[ic] = -INFINITY; continue; }
            orc_dequantize_row(p->type_k, c->k + ic * p->k_nb1 + hk * p->k_nb2, row, D);
            double s = 0;
            for (int64_t d = 0; d < D; d++) s += (double)row[d] * (double)pq[d];
            s *= scale;
            if (p->logit_softcap != 0.0f) s = p->logit_softcap * tanh(s);
            s += mv;
            sc[ic] = s;
            if (s > mx) mx = s;
        }
        double S = 0;
        for (int64_t d = 0; d < D; d++) acc[d] = 0;
        for (int64_t ic = 0; ic < p->n_kv; ic++) {
            if (sc[ic] == -INFINITY) continue;
            const double pw = exp(sc[ic] - mx);
            S += pw;
            orc_dequantize_row(p->type_v, c->v + ic * p->v_nb1 + hk * p->v_nb2, row, D);
            for (int64_t d = 0; d < D; d++) acc[d] += pw * (double)row[d];
        }
        double *out = c->dst64 + (iq1 * p->n_head + h) * D;
        for (int64_t d = 0; d < D; d++) out[d] = acc[d] / S;
    }
    free(row); free(sc); free(acc);
}

int orc_flash_attn_f64(const orc_fattn_params *p, const void *q, const void *k, const void *v,
                       const void *mask, double *dst, int nthreads) {
    if (p->n_head % p->n_head_kv) return -1;
    fa_ctx c = { p, (const uint8_t *)q, (const uint8_t *)k, (const uint8_t *)v, (const uint8_t *)mask, NULL, dst };
    orc_parallel_for(p->n_head * p->N, nthreads, fa_rows_f64, &c);
    return 0;
}
