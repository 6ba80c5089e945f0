// ggml_oracle.c -- TEST INFRASTRUCTURE ONLY (see ggml_oracle.h).  Plain-C restatement of the
// reference's ggml-cpu hot path; every function cites the reference lines it follows.  Nothing
// here is copied: the block layouts are the GGUF file format, the loops are re-derived from it.
// Parity: PINNED against oracle/_ref (the reference compiled in place) by tests/test_oracle.py
// and the committed vectors under tests/golden/.
#include "ggml_oracle.h"
#include <math.h>
#include <string.h>
#include <stdlib.h>

#define QK  32
#define QKK 256

#pragma pack(push, 1)
typedef struct { uint16_t d; uint8_t qs[16]; } q4_0_t;                                           // ggml-common.h:170-175
typedef struct { uint16_t d; int8_t  qs[32]; } q8_0_t;                                           // :219-224
typedef struct { uint16_t d; uint16_t s; int8_t qs[32]; } q8_1_t;                                // :226-237
typedef struct { uint16_t d; uint16_t dmin; uint8_t scales[12]; uint8_t qs[128]; } q4_K_t;        // :295-306
typedef struct { uint16_t d; uint16_t dmin; uint8_t scales[12]; uint8_t qh[32]; uint8_t qs[128]; } q5_K_t;  // :308-320
typedef struct { uint8_t ql[128]; uint8_t qh[64]; int8_t scales[16]; uint16_t d; } q6_K_t;        // :330-336
typedef struct { float d; int8_t qs[256]; int16_t bsums[16]; } q8_K_t;                            // :339-343
#pragma pack(pop)

int64_t orc_blck_size(int type) {
    switch (type) {
        case ORC_F32: case ORC_F16: case ORC_BF16: return 1;
        case ORC_Q4_0: case ORC_Q8_0: case ORC_Q8_1: return 32;
        case ORC_Q4_K: case ORC_Q5_K: case ORC_Q6_K: case ORC_Q8_K: return 256;
    }
    return 0;
}
size_t orc_type_size(int type) {
    switch (type) {
        case ORC_F32: return 4; case ORC_F16: case ORC_BF16: return 2;
        case ORC_Q4_0: return 18; case ORC_Q8_0: return 34; case ORC_Q8_1: return 36;
        case ORC_Q4_K: return 144; case ORC_Q5_K: return 176; case ORC_Q6_K: return 210; case ORC_Q8_K: return 292;
    }
    return 0;
}
size_t orc_row_size(int type, int64_t k) { return (size_t) (k / orc_blck_size(type)) * orc_type_size(type); }

// ---- fp16 <-> fp32, IEEE round-to-nearest-even (what F16C / GGML_FP32_TO_FP16 produce) ----------
static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float    u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

uint16_t orc_fp32_to_fp16(float f) {
    const uint32_t x = f2u(f);
    const uint32_t sign = (x >> 16) & 0x8000u;
    const uint32_t ax = x & 0x7fffffffu;
    if (ax >= 0x7f800000u) return (uint16_t) (sign | (ax > 0x7f800000u ? 0x7e00u : 0x7c00u));   // nan / inf
    if (ax >= 0x477ff000u) return (uint16_t) (sign | 0x7c00u);                                   // rounds to inf (>= 65520)
    if (ax < 0x33000001u)  return (uint16_t) sign;                                               // <= 2^-25 -> 0 (tie to even)
    int32_t e = (int32_t) (ax >> 23) - 127;
    uint32_t m = (ax & 0x7fffffu) | 0x800000u;
    int shift;
    uint32_t base;
    if (e >= -14) { shift = 13; base = (uint32_t) (e + 15) << 10; m &= 0x7fffffu; }   // normal: drop implicit bit
    else          { shift = 13 + (-14 - e); base = 0; }                                 // subnormal: keep it
    uint32_t q = m >> shift;
    const uint32_t rem = m & ((1u << shift) - 1u), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (q & 1u))) q++;
    return (uint16_t) (sign | (base + q));   // a mantissa carry correctly bumps the exponent
}
float orc_fp16_to_fp32(uint16_t h) {
    const uint32_t sign = ((uint32_t) h & 0x8000u) << 16;
    const uint32_t e = (h >> 10) & 0x1f, m = h & 0x3ffu;
    if (e == 0) {
        if (m == 0) return u2f(sign);
        return (sign ? -1.0f : 1.0f) * (float) m * 5.9604644775390625e-8f;   // m * 2^-24, exact
    }
    if (e == 31) return u2f(sign | 0x7f800000u | (m << 13));
    return u2f(sign | ((e + 112) << 23) | (m << 13));
}
void orc_fp32_to_fp16_row(const float * x, uint16_t * y, int64_t n) { for (int64_t i = 0; i < n; ++i) y[i] = orc_fp32_to_fp16(x[i]); }
void orc_fp16_to_fp32_row(const uint16_t * x, float * y, int64_t n) { for (int64_t i = 0; i < n; ++i) y[i] = orc_fp16_to_fp32(x[i]); }
static inline float bf16_to_fp32(uint16_t h) { return u2f((uint32_t) h << 16); }

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
// cvtps2dq semantics: out-of-range and NaN inputs give the "integer indefinite" value 0x80000000
static inline int cvt_x86(float r) { return (r >= -2147483648.0f && r < 2147483648.0f) ? (int) r : (int) 0x80000000u; }

// ---- activation quantizers ----------------------------------------------------------------------
// X86: ggml-cpu/arch/x86/quants.c:388-449 -- id = 127/amax, _mm256_round_ps(NEAREST) = ties-to-even,
//      packs saturate to int8, s = fp16(d * (float) sum_q).
// REF: ggml-quants.c:225-258 -- id = 1/d, roundf (ties away), (int8_t) cast, s = fp16(sum * d).
// MAXPS / MAXSS and the scalar MAX() macro alike: a > b ? a : b, i.e. the SECOND operand whenever one of them is NaN
static inline float mx(float a, float b) { return a > b ? a : b; }
// amax exactly as the AVX2 body computes it (arch/x86/quants.c:402-412): four 8-lane vectors folded with _mm256_max_ps, then
// _mm_max_ps(high half, low half), _mm_max_ps(max4, movehl), _mm_max_ss(max4, movehdup).  For NaN-free input this is the plain
// maximum; with NaNs the result depends on where they sit, and this order reproduces it.
static float amax_x86(const float * x) {
    float m[8];
    for (int j = 0; j < 8; ++j) m[j] = mx(mx(mx(fabsf(x[j]), fabsf(x[8 + j])), fabsf(x[16 + j])), fabsf(x[24 + j]));
    float q[4];
    for (int k = 0; k < 4; ++k) q[k] = mx(m[4 + k], m[k]);
    return mx(mx(q[0], q[2]), mx(q[1], q[3]));
}
void orc_quantize_row_q8_1(const float * x, void * vy, int64_t k, int variant) {
    q8_1_t * y = (q8_1_t *) vy;
    for (int64_t i = 0; i < k / QK; ++i) {
        float amax = 0.0f;
        if (variant == ORC_Q8_1_X86) amax = amax_x86(x + i * QK);
        else for (int j = 0; j < QK; ++j) amax = mx(amax, fabsf(x[i * QK + j]));        // ggml-quants.c:233-236, MAX(amax, fabsf(v))
        const float d = amax / 127.0f;
        y[i].d = orc_fp32_to_fp16(d);
        int sum = 0;
        if (variant == ORC_Q8_1_X86) {
            const float id = amax != 0.0f ? 127.0f / amax : 0.0f;
            uint32_t usum = 0;                                     // the AVX2 body sums the int32 lanes BEFORE the saturating packs,
            for (int j = 0; j < QK; ++j) {                         // and cvtps_epi32 gives 0x80000000 for inf/nan/overflow (a denormal amax)
                const int q32 = cvt_x86(nearbyintf(x[i * QK + j] * id));   // default rounding mode = nearest-even
                y[i].qs[j] = (int8_t) clampi(q32, -128, 127); usum += (uint32_t) q32;
            }
            sum = (int) usum;
            y[i].s = orc_fp32_to_fp16(d * (float) sum);
        } else {
            const float id = d != 0.0f ? 1.0f / d : 0.0f;
            for (int j = 0; j < QK; ++j) {
                const int8_t q = (int8_t) roundf(x[i * QK + j] * id);
                y[i].qs[j] = q; sum += q;
            }
            y[i].s = orc_fp32_to_fp16((float) sum * d);
        }
    }
}
// ggml-cpu/arch/x86/quants.c:290-386 (same arithmetic as the q8_1 x86 body, no sum)
void orc_quantize_row_q8_0(const float * x, void * vy, int64_t k) {
    q8_0_t * y = (q8_0_t *) vy;
    for (int64_t i = 0; i < k / QK; ++i) {
        const float amax = amax_x86(x + i * QK);
        const float d = amax / 127.0f;
        const float id = amax != 0.0f ? 127.0f / amax : 0.0f;
        y[i].d = orc_fp32_to_fp16(d);
        for (int j = 0; j < QK; ++j) y[i].qs[j] = (int8_t) clampi(cvt_x86(nearbyintf(x[i * QK + j] * id)), -128, 127);
    }
}
// ggml-quants.c:2555-2591; nearest_int() there is the 12582912.f magic-number trick = ties-to-even
void orc_quantize_row_q8_K(const float * x, void * vy, int64_t k) {
    q8_K_t * y = (q8_K_t *) vy;
    for (int64_t i = 0; i < k / QKK; ++i, x += QKK) {
        float max = 0.0f, amax = 0.0f;
        for (int j = 0; j < QKK; ++j) { const float ax = fabsf(x[j]); if (ax > amax) { amax = ax; max = x[j]; } }
        if (amax == 0.0f) { y[i].d = 0.0f; memset(y[i].qs, 0, QKK); memset(y[i].bsums, 0, sizeof(y[i].bsums)); continue; }
        const float iscale = -127.0f / max;
        for (int j = 0; j < QKK; ++j) { const int v = (int) nearbyintf(iscale * x[j]); y[i].qs[j] = (int8_t) (v > 127 ? 127 : v); }
        for (int j = 0; j < QKK / 16; ++j) { int s = 0; for (int l = 0; l < 16; ++l) s += y[i].qs[j * 16 + l]; y[i].bsums[j] = (int16_t) s; }
        y[i].d = 1.0f / iscale;
    }
}
// ggml-quants.c:36-71
void orc_quantize_row_q4_0_ref(const float * x, void * vy, int64_t k) {
    q4_0_t * y = (q4_0_t *) vy;
    for (int64_t i = 0; i < k / QK; ++i) {
        float amax = 0.0f, max = 0.0f;
        for (int j = 0; j < QK; ++j) { const float v = x[i * QK + j]; if (amax < fabsf(v)) { amax = fabsf(v); max = v; } }
        const float d = max / -8.0f;
        const float id = d != 0.0f ? 1.0f / d : 0.0f;
        y[i].d = orc_fp32_to_fp16(d);
        for (int j = 0; j < QK / 2; ++j) {
            const float x0 = x[i * QK + j] * id, x1 = x[i * QK + QK / 2 + j] * id;
            int a = (int8_t) (x0 + 8.5f), b = (int8_t) (x1 + 8.5f);
            if (a > 15) a = 15;
            if (b > 15) b = 15;
            y[i].qs[j] = (uint8_t) ((uint8_t) a | ((uint8_t) b << 4));
        }
    }
}
// ggml-quants.c:199-222
void orc_quantize_row_q8_0_ref(const float * x, void * vy, int64_t k) {
    q8_0_t * y = (q8_0_t *) vy;
    for (int64_t i = 0; i < k / QK; ++i) {
        float amax = 0.0f;
        for (int j = 0; j < QK; ++j) { const float a = fabsf(x[i * QK + j]); if (a > amax) amax = a; }
        const float d = amax / 127.0f;
        const float id = d != 0.0f ? 1.0f / d : 0.0f;
        y[i].d = orc_fp32_to_fp16(d);
        for (int j = 0; j < QK; ++j) y[i].qs[j] = (int8_t) roundf(x[i * QK + j] * id);
    }
}

// ---- weight decode ---------------------------------------------------------------------------------
// get_scale_min_k4, ggml-quants.c:703-710: 8 x (6-bit scale, 6-bit min) packed in 12 bytes
static inline void scale_min_k4(int j, const uint8_t * q, int * sc, int * m) {
    if (j < 4) { *sc = q[j] & 63; *m = q[j + 4] & 63; }
    else       { *sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4); *m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}
// integer weight of element e (0..255) of a K-quant block, plus its sub-block index (32 wide for q4_K/q5_K, 16 for q6_K)
static inline int q4_K_q(const q4_K_t * b, int e) { const int g = e >> 6, r = e & 63; const uint8_t v = b->qs[g * 32 + (r & 31)]; return r < 32 ? (v & 0xF) : (v >> 4); }
static inline int q5_K_q(const q5_K_t * b, int e) {
    const int g = e >> 6, r = e & 63, l = r & 31; const uint8_t v = b->qs[g * 32 + l];
    const int hi = (b->qh[l] >> (2 * g + (r >= 32))) & 1;
    return (r < 32 ? (v & 0xF) : (v >> 4)) + 16 * hi;
}
static inline int q6_K_q(const q6_K_t * b, int e) {          // ggml-quants.c:1762-1791
    const int n = e >> 7, r = e & 127, l = r & 31, grp = r >> 5;
    const uint8_t lo = b->ql[n * 64 + (grp & 1) * 32 + l];
    const int nib = grp < 2 ? (lo & 0xF) : (lo >> 4);
    const int hi = (b->qh[n * 32 + l] >> (2 * grp)) & 3;
    return (nib | (hi << 4)) - 32;
}

int orc_dequantize_row(int type, const void * vx, float * y, int64_t k) {
    switch (type) {
        case ORC_F32: memcpy(y, vx, (size_t) k * 4); return 0;
        case ORC_F16: orc_fp16_to_fp32_row((const uint16_t *) vx, y, k); return 0;
        case ORC_BF16: for (int64_t i = 0; i < k; ++i) y[i] = bf16_to_fp32(((const uint16_t *) vx)[i]); return 0;
        case ORC_Q4_0: {                                        // ggml-quants.c:307-324
            const q4_0_t * x = (const q4_0_t *) vx;
            for (int64_t i = 0; i < k / QK; ++i) {
                const float d = orc_fp16_to_fp32(x[i].d);
                for (int j = 0; j < 16; ++j) {
                    y[i * QK + j]      = (float) ((x[i].qs[j] & 0xF) - 8) * d;
                    y[i * QK + j + 16] = (float) ((x[i].qs[j] >> 4) - 8) * d;
                }
            }
            return 0;
        }
        case ORC_Q8_0: {                                        // ggml-quants.c:401-415
            const q8_0_t * x = (const q8_0_t *) vx;
            for (int64_t i = 0; i < k / QK; ++i) { const float d = orc_fp16_to_fp32(x[i].d); for (int j = 0; j < QK; ++j) y[i * QK + j] = (float) x[i].qs[j] * d; }
            return 0;
        }
        case ORC_Q4_K: {                                        // ggml-quants.c:1352-1374
            const q4_K_t * x = (const q4_K_t *) vx;
            for (int64_t i = 0; i < k / QKK; ++i) {
                const float d = orc_fp16_to_fp32(x[i].d), mn = orc_fp16_to_fp32(x[i].dmin);
                for (int e = 0; e < QKK; ++e) { int sc, m; scale_min_k4(e >> 5, x[i].scales, &sc, &m); y[i * QKK + e] = (d * (float) sc) * (float) q4_K_q(&x[i], e) - mn * (float) m; }
            }
            return 0;
        }
        case ORC_Q5_K: {                                        // ggml-quants.c:1554-1580
            const q5_K_t * x = (const q5_K_t *) vx;
            for (int64_t i = 0; i < k / QKK; ++i) {
                const float d = orc_fp16_to_fp32(x[i].d), mn = orc_fp16_to_fp32(x[i].dmin);
                for (int e = 0; e < QKK; ++e) { int sc, m; scale_min_k4(e >> 5, x[i].scales, &sc, &m); y[i * QKK + e] = (d * (float) sc) * (float) q5_K_q(&x[i], e) - mn * (float) m; }
            }
            return 0;
        }
        case ORC_Q6_K: {                                        // ggml-quants.c:1762-1791
            const q6_K_t * x = (const q6_K_t *) vx;
            for (int64_t i = 0; i < k / QKK; ++i) {
                const float d = orc_fp16_to_fp32(x[i].d);
                for (int e = 0; e < QKK; ++e) y[i * QKK + e] = d * (float) x[i].scales[e >> 4] * (float) q6_K_q(&x[i], e);
            }
            return 0;
        }
    }
    return -1;
}

// ---- row dot products ---------------------------------------------------------------------------------
// CPU path (ORC_ACT_CPU): ggml_vec_dot_q4_0_q8_0 / q8_0_q8_0 / q4_K_q8_K / q5_K_q8_K / q6_K_q8_K,
// ggml-cpu/quants.c:115,305,550,625,708 -- integer sums per block, fp32 across blocks.
static float dot_cpu(int type, int64_t K, const void * w, const void * act) {
    float sumf = 0.0f;
    switch (type) {
        case ORC_Q4_0: {
            const q4_0_t * x = (const q4_0_t *) w; const q8_0_t * y = (const q8_0_t *) act;
            for (int64_t i = 0; i < K / QK; ++i) {
                int s = 0;
                for (int j = 0; j < 16; ++j) s += ((x[i].qs[j] & 0xF) - 8) * y[i].qs[j] + ((x[i].qs[j] >> 4) - 8) * y[i].qs[j + 16];
                sumf += (float) s * orc_fp16_to_fp32(x[i].d) * orc_fp16_to_fp32(y[i].d);
            }
            return sumf;
        }
        case ORC_Q8_0: {
            const q8_0_t * x = (const q8_0_t *) w; const q8_0_t * y = (const q8_0_t *) act;
            for (int64_t i = 0; i < K / QK; ++i) {
                int s = 0;
                for (int j = 0; j < QK; ++j) s += x[i].qs[j] * y[i].qs[j];
                sumf += (float) s * (orc_fp16_to_fp32(x[i].d) * orc_fp16_to_fp32(y[i].d));
            }
            return sumf;
        }
        case ORC_Q4_K: case ORC_Q5_K: {
            const q8_K_t * y = (const q8_K_t *) act;
            for (int64_t i = 0; i < K / QKK; ++i) {
                const uint8_t * blk = (const uint8_t *) w + i * orc_type_size(type);
                const q4_K_t * b4 = (const q4_K_t *) blk; const q5_K_t * b5 = (const q5_K_t *) blk;
                int32_t si = 0, smin = 0;
                for (int j = 0; j < 8; ++j) {
                    int sc, m; scale_min_k4(j, b4->scales, &sc, &m);     // same header layout for q4_K and q5_K
                    int s = 0;
                    for (int l = 0; l < 32; ++l) s += (type == ORC_Q4_K ? q4_K_q(b4, j * 32 + l) : q5_K_q(b5, j * 32 + l)) * y[i].qs[j * 32 + l];
                    si += sc * s;
                    smin += m * (y[i].bsums[2 * j] + y[i].bsums[2 * j + 1]);
                }
                sumf += orc_fp16_to_fp32(b4->d) * y[i].d * (float) si;
                sumf -= orc_fp16_to_fp32(b4->dmin) * y[i].d * (float) smin;
            }
            return sumf;
        }
        case ORC_Q6_K: {
            const q6_K_t * x = (const q6_K_t *) w; const q8_K_t * y = (const q8_K_t *) act;
            for (int64_t i = 0; i < K / QKK; ++i) {
                int32_t si = 0;
                for (int j = 0; j < 16; ++j) { int s = 0; for (int l = 0; l < 16; ++l) s += q6_K_q(&x[i], j * 16 + l) * y[i].qs[j * 16 + l]; si += x[i].scales[j] * s; }
                sumf += orc_fp16_to_fp32(x[i].d) * y[i].d * (float) si;
            }
            return sumf;
        }
    }
    return NAN;
}

// B200 GEMV numerics (ORC_ACT_Q8_1): the same integer sums against block_q8_1, with the q4_0 "-8" and the
// K-quant mins folded through s = d*sum(q) (super-llama.cpp_b200/csrc/b200_gemv.cu); the reference's
// own GPU formulas are ggml-cuda/vecdotq.cuh:103-122,231-245,461-484.
static float dot_q8_1(int type, int64_t K, const void * w, const q8_1_t * y) {
    float sumf = 0.0f;
    switch (type) {
        case ORC_Q4_0: {
            const q4_0_t * x = (const q4_0_t *) w;
            for (int64_t i = 0; i < K / QK; ++i) {
                int s = 0;
                for (int j = 0; j < 16; ++j) s += (x[i].qs[j] & 0xF) * y[i].qs[j] + (x[i].qs[j] >> 4) * y[i].qs[j + 16];
                sumf += orc_fp16_to_fp32(x[i].d) * (orc_fp16_to_fp32(y[i].d) * (float) s - 8.0f * orc_fp16_to_fp32(y[i].s));
            }
            return sumf;
        }
        case ORC_Q8_0: {
            const q8_0_t * x = (const q8_0_t *) w;
            for (int64_t i = 0; i < K / QK; ++i) {
                int s = 0;
                for (int j = 0; j < QK; ++j) s += x[i].qs[j] * y[i].qs[j];
                sumf += orc_fp16_to_fp32(x[i].d) * orc_fp16_to_fp32(y[i].d) * (float) s;
            }
            return sumf;
        }
        case ORC_Q4_K: case ORC_Q5_K: {
            for (int64_t i = 0; i < K / QKK; ++i) {
                const uint8_t * blk = (const uint8_t *) w + i * orc_type_size(type);
                const q4_K_t * b4 = (const q4_K_t *) blk; const q5_K_t * b5 = (const q5_K_t *) blk;
                const float d = orc_fp16_to_fp32(b4->d), dmin = orc_fp16_to_fp32(b4->dmin);
                for (int j = 0; j < 8; ++j) {
                    int sc, m; scale_min_k4(j, b4->scales, &sc, &m);
                    const q8_1_t * yb = &y[i * 8 + j];
                    int s = 0;
                    for (int l = 0; l < 32; ++l) s += (type == ORC_Q4_K ? q4_K_q(b4, j * 32 + l) : q5_K_q(b5, j * 32 + l)) * yb->qs[l];
                    sumf += d * (float) sc * orc_fp16_to_fp32(yb->d) * (float) s - dmin * (float) m * orc_fp16_to_fp32(yb->s);
                }
            }
            return sumf;
        }
        case ORC_Q6_K: {
            const q6_K_t * x = (const q6_K_t *) w;
            for (int64_t i = 0; i < K / QKK; ++i) {
                const float d = orc_fp16_to_fp32(x[i].d);
                for (int j = 0; j < 16; ++j) {
                    const q8_1_t * yb = &y[i * 8 + (j >> 1)];
                    int s = 0;
                    for (int l = 0; l < 16; ++l) s += q6_K_q(&x[i], j * 16 + l) * yb->qs[(j & 1) * 16 + l];
                    sumf += d * (float) x[i].scales[j] * orc_fp16_to_fp32(yb->d) * (float) s;
                }
            }
            return sumf;
        }
    }
    return NAN;
}

static int is_float_type(int t) { return t == ORC_F32 || t == ORC_F16 || t == ORC_BF16; }

// one dst column block: rows [0,M) of weight matrix wm against activation column xc
static int mat_col(int type, const uint8_t * wm, int64_t K, int64_t M, const float * xc, float * dcol, int act, void * scratch, float * wrow) {
    const size_t rs = orc_row_size(type, K);
    if (act == ORC_ACT_F64 || is_float_type(type)) {
        // float weights on the CPU: vec_dot_type f16 for f16/bf16 weights rounds the activations (ggml-cpu.c:204-232);
        // we keep the exact product -- the oracle for float weights is only used at NMSE tolerance.
        for (int64_t r = 0; r < M; ++r) {
            if (orc_dequantize_row(type, wm + r * rs, wrow, K)) return -1;
            double s = 0.0;
            for (int64_t i = 0; i < K; ++i) s += (double) wrow[i] * (double) xc[i];
            dcol[r] = (float) s;
        }
        return 0;
    }
    if (act == ORC_ACT_Q8_1) {
        orc_quantize_row_q8_1(xc, scratch, K, ORC_Q8_1_X86);
        for (int64_t r = 0; r < M; ++r) dcol[r] = dot_q8_1(type, K, wm + r * rs, (const q8_1_t *) scratch);
        return 0;
    }
    if (type == ORC_Q4_0 || type == ORC_Q8_0) orc_quantize_row_q8_0(xc, scratch, K);      // vec_dot_type, ggml-cpu.c:217-231
    else                                       orc_quantize_row_q8_K(xc, scratch, K);      // ggml-cpu.c:262-291
    for (int64_t r = 0; r < M; ++r) dcol[r] = dot_cpu(type, K, wm + r * rs, scratch);
    return 0;
}

int orc_mul_mat(int type, const void * w, int64_t K, int64_t M, int64_t ne02, int64_t ne03,
                const float * x, int64_t N, int64_t ne12, int64_t ne13, float * dst, int act) {
    if (orc_blck_size(type) == 0 || K % orc_blck_size(type) != 0 || ne12 % ne02 != 0 || ne13 % ne03 != 0) return -1;
    const size_t rs = orc_row_size(type, K);
    const int64_t r2 = ne12 / ne02, r3 = ne13 / ne03;                          // ggml-cpu.c:1157-1158
    void * scratch = malloc((size_t) K * 2 + 1024);
    float * wrow = (float *) malloc((size_t) K * 4);
    int rc = 0;
    for (int64_t i13 = 0; i13 < ne13 && !rc; ++i13) for (int64_t i12 = 0; i12 < ne12 && !rc; ++i12) {
        const uint8_t * wm = (const uint8_t *) w + ((i13 / r3) * ne02 + (i12 / r2)) * M * rs;
        for (int64_t n = 0; n < N && !rc; ++n) {
            const int64_t col = (i13 * ne12 + i12) * N + n;
            rc = mat_col(type, wm, K, M, x + col * K, dst + col * M, act, scratch, wrow);
        }
    }
    free(scratch); free(wrow);
    return rc;
}

int orc_mul_mat_id(int type, const void * as, int64_t K, int64_t M, int64_t n_expert,
                   const float * b, int64_t ne11, int64_t n_tokens,
                   const int32_t * ids, int64_t n_used, float * dst, int act) {
    if (orc_blck_size(type) == 0 || K % orc_blck_size(type) != 0 || ne11 <= 0 || n_used % ne11 != 0) return -1;
    const size_t rs = orc_row_size(type, K);
    void * scratch = malloc((size_t) K * 2 + 1024);
    float * wrow = (float *) malloc((size_t) K * 4);
    int rc = 0;
    for (int64_t t = 0; t < n_tokens && !rc; ++t) for (int64_t e = 0; e < n_used && !rc; ++e) {
        const int32_t ex = ids[t * n_used + e];
        if (ex < 0 || ex >= n_expert) { rc = -2; break; }                     // reference asserts, ggml-cpu.c:1597
        const uint8_t * wm = (const uint8_t *) as + (size_t) ex * M * rs;
        rc = mat_col(type, wm, K, M, b + (t * ne11 + e % ne11) * K, dst + (t * n_used + e) * M, act, scratch, wrow);
    }
    free(scratch); free(wrow);
    return rc;
}

// ---- flash attention -------------------------------------------------------------------------------------
static inline float rnd16(float f) { return orc_fp16_to_fp32(orc_fp32_to_fp16(f)); }

int orc_flash_attn_ext(const float * q, const uint16_t * k, const uint16_t * v, const uint16_t * mask,
                       const float * sinks, float * dst,
                       int64_t DK, int64_t DV, int64_t N, int64_t H, int64_t S,
                       int64_t KV, int64_t Hk, int64_t Sk, int64_t Nm, int64_t Hm, int64_t Sm,
                       float scale, float max_bias, float logit_softcap, int faithful) {
    if (Hk <= 0 || Sk <= 0 || H % Hk != 0 || S % Sk != 0) return -1;
    const int64_t rk2 = H / Hk, rk3 = S / Sk;
    if (logit_softcap != 0.0f) scale /= logit_softcap;                                         // ops.cpp:8107-8109
    const uint32_t n_head_log2 = 1u << (uint32_t) floor(log2((double) H));
    const float m0 = powf(2.0f, -(max_bias) / n_head_log2), m1 = powf(2.0f, -(max_bias / 2.0f) / n_head_log2);
    float * qf = (float *) malloc((size_t) DK * 4), * acc = (float *) malloc((size_t) DV * 4);
    for (int64_t iq3 = 0; iq3 < S; ++iq3) for (int64_t iq2 = 0; iq2 < H; ++iq2) for (int64_t iq1 = 0; iq1 < N; ++iq1) {
        const uint32_t h = (uint32_t) iq2;
        const float slope = max_bias > 0.0f ? (h < n_head_log2 ? powf(m0, (float) (h + 1)) : powf(m1, (float) (2 * (h - n_head_log2) + 1))) : 1.0f;
        const float * pq = q + ((iq3 * H + iq2) * N + iq1) * DK;
        for (int64_t d = 0; d < DK; ++d) qf[d] = faithful ? rnd16(pq[d]) : pq[d];              // q_to_vec_dot, ops.cpp:8162
        for (int64_t d = 0; d < DV; ++d) acc[d] = 0.0f;
        const uint16_t * kb = k + ((iq3 / rk3) * Hk + (iq2 / rk2)) * KV * DK;
        const uint16_t * vb = v + ((iq3 / rk3) * Hk + (iq2 / rk2)) * KV * DV;
        const uint16_t * mp = mask ? mask + (((iq3 % Sm) * Hm + (iq2 % Hm)) * Nm + iq1) * KV : NULL;
        float Sum = 0.0f, Mx = -INFINITY;
        for (int64_t ic = 0; ic < KV; ++ic) {
            const float mv = mp ? slope * orc_fp16_to_fp32(mp[ic]) : 0.0f;
            if (mv == -INFINITY) continue;                                                     // ops.cpp:8169-8171
            double dot = 0.0;
            for (int64_t d = 0; d < DK; ++d) dot += (double) orc_fp16_to_fp32(kb[ic * DK + d]) * (double) qf[d];
            float s = (float) dot * scale;
            if (logit_softcap != 0.0f) s = logit_softcap * tanhf(s);
            s += mv;
            const float Mold = Mx;
            float ms = 1.0f, vs = 1.0f;
            if (s > Mx) { Mx = s; ms = expf(Mold - Mx); for (int64_t d = 0; d < DV; ++d) acc[d] = faithful ? rnd16(acc[d] * ms) : acc[d] * ms; }
            else        { vs = expf(s - Mx); }
            for (int64_t d = 0; d < DV; ++d) {
                const float t = acc[d] + orc_fp16_to_fp32(vb[ic * DV + d]) * vs;               // fp16 accumulator on the CPU, ops.cpp:8193-8207
                acc[d] = faithful ? rnd16(t) : t;
            }
            Sum = Sum * ms + vs;
        }
        if (sinks) {                                                                           // ops.cpp:8241-8255
            const float s = sinks[h];
            float ms = 1.0f, vs = 1.0f;
            if (s > Mx) { ms = expf(Mx - s); for (int64_t d = 0; d < DV; ++d) acc[d] *= ms; }
            else        { vs = expf(s - Mx); }
            Sum = Sum * ms + vs;
        }
        const float Sinv = Sum == 0.0f ? 0.0f : 1.0f / Sum;
        float * o = dst + ((iq3 * N + iq1) * H + iq2) * DV;                                    // permuted store, ops.cpp:8270
        for (int64_t d = 0; d < DV; ++d) o[d] = acc[d] * Sinv;
    }
    free(qf); free(acc);
    return 0;
}

// ggml-cuda.cu:777-803: tensor_split[i] = cumulative fraction at which device i starts (NULL = even)
void orc_split_row_range(int64_t nrows, const float * tensor_split, int world, int rank, int64_t rounding,
                         int64_t * row_low, int64_t * row_high) {
    float lo_f = tensor_split ? tensor_split[rank] : (float) rank / (float) world;
    int64_t lo = rank == 0 ? 0 : (int64_t) ((float) nrows * lo_f);
    lo -= lo % rounding;
    int64_t hi;
    if (rank == world - 1) hi = nrows;
    else {
        float hi_f = tensor_split ? tensor_split[rank + 1] : (float) (rank + 1) / (float) world;
        hi = (int64_t) ((float) nrows * hi_f); hi -= hi % rounding;
    }
    *row_low = lo; *row_high = hi;
}
