/*
 * aios_oracle.c — CPU ORACLE (test infrastructure, NOT product code).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may build, load or call this file.  The product (libaios_b200.so) never links it
 * and has no CPU fallback.
 *
 * PARITY UNPINNED.  The reference (MohaMehrzad/aiOS, `runtime` crate) holds no inference
 * arithmetic; it proxies to an un-vendored, un-pinned llama.cpp `llama-server`
 * (reference: scripts/build-llamacpp.sh:89, runtime/src/model_manager.rs:187-204,
 * runtime/src/inference.rs:100,136-142).  llama.cpp's source is not on this machine, so
 * this file restates the *published* ggml CPU algorithm from its documented formats:
 *   - block formats and dequantisation: gguf-py quants.py (llama.cpp's own Python package,
 *     gguf 0.19.0): Q8_0 :396-401, Q4_K :479-522, Q5_K :527-549, Q6_K :554-572 — the
 *     dequantisers here are checked BIT-EXACTLY against that package in tests/.
 *   - activation quantisation (Q8_K / Q8_0), integer block dot products, f16 KV cache,
 *     NORM-style RoPE, RMSNorm, SwiGLU: SURVEY.md Appendix C (ggml "generic" scalar
 *     kernels; summation order = the generic kernel's).
 * None of the reference's own tests pin a logit/token, so there is no
 * golden vector from the reference to check against; tests/golden/ holds vectors made
 * with gguf-py and HF transformers fp32 by the committed generator script.
 *
 * Build: see oracle/Makefile (gcc -O2 -fopenmp -ffp-contract=off -shared).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define QK_K 256
#define QK8_0 32

enum { T_F32 = 0, T_F16 = 1, T_Q8_0 = 8, T_Q4_K = 12, T_Q5_K = 13, T_Q6_K = 14 };

/* ---------------------------------------------------------------- fp16 <-> fp32 (IEEE, RNE) */
static inline float h2f(uint16_t h) {
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1f;
    uint32_t man = h & 0x3ffu;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) {
            bits = sign;
        } else { /* subnormal: normalise */
            int e = -1;
            do { man <<= 1; e++; } while ((man & 0x400u) == 0);
            man &= 0x3ffu;
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | (man << 13);
        }
    } else if (exp == 31) {
        bits = sign | 0x7f800000u | (man << 13);
    } else {
        bits = sign | ((exp + 112) << 23) | (man << 13);
    }
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

static inline uint16_t f2h(float f) {
    uint32_t x;
    memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t absx = x & 0x7fffffffu;
    if (absx >= 0x7f800000u) { /* inf / nan */
        return (uint16_t)(sign | 0x7c00u | ((absx > 0x7f800000u) ? 0x200u : 0));
    }
    if (absx >= 0x477ff000u) { /* rounds to >= 65520 -> inf */
        return (uint16_t)(sign | 0x7c00u);
    }
    if (absx < 0x33000001u) { /* < 2^-25 (or exactly 2^-25 -> ties to even 0) */
        return (uint16_t)sign;
    }
    int32_t e = (int32_t)(absx >> 23) - 127;
    uint32_t m = (absx & 0x7fffffu) | 0x800000u;
    if (e < -14) { /* subnormal half */
        int shift = (-14 - e) + 13; /* 14..24 */
        uint32_t q = m >> shift;
        uint32_t rem = m & ((1u << shift) - 1);
        uint32_t half = 1u << (shift - 1);
        if (rem > half || (rem == half && (q & 1))) q++;
        return (uint16_t)(sign | q);
    }
    uint32_t q = ((uint32_t)(e + 15) << 10) | ((m >> 13) & 0x3ffu);
    uint32_t rem = m & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (q & 1))) q++;
    return (uint16_t)(sign | q);
}

float oracle_h2f(uint16_t h) { return h2f(h); }
uint16_t oracle_f2h(float f) { return f2h(f); }

/* ---------------------------------------------------------------- block layouts */
#pragma pack(push, 1)
typedef struct { uint16_t d; int8_t qs[QK8_0]; } block_q8_0;                              /* 34 B  */
typedef struct { uint16_t d; uint16_t dmin; uint8_t scales[12]; uint8_t qs[QK_K / 2]; } block_q4_K; /* 144 B */
typedef struct { uint16_t d; uint16_t dmin; uint8_t scales[12]; uint8_t qh[QK_K / 8]; uint8_t qs[QK_K / 2]; } block_q5_K; /* 176 B */
typedef struct { uint8_t ql[QK_K / 2]; uint8_t qh[QK_K / 4]; int8_t scales[QK_K / 16]; uint16_t d; } block_q6_K; /* 210 B */
typedef struct { float d; int8_t qs[QK_K]; int16_t bsums[QK_K / 16]; } block_q8_K;       /* 292 B */
#pragma pack(pop)

static size_t type_block_bytes(int type) {
    switch (type) {
        case T_F32: return 4;
        case T_F16: return 2;
        case T_Q8_0: return sizeof(block_q8_0);
        case T_Q4_K: return sizeof(block_q4_K);
        case T_Q5_K: return sizeof(block_q5_K);
        case T_Q6_K: return sizeof(block_q6_K);
    }
    return 0;
}
static int type_block_elems(int type) {
    switch (type) {
        case T_F32: case T_F16: return 1;
        case T_Q8_0: return QK8_0;
        case T_Q4_K: case T_Q5_K: case T_Q6_K: return QK_K;
    }
    return 0;
}
static size_t row_bytes(int type, int64_t k) { return (size_t)(k / type_block_elems(type)) * type_block_bytes(type); }

/* 6-bit (scale, min) unpack — gguf-py quants.py:479-501 (Q4_K.get_scale_min) */
static inline void get_scale_min_k4(int j, const uint8_t *q, uint8_t *d, uint8_t *m) {
    if (j < 4) {
        *d = q[j] & 63;
        *m = q[j + 4] & 63;
    } else {
        *d = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4);
        *m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4);
    }
}

/* ---------------------------------------------------------------- dequantize_row_* (bit-exact tier)
 * Each product is rounded to f32 on its own; -ffp-contract=off keeps `d*q - m` un-fused,
 * matching gguf-py (numpy has no FMA).  quants.py:396-401, :504-522, :527-549, :554-572. */
static void dequantize_row_q8_0(const block_q8_0 *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK8_0; i++) {
        const float d = h2f(x[i].d);
        for (int j = 0; j < QK8_0; j++) y[i * QK8_0 + j] = (float)x[i].qs[j] * d;
    }
}

static void dequantize_row_q4_K(const block_q4_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const float d = h2f(x[i].d), dmin = h2f(x[i].dmin);
        const uint8_t *q = x[i].qs;
        for (int p = 0; p < 4; p++) {
            uint8_t sc, m;
            get_scale_min_k4(2 * p, x[i].scales, &sc, &m);
            const float d1 = d * (float)sc, m1 = dmin * (float)m;
            get_scale_min_k4(2 * p + 1, x[i].scales, &sc, &m);
            const float d2 = d * (float)sc, m2 = dmin * (float)m;
            for (int l = 0; l < 32; l++) *y++ = d1 * (float)(q[l] & 0xF) - m1;
            for (int l = 0; l < 32; l++) *y++ = d2 * (float)(q[l] >> 4) - m2;
            q += 32;
        }
    }
}

static void dequantize_row_q5_K(const block_q5_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const float d = h2f(x[i].d), dmin = h2f(x[i].dmin);
        const uint8_t *ql = x[i].qs, *qh = x[i].qh;
        for (int p = 0; p < 4; p++) {
            uint8_t sc, m;
            get_scale_min_k4(2 * p, x[i].scales, &sc, &m);
            const float d1 = d * (float)sc, m1 = dmin * (float)m;
            get_scale_min_k4(2 * p + 1, x[i].scales, &sc, &m);
            const float d2 = d * (float)sc, m2 = dmin * (float)m;
            for (int l = 0; l < 32; l++)
                *y++ = d1 * (float)((ql[l] & 0xF) | (((qh[l] >> (2 * p)) & 1) << 4)) - m1;
            for (int l = 0; l < 32; l++)
                *y++ = d2 * (float)((ql[l] >> 4) | (((qh[l] >> (2 * p + 1)) & 1) << 4)) - m2;
            ql += 32;
        }
    }
}

static void dequantize_row_q6_K(const block_q6_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const float d = h2f(x[i].d);
        const uint8_t *ql = x[i].ql, *qh = x[i].qh;
        const int8_t *sc = x[i].scales;
        for (int h = 0; h < 2; h++) {
            for (int l = 0; l < 32; l++) {
                const int is = l / 16;
                const int8_t q1 = (int8_t)((ql[l] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                const int8_t q2 = (int8_t)((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                const int8_t q3 = (int8_t)((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                const int8_t q4 = (int8_t)((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                y[l + 0] = (d * (float)sc[is + 0]) * (float)q1;
                y[l + 32] = (d * (float)sc[is + 2]) * (float)q2;
                y[l + 64] = (d * (float)sc[is + 4]) * (float)q3;
                y[l + 96] = (d * (float)sc[is + 6]) * (float)q4;
            }
            y += 128; ql += 64; qh += 32; sc += 8;
        }
    }
}

/* public: dequantise n elements (n multiple of the block size) of `type` */
int oracle_dequantize(int type, const void *blocks, int64_t n, float *out) {
    switch (type) {
        case T_F32: memcpy(out, blocks, (size_t)n * 4); return 0;
        case T_F16: for (int64_t i = 0; i < n; i++) out[i] = h2f(((const uint16_t *)blocks)[i]); return 0;
        case T_Q8_0: dequantize_row_q8_0((const block_q8_0 *)blocks, out, n); return 0;
        case T_Q4_K: dequantize_row_q4_K((const block_q4_K *)blocks, out, n); return 0;
        case T_Q5_K: dequantize_row_q5_K((const block_q5_K *)blocks, out, n); return 0;
        case T_Q6_K: dequantize_row_q6_K((const block_q6_K *)blocks, out, n); return 0;
    }
    return -1;
}

/* ---------------------------------------------------------------- activation quantisation
 * SURVEY.md Appendix C.1 (ggml quantize_row_q8_K / quantize_row_q8_0 reference versions). */
static inline int nearest_int(float fval) { /* round-to-nearest-even via the 1.5*2^23 trick */
    float val = fval + 12582912.f;
    int i;
    memcpy(&i, &val, sizeof(int));
    return (i & 0x007fffff) - 0x00400000;
}

void oracle_quantize_row_q8_K(const float *x, void *vy, int64_t k) {
    block_q8_K *y = (block_q8_K *)vy;
    for (int64_t i = 0; i < k / QK_K; i++) {
        float max = 0, amax = 0;
        for (int j = 0; j < QK_K; j++) {
            const float ax = fabsf(x[j]);
            if (ax > amax) { amax = ax; max = x[j]; }
        }
        if (!amax) {
            y[i].d = 0;
            memset(y[i].qs, 0, QK_K);
            memset(y[i].bsums, 0, sizeof(y[i].bsums));
            x += QK_K;
            continue;
        }
        const float iscale = -127.f / max;
        for (int j = 0; j < QK_K; j++) {
            int v = nearest_int(iscale * x[j]);
            y[i].qs[j] = (int8_t)(v < 127 ? v : 127);
        }
        for (int j = 0; j < QK_K / 16; j++) {
            int sum = 0;
            for (int ii = 0; ii < 16; ii++) sum += y[i].qs[j * 16 + ii];
            y[i].bsums[j] = (int16_t)sum;
        }
        y[i].d = 1 / iscale;
        x += QK_K;
    }
}

void oracle_quantize_row_q8_0(const float *x, void *vy, int64_t k) {
    block_q8_0 *y = (block_q8_0 *)vy;
    for (int64_t i = 0; i < k / QK8_0; i++) {
        float amax = 0;
        for (int j = 0; j < QK8_0; j++) {
            const float v = fabsf(x[i * QK8_0 + j]);
            if (v > amax) amax = v;
        }
        const float d = amax / 127;
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d = f2h(d);
        for (int j = 0; j < QK8_0; j++) y[i].qs[j] = (int8_t)roundf(x[i * QK8_0 + j] * id);
    }
}

/* ---------------------------------------------------------------- integer block dot products
 * ggml's arithmetic: integer dot products inside a block, float
 * scaling by d_weight*d_activation, accumulation across blocks.  ggml does not fix the order of
 * that last accumulation (its generic, AVX2, AVX-512 and NEON kernels all differ), so this
 * restatement picks an ORDER-FREE one: every 64-weight unit (32 for Q8_0) contributes one f32
 * term, and the terms are summed in double and rounded to f32 once.  A double sum of a few
 * hundred f32 terms rounds to the same f32 in any order (except with probability ~1e-7), which
 * lets a parallel implementation be compared bit for bit instead of through a chaotic tolerance
 * (one flipped Q8 rounding re-randomises the quantisation noise of every later layer). */
static float vec_dot_q4_K_q8_K(int64_t n, const block_q4_K *x, const block_q8_K *y) {
    const int64_t nb = n / QK_K;
    double acc = 0.0;
    for (int64_t i = 0; i < nb; i++) {
        const float d = h2f(x[i].d), dmin = h2f(x[i].dmin), d8 = y[i].d;
        const float fd = d * d8, fm = dmin * d8;
        for (int p = 0; p < 4; p++) { /* unit p: sub-blocks 2p (low nibbles) and 2p+1 (high nibbles) */
            uint8_t sc_lo, m_lo, sc_hi, m_hi;
            get_scale_min_k4(2 * p, x[i].scales, &sc_lo, &m_lo);
            get_scale_min_k4(2 * p + 1, x[i].scales, &sc_hi, &m_hi);
            const uint8_t *q4 = x[i].qs + 32 * p;
            const int8_t *q8 = y[i].qs + 64 * p;
            int dlo = 0, dhi = 0;
            for (int l = 0; l < 32; l++) {
                dlo += (q4[l] & 0xF) * q8[l];
                dhi += (q4[l] >> 4) * q8[l + 32];
            }
            const int bs_lo = y[i].bsums[4 * p] + y[i].bsums[4 * p + 1];
            const int bs_hi = y[i].bsums[4 * p + 2] + y[i].bsums[4 * p + 3];
            const int isum = sc_lo * dlo + sc_hi * dhi;
            const int msum = m_lo * bs_lo + m_hi * bs_hi;
            const float term = fmaf(fd, (float)isum, -(fm * (float)msum));
            acc += (double)term;
        }
    }
    return (float)acc;
}

static float vec_dot_q5_K_q8_K(int64_t n, const block_q5_K *x, const block_q8_K *y) {
    const int64_t nb = n / QK_K;
    double acc = 0.0;
    for (int64_t i = 0; i < nb; i++) {
        const float d = h2f(x[i].d), dmin = h2f(x[i].dmin), d8 = y[i].d;
        const float fd = d * d8, fm = dmin * d8;
        for (int p = 0; p < 4; p++) {
            uint8_t sc_lo, m_lo, sc_hi, m_hi;
            get_scale_min_k4(2 * p, x[i].scales, &sc_lo, &m_lo);
            get_scale_min_k4(2 * p + 1, x[i].scales, &sc_hi, &m_hi);
            const uint8_t *q4 = x[i].qs + 32 * p, *qh = x[i].qh;
            const int8_t *q8 = y[i].qs + 64 * p;
            int dlo = 0, dhi = 0;
            for (int l = 0; l < 32; l++) {
                dlo += ((q4[l] & 0xF) | (((qh[l] >> (2 * p)) & 1) << 4)) * q8[l];
                dhi += ((q4[l] >> 4) | (((qh[l] >> (2 * p + 1)) & 1) << 4)) * q8[l + 32];
            }
            const int bs_lo = y[i].bsums[4 * p] + y[i].bsums[4 * p + 1];
            const int bs_hi = y[i].bsums[4 * p + 2] + y[i].bsums[4 * p + 3];
            const int isum = sc_lo * dlo + sc_hi * dhi;
            const int msum = m_lo * bs_lo + m_hi * bs_hi;
            const float term = fmaf(fd, (float)isum, -(fm * (float)msum));
            acc += (double)term;
        }
    }
    return (float)acc;
}

static float vec_dot_q6_K_q8_K(int64_t n, const block_q6_K *x, const block_q8_K *y) {
    const int64_t nb = n / QK_K;
    double acc = 0.0;
    for (int64_t i = 0; i < nb; i++) {
        const float fd = h2f(x[i].d) * y[i].d;
        for (int h = 0; h < 2; h++) {
            for (int t = 0; t < 2; t++) { /* unit (h,t): elements 128h + 16t + 32k + [0,16), k = 0..3 */
                const uint8_t *ql = x[i].ql + 64 * h + 16 * t, *qh = x[i].qh + 32 * h + 16 * t;
                const int8_t *q8 = y[i].qs + 128 * h + 16 * t;
                int dk[4] = {0, 0, 0, 0};
                for (int l = 0; l < 16; l++) {
                    dk[0] += ((ql[l] & 0xF) | (((qh[l] >> 0) & 3) << 4)) * q8[l];
                    dk[1] += ((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) * q8[l + 32];
                    dk[2] += ((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) * q8[l + 64];
                    dk[3] += ((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) * q8[l + 96];
                }
                int isum = 0;
                for (int k = 0; k < 4; k++) {
                    const int g = 8 * h + t + 2 * k; /* 16-wide scale group */
                    isum += x[i].scales[g] * (dk[k] - 32 * y[i].bsums[g]);
                }
                acc += (double)(fd * (float)isum);
            }
        }
    }
    return (float)acc;
}

static float vec_dot_q8_0_q8_0(int64_t n, const block_q8_0 *x, const block_q8_0 *y) {
    const int64_t nb = n / QK8_0;
    double acc = 0.0;
    for (int64_t ib = 0; ib < nb; ib++) {
        int sumi = 0;
        for (int j = 0; j < QK8_0; j++) sumi += x[ib].qs[j] * y[ib].qs[j];
        acc += (double)((float)sumi * (h2f(x[ib].d) * h2f(y[ib].d)));
    }
    return (float)acc;
}


static float vec_dot_f32(int64_t n, const float *x, const float *y) {
    float s = 0;
    for (int64_t i = 0; i < n; i++) s += x[i] * y[i];
    return s;
}

/* ---------------------------------------------------------------- AVX2 dot kernels (timed CPU baseline)
 * The integer sub-block dots (exact in any order) are vectorised; everything that rounds — the one f32 term per
 * 64-weight unit (32-weight block for Q8_0) and the double sum of the terms — is the generic kernels' code, operation
 * by operation.  The results are therefore BIT-IDENTICAL to the generic kernels (tests/test_cpu.py checks this on
 * every type), which is what lets the fast path serve as the parity oracle at Mistral-7B size.
 * (Round 1's AVX2 kernels accumulated per-block f32 partial sums like ggml's own AVX2 path; on a 32-layer random-init
 * model that last-bit difference grows to ~0.4 of the logit std, so they could not pin anything.)
 * Selected with oracle_set_simd(1). */
static int g_simd = 0;
void oracle_set_simd(int on) { g_simd = on; }
#if defined(__AVX2__)
#include <immintrin.h>
static inline int hsum8_i32(__m256i v) {
    __m128i s = _mm_add_epi32(_mm256_castsi256_si128(v), _mm256_extracti128_si256(v, 1));
    s = _mm_add_epi32(s, _mm_shuffle_epi32(s, 0x4e));
    s = _mm_add_epi32(s, _mm_shuffle_epi32(s, 0xb1));
    return _mm_cvtsi128_si32(s);
}
/* u8 x i8 -> i32 lanes (8 lanes, each the sum of 4 products; |pair sums| <= 2*128*127 < 2^15: no saturation) */
static inline __m256i dot_u8_i8(__m256i u, __m256i s) {
    const __m256i p16 = _mm256_maddubs_epi16(u, s);
    return _mm256_madd_epi16(p16, _mm256_set1_epi16(1));
}
/* horizontal sums of eight 8-lane vectors at once: returns lanes {sum(v0), ..., sum(v7)} */
static inline __m256i hsum8x8(__m256i v0, __m256i v1, __m256i v2, __m256i v3, __m256i v4, __m256i v5, __m256i v6, __m256i v7) {
    const __m256i g0 = _mm256_hadd_epi32(_mm256_hadd_epi32(v0, v1), _mm256_hadd_epi32(v2, v3)); /* [s0..s3 lo | s0..s3 hi] */
    const __m256i g1 = _mm256_hadd_epi32(_mm256_hadd_epi32(v4, v5), _mm256_hadd_epi32(v6, v7));
    const __m128i a = _mm_add_epi32(_mm256_castsi256_si128(g0), _mm256_extracti128_si256(g0, 1));
    const __m128i b = _mm_add_epi32(_mm256_castsi256_si128(g1), _mm256_extracti128_si256(g1, 1));
    return _mm256_set_m128i(b, a);
}
static float vec_dot_q4_K_q8_K_avx2(int64_t n, const block_q4_K *x, const block_q8_K *y) {
    const int64_t nb = n / QK_K;
    const __m256i m4 = _mm256_set1_epi8(0xF);
    double acc = 0.0;
    for (int64_t i = 0; i < nb; i++) {
        const float d = h2f(x[i].d), dmin = h2f(x[i].dmin), d8 = y[i].d;
        const float fd = d * d8, fm = dmin * d8;
        __m256i v[8];
        for (int p = 0; p < 4; p++) {
            const __m256i q4 = _mm256_loadu_si256((const __m256i *)(x[i].qs + 32 * p));
            const __m256i a0 = _mm256_loadu_si256((const __m256i *)(y[i].qs + 64 * p));
            const __m256i a1 = _mm256_loadu_si256((const __m256i *)(y[i].qs + 64 * p + 32));
            v[2 * p] = dot_u8_i8(_mm256_and_si256(q4, m4), a0);
            v[2 * p + 1] = dot_u8_i8(_mm256_and_si256(_mm256_srli_epi16(q4, 4), m4), a1);
        }
        int32_t dots[8]; /* dots[2p] = low-nibble dot of unit p, dots[2p+1] = high-nibble dot */
        _mm256_storeu_si256((__m256i *)dots, hsum8x8(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]));
        for (int p = 0; p < 4; p++) {
            uint8_t sc_lo, m_lo, sc_hi, m_hi;
            get_scale_min_k4(2 * p, x[i].scales, &sc_lo, &m_lo);
            get_scale_min_k4(2 * p + 1, x[i].scales, &sc_hi, &m_hi);
            const int bs_lo = y[i].bsums[4 * p] + y[i].bsums[4 * p + 1];
            const int bs_hi = y[i].bsums[4 * p + 2] + y[i].bsums[4 * p + 3];
            const int isum = sc_lo * dots[2 * p] + sc_hi * dots[2 * p + 1];
            const int msum = m_lo * bs_lo + m_hi * bs_hi;
            const float term = fmaf(fd, (float)isum, -(fm * (float)msum));
            acc += (double)term;
        }
    }
    return (float)acc;
}
static float vec_dot_q6_K_q8_K_avx2(int64_t n, const block_q6_K *x, const block_q8_K *y) {
    const int64_t nb = n / QK_K;
    const __m256i m4 = _mm256_set1_epi8(0xF), m2 = _mm256_set1_epi8(0x30);
    double acc = 0.0;
    for (int64_t i = 0; i < nb; i++) {
        const float fd = h2f(x[i].d) * y[i].d;
        for (int h = 0; h < 2; h++) {
            const __m256i qla = _mm256_loadu_si256((const __m256i *)(x[i].ql + 64 * h));
            const __m256i qlb = _mm256_loadu_si256((const __m256i *)(x[i].ql + 64 * h + 32));
            const __m256i qh = _mm256_loadu_si256((const __m256i *)(x[i].qh + 32 * h));
            __m256i q[4], v[4];
            q[0] = _mm256_or_si256(_mm256_and_si256(qla, m4), _mm256_and_si256(_mm256_slli_epi16(qh, 4), m2));
            q[1] = _mm256_or_si256(_mm256_and_si256(qlb, m4), _mm256_and_si256(_mm256_slli_epi16(qh, 2), m2));
            q[2] = _mm256_or_si256(_mm256_and_si256(_mm256_srli_epi16(qla, 4), m4), _mm256_and_si256(qh, m2));
            q[3] = _mm256_or_si256(_mm256_and_si256(_mm256_srli_epi16(qlb, 4), m4), _mm256_and_si256(_mm256_srli_epi16(qh, 2), m2));
            for (int g = 0; g < 4; g++)  /* 8 lanes: lanes 0-3 = first 16 weights of the 32-group, 4-7 = second 16 */
                v[g] = dot_u8_i8(q[g], _mm256_loadu_si256((const __m256i *)(y[i].qs + 128 * h + 32 * g)));
            /* [d0(g=0..3) | d1(g=0..3)]: d0 = dot of the first 16 (t = 0), d1 = of the second 16 (t = 1) */
            int32_t dd[8];
            _mm256_storeu_si256((__m256i *)dd, _mm256_hadd_epi32(_mm256_hadd_epi32(v[0], v[1]), _mm256_hadd_epi32(v[2], v[3])));
            int isum0 = 0, isum1 = 0; /* units (h, t = 0) and (h, t = 1) */
            for (int g = 0; g < 4; g++) {
                const int is = 8 * h + 2 * g;
                isum0 += x[i].scales[is] * (dd[g] - 32 * y[i].bsums[is]);
                isum1 += x[i].scales[is + 1] * (dd[4 + g] - 32 * y[i].bsums[is + 1]);
            }
            acc += (double)(fd * (float)isum0);
            acc += (double)(fd * (float)isum1);
        }
    }
    return (float)acc;
}
static float vec_dot_q8_0_q8_0_avx2(int64_t n, const block_q8_0 *x, const block_q8_0 *y) {
    const int64_t nb = n / QK8_0;
    double acc = 0.0;
    int64_t ib = 0;
    for (; ib + 8 <= nb; ib += 8) {
        __m256i v[8];
        for (int j = 0; j < 8; j++) {
            const __m256i qx = _mm256_loadu_si256((const __m256i *)x[ib + j].qs);
            const __m256i qy = _mm256_loadu_si256((const __m256i *)y[ib + j].qs);
            v[j] = dot_u8_i8(_mm256_sign_epi8(qx, qx), _mm256_sign_epi8(qy, qx));
        }
        int32_t sums[8];
        _mm256_storeu_si256((__m256i *)sums, hsum8x8(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]));
        for (int j = 0; j < 8; j++) acc += (double)((float)sums[j] * (h2f(x[ib + j].d) * h2f(y[ib + j].d)));
    }
    for (; ib < nb; ib++) {
        const __m256i qx = _mm256_loadu_si256((const __m256i *)x[ib].qs);
        const __m256i qy = _mm256_loadu_si256((const __m256i *)y[ib].qs);
        const int sumi = hsum8_i32(dot_u8_i8(_mm256_sign_epi8(qx, qx), _mm256_sign_epi8(qy, qx)));
        acc += (double)((float)sumi * (h2f(x[ib].d) * h2f(y[ib].d)));
    }
    return (float)acc;
}
#define HAVE_AVX2_DOTS 1
#else
#define HAVE_AVX2_DOTS 0
#endif

/* one row dot: x row of `type` against pre-quantised activation `yq` (or f32 `yf`) */
static float row_dot(int type, int64_t k, const void *row, const void *yq, const float *yf) {
#if HAVE_AVX2_DOTS
    if (g_simd) {
        switch (type) {
            case T_Q4_K: return vec_dot_q4_K_q8_K_avx2(k, (const block_q4_K *)row, (const block_q8_K *)yq);
            case T_Q6_K: return vec_dot_q6_K_q8_K_avx2(k, (const block_q6_K *)row, (const block_q8_K *)yq);
            case T_Q8_0: return vec_dot_q8_0_q8_0_avx2(k, (const block_q8_0 *)row, (const block_q8_0 *)yq);
            default: break;
        }
    }
#endif
    switch (type) {
        case T_Q4_K: return vec_dot_q4_K_q8_K(k, (const block_q4_K *)row, (const block_q8_K *)yq);
        case T_Q5_K: return vec_dot_q5_K_q8_K(k, (const block_q5_K *)row, (const block_q8_K *)yq);
        case T_Q6_K: return vec_dot_q6_K_q8_K(k, (const block_q6_K *)row, (const block_q8_K *)yq);
        case T_Q8_0: return vec_dot_q8_0_q8_0(k, (const block_q8_0 *)row, (const block_q8_0 *)yq);
        case T_F32: return vec_dot_f32(k, (const float *)row, yf);
    }
    return NAN;
}

/* y[N] = W[N,K] . x[K]   (ggml mul_mat with src1 quantised to the weight type's vec_dot_type) */
int oracle_matvec(int type, const void *w, int64_t n_rows, int64_t k, const float *x, float *y) {
    void *yq = NULL;
    if (type == T_Q4_K || type == T_Q5_K || type == T_Q6_K) {
        yq = malloc((size_t)(k / QK_K) * sizeof(block_q8_K));
        oracle_quantize_row_q8_K(x, yq, k);
    } else if (type == T_Q8_0) {
        yq = malloc((size_t)(k / QK8_0) * sizeof(block_q8_0));
        oracle_quantize_row_q8_0(x, yq, k);
    } else if (type != T_F32) {
        return -1;
    }
    const size_t rb = row_bytes(type, k);
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < n_rows; r++) y[r] = row_dot(type, k, (const char *)w + (size_t)r * rb, yq, x);
    free(yq);
    return 0;
}

/* ---------------------------------------------------------------- model (arch "llama") */
typedef struct { int32_t type; int64_t ne0, ne1; const void *data; } otensor;
typedef struct {
    otensor attn_norm, wq, wk, wv, wo, ffn_norm, w_gate, w_up, w_down;
} olayer;

typedef struct {
    int32_t n_layer, n_embd, n_head, n_head_kv, head_dim, n_ff, n_vocab, n_ctx, n_rot;
    float rms_eps, rope_freq_base;
} oracle_hparams;

typedef struct oracle_model {
    oracle_hparams hp;
    otensor tok_embd, output_norm, output;
    olayer *layers;
    uint16_t *k_cache, *v_cache; /* f16 [L][n_ctx][n_head_kv*head_dim] */
    /* scratch */
    float *x, *xn, *q, *k, *v, *att_out, *scores, *gate, *up, *h, *tmp;
    float *logits;
} oracle_model;

oracle_model *oracle_model_new(const oracle_hparams *hp) {
    oracle_model *m = (oracle_model *)calloc(1, sizeof(*m));
    m->hp = *hp;
    m->layers = (olayer *)calloc((size_t)hp->n_layer, sizeof(olayer));
    const size_t kvd = (size_t)hp->n_head_kv * hp->head_dim;
    m->k_cache = (uint16_t *)calloc((size_t)hp->n_layer * hp->n_ctx * kvd, 2);
    m->v_cache = (uint16_t *)calloc((size_t)hp->n_layer * hp->n_ctx * kvd, 2);
    const size_t qd = (size_t)hp->n_head * hp->head_dim;
    m->x = (float *)calloc((size_t)hp->n_embd, 4);
    m->xn = (float *)calloc((size_t)hp->n_embd, 4);
    m->q = (float *)calloc(qd, 4);
    m->k = (float *)calloc(kvd, 4);
    m->v = (float *)calloc(kvd, 4);
    m->att_out = (float *)calloc(qd, 4);
    m->scores = (float *)calloc((size_t)hp->n_head * hp->n_ctx, 4);
    m->gate = (float *)calloc((size_t)hp->n_ff, 4);
    m->up = (float *)calloc((size_t)hp->n_ff, 4);
    m->h = (float *)calloc((size_t)hp->n_ff, 4);
    m->tmp = (float *)calloc((size_t)hp->n_embd, 4);
    m->logits = (float *)calloc((size_t)hp->n_vocab, 4);
    return m;
}

void oracle_model_free(oracle_model *m) {
    if (!m) return;
    free(m->layers); free(m->k_cache); free(m->v_cache);
    free(m->x); free(m->xn); free(m->q); free(m->k); free(m->v); free(m->att_out);
    free(m->scores); free(m->gate); free(m->up); free(m->h); free(m->tmp); free(m->logits);
    free(m);
}

/* name follows GGUF tensor names for arch llama (gguf-py constants.py:1550-1569) */
int oracle_model_set_tensor(oracle_model *m, const char *name, int type, int64_t ne0, int64_t ne1, const void *data) {
    otensor t = {type, ne0, ne1, data};
    if (!strcmp(name, "token_embd.weight")) { m->tok_embd = t; return 0; }
    if (!strcmp(name, "output_norm.weight")) { m->output_norm = t; return 0; }
    if (!strcmp(name, "output.weight")) { m->output = t; return 0; }
    int il;
    char rest[64];
    if (sscanf(name, "blk.%d.%63s", &il, rest) == 2 && il >= 0 && il < m->hp.n_layer) {
        olayer *l = &m->layers[il];
        if (!strcmp(rest, "attn_norm.weight")) l->attn_norm = t;
        else if (!strcmp(rest, "attn_q.weight")) l->wq = t;
        else if (!strcmp(rest, "attn_k.weight")) l->wk = t;
        else if (!strcmp(rest, "attn_v.weight")) l->wv = t;
        else if (!strcmp(rest, "attn_output.weight")) l->wo = t;
        else if (!strcmp(rest, "ffn_norm.weight")) l->ffn_norm = t;
        else if (!strcmp(rest, "ffn_gate.weight")) l->w_gate = t;
        else if (!strcmp(rest, "ffn_up.weight")) l->w_up = t;
        else if (!strcmp(rest, "ffn_down.weight")) l->w_down = t;
        else return -1;
        return 0;
    }
    return -1;
}

/* expf restated with fixed operations so that a GPU can reproduce it bit for bit (glibc's and
 * CUDA's expf differ in the last ulp; ggml's own vectorised paths use their own polynomial too):
 * n = rint(x*log2e), r = x - n*ln2 (two-step Cody-Waite), degree-6 Horner with fmaf, scale by 2^n. */
float oracle_expf(float x) {
    if (x < -87.0f) return 0.0f;
    if (x > 88.0f) return INFINITY;
    const float t = x * 1.44269504088896341f;
    const float n = rintf(t);
    float r = fmaf(n, -0.693145751953125f, x);
    r = fmaf(n, -1.42860682030941723212e-6f, r);
    float p = 1.0f / 720.0f;
    p = fmaf(p, r, 1.0f / 120.0f);
    p = fmaf(p, r, 1.0f / 24.0f);
    p = fmaf(p, r, 1.0f / 6.0f);
    p = fmaf(p, r, 0.5f);
    p = fmaf(p, r, 1.0f);
    p = fmaf(p, r, 1.0f);
    const int ni = (int)n; /* [-126, 127] after the clamps above */
    uint32_t bits = (uint32_t)(ni + 127) << 23;
    float s;
    memcpy(&s, &bits, 4);
    return p * s;
}

/* RMSNorm: sum of squares in double (ggml_float), scale = 1/sqrtf(mean+eps); then the
 * separate elementwise multiply by the norm weight (Appendix C.4). */
static void rms_norm_mul(const float *x, const float *w, float *y, int n, float eps) {
    double sum = 0.0;
    for (int i = 0; i < n; i++) sum += (double)(x[i] * x[i]);
    const float mean = (float)(sum / n);
    const float scale = 1.0f / sqrtf(mean + eps);
    for (int i = 0; i < n; i++) y[i] = (x[i] * scale) * w[i];
}

/* RoPE, NORM flavour: adjacent pairs, theta_i = pos * base^(-2i/n_rot) built by repeated
 * multiplication in f32 (Appendix C.3). */
void oracle_rope_table(int pos, int n_rot, float freq_base, float *cos_out, float *sin_out) {
    const float theta_scale = powf(freq_base, -2.0f / (float)n_rot);
    float theta = (float)pos;
    for (int i = 0; i < n_rot / 2; i++) {
        cos_out[i] = cosf(theta);
        sin_out[i] = sinf(theta);
        theta *= theta_scale;
    }
}
static void rope_inplace(float *v, int n_heads, int head_dim, int n_rot, const float *c, const float *s) {
    for (int h = 0; h < n_heads; h++) {
        float *p = v + (size_t)h * head_dim;
        for (int i = 0; i < n_rot / 2; i++) {
            const float x0 = p[2 * i], x1 = p[2 * i + 1];
            p[2 * i] = x0 * c[i] - x1 * s[i];
            p[2 * i + 1] = x0 * s[i] + x1 * c[i];
        }
    }
}

static void embed_row(const otensor *t, int tok, float *out) {
    const size_t rb = row_bytes(t->type, t->ne0);
    oracle_dequantize(t->type, (const char *)t->data + (size_t)tok * rb, t->ne0, out);
}

/* one token through the whole stack at position `pos`; logits (if non-NULL) get n_vocab floats */
static int eval_token(oracle_model *m, int tok, int pos, float *logits) {
    const oracle_hparams *hp = &m->hp;
    const int E = hp->n_embd, H = hp->n_head, G = hp->n_head_kv, D = hp->head_dim, FF = hp->n_ff;
    const int kvd = G * D, rep = H / G;
    if (pos >= hp->n_ctx || tok < 0 || tok >= hp->n_vocab) return -1;
    float cs[512], sn[512];
    oracle_rope_table(pos, hp->n_rot, hp->rope_freq_base, cs, sn);
    embed_row(&m->tok_embd, tok, m->x);
    const float kq_scale = 1.0f / sqrtf((float)D);
    for (int il = 0; il < hp->n_layer; il++) {
        const olayer *l = &m->layers[il];
        rms_norm_mul(m->x, (const float *)l->attn_norm.data, m->xn, E, hp->rms_eps);
        oracle_matvec(l->wq.type, l->wq.data, (int64_t)H * D, E, m->xn, m->q);
        oracle_matvec(l->wk.type, l->wk.data, kvd, E, m->xn, m->k);
        oracle_matvec(l->wv.type, l->wv.data, kvd, E, m->xn, m->v);
        rope_inplace(m->q, H, D, hp->n_rot, cs, sn);
        rope_inplace(m->k, G, D, hp->n_rot, cs, sn);
        uint16_t *kc = m->k_cache + ((size_t)il * hp->n_ctx) * kvd;
        uint16_t *vc = m->v_cache + ((size_t)il * hp->n_ctx) * kvd;
        for (int i = 0; i < kvd; i++) {
            kc[(size_t)pos * kvd + i] = f2h(m->k[i]);
            vc[(size_t)pos * kvd + i] = f2h(m->v[i]);
        }
        const int n_kv = pos + 1;
#pragma omp parallel for schedule(static)
        for (int h = 0; h < H; h++) {
            const int g = h / rep;
            float *sc = m->scores + (size_t)h * hp->n_ctx;
            float qh[512];
            for (int d = 0; d < D; d++) qh[d] = h2f(f2h(m->q[(size_t)h * D + d])); /* q -> f16 (vec_dot_type of an f16 K) */
            float mx = -INFINITY;
            for (int j = 0; j < n_kv; j++) {
                const uint16_t *kr = kc + (size_t)j * kvd + (size_t)g * D;
                double sd = 0; /* f16 x f16 products are exact in f32; summed order-free in double */
                for (int d = 0; d < D; d++) sd += (double)(qh[d] * h2f(kr[d]));
                float s = (float)sd;
                s *= kq_scale;
                sc[j] = s;
                if (s > mx) mx = s;
            }
            double sum = 0.0;
            for (int j = 0; j < n_kv; j++) {
                const float e = oracle_expf(sc[j] - mx);
                sc[j] = e;
                sum += (double)e;
            }
            const float inv = (float)(1.0 / sum);
            float *o = m->att_out + (size_t)h * D;
            double od[512];
            for (int d = 0; d < D; d++) od[d] = 0;
            for (int j = 0; j < n_kv; j++) {
                const float p = h2f(f2h(sc[j] * inv)); /* P -> f16 before the f16 V mat-mul */
                const uint16_t *vr = vc + (size_t)j * kvd + (size_t)g * D;
                for (int d = 0; d < D; d++) od[d] += (double)(p * h2f(vr[d])); /* exact product, order-free sum */
            }
            for (int d = 0; d < D; d++) o[d] = (float)od[d];
        }
        oracle_matvec(l->wo.type, l->wo.data, E, (int64_t)H * D, m->att_out, m->tmp);
        for (int i = 0; i < E; i++) m->x[i] += m->tmp[i];
        rms_norm_mul(m->x, (const float *)l->ffn_norm.data, m->xn, E, hp->rms_eps);
        oracle_matvec(l->w_gate.type, l->w_gate.data, FF, E, m->xn, m->gate);
        oracle_matvec(l->w_up.type, l->w_up.data, FF, E, m->xn, m->up);
        for (int i = 0; i < FF; i++) {
            const float gx = m->gate[i];
            m->h[i] = (gx / (1.0f + oracle_expf(-gx))) * m->up[i];
        }
        oracle_matvec(l->w_down.type, l->w_down.data, E, FF, m->h, m->tmp);
        for (int i = 0; i < E; i++) m->x[i] += m->tmp[i];
    }
    if (logits) {
        rms_norm_mul(m->x, (const float *)m->output_norm.data, m->xn, E, hp->rms_eps);
        const otensor *out = m->output.data ? &m->output : &m->tok_embd; /* tied embeddings */
        oracle_matvec(out->type, out->data, hp->n_vocab, E, m->xn, logits);
    }
    return 0;
}

/* evaluate n tokens at positions n_past..n_past+n-1; logits of the LAST token -> logits_out */
int oracle_eval(oracle_model *m, const int32_t *tokens, int n, int n_past, float *logits_out) {
    for (int i = 0; i < n; i++) {
        const int last = (i == n - 1);
        int rc = eval_token(m, tokens[i], n_past + i, last ? (logits_out ? logits_out : m->logits) : NULL);
        if (rc) return rc;
    }
    return 0;
}

/* logits of every token (for per-position parity checks): logits_out[n][n_vocab] */
int oracle_eval_all(oracle_model *m, const int32_t *tokens, int n, int n_past, float *logits_out) {
    for (int i = 0; i < n; i++) {
        int rc = eval_token(m, tokens[i], n_past + i, logits_out + (size_t)i * m->hp.n_vocab);
        if (rc) return rc;
    }
    return 0;
}

static int argmax_first(const float *v, int n, float *margin) {
    int best = 0;
    for (int i = 1; i < n; i++) if (v[i] > v[best]) best = i; /* lowest index wins ties */
    if (margin) {
        float second = -INFINITY;
        for (int i = 0; i < n; i++) if (i != best && v[i] > second) second = v[i];
        *margin = v[best] - second;
    }
    return best;
}

/* greedy: prefill prompt, then n_gen argmax tokens; margins[i] = top1-top2 logit gap at step i */
int oracle_greedy(oracle_model *m, const int32_t *prompt, int n_prompt, int n_gen, int32_t *out, float *margins) {
    int rc = oracle_eval(m, prompt, n_prompt, 0, m->logits);
    if (rc) return rc;
    int pos = n_prompt;
    for (int i = 0; i < n_gen; i++) {
        float mg;
        const int t = argmax_first(m->logits, m->hp.n_vocab, &mg);
        out[i] = t;
        if (margins) margins[i] = mg;
        if (i + 1 < n_gen) {
            rc = eval_token(m, t, pos++, m->logits);
            if (rc) return rc;
        }
    }
    return 0;
}

/* ---------------------------------------------------------------- sampler chain
 * llama-server's default chain, of which the reference sets only `temperature` and `max_tokens`
 * (runtime/src/inference.rs:103-122): top-k -> top-p -> min-p -> temperature -> soft-max -> draw.
 * llama.cpp's own draw (std::mt19937 + std::discrete_distribution) is neither specified nor stable across
 * standard libraries; the draw here is an inverse-CDF walk over the surviving candidates with a splitmix64
 * uniform keyed by (seed, counter).  Every float step is a single rounded f32 operation in a fixed order
 * (this file is built with -ffp-contract=off), so a GPU can reproduce the candidate set, both cut-offs and
 * the drawn token bit for bit.
 *
 *   candidates  = allowed tokens (mask bit set, or all when mask == NULL) sorted by (logit descending, index
 *                 ascending); K = min(top_k > 0 ? top_k : 128, 128, #allowed)      [128 = candidate cap]
 *   p_i         = expf(l_i - l_0), sum = p_0 + p_1 + ... (left to right)
 *   top-p       : n = smallest i+1 with (p_0/sum + ... + p_i/sum) >= top_p          (top_p >= 1: off)
 *   min-p       : n = first i in [1, n) with p_i < min_p * p_0, if any               (min_p <= 0: off)
 *   temperature : w_i = expf((l_i - l_0) / temperature), i < n; sum2 left to right
 *   draw        : u = uniform(seed, counter) * sum2; first i with u < w_0 + ... + w_i (else n - 1)
 * Returns the token; cand[0..K) = sorted candidate ids, counts[3] = {K, n after top-p, n after min-p}.
 * temperature <= 0 is greedy (arg-max over the allowed tokens, lowest index wins). */
float oracle_uniform(uint64_t seed, uint64_t ctr) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (ctr + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (float)(z >> 40) * (1.0f / 16777216.0f);
}

#define ORACLE_SAMPLE_MAX_K 128
int oracle_sample(const float *logits, int n_vocab, const uint8_t *mask, float temperature, int top_k, float top_p,
                  float min_p, uint64_t seed, uint64_t counter, int32_t *cand, int32_t *counts) {
    int K = top_k > 0 ? top_k : ORACLE_SAMPLE_MAX_K;
    if (K > ORACLE_SAMPLE_MAX_K) K = ORACLE_SAMPLE_MAX_K;
    if (!(temperature > 0.f)) K = 1;
    int idx[ORACLE_SAMPLE_MAX_K];
    float val[ORACLE_SAMPLE_MAX_K];
    int n = 0;
    /* insertion into a sorted list of at most K: (logit descending, index ascending) */
    for (int i = 0; i < n_vocab; i++) {
        if (mask && !((mask[i >> 3] >> (i & 7)) & 1)) continue;
        const float v = logits[i];
        if (n == K && !(v > val[n - 1])) continue; /* ties keep the earlier index */
        int j = n < K ? n : K - 1;
        while (j > 0 && v > val[j - 1]) { val[j] = val[j - 1]; idx[j] = idx[j - 1]; j--; }
        val[j] = v; idx[j] = i;
        if (n < K) n++;
    }
    if (n == 0) return -1;
    K = n;
    if (cand) for (int i = 0; i < K; i++) cand[i] = idx[i];
    if (counts) counts[0] = counts[1] = counts[2] = K;
    if (!(temperature > 0.f)) return idx[0];
    float p[ORACLE_SAMPLE_MAX_K];
    float sum = 0.f;
    for (int i = 0; i < K; i++) { p[i] = oracle_expf(val[i] - val[0]); sum = sum + p[i]; }
    n = K;
    if (top_p < 1.f) {
        float c = 0.f;
        for (int i = 0; i < K; i++) { c = c + p[i] / sum; if (c >= top_p) { n = i + 1; break; } }
    }
    if (counts) counts[1] = n;
    if (min_p > 0.f) {
        int n2 = 1;
        while (n2 < n && p[n2] >= min_p * p[0]) n2++;
        n = n2;
    }
    if (counts) counts[2] = n;
    float sum2 = 0.f;
    for (int i = 0; i < n; i++) { p[i] = oracle_expf((val[i] - val[0]) / temperature); sum2 = sum2 + p[i]; }
    const float u = oracle_uniform(seed, counter) * sum2;
    float c = 0.f;
    int pick = n - 1;
    for (int i = 0; i < n; i++) { c = c + p[i]; if (u < c) { pick = i; break; } }
    return idx[pick];
}

const uint16_t *oracle_k_cache(const oracle_model *m) { return m->k_cache; }
const uint16_t *oracle_v_cache(const oracle_model *m) { return m->v_cache; }
int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
