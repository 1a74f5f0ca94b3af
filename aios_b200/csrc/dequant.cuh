// dequant.cuh — bit-exact single-element dequantisation on GGUF's natural block layout.
// See dequant.cu for the bit-exactness notes (__fmul_rn / __fsub_rn: never FMA-contracted).
#pragma once
#include "common.cuh"
#include "layout.cuh"

namespace aios {

__device__ __forceinline__ void scale_min_k4(int j, const uint8_t *q, int &sc, int &m) {
    // gguf-py quants.py:479-501
    if (j < 4) {
        sc = q[j] & 63;
        m = q[j + 4] & 63;
    } else {
        sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4);
        m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4);
    }
}

// one element of a natural-layout row; `row` points at the first block of the row
__device__ __forceinline__ float dequant_elem(int type, const uint8_t *row, int64_t i) {
    switch (type) {
        case T_F32: return reinterpret_cast<const float *>(row)[i];
        case T_F16: return __half2float(reinterpret_cast<const __half *>(row)[i]);
        case T_Q8_0: {
            const uint8_t *b = row + (i / 32) * 34;
            const float d = h2f_bits((uint16_t)(b[0] | (b[1] << 8)));
            return __fmul_rn((float)(int8_t)b[2 + (i % 32)], d);
        }
        case T_Q4_K: {
            const uint8_t *b = row + (i / 256) * 144;
            const int e = (int)(i % 256), j = e / 32, l = e % 32;
            const float d = h2f_bits((uint16_t)(b[0] | (b[1] << 8))), dmin = h2f_bits((uint16_t)(b[2] | (b[3] << 8)));
            int sc, m;
            scale_min_k4(j, b + 4, sc, m);
            const uint8_t qb = b[16 + 32 * (j / 2) + l];
            const int q = (j & 1) ? (qb >> 4) : (qb & 0xF);
            return __fsub_rn(__fmul_rn(__fmul_rn(d, (float)sc), (float)q), __fmul_rn(dmin, (float)m));
        }
        case T_Q5_K: {
            const uint8_t *b = row + (i / 256) * 176;
            const int e = (int)(i % 256), j = e / 32, l = e % 32;
            const float d = h2f_bits((uint16_t)(b[0] | (b[1] << 8))), dmin = h2f_bits((uint16_t)(b[2] | (b[3] << 8)));
            int sc, m;
            scale_min_k4(j, b + 4, sc, m);
            const uint8_t qb = b[48 + 32 * (j / 2) + l];
            const int hb = (b[16 + l] >> j) & 1;
            const int q = ((j & 1) ? (qb >> 4) : (qb & 0xF)) | (hb << 4);
            return __fsub_rn(__fmul_rn(__fmul_rn(d, (float)sc), (float)q), __fmul_rn(dmin, (float)m));
        }
        case T_Q6_K: {
            const uint8_t *b = row + (i / 256) * 210;
            const int e = (int)(i % 256), h = e / 128, r = e % 128, k = r / 32, l = r % 32;
            const uint8_t *ql = b + 64 * h, *qh = b + 128 + 32 * h;
            const int8_t *sc = reinterpret_cast<const int8_t *>(b + 192) + 8 * h;
            const float d = h2f_bits((uint16_t)(b[208] | (b[209] << 8)));
            const uint8_t lo = ql[l + 32 * (k & 1)];
            const int q4 = (k >= 2) ? (lo >> 4) : (lo & 0xF);
            const int q = (int)(int8_t)(q4 | (((qh[l] >> (2 * k)) & 3) << 4)) - 32;
            return __fmul_rn(__fmul_rn(d, (float)sc[l / 16 + 2 * k]), (float)q);
        }
    }
    return 0.f;
}

}  // namespace aios
