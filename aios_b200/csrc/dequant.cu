// dequant.cu — (1) bit-exact dequantize_row_* kernels on GGUF's natural block layout
// (test hook + embedding-row gather), (2) the GPU-side re-tiling of natural blocks into
// the aligned planes of layout.cuh.
//
// Bit-exactness: ggml / gguf-py compute  (d*sc)*q - (dmin*m)  with every product rounded
// to f32 on its own (gguf-py quants.py:504-522; numpy has no FMA).  nvcc contracts a*b-c
// into an FMA by default, so the arithmetic here is spelled with __fmul_rn / __fsub_rn,
// which are never contracted.
#include "common.cuh"
#include "kernels.h"
#include "layout.cuh"
#include "dequant.cuh"

namespace aios {

__global__ void dequantize_rows_kernel(int type, const uint8_t *blocks, int64_t n, float *out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = dequant_elem(type, blocks, i);
}

void launch_dequantize(int type, const void *d_blocks, int64_t n, float *d_out, cudaStream_t st) {
    const int threads = 256;
    const int64_t grid = (n + threads - 1) / threads;
    dequantize_rows_kernel<<<(unsigned)grid, threads, 0, st>>>(type, (const uint8_t *)d_blocks, n, d_out);
}

// embedding gather: x[0..k) = dequant(table row *tok_ptr); the token id is read on the device
__global__ void embed_row_kernel(int type, const uint8_t *table, int64_t row_bytes, int k, const int32_t *tok_ptr,
                                 float *x, int32_t *pos_ptr, int pos_value, int32_t *cur_token) {
    const int tok = *tok_ptr;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (pos_ptr) *pos_ptr = pos_value;      // prompt feeding: the host names the position
        if (cur_token) *cur_token = tok;
    }
    const uint8_t *row = table + (int64_t)tok * row_bytes;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < k; i += gridDim.x * blockDim.x) x[i] = dequant_elem(type, row, i);
}

void launch_embed_row(int type, const void *d_table, int64_t row_bytes, int k, const int32_t *d_tok, float *d_x,
                      int32_t *d_pos, int pos_value, int32_t *d_cur_token, cudaStream_t st) {
    const int threads = 256;
    embed_row_kernel<<<(k + threads - 1) / threads, threads, 0, st>>>(type, (const uint8_t *)d_table, row_bytes, k, d_tok,
                                                                       d_x, d_pos, pos_value, d_cur_token);
}

// batched gather (continuous batching / prompt chunks): x[t][0..k) = dequant(table row toks[t]), t = blockIdx.y
__global__ void embed_rows_kernel(int type, const uint8_t *table, int64_t row_bytes, int k, const int32_t *toks, float *x,
                                  int64_t ldx) {
    pdl_wait();
    pdl_launch_dependents();
    const int t = blockIdx.y;
    const uint8_t *row = table + (int64_t)toks[t] * row_bytes;
    float *xt = x + (size_t)t * ldx;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < k; i += gridDim.x * blockDim.x) xt[i] = dequant_elem(type, row, i);
}
void launch_embed_rows(int type, const void *d_table, int64_t row_bytes, int k, const int32_t *d_toks, int T, float *d_x,
                       int64_t ldx, cudaStream_t st) {
    const int threads = 256;
    launch_k(embed_rows_kernel, dim3((unsigned)((k + threads - 1) / threads), (unsigned)T), dim3(threads), 0, st, type,
             (const uint8_t *)d_table, row_bytes, k, d_toks, d_x, ldx);
}

// ---------------------------------------------------------------- natural blocks -> planes
// One thread per (row, unit).  Raw blocks are only 2-byte aligned (Q6_K, Q8_0), so they are
// read bytewise; this runs once per tensor at load time.
__device__ __forceinline__ uint4 load16_bytes(const uint8_t *p) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; i++)
        w[i] = (uint32_t)p[4 * i] | ((uint32_t)p[4 * i + 1] << 8) | ((uint32_t)p[4 * i + 2] << 16) |
               ((uint32_t)p[4 * i + 3] << 24);
    return make_uint4(w[0], w[1], w[2], w[3]);
}

__global__ void repack_q4k_kernel(const uint8_t *raw, int64_t n_rows, int k, uint8_t *qsa, uint8_t *qsb, uint8_t *meta,
                                  uint8_t *qh5, int is_q5) {
    const int U = k / 64, nb = k / 256;
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_rows * U) return;
    const int64_t row = idx / U;
    const int u = (int)(idx % U), b = u / 4, p = u % 4;
    const int bbytes = is_q5 ? 176 : 144;
    const uint8_t *blk = raw + (row * nb + b) * bbytes;
    const uint8_t *qs = blk + (is_q5 ? 48 : 16);
    reinterpret_cast<uint4 *>(qsa)[idx] = load16_bytes(qs + 32 * p);
    reinterpret_cast<uint4 *>(qsb)[idx] = load16_bytes(qs + 32 * p + 16);
    if (p == 0) {
        // META word j (one per unit): sc(2j) | sc(2j+1)<<6 | m(2j)<<12 | m(2j+1)<<18 | byte j of (d | dmin<<16) << 24.
        // The 6-bit fields are unpacked here once (gguf-py quants.py:479-501) so the GEMV needs no bit surgery;
        // the block still costs exactly 16 bytes.
        const uint32_t ddm = (uint32_t)blk[0] | ((uint32_t)blk[1] << 8) | ((uint32_t)blk[2] << 16) | ((uint32_t)blk[3] << 24);
        uint32_t w[4];
        for (int j = 0; j < 4; j++) {
            int s0, m0, s1, m1;
            scale_min_k4(2 * j, blk + 4, s0, m0);
            scale_min_k4(2 * j + 1, blk + 4, s1, m1);
            w[j] = (uint32_t)s0 | ((uint32_t)s1 << 6) | ((uint32_t)m0 << 12) | ((uint32_t)m1 << 18) |
                   (((ddm >> (8 * j)) & 0xffu) << 24);
        }
        reinterpret_cast<uint4 *>(meta)[row * nb + b] = make_uint4(w[0], w[1], w[2], w[3]);
    }
    if (is_q5) {
        const uint8_t *qh = blk + 16;
        uint32_t lo = 0, hi = 0;
        for (int l = 0; l < 32; l++) {
            lo |= (uint32_t)((qh[l] >> (2 * p)) & 1) << l;
            hi |= (uint32_t)((qh[l] >> (2 * p + 1)) & 1) << l;
        }
        reinterpret_cast<uint2 *>(qh5)[idx] = make_uint2(lo, hi);
    }
}

__global__ void repack_q6k_kernel(const uint8_t *raw, int64_t n_rows, int k, uint8_t *qla, uint8_t *qlb, uint8_t *qh,
                                  uint8_t *sc, uint8_t *dpl) {
    const int U = k / 64, nb = k / 256;
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_rows * U) return;
    const int64_t row = idx / U;
    const int u = (int)(idx % U), b = u / 4, h = (u % 4) / 2, t = u % 2;
    const uint8_t *blk = raw + (row * nb + b) * 210;
    reinterpret_cast<uint4 *>(qla)[idx] = load16_bytes(blk + 64 * h + 16 * t);
    reinterpret_cast<uint4 *>(qlb)[idx] = load16_bytes(blk + 64 * h + 32 + 16 * t);
    reinterpret_cast<uint4 *>(qh)[idx] = load16_bytes(blk + 128 + 32 * h + 16 * t);
    const uint8_t *s = blk + 192 + 8 * h + t;
    reinterpret_cast<uint32_t *>(sc)[idx] =
        (uint32_t)s[0] | ((uint32_t)s[2] << 8) | ((uint32_t)s[4] << 16) | ((uint32_t)s[6] << 24);
    if (u % 4 == 0) reinterpret_cast<uint16_t *>(dpl)[row * nb + b] = (uint16_t)(blk[208] | (blk[209] << 8));
}

__global__ void repack_q80_kernel(const uint8_t *raw, int64_t n_rows, int k, uint8_t *qsa, uint8_t *qsb, uint8_t *dpl) {
    const int U8 = k / 32;
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_rows * U8) return;
    const uint8_t *blk = raw + idx * 34;
    reinterpret_cast<uint4 *>(qsa)[idx] = load16_bytes(blk + 2);
    reinterpret_cast<uint4 *>(qsb)[idx] = load16_bytes(blk + 18);
    reinterpret_cast<uint16_t *>(dpl)[idx] = (uint16_t)(blk[0] | (blk[1] << 8));
}

void launch_repack(int type, const void *d_raw, int64_t n_rows, int k, uint8_t *const planes[5], cudaStream_t st) {
    const int threads = 256;
    if (type == T_Q4_K || type == T_Q5_K) {
        const int64_t n = n_rows * (k / 64);
        repack_q4k_kernel<<<(unsigned)((n + threads - 1) / threads), threads, 0, st>>>(
            (const uint8_t *)d_raw, n_rows, k, planes[0], planes[1], planes[2], planes[3], type == T_Q5_K);
    } else if (type == T_Q6_K) {
        const int64_t n = n_rows * (k / 64);
        repack_q6k_kernel<<<(unsigned)((n + threads - 1) / threads), threads, 0, st>>>(
            (const uint8_t *)d_raw, n_rows, k, planes[0], planes[1], planes[2], planes[3], planes[4]);
    } else if (type == T_Q8_0) {
        const int64_t n = n_rows * (k / 32);
        repack_q80_kernel<<<(unsigned)((n + threads - 1) / threads), threads, 0, st>>>((const uint8_t *)d_raw, n_rows, k,
                                                                                        planes[0], planes[1], planes[2]);
    }
}

}  // namespace aios
