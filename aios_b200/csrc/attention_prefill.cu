// attention_prefill.cu — causal attention for a whole prompt (BASELINE.json configs[2]) and the
// RoPE + KV-append step in front of it.  Part of the tensor-core prefill path (prefill_tc.cu): float arithmetic with a
// stated tolerance, not the decode path's bit-exact formulation.
//
// rope_kv_rows_kernel: for T new tokens at positions pos0.., rotate q and k (NORM-style RoPE, adjacent pairs, the
// host-built cos/sin table — the same operations as the decode path), round q, k, v to f16, append k / v to the
// sequence's KV range (position-major pages) and leave q as f16 [T][n_head*hd] for the attention kernel.
//
// attn_prefill_kernel: flash attention, one CTA = 64 queries of one head (4 warps x 16 rows), keys and values streamed
// from the KV range in tiles of 64 positions (cp.async, double buffered), S = Q.K^T and O += P.V on the tensor cores
// (mma.sync m16n8k16, f16 operands, f32 accumulation), online soft-max in f32 registers.  Queries see the cached
// positions before the chunk plus the chunk's own positions up to themselves (causal).  The four query heads of a GQA
// group are separate CTAs that read the same K/V rows (L2 hits).
#include <cuda_fp16.h>

#include "common.cuh"
#include "kernels.h"

namespace aios {

struct RopeRowsParams {
    const float *qkv;       // [T][lq] raw projections, lq = qd + 2*kvd
    __half *qh;             // [T][qd] out
    uint16_t *k_cache, *v_cache;  // this layer; row of position j at + j * kv_stride
    const float *cos, *sin; // [n_ctx][n_rot/2]
    int32_t T, pos0, n_head, n_kv, hd, n_rot, kv_stride;
};

__global__ void rope_kv_rows_kernel(const RopeRowsParams p) {
    const int t = blockIdx.x;
    const int qd = p.n_head * p.hd, kvd = p.n_kv * p.hd, lq = qd + 2 * kvd;
    const float *row = p.qkv + (size_t)t * lq;
    const int pos = p.pos0 + t;
    const float *cs = p.cos + (size_t)pos * (p.n_rot / 2), *sn = p.sin + (size_t)pos * (p.n_rot / 2);
    // pairs (2i, 2i+1) of q | k | v
    for (int i = threadIdx.x; i < lq / 2; i += blockDim.x) {
        const int e = 2 * i;
        float x0 = row[e], x1 = row[e + 1];
        if (e < qd + kvd) {  // q and k are rotated
            const int d = (e < qd ? e : e - qd) % p.hd;
            if (d < p.n_rot) {
                const float c = cs[d >> 1], s = sn[d >> 1];
                const float r0 = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, s));
                const float r1 = __fadd_rn(__fmul_rn(x0, s), __fmul_rn(x1, c));
                x0 = r0; x1 = r1;
            }
        }
        const __half2 h = __floats2half2_rn(x0, x1);
        if (e < qd) *reinterpret_cast<__half2 *>(p.qh + (size_t)t * qd + e) = h;
        else if (e < qd + kvd) *reinterpret_cast<__half2 *>(p.k_cache + (size_t)pos * p.kv_stride + (e - qd)) = h;
        else *reinterpret_cast<__half2 *>(p.v_cache + (size_t)pos * p.kv_stride + (e - qd - kvd)) = h;
    }
}

void launch_rope_kv_rows(const float *qkv, __half *qh, uint16_t *k_cache, uint16_t *v_cache, const float *cos, const float *sin, int T,
                         int pos0, int n_head, int n_kv, int hd, int n_rot, int kv_stride, cudaStream_t st) {
    RopeRowsParams p{qkv, qh, k_cache, v_cache, cos, sin, T, pos0, n_head, n_kv, hd, n_rot, kv_stride};
    rope_kv_rows_kernel<<<T, 256, 0, st>>>(p);
    CUDA_CHECK(cudaGetLastError());
}

// ---------------------------------------------------------------- flash attention (prefill)
constexpr int FA_BM = 64, FA_BN = 64, FA_THREADS = 128;

struct FaParams {
    const __half *qh;             // [T][n_head*hd]
    __half *out;                  // [T][n_head*hd]
    const uint16_t *k_cache, *v_cache;
    int32_t T, pos0, n_head, n_kv, kv_stride;
    float scale_log2;             // (1/sqrt(hd)) * log2(e)
};

__device__ __forceinline__ void mma_f16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void *smem_row) {
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(smem_row);
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], const void *smem_row) {
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(smem_row);
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t *>(&h);
}

template <int HD>
__global__ void __launch_bounds__(FA_THREADS) attn_prefill_kernel(const FaParams p) {
    constexpr int LDS = HD + 8;               // padded row (halves): ldmatrix rows land in different banks
    constexpr int KS = HD / 16;               // k-steps of the S = Q.K^T product
    constexpr int NT_O = HD / 8;              // n-tiles of the output
    extern __shared__ __align__(16) uint8_t fa_smem[];
    __half *ks = reinterpret_cast<__half *>(fa_smem);        // [2][FA_BN][LDS]
    __half *vs = ks + 2 * FA_BN * LDS;                       // [2][FA_BN][LDS]
    const int head = blockIdx.y, g = head / (p.n_head / p.n_kv);
    const int m0 = blockIdx.x * FA_BM;                       // first query row (token index inside the chunk)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gq = lane >> 2, qq = lane & 3;
    const int qd = p.n_head * HD;
    const int row_a = m0 + warp * 16 + gq, row_b = row_a + 8;    // my two query rows
    // keys this CTA needs: positions [0, pos0 + min(T, m0 + FA_BM))
    const int n_keys = p.pos0 + min(p.T, m0 + FA_BM);
    const int n_tiles = (n_keys + FA_BN - 1) / FA_BN;

    // Q fragments (registers, whole kernel)
    uint32_t qf[KS][4];
    {
        const __half *qa = p.qh + (size_t)min(row_a, p.T - 1) * qd + head * HD;
        const __half *qb = p.qh + (size_t)min(row_b, p.T - 1) * qd + head * HD;
#pragma unroll
        for (int k = 0; k < KS; k++) {
            qf[k][0] = *reinterpret_cast<const uint32_t *>(qa + 16 * k + 2 * qq);
            qf[k][1] = *reinterpret_cast<const uint32_t *>(qb + 16 * k + 2 * qq);
            qf[k][2] = *reinterpret_cast<const uint32_t *>(qa + 16 * k + 8 + 2 * qq);
            qf[k][3] = *reinterpret_cast<const uint32_t *>(qb + 16 * k + 8 + 2 * qq);
        }
    }
    auto load_tile = [&](int tile, int buf) {
        const int j0 = tile * FA_BN;
        constexpr int CPR = HD / 8;  // 16-byte chunks per row
        for (int c = threadIdx.x; c < FA_BN * CPR; c += FA_THREADS) {
            const int r = c / CPR, ch = c % CPR;
            const int j = min(j0 + r, n_keys - 1);  // rows beyond the last key repeat it (masked below)
            const uint32_t dk = (uint32_t)__cvta_generic_to_shared(ks + ((size_t)buf * FA_BN + r) * LDS + ch * 8);
            const uint32_t dv = (uint32_t)__cvta_generic_to_shared(vs + ((size_t)buf * FA_BN + r) * LDS + ch * 8);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dk), "l"(p.k_cache + (size_t)j * p.kv_stride + g * HD + ch * 8) : "memory");
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dv), "l"(p.v_cache + (size_t)j * p.kv_stride + g * HD + ch * 8) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    float o[NT_O][4];
#pragma unroll
    for (int i = 0; i < NT_O; i++) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m_a = -INFINITY, m_b = -INFINITY, l_a = 0.f, l_b = 0.f;
    const int pos_a = p.pos0 + row_a, pos_b = p.pos0 + row_b;  // absolute positions: key j is visible iff j <= pos

    load_tile(0, 0);
    for (int tile = 0; tile < n_tiles; tile++) {
        const int buf = tile & 1;
        if (tile + 1 < n_tiles) {
            load_tile(tile + 1, buf ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        const __half *kt = ks + (size_t)buf * FA_BN * LDS, *vt = vs + (size_t)buf * FA_BN * LDS;
        const int j0 = tile * FA_BN;
        // ---- S = Q . K^T for 16 rows x 64 keys
        float s[FA_BN / 8][4];
#pragma unroll
        for (int n = 0; n < FA_BN / 8; n++) s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.f;
#pragma unroll
        for (int k = 0; k < KS; k++) {
#pragma unroll
            for (int n2 = 0; n2 < FA_BN / 16; n2++) {
                // matrices: (keys 16*n2 + 0..7, dims 16k + 0..7), (same keys, dims + 8), (keys + 8, dims), (keys + 8, dims + 8)
                uint32_t b[4];
                const int mi = lane >> 3, r8 = lane & 7;
                ldsm_x4(b, kt + (size_t)(16 * n2 + (mi >> 1) * 8 + r8) * LDS + 16 * k + (mi & 1) * 8);
                mma_f16(s[2 * n2], qf[k], b[0], b[1]);
                mma_f16(s[2 * n2 + 1], qf[k], b[2], b[3]);
            }
        }
        // ---- mask (causal + keys beyond n_keys), scale, online soft-max
        float mx_a = m_a, mx_b = m_b;
#pragma unroll
        for (int n = 0; n < FA_BN / 8; n++) {
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int j = j0 + 8 * n + 2 * qq + e;
                s[n][e] = j <= pos_a ? s[n][e] * p.scale_log2 : -INFINITY;
                s[n][2 + e] = j <= pos_b ? s[n][2 + e] * p.scale_log2 : -INFINITY;
                mx_a = fmaxf(mx_a, s[n][e]);
                mx_b = fmaxf(mx_b, s[n][2 + e]);
            }
        }
        mx_a = fmaxf(mx_a, __shfl_xor_sync(0xffffffffu, mx_a, 1)); mx_a = fmaxf(mx_a, __shfl_xor_sync(0xffffffffu, mx_a, 2));
        mx_b = fmaxf(mx_b, __shfl_xor_sync(0xffffffffu, mx_b, 1)); mx_b = fmaxf(mx_b, __shfl_xor_sync(0xffffffffu, mx_b, 2));
        // (a row's first visible key is position 0, which is in tile 0: the running max is finite from the first tile on)
        const float corr_a = exp2f(m_a - mx_a), corr_b = exp2f(m_b - mx_b);
        m_a = mx_a; m_b = mx_b;
        float sum_a = 0.f, sum_b = 0.f;
        uint32_t pf[FA_BN / 16][4];  // P as A fragments (16 rows x 16 keys per k-step)
#pragma unroll
        for (int n = 0; n < FA_BN / 8; n++) {
            const float p0 = exp2f(s[n][0] - m_a), p1 = exp2f(s[n][1] - m_a), p2 = exp2f(s[n][2] - m_b), p3 = exp2f(s[n][3] - m_b);
            sum_a += p0 + p1; sum_b += p2 + p3;
            pf[n >> 1][(n & 1) * 2 + 0] = pack_h2(p0, p1);
            pf[n >> 1][(n & 1) * 2 + 1] = pack_h2(p2, p3);
        }
        l_a = l_a * corr_a + sum_a; l_b = l_b * corr_b + sum_b;
#pragma unroll
        for (int i = 0; i < NT_O; i++) { o[i][0] *= corr_a; o[i][1] *= corr_a; o[i][2] *= corr_b; o[i][3] *= corr_b; }
        // ---- O += P . V
#pragma unroll
        for (int kk = 0; kk < FA_BN / 16; kk++) {
#pragma unroll
            for (int n2 = 0; n2 < NT_O / 2; n2++) {
                // transposed 8x8 tiles of V: (keys 16kk + 0..7, dims 16*n2 + 0..7), (keys + 8, dims), (keys, dims + 8), (keys + 8, dims + 8)
                uint32_t b[4];
                const int mi = lane >> 3, r8 = lane & 7;
                ldsm_x4_trans(b, vt + (size_t)(16 * kk + (mi & 1) * 8 + r8) * LDS + 16 * n2 + (mi >> 1) * 8);
                mma_f16(o[2 * n2], pf[kk], b[0], b[1]);
                mma_f16(o[2 * n2 + 1], pf[kk], b[2], b[3]);
            }
        }
        __syncthreads();  // the buffer is refilled two iterations from now
    }
    // row sums live in the 4 lanes of a quad
    l_a += __shfl_xor_sync(0xffffffffu, l_a, 1); l_a += __shfl_xor_sync(0xffffffffu, l_a, 2);
    l_b += __shfl_xor_sync(0xffffffffu, l_b, 1); l_b += __shfl_xor_sync(0xffffffffu, l_b, 2);
    const float inv_a = 1.f / l_a, inv_b = 1.f / l_b;
#pragma unroll
    for (int i = 0; i < NT_O; i++) {
        const int d = 8 * i + 2 * qq;
        if (row_a < p.T) *reinterpret_cast<uint32_t *>(p.out + (size_t)row_a * qd + head * HD + d) = pack_h2(o[i][0] * inv_a, o[i][1] * inv_a);
        if (row_b < p.T) *reinterpret_cast<uint32_t *>(p.out + (size_t)row_b * qd + head * HD + d) = pack_h2(o[i][2] * inv_b, o[i][3] * inv_b);
    }
}

void launch_attn_prefill(const __half *qh, __half *out, const uint16_t *k_cache, const uint16_t *v_cache, int T, int pos0, int n_head,
                         int n_kv, int hd, int kv_stride, cudaStream_t st) {
    FaParams p{qh, out, k_cache, v_cache, T, pos0, n_head, n_kv, kv_stride, (1.0f / sqrtf((float)hd)) * 1.4426950408889634f};
    const dim3 grid((T + FA_BM - 1) / FA_BM, n_head);
    const size_t smem = (size_t)4 * FA_BN * (hd + 8) * 2;
    if (hd == 128) {
        static PerDeviceOnce once;
        once.run([] { CUDA_CHECK(cudaFuncSetAttribute(attn_prefill_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024)); });
        attn_prefill_kernel<128><<<grid, FA_THREADS, smem, st>>>(p);
    } else if (hd == 64) {
        static PerDeviceOnce once;
        once.run([] { CUDA_CHECK(cudaFuncSetAttribute(attn_prefill_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024)); });
        attn_prefill_kernel<64><<<grid, FA_THREADS, smem, st>>>(p);
    } else {
        throw CudaError("prefill attention: head_dim 64 and 128 are built");
    }
    CUDA_CHECK(cudaGetLastError());
}

}  // namespace aios
