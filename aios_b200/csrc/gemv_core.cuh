// gemv_core.cuh — device code shared by the per-op GEMV kernel (gemv.cu) and the persistent fused
// decode kernel (pfd.cu): the shared-memory activation image, the Q8_K / Q8_0 activation quantiser,
// and the per-unit integer dot products on the weight planes of layout.cuh.
//
// Numerics: oracle/aios_oracle.c — see gemv.cu's header.
#pragma once
#include "common.cuh"
#include "kernels.h"
#include "layout.cuh"

namespace aios {

#ifndef GEMV_CORE_WARPS
#error "define GEMV_CORE_WARPS (warps per CTA taking part in the activation prologue) before including gemv_core.cuh"
#endif
// barrier over the GEMV_WARPS warps that take part in the activation prologue: the whole CTA by default; a kernel
// with extra (producer) warps overrides it with a named barrier
#ifndef GEMV_CORE_SYNC
#define GEMV_CORE_SYNC() __syncthreads()
#endif
constexpr int GEMV_WARPS = GEMV_CORE_WARPS;
constexpr int GEMV_THREADS = GEMV_WARPS * 32;
constexpr int GEMV_R = 2;  // rows per warp per pass

// ---------------------------------------------------------------- shared-memory activation image
// act[4][U] uint4 (Q8_0: act[2][U8]) | aux (bsums) | d8
struct ActSmem {
    uint4 *act;
    int32_t *bs32;  // Q4_K/Q5_K: [nb*8] 32-wide sums of q8
    int16_t *bs16;  // Q6_K: [U*4] 16-wide sums, permuted per unit
    float *d8;      // per 256-block (Q8_K) or per 32-block (Q8_0) activation scale
};

__device__ __forceinline__ ActSmem carve(uint8_t *smem, int k) {
    ActSmem s;
    s.act = reinterpret_cast<uint4 *>(smem);
    uint8_t *aux = smem + k;
    s.bs32 = reinterpret_cast<int32_t *>(aux);
    s.bs16 = reinterpret_cast<int16_t *>(aux);
    s.d8 = reinterpret_cast<float *>(aux + k / 8);
    return s;
}

__host__ __device__ inline size_t act_image_bytes(int k) { return (size_t)k + (size_t)k / 8 + (size_t)k / 8 + 64; }

enum : int { PERM_Q4 = 0, PERM_Q6 = 1, PERM_Q80 = 2 };

__device__ __forceinline__ double block_sum(double v, double *red) {
    v = warp_sum_f64(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) red[warp] = v;
    GEMV_CORE_SYNC();
    double t = 0.0;
#pragma unroll
    for (int i = 0; i < GEMV_WARPS; i++) t += red[i];
    GEMV_CORE_SYNC();
    return t;
}

__device__ __forceinline__ void load8(const float *p, float (&v)[8]) {
    const float4 a = reinterpret_cast<const float4 *>(p)[0], b = reinterpret_cast<const float4 *>(p)[1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

// quantise the 8 values this lane holds of 256-slab b and store them in the lane order of the planes.
// Follows quantize_row_q8_K / quantize_row_q8_0 (oracle/aios_oracle.c).
template <int PERM>
__device__ __forceinline__ void quant_slab(const float (&v)[8], int b, int k, const ActSmem &s) {
    const int lane = threadIdx.x & 31;
    int q[8];
    if (PERM == PERM_Q80) {
        float am = 0.f;
#pragma unroll
        for (int i = 0; i < 8; i++) am = fmaxf(am, fabsf(v[i]));
        am = fmaxf(am, __shfl_xor_sync(0xffffffffu, am, 1));
        am = fmaxf(am, __shfl_xor_sync(0xffffffffu, am, 2));
        const float d = __fdiv_rn(am, 127.f);
        const float id = d != 0.f ? __fdiv_rn(1.0f, d) : 0.f;
#pragma unroll
        for (int i = 0; i < 8; i++) q[i] = (int)roundf(__fmul_rn(v[i], id));
        const int u8 = b * 8 + (lane >> 2);
        if ((lane & 3) == 0) s.d8[u8] = __half2float(__float2half_rn(d));
        const int U8 = k / 32;
        const int kk = (lane & 3) >> 1;
        uint2 pk;
        pk.x = (q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((q[3] & 0xff) << 24);
        pk.y = (q[4] & 0xff) | ((q[5] & 0xff) << 8) | ((q[6] & 0xff) << 16) | ((q[7] & 0xff) << 24);
        *reinterpret_cast<uint2 *>(reinterpret_cast<uint8_t *>(s.act + kk * U8 + u8) + (lane & 1) * 8) = pk;
        return;
    }
    // amax with ggml's "first element whose |x| is strictly larger" rule
    // (the signed value travels with its index through the reduction: selecting v[idx & 7] afterwards makes the
    // compiler spill v[] to local memory for an indexed load)
    float mv = 0.f;
    int ai = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        if (fabsf(v[i]) > fabsf(mv)) { mv = v[i]; ai = i; }
    }
    int idx = lane * 8 + ai;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float mv2 = __shfl_xor_sync(0xffffffffu, mv, o);
        const int idx2 = __shfl_xor_sync(0xffffffffu, idx, o);
        const float a = fabsf(mv), a2 = fabsf(mv2);
        if (a2 > a || (a2 == a && idx2 < idx)) { mv = mv2; idx = idx2; }
    }
    const float am = fabsf(mv);
    float d = 0.f;
    if (am == 0.f) {
#pragma unroll
        for (int i = 0; i < 8; i++) q[i] = 0;
    } else {
        const float iscale = __fdiv_rn(-127.f, mv);
#pragma unroll
        for (int i = 0; i < 8; i++) q[i] = min(127, __float2int_rn(__fmul_rn(iscale, v[i])));
        d = __fdiv_rn(1.0f, iscale);
    }
    if (lane == 0) s.d8[b] = d;
    int s8 = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s8 += q[i];
    const int s16 = s8 + __shfl_xor_sync(0xffffffffu, s8, 1);
    const int s32 = s16 + __shfl_xor_sync(0xffffffffu, s16, 2);
    const int U = k / 64;
    uint2 pk;
    pk.x = (q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((q[3] & 0xff) << 24);
    pk.y = (q[4] & 0xff) | ((q[5] & 0xff) << 8) | ((q[6] & 0xff) << 16) | ((q[7] & 0xff) << 24);
    if (PERM == PERM_Q4) {
        const int unit = b * 4 + (lane >> 3), kk = (lane & 7) >> 1;
        *reinterpret_cast<uint2 *>(reinterpret_cast<uint8_t *>(s.act + kk * U + unit) + (lane & 1) * 8) = pk;
        if ((lane & 3) == 0) s.bs32[b * 8 + (lane >> 2)] = s32;
    } else {
        const int h = lane >> 4, kk = (lane & 15) >> 2, t = (lane >> 1) & 1;
        const int unit = b * 4 + h * 2 + t;
        *reinterpret_cast<uint2 *>(reinterpret_cast<uint8_t *>(s.act + kk * U + unit) + (lane & 1) * 8) = pk;
        if ((lane & 1) == 0) s.bs16[unit * 4 + kk] = (int16_t)s16;
    }
}

__device__ __forceinline__ void norm8(float (&v)[8], const float *w, float scale) {
    float wv[8];
    load8(w, wv);
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = __fmul_rn(__fmul_rn(v[i], scale), wv[i]);
}

__device__ __forceinline__ double sumsq8(const float (&v)[8]) {
    double s = 0.0;  // f32 squares summed in double, like ggml_rms_norm (order-free once rounded to f32)
#pragma unroll
    for (int i = 0; i < 8; i++) s += (double)__fmul_rn(v[i], v[i]);
    return s;
}

// Quantise x[0..k) (optionally RMS-normed) into shared memory.  Warp w owns slabs w, w+8, ...
template <int PERM>
__device__ void prep_activations(const float *__restrict__ x, const float *__restrict__ norm_w, float eps, int k,
                                 const ActSmem &s, double *red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nslab = k / 256;
    if (nslab <= GEMV_WARPS) {
        // one slab per warp (K <= 4096 with 16 warps): x and the norm weights are requested together, so the norm
        // weights do not cost a second trip to L2 after the sum of squares
        float v0[8], w0[8];
#pragma unroll
        for (int i = 0; i < 8; i++) v0[i] = 0.f;
        if (warp < nslab) {
            load8(x + warp * 256 + lane * 8, v0);
            if (norm_w) load8(norm_w + warp * 256 + lane * 8, w0);
        }
        if (norm_w) {
            const double ss = block_sum(sumsq8(v0), red);
            const float mean = (float)(ss / (double)k);
            const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
            if (warp < nslab) {
#pragma unroll
                for (int i = 0; i < 8; i++) v0[i] = __fmul_rn(__fmul_rn(v0[i], scale), w0[i]);
            }
        }
        if (warp < nslab) quant_slab<PERM>(v0, warp, k, s);
    } else if (nslab <= 2 * GEMV_WARPS) {
        // each lane keeps its (up to) 16 values in registers — one trip to L2
        const int b0 = warp, b1 = warp + GEMV_WARPS;
        float v0[8], v1[8];
#pragma unroll
        for (int i = 0; i < 8; i++) v0[i] = v1[i] = 0.f;
        if (b0 < nslab) load8(x + b0 * 256 + lane * 8, v0);
        if (b1 < nslab) load8(x + b1 * 256 + lane * 8, v1);
        if (norm_w) {
            const double ss = block_sum(sumsq8(v0) + sumsq8(v1), red);
            const float mean = (float)(ss / (double)k);
            const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
            if (b0 < nslab) norm8(v0, norm_w + b0 * 256 + lane * 8, scale);
            if (b1 < nslab) norm8(v1, norm_w + b1 * 256 + lane * 8, scale);
        }
        if (b0 < nslab) quant_slab<PERM>(v0, b0, k, s);
        if (b1 < nslab) quant_slab<PERM>(v1, b1, k, s);
    } else {
        float scale = 1.f;
        if (norm_w) {
            double ss = 0.0;
            for (int b = warp; b < nslab; b += GEMV_WARPS) {
                float v[8];
                load8(x + b * 256 + lane * 8, v);
                ss += sumsq8(v);
            }
            ss = block_sum(ss, red);
            const float mean = (float)(ss / (double)k);
            scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
        }
        // (software-pipelining this loop — next pair requested before the current one is quantised — was tried: with
        // the held weight images it spills in every GEMV kernel, 1.72 -> 1.76 ms/token)
        for (int b = warp; b < nslab; b += 2 * GEMV_WARPS) {
            const int b1 = b + GEMV_WARPS;
            float v0[8], v1[8];
            load8(x + b * 256 + lane * 8, v0);
            if (b1 < nslab) load8(x + b1 * 256 + lane * 8, v1);
            if (norm_w) {
                norm8(v0, norm_w + b * 256 + lane * 8, scale);
                if (b1 < nslab) norm8(v1, norm_w + b1 * 256 + lane * 8, scale);
            }
            quant_slab<PERM>(v0, b, k, s);
            if (b1 < nslab) quant_slab<PERM>(v1, b1, k, s);
        }
    }
    GEMV_CORE_SYNC();
}

// The norm weights are weights: a kernel can fetch them BEFORE its dependency wait (K <= 4096: 16 per lane) ...
__device__ __forceinline__ void preload_norm(const float *__restrict__ norm_w, int k, float (&nw0)[8], float (&nw1)[8]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nslab = k / 256;
    if (norm_w && nslab <= 2 * GEMV_WARPS) {
        if (warp < nslab) load8(norm_w + warp * 256 + lane * 8, nw0);
        if (warp + GEMV_WARPS < nslab) load8(norm_w + (warp + GEMV_WARPS) * 256 + lane * 8, nw1);
    }
}

// ... and hand them to this variant of prep_activations.  Same values; two differences in latency: no trip to L2 for
// the norm weights after the wait, and for a big un-normed K (ffn_down) four slabs of x in flight per lane, not two.
template <int PERM>
__device__ void prep_activations_pre(const float *__restrict__ x, const float *__restrict__ norm_w, float eps, int k,
                                     const ActSmem &s, double *red, const float (&nw0)[8], const float (&nw1)[8]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nslab = k / 256;
    if (nslab <= 2 * GEMV_WARPS) {
        const int b0 = warp, b1 = warp + GEMV_WARPS;
        float v0[8], v1[8];
#pragma unroll
        for (int i = 0; i < 8; i++) v0[i] = v1[i] = 0.f;
        if (b0 < nslab) load8(x + b0 * 256 + lane * 8, v0);
        if (b1 < nslab) load8(x + b1 * 256 + lane * 8, v1);
        if (norm_w) {
            const double ss = block_sum(sumsq8(v0) + sumsq8(v1), red);
            const float mean = (float)(ss / (double)k);
            const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
#pragma unroll
            for (int i = 0; i < 8; i++) {
                v0[i] = __fmul_rn(__fmul_rn(v0[i], scale), nw0[i]);
                v1[i] = __fmul_rn(__fmul_rn(v1[i], scale), nw1[i]);
            }
        }
        if (b0 < nslab) quant_slab<PERM>(v0, b0, k, s);
        if (b1 < nslab) quant_slab<PERM>(v1, b1, k, s);
        GEMV_CORE_SYNC();
    } else if (norm_w) {
        prep_activations<PERM>(x, norm_w, eps, k, s, red);
    } else {
        for (int b = warp; b < nslab; b += 4 * GEMV_WARPS) {
            float v[4][8];
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (b + j * GEMV_WARPS < nslab) load8(x + (b + j * GEMV_WARPS) * 256 + lane * 8, v[j]);
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (b + j * GEMV_WARPS < nslab) quant_slab<PERM>(v[j], b + j * GEMV_WARPS, k, s);
        }
        GEMV_CORE_SYNC();
    }
}

// ---------------------------------------------------------------- weight registers of one pass
struct RowPtrs {
    const uint8_t *p[5];  // row starts inside each plane
};

template <int TYPE>
__device__ __forceinline__ RowPtrs row_ptrs(const WPlanes &w, int64_t row) {
    RowPtrs r;
    const int64_t U = w.k / 64, nb = w.k / 256, U8 = w.k / 32;
    if (TYPE == T_Q4_K || TYPE == T_Q5_K) {
        r.p[0] = w.p[0] + row * U * 16; r.p[1] = w.p[1] + row * U * 16; r.p[2] = w.p[2] + row * nb * 16;
        r.p[3] = TYPE == T_Q5_K ? w.p[3] + row * U * 8 : nullptr; r.p[4] = nullptr;
    } else if (TYPE == T_Q6_K) {
        r.p[0] = w.p[0] + row * U * 16; r.p[1] = w.p[1] + row * U * 16; r.p[2] = w.p[2] + row * U * 16;
        r.p[3] = w.p[3] + row * U * 4; r.p[4] = w.p[4] + row * nb * 2;
    } else {
        r.p[0] = w.p[0] + row * U8 * 16; r.p[1] = w.p[1] + row * U8 * 16; r.p[2] = w.p[2] + row * U8 * 2;
        r.p[3] = r.p[4] = nullptr;
    }
    return r;
}

// lane-iterations per pass: a pass covers ITS*32 units of each of the GEMV_R rows
template <int TYPE> struct PassCfg { static constexpr int ITS = 2; };
template <> struct PassCfg<T_Q8_0> { static constexpr int ITS = 4; };

template <int TYPE>
struct PassRegs {
    static constexpr int ITS = PassCfg<TYPE>::ITS;
    uint4 a[GEMV_R][ITS];   // QSa / QLa / QSa
    uint4 b[GEMV_R][ITS];   // QSb / QLb / QSb
    uint4 c[GEMV_R][ITS];   // META (Q4_K/Q5_K) / QH (Q6_K)
    uint32_t e[GEMV_R][ITS];  // SC word (Q6_K) / d bits (Q8_0) / qh5 low (Q5_K)
    uint32_t f[GEMV_R][ITS];  // d bits (Q6_K) / qh5 high (Q5_K)
};

// dp4a with an unsigned first operand (bytes 0..255) and a signed second one
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// FULL: the pass is known to be entirely in range (no per-lane predicates)
template <int TYPE, bool FULL, typename Regs>
__device__ __forceinline__ void load_pass(Regs &g, const RowPtrs (&rp)[GEMV_R], int u0, int nu, uint64_t pol) {
    const int lane = threadIdx.x & 31;
    constexpr int ITS = Regs::ITS;
#pragma unroll
    for (int it = 0; it < ITS; it++) {
        const int u = u0 + it * 32 + lane;
        if (FULL || u < nu) {
#pragma unroll
            for (int r = 0; r < GEMV_R; r++) {
                g.a[r][it] = ldg_stream(rp[r].p[0] + (size_t)u * 16, pol);
                g.b[r][it] = ldg_stream(rp[r].p[1] + (size_t)u * 16, pol);
                if (TYPE == T_Q4_K || TYPE == T_Q5_K) {
                    g.c[r][it] = ldg_stream(rp[r].p[2] + (size_t)(u >> 2) * 16, pol);
                    if (TYPE == T_Q5_K) {
                        const uint2 h = *reinterpret_cast<const uint2 *>(rp[r].p[3] + (size_t)u * 8);
                        g.e[r][it] = h.x; g.f[r][it] = h.y;
                    }
                } else if (TYPE == T_Q6_K) {
                    g.c[r][it] = ldg_stream(rp[r].p[2] + (size_t)u * 16, pol);
                    g.e[r][it] = ldg_stream_u32(rp[r].p[3] + (size_t)u * 4);
                    g.f[r][it] = ldg_stream_u16(rp[r].p[4] + (size_t)(u >> 2) * 2);
                } else {
                    g.e[r][it] = ldg_stream_u16(rp[r].p[2] + (size_t)u * 2);
                }
            }
        }
    }
}

template <int TYPE, bool FULL, typename Regs>
__device__ __forceinline__ void compute_pass(const Regs &g, int u0, int nu, int k, const ActSmem &s,
                                             double (&acc)[GEMV_R]) {
    const int lane = threadIdx.x & 31;
    constexpr int ITS = Regs::ITS;
#pragma unroll
    for (int it = 0; it < ITS; it++) {
        const int u = u0 + it * 32 + lane;
        if (FULL || u < nu) {
            if (TYPE == T_Q4_K || TYPE == T_Q5_K) {
                const int U = k / 64;
                const uint4 a0 = s.act[u], a1 = s.act[U + u], a2 = s.act[2 * U + u], a3 = s.act[3 * U + u];
                const int2 bs = *reinterpret_cast<const int2 *>(s.bs32 + 2 * u);
                const float d8 = s.d8[u >> 2];
                const uint32_t A0[4] = {a0.x, a0.y, a0.z, a0.w}, A1[4] = {a1.x, a1.y, a1.z, a1.w};
                const uint32_t A2[4] = {a2.x, a2.y, a2.z, a2.w}, A3[4] = {a3.x, a3.y, a3.z, a3.w};
                const int p = lane & 3;  // unit within its block (u0 is a multiple of 4)
#pragma unroll
                for (int r = 0; r < GEMV_R; r++) {
                    const uint32_t QA[4] = {g.a[r][it].x, g.a[r][it].y, g.a[r][it].z, g.a[r][it].w};
                    const uint32_t QB[4] = {g.b[r][it].x, g.b[r][it].y, g.b[r][it].z, g.b[r][it].w};
                    int dlo = 0, dhi = 0;
                    if (TYPE == T_Q4_K) {
                        int dhi16 = 0;  // high nibbles stay in place (value*16): one mask, no shift
#pragma unroll
                        for (int w = 0; w < 4; w++) {
                            dlo = __dp4a((int)(QA[w] & 0x0F0F0F0Fu), (int)A0[w], dlo);
                            dlo = __dp4a((int)(QB[w] & 0x0F0F0F0Fu), (int)A1[w], dlo);
                            dhi16 = dp4a_us(QA[w] & 0xF0F0F0F0u, A2[w], dhi16);
                            dhi16 = dp4a_us(QB[w] & 0xF0F0F0F0u, A3[w], dhi16);
                        }
                        dhi = dhi16 >> 4;  // exact: a multiple of 16
                    } else {
#pragma unroll
                        for (int w = 0; w < 4; w++) {
                            uint32_t la = QA[w] & 0x0F0F0F0Fu, ha = (QA[w] >> 4) & 0x0F0F0F0Fu;
                            uint32_t lb = QB[w] & 0x0F0F0F0Fu, hb = (QB[w] >> 4) & 0x0F0F0F0Fu;
                            // 5th bits: bit (4w+i) of e -> byte i of word w (low sub-block; QSa: l = 4w+i, QSb: l = 16+4w+i)
                            const uint32_t xa = (g.e[r][it] >> (4 * w)) & 0xFu, xb = (g.e[r][it] >> (16 + 4 * w)) & 0xFu;
                            const uint32_t ya = (g.f[r][it] >> (4 * w)) & 0xFu, yb = (g.f[r][it] >> (16 + 4 * w)) & 0xFu;
                            la |= ((xa * 0x00204081u) & 0x01010101u) << 4;
                            lb |= ((xb * 0x00204081u) & 0x01010101u) << 4;
                            ha |= ((ya * 0x00204081u) & 0x01010101u) << 4;
                            hb |= ((yb * 0x00204081u) & 0x01010101u) << 4;
                            dlo = __dp4a((int)la, (int)A0[w], dlo);
                            dlo = __dp4a((int)lb, (int)A1[w], dlo);
                            dhi = __dp4a((int)ha, (int)A2[w], dhi);
                            dhi = __dp4a((int)hb, (int)A3[w], dhi);
                        }
                    }
                    // META word of my unit: 4 x 6-bit fields; d|dmin is spread over the top bytes of the 4 words
                    const uint4 mt = g.c[r][it];
                    const uint32_t mw = (p & 2) ? ((p & 1) ? mt.w : mt.z) : ((p & 1) ? mt.y : mt.x);
                    const uint32_t ddm = __byte_perm(__byte_perm(mt.x, mt.y, 0x0073), __byte_perm(mt.z, mt.w, 0x0073), 0x5410);
                    const float2 dd = __half22float2(*reinterpret_cast<const __half2 *>(&ddm));
                    const int sc_lo = mw & 63, sc_hi = (mw >> 6) & 63, m_lo = (mw >> 12) & 63, m_hi = (mw >> 18) & 63;
                    const int isum = sc_lo * dlo + sc_hi * dhi;
                    const int msum = m_lo * bs.x + m_hi * bs.y;
                    // one f32 term per unit, summed in double (order-free; oracle: vec_dot_q4_K_q8_K)
                    const float term = __fmaf_rn(__fmul_rn(dd.x, d8), (float)isum, -__fmul_rn(__fmul_rn(dd.y, d8), (float)msum));
                    acc[r] += (double)term;
                }
            } else if (TYPE == T_Q6_K) {
                const int U = k / 64;
                const uint4 a0 = s.act[u], a1 = s.act[U + u], a2 = s.act[2 * U + u], a3 = s.act[3 * U + u];
                const uint2 bsp = *reinterpret_cast<const uint2 *>(s.bs16 + 4 * u);
                const int b0 = (int)(int16_t)(bsp.x & 0xffff), b1 = (int)(int16_t)(bsp.x >> 16);
                const int b2 = (int)(int16_t)(bsp.y & 0xffff), b3 = (int)(int16_t)(bsp.y >> 16);
                const float d8 = s.d8[u >> 2];
                const uint32_t A0[4] = {a0.x, a0.y, a0.z, a0.w}, A1[4] = {a1.x, a1.y, a1.z, a1.w};
                const uint32_t A2[4] = {a2.x, a2.y, a2.z, a2.w}, A3[4] = {a3.x, a3.y, a3.z, a3.w};
#pragma unroll
                for (int r = 0; r < GEMV_R; r++) {
                    const uint32_t LA[4] = {g.a[r][it].x, g.a[r][it].y, g.a[r][it].z, g.a[r][it].w};
                    const uint32_t LB[4] = {g.b[r][it].x, g.b[r][it].y, g.b[r][it].z, g.b[r][it].w};
                    const uint32_t HH[4] = {g.c[r][it].x, g.c[r][it].y, g.c[r][it].z, g.c[r][it].w};
                    int d0 = 0, d1 = 0, d2 = 0, d3 = 0;
#pragma unroll
                    for (int w = 0; w < 4; w++) {
                        const uint32_t h = HH[w];
                        const uint32_t q0 = (LA[w] & 0x0F0F0F0Fu) | ((h << 4) & 0x30303030u);
                        const uint32_t q1 = (LB[w] & 0x0F0F0F0Fu) | ((h << 2) & 0x30303030u);
                        const uint32_t q2 = ((LA[w] >> 4) & 0x0F0F0F0Fu) | (h & 0x30303030u);
                        const uint32_t q3 = ((LB[w] >> 4) & 0x0F0F0F0Fu) | ((h >> 2) & 0x30303030u);
                        d0 = __dp4a((int)q0, (int)A0[w], d0);
                        d1 = __dp4a((int)q1, (int)A1[w], d1);
                        d2 = __dp4a((int)q2, (int)A2[w], d2);
                        d3 = __dp4a((int)q3, (int)A3[w], d3);
                    }
                    const uint32_t scw = g.e[r][it];
                    const int s0 = (int)(int8_t)(scw & 0xff), s1 = (int)(int8_t)((scw >> 8) & 0xff);
                    const int s2 = (int)(int8_t)((scw >> 16) & 0xff), s3 = (int)(int8_t)(scw >> 24);
                    const int isum = s0 * (d0 - 32 * b0) + s1 * (d1 - 32 * b1) + s2 * (d2 - 32 * b2) + s3 * (d3 - 32 * b3);
                    acc[r] += (double)__fmul_rn(__fmul_rn(h2f_bits((uint16_t)g.f[r][it]), d8), (float)isum);
                }
            } else {
                const int U8 = k / 32;
                const uint4 a0 = s.act[u], a1 = s.act[U8 + u];
                const float d8 = s.d8[u];
#pragma unroll
                for (int r = 0; r < GEMV_R; r++) {
                    int d = 0;
                    d = __dp4a((int)g.a[r][it].x, (int)a0.x, d); d = __dp4a((int)g.a[r][it].y, (int)a0.y, d);
                    d = __dp4a((int)g.a[r][it].z, (int)a0.z, d); d = __dp4a((int)g.a[r][it].w, (int)a0.w, d);
                    d = __dp4a((int)g.b[r][it].x, (int)a1.x, d); d = __dp4a((int)g.b[r][it].y, (int)a1.y, d);
                    d = __dp4a((int)g.b[r][it].z, (int)a1.z, d); d = __dp4a((int)g.b[r][it].w, (int)a1.w, d);
                    acc[r] += (double)__fmul_rn((float)d, __fmul_rn(h2f_bits((uint16_t)g.e[r][it]), d8));
                }
            }
        }
    }
}

// sum acc[0] and acc[1] over the warp: after the call every lane of the lower half-warp holds the
// total of row 0 and every lane of the upper half holds row 1 (5 double shuffles instead of 10)
__device__ __forceinline__ double warp_sum2_f64(double a0, double a1) {
    const int lane = threadIdx.x & 31;
    const bool hi = lane & 16;
    const double give = hi ? a0 : a1, keep = hi ? a1 : a0;
    double v = keep + __shfl_xor_sync(0xffffffffu, give, 16);
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// The activation image was built by an earlier launch (quant_rows_kernel) — copy it into shared memory.
__device__ __forceinline__ void copy_activation_image(const uint8_t *__restrict__ image, int k, uint8_t *smem) {
    const int n16 = (int)(act_image_bytes(k) / 16);
    for (int i = threadIdx.x; i < n16; i += GEMV_THREADS)
        reinterpret_cast<uint4 *>(smem)[i] = __ldcg(reinterpret_cast<const uint4 *>(image) + i);
    GEMV_CORE_SYNC();
}

// EPI_QKV_ROPE: the warp holds the dots of rows row0 (lanes 0-15) and row0+1 (lanes 16-31) of segment seg — an
// adjacent (even, odd) pair, which is what GGML's NORM-style RoPE rotates.  Same operations, in the same order, as
// the attention kernels' own prologue (attention_split.cu) and the oracle: rotate (q, k), round to f16.
__device__ __forceinline__ void qkv_rope_store(const GemvParams &p, int seg, int row0, double tot, int n_past) {
    const float x0 = (float)__shfl_sync(0xffffffffu, tot, 0), x1 = (float)__shfl_sync(0xffffffffu, tot, 16);
    if ((threadIdx.x & 31) != 0) return;
    const RopeKv &r = p.rope;
    float a = x0, b = x1;
    if (seg < 2) {
        const int d = row0 % r.hd;
        if (d < r.n_rot) {
            const int i = d >> 1;
            const float c = r.cos[(size_t)n_past * (r.n_rot / 2) + i], s = r.sin[(size_t)n_past * (r.n_rot / 2) + i];
            a = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, s));
            b = __fadd_rn(__fmul_rn(x0, s), __fmul_rn(x1, c));
        }
    }
    const uint16_t ha = f2h_bits(a), hb = f2h_bits(b);
    if (seg == 0) {
        *reinterpret_cast<float2 *>(p.out[0] + row0) = make_float2(h2f_bits(ha), h2f_bits(hb));
    } else {
        uint16_t *dst = (seg == 1 ? r.k_cache : r.v_cache) + (size_t)n_past * r.kv_stride + row0;
        *reinterpret_cast<uint32_t *>(dst) = (uint32_t)ha | ((uint32_t)hb << 16);
    }
}

// order-preserving float -> u32 so that packed (value, ~index) keys can be atomicMax-ed
__device__ __forceinline__ unsigned long long argmax_key(float v, int idx) {
    uint32_t b = __float_as_uint(v);
    b = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
    return ((unsigned long long)b << 32) | (uint32_t)(0x7fffffff - idx);
}

}  // namespace aios
