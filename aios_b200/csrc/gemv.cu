// gemv.cu — the decode hot loop's kernel: y[N] = dq(W[N,K]) . x[K] for Q4_K / Q5_K / Q6_K / Q8_0
// weights at batch 1, HBM-bandwidth bound.
//
// Numerics mirror llama.cpp's CPU path (oracle/aios_oracle.c): the
// input vector is quantised to Q8_K (K-quants) or Q8_0 (Q8_0 weights), dot products are
// integer inside a block (dp4a) and float across blocks.  Only the float summation order
// differs from the oracle's generic kernel.
//
// One launch = [optional RMSNorm] + activation quantisation (done redundantly by every CTA
// into shared memory: K bytes, so no separate launch and no HBM round trip) + the weight
// stream (128-bit coalesced no-allocate loads from the planes of layout.cuh, two rows and
// two lane-iterations = up to 12 loads in flight per lane) + warp-shuffle reduction + a fused
// epilogue (store / residual add / SwiGLU).  Up to three matrices that share the input
// (attn_q/k/v; ffn_gate/up) ride in one launch; CTAs are split between them by byte share.
#include "common.cuh"
#include "kernels.h"
#include "layout.cuh"

namespace aios {

constexpr int GEMV_THREADS = 256;
constexpr int GEMV_WARPS = GEMV_THREADS / 32;
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

size_t gemv_smem_bytes(int k) { return (size_t)k + (size_t)k / 8 + (size_t)k / 8 + 64; }

enum : int { PERM_Q4 = 0, PERM_Q6 = 1, PERM_Q80 = 2 };

__device__ __forceinline__ double block_sum(double v, double *red) {
    v = warp_sum_f64(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int i = 0; i < GEMV_WARPS; i++) t += red[i];
    __syncthreads();
    return t;
}

// Quantise x[0..k) (optionally RMS-normed) into shared memory in the lane order the weight
// planes expect.  Follows quantize_row_q8_K / quantize_row_q8_0 (oracle/aios_oracle.c).
template <int PERM>
__device__ void prep_activations(const float *__restrict__ x, const float *__restrict__ norm_w, float eps, int k,
                                 const ActSmem &s, double *red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float scale = 1.f;
    if (norm_w) {
        // sum of squares in double like ggml (order-free once rounded to f32); x*x itself is an f32 product
        double ss = 0.0;
        for (int i = threadIdx.x * 4; i < k; i += GEMV_THREADS * 4) {
            const float4 v = *reinterpret_cast<const float4 *>(x + i);
            ss += (double)__fmul_rn(v.x, v.x) + (double)__fmul_rn(v.y, v.y) + (double)__fmul_rn(v.z, v.z) +
                  (double)__fmul_rn(v.w, v.w);
        }
        ss = block_sum(ss, red);
        const float mean = (float)(ss / (double)k);
        scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
    }
    const int nslab = k / 256;
    for (int b = warp; b < nslab; b += GEMV_WARPS) {
        float v[8];
        {
            const float4 *xp = reinterpret_cast<const float4 *>(x + b * 256 + lane * 8);
            const float4 v0 = xp[0], v1 = xp[1];
            v[0] = v0.x; v[1] = v0.y; v[2] = v0.z; v[3] = v0.w;
            v[4] = v1.x; v[5] = v1.y; v[6] = v1.z; v[7] = v1.w;
        }
        if (norm_w) {
            const float4 *wp = reinterpret_cast<const float4 *>(norm_w + b * 256 + lane * 8);
            const float4 w0 = wp[0], w1 = wp[1];
            const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = __fmul_rn(__fmul_rn(v[i], scale), w[i]);
        }
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
        } else {
            // amax with ggml's "first element whose |x| is strictly larger" rule
            float am = 0.f;
            int ai = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const float ax = fabsf(v[i]);
                if (ax > am) { am = ax; ai = i; }
            }
            int idx = lane * 8 + ai;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float am2 = __shfl_xor_sync(0xffffffffu, am, o);
                const int idx2 = __shfl_xor_sync(0xffffffffu, idx, o);
                if (am2 > am || (am2 == am && idx2 < idx)) { am = am2; idx = idx2; }
            }
            float cand = 0.f;
#pragma unroll
            for (int i = 0; i < 8; i++) if (i == (idx & 7)) cand = v[i];
            const float mv = __shfl_sync(0xffffffffu, cand, idx >> 3);
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
    }
    __syncthreads();
}

// ---------------------------------------------------------------- per-type row dot products
struct RowPtrs {
    const uint8_t *p[5];  // row starts inside each plane
};

__device__ __forceinline__ RowPtrs row_ptrs(const WPlanes &w, int64_t row) {
    RowPtrs r;
    const int64_t U = w.k / 64, nb = w.k / 256, U8 = w.k / 32;
    switch (w.type) {
        case T_Q4_K:
        case T_Q5_K:
            r.p[0] = w.p[0] + row * U * 16; r.p[1] = w.p[1] + row * U * 16; r.p[2] = w.p[2] + row * nb * 16;
            r.p[3] = w.p[3] ? w.p[3] + row * U * 8 : nullptr; r.p[4] = nullptr;
            break;
        case T_Q6_K:
            r.p[0] = w.p[0] + row * U * 16; r.p[1] = w.p[1] + row * U * 16; r.p[2] = w.p[2] + row * U * 16;
            r.p[3] = w.p[3] + row * U * 4; r.p[4] = w.p[4] + row * nb * 2;
            break;
        default:  // Q8_0
            r.p[0] = w.p[0] + row * U8 * 16; r.p[1] = w.p[1] + row * U8 * 16; r.p[2] = w.p[2] + row * U8 * 2;
            r.p[3] = r.p[4] = nullptr;
            break;
    }
    return r;
}

// 6-bit scale/min of sub-blocks 2p and 2p+1 from the 12 packed bytes (w0,w1,w2)
__device__ __forceinline__ void q4k_unit_scales(const uint4 &mt, int p, int &sc_lo, int &sc_hi, int &m_lo, int &m_hi) {
    const int sh = (p & 1) * 16;
    const uint32_t t0 = mt.y >> sh, t1 = mt.z >> sh, t2 = mt.w >> sh;
    if (p < 2) {
        sc_lo = t0 & 63; sc_hi = (t0 >> 8) & 63;
        m_lo = t1 & 63;  m_hi = (t1 >> 8) & 63;
    } else {
        sc_lo = (t2 & 0xF) | ((t0 >> 2) & 0x30);
        sc_hi = ((t2 >> 8) & 0xF) | ((t0 >> 10) & 0x30);
        m_lo = ((t2 >> 4) & 0xF) | ((t1 >> 2) & 0x30);
        m_hi = ((t2 >> 12) & 0xF) | ((t1 >> 10) & 0x30);
    }
}

template <bool Q5>
__device__ __forceinline__ void dot_q4k(const RowPtrs (&rp)[GEMV_R], int k, const ActSmem &s, double (&acc)[GEMV_R]) {
    const int lane = threadIdx.x & 31;
    const int U = k / 64;
    for (int u0 = 0; u0 < U; u0 += 64) {
        uint4 qa[GEMV_R][2], qb[GEMV_R][2], mt[GEMV_R][2];
        uint2 qh[GEMV_R][2];
#pragma unroll
        for (int it = 0; it < 2; it++) {
            const int u = u0 + it * 32 + lane;
            if (u < U) {
#pragma unroll
                for (int r = 0; r < GEMV_R; r++) {
                    qa[r][it] = ldg_stream(rp[r].p[0] + (size_t)u * 16);
                    qb[r][it] = ldg_stream(rp[r].p[1] + (size_t)u * 16);
                    mt[r][it] = ldg_stream(rp[r].p[2] + (size_t)(u >> 2) * 16);
                    if (Q5) qh[r][it] = *reinterpret_cast<const uint2 *>(rp[r].p[3] + (size_t)u * 8);
                }
            }
        }
#pragma unroll
        for (int it = 0; it < 2; it++) {
            const int u = u0 + it * 32 + lane;
            if (u < U) {
                const uint4 a0 = s.act[u], a1 = s.act[U + u], a2 = s.act[2 * U + u], a3 = s.act[3 * U + u];
                const int2 bs = *reinterpret_cast<const int2 *>(s.bs32 + 2 * u);
                const float d8 = s.d8[u >> 2];
                const uint32_t A0[4] = {a0.x, a0.y, a0.z, a0.w}, A1[4] = {a1.x, a1.y, a1.z, a1.w};
                const uint32_t A2[4] = {a2.x, a2.y, a2.z, a2.w}, A3[4] = {a3.x, a3.y, a3.z, a3.w};
#pragma unroll
                for (int r = 0; r < GEMV_R; r++) {
                    const uint32_t QA[4] = {qa[r][it].x, qa[r][it].y, qa[r][it].z, qa[r][it].w};
                    const uint32_t QB[4] = {qb[r][it].x, qb[r][it].y, qb[r][it].z, qb[r][it].w};
                    int dlo = 0, dhi = 0;
#pragma unroll
                    for (int w = 0; w < 4; w++) {
                        uint32_t la = QA[w] & 0x0F0F0F0Fu, ha = (QA[w] >> 4) & 0x0F0F0F0Fu;
                        uint32_t lb = QB[w] & 0x0F0F0F0Fu, hb = (QB[w] >> 4) & 0x0F0F0F0Fu;
                        if (Q5) {
                            // 5th bits: bit (4w+i) of qh.x -> byte i of word w (low sub-block, QSa: l = 4w+i;
                            // QSb: l = 16+4w+i); same for qh.y (high sub-block)
                            const uint32_t xa = (qh[r][it].x >> (4 * w)) & 0xFu, xb = (qh[r][it].x >> (16 + 4 * w)) & 0xFu;
                            const uint32_t ya = (qh[r][it].y >> (4 * w)) & 0xFu, yb = (qh[r][it].y >> (16 + 4 * w)) & 0xFu;
                            la |= ((xa * 0x00204081u) & 0x01010101u) << 4;
                            lb |= ((xb * 0x00204081u) & 0x01010101u) << 4;
                            ha |= ((ya * 0x00204081u) & 0x01010101u) << 4;
                            hb |= ((yb * 0x00204081u) & 0x01010101u) << 4;
                        }
                        dlo = __dp4a((int)la, (int)A0[w], dlo);
                        dlo = __dp4a((int)lb, (int)A1[w], dlo);
                        dhi = __dp4a((int)ha, (int)A2[w], dhi);
                        dhi = __dp4a((int)hb, (int)A3[w], dhi);
                    }
                    int sc_lo, sc_hi, m_lo, m_hi;
                    q4k_unit_scales(mt[r][it], u & 3, sc_lo, sc_hi, m_lo, m_hi);
                    const float d = h2f_bits((uint16_t)(mt[r][it].x & 0xffff));
                    const float dmin = h2f_bits((uint16_t)(mt[r][it].x >> 16));
                    const int isum = sc_lo * dlo + sc_hi * dhi;
                    const int msum = m_lo * bs.x + m_hi * bs.y;
                    // one f32 term per unit, summed in double (order-free; oracle: vec_dot_q4_K_q8_K)
                    const float term = __fmaf_rn(__fmul_rn(d, d8), (float)isum, -__fmul_rn(__fmul_rn(dmin, d8), (float)msum));
                    acc[r] += (double)term;
                }
            }
        }
    }
}

__device__ __forceinline__ void dot_q6k(const RowPtrs (&rp)[GEMV_R], int k, const ActSmem &s, double (&acc)[GEMV_R]) {
    const int lane = threadIdx.x & 31;
    const int U = k / 64;
    for (int u0 = 0; u0 < U; u0 += 64) {
        uint4 la[GEMV_R][2], lb[GEMV_R][2], hh[GEMV_R][2];
        uint32_t sc[GEMV_R][2];
        uint16_t dd[GEMV_R][2];
#pragma unroll
        for (int it = 0; it < 2; it++) {
            const int u = u0 + it * 32 + lane;
            if (u < U) {
#pragma unroll
                for (int r = 0; r < GEMV_R; r++) {
                    la[r][it] = ldg_stream(rp[r].p[0] + (size_t)u * 16);
                    lb[r][it] = ldg_stream(rp[r].p[1] + (size_t)u * 16);
                    hh[r][it] = ldg_stream(rp[r].p[2] + (size_t)u * 16);
                    sc[r][it] = ldg_stream_u32(rp[r].p[3] + (size_t)u * 4);
                    dd[r][it] = ldg_stream_u16(rp[r].p[4] + (size_t)(u >> 2) * 2);
                }
            }
        }
#pragma unroll
        for (int it = 0; it < 2; it++) {
            const int u = u0 + it * 32 + lane;
            if (u < U) {
                const uint4 a0 = s.act[u], a1 = s.act[U + u], a2 = s.act[2 * U + u], a3 = s.act[3 * U + u];
                const uint2 bsp = *reinterpret_cast<const uint2 *>(s.bs16 + 4 * u);
                const int b0 = (int)(int16_t)(bsp.x & 0xffff), b1 = (int)(int16_t)(bsp.x >> 16);
                const int b2 = (int)(int16_t)(bsp.y & 0xffff), b3 = (int)(int16_t)(bsp.y >> 16);
                const float d8 = s.d8[u >> 2];
                const uint32_t A0[4] = {a0.x, a0.y, a0.z, a0.w}, A1[4] = {a1.x, a1.y, a1.z, a1.w};
                const uint32_t A2[4] = {a2.x, a2.y, a2.z, a2.w}, A3[4] = {a3.x, a3.y, a3.z, a3.w};
#pragma unroll
                for (int r = 0; r < GEMV_R; r++) {
                    const uint32_t LA[4] = {la[r][it].x, la[r][it].y, la[r][it].z, la[r][it].w};
                    const uint32_t LB[4] = {lb[r][it].x, lb[r][it].y, lb[r][it].z, lb[r][it].w};
                    const uint32_t HH[4] = {hh[r][it].x, hh[r][it].y, hh[r][it].z, hh[r][it].w};
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
                    const uint32_t scw = sc[r][it];
                    const int s0 = (int)(int8_t)(scw & 0xff), s1 = (int)(int8_t)((scw >> 8) & 0xff);
                    const int s2 = (int)(int8_t)((scw >> 16) & 0xff), s3 = (int)(int8_t)(scw >> 24);
                    const int isum = s0 * (d0 - 32 * b0) + s1 * (d1 - 32 * b1) + s2 * (d2 - 32 * b2) + s3 * (d3 - 32 * b3);
                    acc[r] += (double)__fmul_rn(__fmul_rn(h2f_bits(dd[r][it]), d8), (float)isum);
                }
            }
        }
    }
}

__device__ __forceinline__ void dot_q80(const RowPtrs (&rp)[GEMV_R], int k, const ActSmem &s, double (&acc)[GEMV_R]) {
    const int lane = threadIdx.x & 31;
    const int U8 = k / 32;
    for (int u0 = 0; u0 < U8; u0 += 128) {
        uint4 qa[GEMV_R][4], qb[GEMV_R][4];
        uint16_t dd[GEMV_R][4];
#pragma unroll
        for (int it = 0; it < 4; it++) {
            const int u = u0 + it * 32 + lane;
            if (u < U8) {
#pragma unroll
                for (int r = 0; r < GEMV_R; r++) {
                    qa[r][it] = ldg_stream(rp[r].p[0] + (size_t)u * 16);
                    qb[r][it] = ldg_stream(rp[r].p[1] + (size_t)u * 16);
                    dd[r][it] = ldg_stream_u16(rp[r].p[2] + (size_t)u * 2);
                }
            }
        }
#pragma unroll
        for (int it = 0; it < 4; it++) {
            const int u = u0 + it * 32 + lane;
            if (u < U8) {
                const uint4 a0 = s.act[u], a1 = s.act[U8 + u];
                const float d8 = s.d8[u];
#pragma unroll
                for (int r = 0; r < GEMV_R; r++) {
                    int d = 0;
                    d = __dp4a((int)qa[r][it].x, (int)a0.x, d); d = __dp4a((int)qa[r][it].y, (int)a0.y, d);
                    d = __dp4a((int)qa[r][it].z, (int)a0.z, d); d = __dp4a((int)qa[r][it].w, (int)a0.w, d);
                    d = __dp4a((int)qb[r][it].x, (int)a1.x, d); d = __dp4a((int)qb[r][it].y, (int)a1.y, d);
                    d = __dp4a((int)qb[r][it].z, (int)a1.z, d); d = __dp4a((int)qb[r][it].w, (int)a1.w, d);
                    acc[r] += (double)__fmul_rn((float)d, __fmul_rn(h2f_bits(dd[r][it]), d8));
                }
            }
        }
    }
}

// out-of-line copies for launches that mix weight types (keeps each path's register allocation apart)
__device__ __noinline__ void dot_q4k_nl(const RowPtrs (&rp)[GEMV_R], int k, const ActSmem &s, double (&acc)[GEMV_R]) { dot_q4k<false>(rp, k, s, acc); }
__device__ __noinline__ void dot_q5k_nl(const RowPtrs (&rp)[GEMV_R], int k, const ActSmem &s, double (&acc)[GEMV_R]) { dot_q4k<true>(rp, k, s, acc); }
__device__ __noinline__ void dot_q6k_nl(const RowPtrs (&rp)[GEMV_R], int k, const ActSmem &s, double (&acc)[GEMV_R]) { dot_q6k(rp, k, s, acc); }
__device__ __noinline__ void dot_q80_nl(const RowPtrs (&rp)[GEMV_R], int k, const ActSmem &s, double (&acc)[GEMV_R]) { dot_q80(rp, k, s, acc); }

template <int TYPE>
__device__ __forceinline__ void dot_rows(int type, const RowPtrs (&rp)[GEMV_R], int k, const ActSmem &s,
                                         double (&acc)[GEMV_R]) {
    if (TYPE == T_Q4_K) { dot_q4k<false>(rp, k, s, acc); return; }
    if (TYPE == T_Q5_K) { dot_q4k<true>(rp, k, s, acc); return; }
    if (TYPE == T_Q6_K) { dot_q6k(rp, k, s, acc); return; }
    if (TYPE == T_Q8_0) { dot_q80(rp, k, s, acc); return; }
    switch (type) {
        case T_Q4_K: dot_q4k_nl(rp, k, s, acc); break;
        case T_Q5_K: dot_q5k_nl(rp, k, s, acc); break;
        case T_Q6_K: dot_q6k_nl(rp, k, s, acc); break;
        default: dot_q80_nl(rp, k, s, acc); break;
    }
}

// order-preserving float -> u32 so that packed (value, ~index) keys can be atomicMax-ed
__device__ __forceinline__ unsigned long long argmax_key(float v, int idx) {
    uint32_t b = __float_as_uint(v);
    b = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
    return ((unsigned long long)b << 32) | (uint32_t)(0x7fffffff - idx);
}

// TYPE = the ggml type of every segment, or -1 for a launch that mixes types (runtime switch)
template <int TYPE>
__global__ void __launch_bounds__(GEMV_THREADS, 2) gemv_kernel(const GemvParams p) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ double red[GEMV_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // which segment does this CTA serve?
    int seg = 0;
    if (p.epi != EPI_SWIGLU) {
        while (seg + 1 < p.n_seg && (int)blockIdx.x >= p.cta_begin[seg + 1]) seg++;
    }
    const WPlanes &W = p.seg[seg];
    const int cta_local = (int)blockIdx.x - p.cta_begin[seg];
    const int cta_count = p.cta_begin[seg + 1] - p.cta_begin[seg];
    const ActSmem s = carve(smem, p.k);
    const int wtype = TYPE >= 0 ? TYPE : W.type;
    if (wtype == T_Q6_K) prep_activations<PERM_Q6>(p.x, p.norm_w, p.eps, p.k, s, red);
    else if (wtype == T_Q8_0) prep_activations<PERM_Q80>(p.x, p.norm_w, p.eps, p.k, s, red);
    else prep_activations<PERM_Q4>(p.x, p.norm_w, p.eps, p.k, s, red);

    const int N = W.n_rows;
    if (p.epi == EPI_SWIGLU) {
        // one warp = one output row: gate row from seg[0] and up row from seg[1]
        const WPlanes &W1 = p.seg[1];
        for (int row = (cta_local * GEMV_WARPS + warp); row < N; row += cta_count * GEMV_WARPS) {
            RowPtrs rp[GEMV_R] = {row_ptrs(W, row), row_ptrs(W1, row)};
            double acc[GEMV_R] = {0.0, 0.0};
            dot_rows<TYPE>(wtype, rp, p.k, s, acc);
            const float g = (float)warp_sum_f64(acc[0]), u = (float)warp_sum_f64(acc[1]);
            if (lane == 0) p.out[0][row] = __fmul_rn(__fdiv_rn(g, __fadd_rn(1.0f, expf_det(-g))), u);
        }
        return;
    }
    float *out = p.out[seg];
    unsigned long long best = 0ull;
    for (int row0 = (cta_local * GEMV_WARPS + warp) * GEMV_R; row0 < N; row0 += cta_count * GEMV_WARPS * GEMV_R) {
        RowPtrs rp[GEMV_R];
#pragma unroll
        for (int r = 0; r < GEMV_R; r++) rp[r] = row_ptrs(W, min(row0 + r, N - 1));
        double acc[GEMV_R];
#pragma unroll
        for (int r = 0; r < GEMV_R; r++) acc[r] = 0.0;
        dot_rows<TYPE>(wtype, rp, p.k, s, acc);
#pragma unroll
        for (int r = 0; r < GEMV_R; r++) {
            const float v = (float)warp_sum_f64(acc[r]);
            const int row = row0 + r;
            if (lane == 0 && row < N) {
                if (p.epi == EPI_ADD) out[row] = __fadd_rn(out[row], v);
                else out[row] = v;
                if (p.argmax_key) {
                    const unsigned long long key = argmax_key(v, row);
                    best = key > best ? key : best;
                }
            }
        }
    }
    if (p.argmax_key && lane == 0 && best) atomicMax(p.argmax_key, best);
}

static size_t seg_row_bytes(const WPlanes &w) { return tensor_bytes(w.type, 1, w.k); }

int gemv_plan_grid(GemvParams &p, int n_sms) {
    const int target = n_sms * 2;  // two resident CTAs per SM (launch bounds)
    if (p.epi == EPI_SWIGLU) {
        const int groups = (p.seg[0].n_rows + GEMV_WARPS - 1) / GEMV_WARPS;
        const int g = groups < target ? groups : target;
        p.cta_begin[0] = 0;
        p.cta_begin[1] = g;
        p.cta_begin[2] = p.cta_begin[3] = g;
        return g;
    }
    double total = 0;
    int groups[3] = {0, 0, 0};
    int total_groups = 0;
    for (int i = 0; i < p.n_seg; i++) {
        total += (double)seg_row_bytes(p.seg[i]) * p.seg[i].n_rows;
        groups[i] = (p.seg[i].n_rows + GEMV_WARPS * GEMV_R - 1) / (GEMV_WARPS * GEMV_R);
        total_groups += groups[i];
    }
    const int grid_target = total_groups < target ? total_groups : target;
    int begin = 0;
    for (int i = 0; i < p.n_seg; i++) {
        p.cta_begin[i] = begin;
        double share = (double)seg_row_bytes(p.seg[i]) * p.seg[i].n_rows / total;
        int c = (int)(share * grid_target + 0.5);
        if (c < 1) c = 1;
        if (c > groups[i]) c = groups[i];
        begin += c;
    }
    for (int i = p.n_seg; i < 4; i++) p.cta_begin[i] = begin;
    return begin;
}

void gemv_init() {
    static bool done = false;
    if (done) return;
    CUDA_CHECK(cudaFuncSetAttribute(gemv_kernel<T_Q4_K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CUDA_CHECK(cudaFuncSetAttribute(gemv_kernel<T_Q5_K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CUDA_CHECK(cudaFuncSetAttribute(gemv_kernel<T_Q6_K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CUDA_CHECK(cudaFuncSetAttribute(gemv_kernel<T_Q8_0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CUDA_CHECK(cudaFuncSetAttribute(gemv_kernel<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    done = true;
}

void launch_gemv(const GemvParams &p, int grid, cudaStream_t st) {
    int t = p.seg[0].type;
    for (int i = 1; i < p.n_seg; i++) if (p.seg[i].type != t) t = -1;
    const size_t sm = gemv_smem_bytes(p.k);
    switch (t) {
        case T_Q4_K: gemv_kernel<T_Q4_K><<<grid, GEMV_THREADS, sm, st>>>(p); break;
        case T_Q5_K: gemv_kernel<T_Q5_K><<<grid, GEMV_THREADS, sm, st>>>(p); break;
        case T_Q6_K: gemv_kernel<T_Q6_K><<<grid, GEMV_THREADS, sm, st>>>(p); break;
        case T_Q8_0: gemv_kernel<T_Q8_0><<<grid, GEMV_THREADS, sm, st>>>(p); break;
        default: gemv_kernel<-1><<<grid, GEMV_THREADS, sm, st>>>(p); break;
    }
}

}  // namespace aios
