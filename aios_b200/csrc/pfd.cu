// pfd.cu — the persistent fused decode kernel: ONE cooperative launch generates n tokens.
//
// Why: at batch 1 a Mistral-7B token is 177 dependent kernels of ~4 us of HBM time each; launched one
// by one, every kernel pays a drain + launch + activation prologue + one exposed memory latency, and
// HBM idles for more than half of the step (profiles/r01_v1_summary.md).  Here one CTA per SM stays
// resident for the whole decode loop and walks a host-built table of phases
//   [QKV gemv] [attention] [O gemv + residual] [gate/up gemv + SwiGLU] [down gemv + residual] x L, [lm_head]
// separated by grid barriers (a global counter; the launch is cooperative, so all CTAs are resident).
// Weights never depend on activations, so every warp keeps the NEXT item's weights in flight in a
// second register image while it computes the current one — across phase boundaries and grid barriers
// too.  HBM therefore keeps streaming during the barrier and the activation prologue of the next phase.
//
// Arithmetic is the shared gemv_core.cuh code: results are bit-identical to the per-op kernels and
// to the oracle (order-free double sums).  Greedy sampling is fused (arg-max key updated by the lm_head
// epilogue; lowest index wins ties); other samplers use the per-op path.
#define GEMV_CORE_WARPS 8
#include "gemv_core.cuh"
#include "dequant.cuh"
#include "pfd.h"

#include <stdio.h>
#include <stdlib.h>

namespace aios {

constexpr int PFD_WARPS = GEMV_CORE_WARPS;
constexpr int PFD_THREADS = PFD_WARPS * 32;
constexpr int PFD_ATT_BATCH = 4;
constexpr int PFD_MAXREP = 8;

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// ---------------------------------------------------------------- barriers
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_add(unsigned *p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// every participating CTA arrives once on a monotonically increasing counter and waits for `target`
__device__ __forceinline__ void counter_barrier(unsigned *counter, unsigned target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        red_release_add(counter, 1u);  // release: this CTA's writes (ordered by the bar.sync above) are visible first
        // poll with a plain volatile load: an acquire load would invalidate this SM's L1 on every iteration
        while (*reinterpret_cast<volatile unsigned *>(counter) < target) { }
        __threadfence();               // one acquire-side fence once the barrier has opened
    }
    __syncthreads();
}

// ---------------------------------------------------------------- item cursor
struct Item {
    int tok, ph;      // token index within the launch, phase index (always a GEMV phase)
    int unit, chunk;  // work unit (row pair) inside the phase; K-chunk (pass) inside the unit
    int ue, nchunks;  // end of this CTA's unit range in the phase; passes per unit
    bool valid;
};

__device__ __forceinline__ int unit_elems(int type) { return type == T_Q8_0 ? 32 : 64; }

// this CTA's contiguous share of a phase's work units (32-bit: units_total * gridDim.x < 2^31)
__device__ __forceinline__ void cta_range(const PfdPhase &P, int &ub, int &ue) {
    const unsigned tot = (unsigned)P.units_total, g = gridDim.x, b = blockIdx.x;
    ub = (int)(tot * b / g);
    ue = (int)(tot * (b + 1) / g);
}

// passes per work unit: a pass covers 2 lane-iterations x 32 lanes = 64 units of each row.  All segments of
// a phase share k; the host refuses to mix Q8_0 (32-weight units) with K-quants inside one phase.
__device__ __forceinline__ int n_chunks(const PfdPhase &P) {
    const int nu = P.k / unit_elems(P.seg[0].type);
    return (nu + 63) / 64;
}

// first item of this warp at or after phase ph of token tok
__device__ __forceinline__ Item first_item_from(const PfdParams &p, int tok, int ph) {
    const int warp = threadIdx.x >> 5;
    Item it;
    it.valid = false; it.tok = tok; it.ph = ph; it.unit = 0; it.chunk = 0;
    while (tok < p.n_tokens) {
        for (; ph < p.n_phases; ph++) {
            const PfdPhase &P = p.phases[ph];
            if (P.kind != PH_GEMV) continue;
            int ub, ue;
            cta_range(P, ub, ue);
            if (ub + warp < ue) {
                it.valid = true; it.tok = tok; it.ph = ph; it.unit = ub + warp; it.chunk = 0;
                it.ue = ue; it.nchunks = n_chunks(P);
                return it;
            }
        }
        tok++; ph = 0;
    }
    it.tok = tok;
    return it;
}

__device__ __forceinline__ Item next_item(const PfdParams &p, const Item &c) {
    Item n = c;
    if (c.chunk + 1 < c.nchunks) { n.chunk = c.chunk + 1; return n; }
    if (c.unit + PFD_WARPS < c.ue) { n.unit = c.unit + PFD_WARPS; n.chunk = 0; return n; }
    return first_item_from(p, c.tok, c.ph + 1);
}

// rows of a work unit -> plane pointers
template <int TYPE>
__device__ __forceinline__ void rows_typed(const PfdPhase &P, int unit, int seg, int row0, RowPtrs (&rp)[GEMV_R]) {
    if (P.epi == EPI_SWIGLU) {
        rp[0] = row_ptrs<TYPE>(P.seg[0], row0);
        rp[1] = row_ptrs<TYPE>(P.seg[1], row0);
    } else {
        const WPlanes &W = P.seg[seg];
        rp[0] = row_ptrs<TYPE>(W, row0);
        rp[1] = row_ptrs<TYPE>(W, min(row0 + 1, W.n_rows - 1));
    }
}

__device__ __forceinline__ void unit_to_rows(const PfdPhase &P, int unit, int &seg, int &row0) {
    if (P.epi == EPI_SWIGLU) { seg = 0; row0 = unit; return; }
    seg = 0;
    while (seg + 1 < P.n_seg && unit >= P.unit_begin[seg + 1]) seg++;
    row0 = (unit - P.unit_begin[seg]) * 2;
}

// one register image for every weight type (fields a type does not use stay dead)
struct AnyRegs {
    uint4 a[GEMV_R][2], b[GEMV_R][2], c[GEMV_R][2];
    uint32_t e[GEMV_R][2], f[GEMV_R][2];
};

template <int TYPE>
__device__ __forceinline__ void pfd_load(AnyRegs &g, const RowPtrs (&rp)[GEMV_R], int u0, int nu) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int it = 0; it < 2; it++) {
        const int u = u0 + it * 32 + lane;
        if (u < nu) {
#pragma unroll
            for (int r = 0; r < GEMV_R; r++) {
                g.a[r][it] = ldg_stream(rp[r].p[0] + (size_t)u * 16);
                g.b[r][it] = ldg_stream(rp[r].p[1] + (size_t)u * 16);
                if (TYPE == T_Q4_K || TYPE == T_Q5_K) {
                    g.c[r][it] = ldg_stream(rp[r].p[2] + (size_t)(u >> 2) * 16);
                    if (TYPE == T_Q5_K) {
                        const uint2 h = *reinterpret_cast<const uint2 *>(rp[r].p[3] + (size_t)u * 8);
                        g.e[r][it] = h.x; g.f[r][it] = h.y;
                    }
                } else if (TYPE == T_Q6_K) {
                    g.c[r][it] = ldg_stream(rp[r].p[2] + (size_t)u * 16);
                    g.e[r][it] = ldg_stream_u32(rp[r].p[3] + (size_t)u * 4);
                    g.f[r][it] = ldg_stream_u16(rp[r].p[4] + (size_t)(u >> 2) * 2);
                } else {
                    g.e[r][it] = ldg_stream_u16(rp[r].p[2] + (size_t)u * 2);
                }
            }
        }
    }
}

__device__ __forceinline__ void load_item(const PfdParams &p, const Item &it, AnyRegs &g) {
    const PfdPhase &P = p.phases[it.ph];
    int seg, row0;
    unit_to_rows(P, it.unit, seg, row0);
    const int type = P.seg[seg].type;
    const int nu = P.k / unit_elems(type);
    const int u0 = it.chunk * 64;
    RowPtrs rp[GEMV_R];
    switch (type) {
        case T_Q4_K: rows_typed<T_Q4_K>(P, it.unit, seg, row0, rp); pfd_load<T_Q4_K>(g, rp, u0, nu); break;
        case T_Q5_K: rows_typed<T_Q5_K>(P, it.unit, seg, row0, rp); pfd_load<T_Q5_K>(g, rp, u0, nu); break;
        case T_Q6_K: rows_typed<T_Q6_K>(P, it.unit, seg, row0, rp); pfd_load<T_Q6_K>(g, rp, u0, nu); break;
        default: rows_typed<T_Q8_0>(P, it.unit, seg, row0, rp); pfd_load<T_Q8_0>(g, rp, u0, nu); break;
    }
}

template <int TYPE>
__device__ __forceinline__ void pfd_compute(const AnyRegs &g, int u0, int nu, int k, const ActSmem &s, double (&acc)[GEMV_R]) {
    if (TYPE == T_Q8_0) {
        const int lane = threadIdx.x & 31;
        const int U8 = k / 32;
#pragma unroll
        for (int it = 0; it < 2; it++) {
            const int u = u0 + it * 32 + lane;
            if (u < nu) {
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
    } else {
        PassRegs<TYPE> t;  // same field names; ITS == 2 for the K-quants
#pragma unroll
        for (int r = 0; r < GEMV_R; r++)
#pragma unroll
            for (int it = 0; it < 2; it++) {
                t.a[r][it] = g.a[r][it]; t.b[r][it] = g.b[r][it]; t.c[r][it] = g.c[r][it];
                t.e[r][it] = g.e[r][it]; t.f[r][it] = g.f[r][it];
            }
        compute_pass<TYPE, false>(t, u0, nu, k, s, acc);
    }
}

struct PfdSmem {
    uint8_t *act_a;    // activation image in Q4_K/Q5_K lane order, or the Q8_0 image
    uint8_t *act_b;    // activation image in Q6_K lane order
    float *scores;     // attention: [rep][per_cap]
    double *att_part;  // attention: [PFD_WARPS][rep*hd]
    double *red;       // [PFD_WARPS]
};

__device__ __forceinline__ void compute_item(const PfdPhase &P, const Item &it, const AnyRegs &g, const PfdSmem &sm,
                                             double (&acc)[GEMV_R]) {
    int seg, row0;
    unit_to_rows(P, it.unit, seg, row0);
    const int type = P.seg[seg].type;
    const int nu = P.k / unit_elems(type);
    const int u0 = it.chunk * 64;
    switch (type) {
        case T_Q4_K: pfd_compute<T_Q4_K>(g, u0, nu, P.k, carve(sm.act_a, P.k), acc); break;
        case T_Q5_K: pfd_compute<T_Q5_K>(g, u0, nu, P.k, carve(sm.act_a, P.k), acc); break;
        case T_Q6_K: pfd_compute<T_Q6_K>(g, u0, nu, P.k, carve(sm.act_b, P.k), acc); break;
        default: pfd_compute<T_Q8_0>(g, u0, nu, P.k, carve(sm.act_a, P.k), acc); break;
    }
}

// reduction + fused epilogue of a finished work unit (same code path as gemv_kernel's)
__device__ __forceinline__ void finish_unit(const PfdParams &p, const PfdPhase &P, const Item &it, double (&acc)[GEMV_R],
                                            unsigned long long &best) {
    const int lane = threadIdx.x & 31;
    int seg, row0;
    unit_to_rows(P, it.unit, seg, row0);
    const double tot = warp_sum2_f64(acc[0], acc[1]);  // lanes 0-15: row 0, lanes 16-31: row 1
    float *out = P.out[seg];
    if (P.epi == EPI_SWIGLU) {
        const float gte = (float)__shfl_sync(0xffffffffu, tot, 0), up = (float)__shfl_sync(0xffffffffu, tot, 16);
        if (lane == 0) out[row0] = __fmul_rn(__fdiv_rn(gte, __fadd_rn(1.0f, expf_det(-gte))), up);
    } else if ((lane & 15) == 0) {
        const int row = row0 + (lane >> 4);
        if (row < P.seg[seg].n_rows) {
            const float v = (float)tot;
            if (P.epi == EPI_ADD) out[row] = __fadd_rn(__ldcg(out + row), v);
            else out[row] = v;
            if (P.argmax) {
                const unsigned long long key = argmax_key(v, row);
                best = key > best ? key : best;
            }
        }
    }
    acc[0] = 0.0;
    acc[1] = 0.0;
}

// ---------------------------------------------------------------- activation prologue of a GEMV phase
__device__ __forceinline__ void src_load8(const PfdParams &p, const PfdPhase &P, int token, int idx, float (&v)[8]) {
    if (P.x64) {
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = (float)__ldcg(P.x64 + idx + i);
    } else if (P.embed) {
        const uint8_t *row = p.tok_embd + (int64_t)token * p.tok_embd_row_bytes;
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = dequant_elem(p.tok_embd_type, row, idx + i);
    } else {
        const float4 a = __ldcg(reinterpret_cast<const float4 *>(P.x + idx));
        const float4 b = __ldcg(reinterpret_cast<const float4 *>(P.x + idx) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    }
}

__device__ void pfd_prologue(const PfdParams &p, const PfdPhase &P, const PfdSmem &sm, int token) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int k = P.k, nslab = k / 256;
    const ActSmem sa = carve(sm.act_a, k), sb = carve(sm.act_b, k);
    auto emit = [&](float (&v)[8], int b, float scale) {
        if (P.embed && blockIdx.x == 0) {  // one CTA publishes the residual stream's starting value
            float4 *dst = reinterpret_cast<float4 *>(p.x + b * 256 + lane * 8);
            dst[0] = make_float4(v[0], v[1], v[2], v[3]);
            dst[1] = make_float4(v[4], v[5], v[6], v[7]);
        }
        if (P.norm_w) norm8(v, P.norm_w + b * 256 + lane * 8, scale);
        if (P.perm_mask & 1) quant_slab<PERM_Q4>(v, b, k, sa);
        if (P.perm_mask & 2) quant_slab<PERM_Q6>(v, b, k, sb);
        if (P.perm_mask & 4) quant_slab<PERM_Q80>(v, b, k, sa);
    };
    if (nslab <= 2 * PFD_WARPS) {
        // k <= 512*warps: each lane keeps its (up to) 16 values in registers — one trip to L2 per phase
        const int b0 = warp, b1 = warp + PFD_WARPS;
        float v0[8], v1[8];
#pragma unroll
        for (int i = 0; i < 8; i++) v0[i] = v1[i] = 0.f;
        if (b0 < nslab) src_load8(p, P, token, b0 * 256 + lane * 8, v0);
        if (b1 < nslab) src_load8(p, P, token, b1 * 256 + lane * 8, v1);
        float scale = 1.f;
        if (P.norm_w) {
            const double ss = block_sum(sumsq8(v0) + sumsq8(v1), sm.red);
            const float mean = (float)(ss / (double)k);
            scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, P.eps)));
        }
        if (b0 < nslab) emit(v0, b0, scale);
        if (b1 < nslab) emit(v1, b1, scale);
    } else {
        float scale = 1.f;
        if (P.norm_w) {
            double ss = 0.0;
            for (int b = warp; b < nslab; b += PFD_WARPS) {
                float v[8];
                src_load8(p, P, token, b * 256 + lane * 8, v);
                ss += sumsq8(v);
            }
            ss = block_sum(ss, sm.red);
            const float mean = (float)(ss / (double)k);
            scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, P.eps)));
        }
        // three slabs per trip to L2
        for (int b = warp; b < nslab; b += 3 * PFD_WARPS) {
            const int b1 = b + PFD_WARPS, b2 = b + 2 * PFD_WARPS;
            float v0[8], v1[8], v2[8];
            src_load8(p, P, token, b * 256 + lane * 8, v0);
            if (b1 < nslab) src_load8(p, P, token, b1 * 256 + lane * 8, v1);
            if (b2 < nslab) src_load8(p, P, token, b2 * 256 + lane * 8, v2);
            emit(v0, b, scale);
            if (b1 < nslab) emit(v1, b1, scale);
            if (b2 < nslab) emit(v2, b2, scale);
        }
    }
    __syncthreads();
}

// ---------------------------------------------------------------- attention phase
template <int EPL>
__device__ __forceinline__ void att_load_row(const uint16_t *p, float (&f)[EPL]) {
    if constexpr (EPL == 4) {
        const uint2 v = __ldcg(reinterpret_cast<const uint2 *>(p));
        f[0] = h2f_bits((uint16_t)(v.x & 0xffff)); f[1] = h2f_bits((uint16_t)(v.x >> 16));
        f[2] = h2f_bits((uint16_t)(v.y & 0xffff)); f[3] = h2f_bits((uint16_t)(v.y >> 16));
    } else {
        const uint32_t v = __ldcg(reinterpret_cast<const uint32_t *>(p));
        f[0] = h2f_bits((uint16_t)(v & 0xffff)); f[1] = h2f_bits((uint16_t)(v >> 16));
    }
}

// Same arithmetic as attention.cu (oracle-exact).  The cross-CTA exchanges (row max, row sum) go through
// global memory and per-kv-head counter barriers instead of a cluster; the partial outputs are added
// into global f64 accumulators (order-free), which must be zero on entry.
template <int EPL>
__device__ void pfd_attention(const PfdParams &p, const PfdPhase &P, const PfdSmem &sm, int pos, int epoch) {
    constexpr int HD = EPL * 32;
    const int grp = p.att_group;  // CTAs per kv-head
    if ((int)blockIdx.x >= p.n_kv * grp) return;  // (uniform per CTA)
    const int g = blockIdx.x / grp, rank = blockIdx.x % grp;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rep = p.n_head / p.n_kv;
    const int kvd = p.n_kv * HD;
    const int n_kv_pos = pos + 1;
    const int per = (n_kv_pos + grp - 1) / grp;
    const int j0 = min(n_kv_pos, rank * per), j1 = min(n_kv_pos, j0 + per);
    const int cnt = j1 - j0;
    const int per_cap = p.att_per_cap;
    float *sc = sm.scores;
    const float *cs = p.rope_cos + (size_t)pos * (p.n_rot / 2);
    const float *sn = p.rope_sin + (size_t)pos * (p.n_rot / 2);
    const float *qkv = p.qkv;

    __shared__ float sh_wmax[PFD_WARPS][PFD_MAXREP];
    __shared__ double sh_wsum[PFD_WARPS];
    __shared__ float sh_gmax[PFD_MAXREP], sh_ginv[PFD_MAXREP];

    if (pos >= j0 && pos < j1) {
        const float *kraw = qkv + (size_t)p.n_head * HD + (size_t)g * HD;
        const float *vraw = qkv + (size_t)p.n_head * HD + (size_t)kvd + (size_t)g * HD;
        uint16_t *kc = P.k_cache + (size_t)pos * kvd + (size_t)g * HD;
        uint16_t *vc = P.v_cache + (size_t)pos * kvd + (size_t)g * HD;
        for (int i = threadIdx.x; i < HD / 2; i += PFD_THREADS) {
            float x0 = __ldcg(kraw + 2 * i), x1 = __ldcg(kraw + 2 * i + 1);
            if (2 * i < p.n_rot) {
                const float c = cs[i], s = sn[i];
                const float r0 = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, s));
                const float r1 = __fadd_rn(__fmul_rn(x0, s), __fmul_rn(x1, c));
                x0 = r0; x1 = r1;
            }
            kc[2 * i] = f2h_bits(x0);
            kc[2 * i + 1] = f2h_bits(x1);
            vc[2 * i] = f2h_bits(__ldcg(vraw + 2 * i));
            vc[2 * i + 1] = f2h_bits(__ldcg(vraw + 2 * i + 1));
        }
        __syncthreads();
    }

    float q[PFD_MAXREP][EPL];
#pragma unroll
    for (int r = 0; r < PFD_MAXREP; r++) {
        if (r < rep) {
            const float *qraw = qkv + (size_t)(g * rep + r) * HD + lane * EPL;
#pragma unroll
            for (int e = 0; e < EPL; e += 2) {
                float x0 = __ldcg(qraw + e), x1 = __ldcg(qraw + e + 1);
                const int i = (lane * EPL + e) / 2;
                if (2 * i < p.n_rot) {
                    const float c = cs[i], s = sn[i];
                    const float r0 = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, s));
                    const float r1 = __fadd_rn(__fmul_rn(x0, s), __fmul_rn(x1, c));
                    x0 = r0; x1 = r1;
                }
                q[r][e] = __half2float(__float2half_rn(x0));
                q[r][e + 1] = __half2float(__float2half_rn(x1));
            }
        }
    }

    // ---- scores of my positions + local row max
    float wmax[PFD_MAXREP];
#pragma unroll
    for (int r = 0; r < PFD_MAXREP; r++) wmax[r] = -INFINITY;
    for (int jb = j0 + warp * PFD_ATT_BATCH; jb < j1; jb += PFD_WARPS * PFD_ATT_BATCH) {
        float kf[PFD_ATT_BATCH][EPL];
#pragma unroll
        for (int t = 0; t < PFD_ATT_BATCH; t++) {
            const int j = min(jb + t, j1 - 1);
            att_load_row<EPL>(P.k_cache + (size_t)j * kvd + (size_t)g * HD + lane * EPL, kf[t]);
        }
#pragma unroll
        for (int t = 0; t < PFD_ATT_BATCH; t++) {
            if (jb + t < j1) {
#pragma unroll
                for (int r = 0; r < PFD_MAXREP; r++) {
                    if (r < rep) {
                        double sd = 0.0;
#pragma unroll
                        for (int e = 0; e < EPL; e++) sd += (double)__fmul_rn(q[r][e], kf[t][e]);
                        const float sdot = __fmul_rn((float)warp_sum_f64(sd), p.att_scale);
                        if (lane == 0) sc[(size_t)r * per_cap + (jb + t - j0)] = sdot;
                        wmax[r] = fmaxf(wmax[r], sdot);
                    }
                }
            }
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int r = 0; r < PFD_MAXREP; r++) sh_wmax[warp][r] = wmax[r];
    }
    __syncthreads();
    if ((int)threadIdx.x < rep) {
        float mx = -INFINITY;
        for (int w = 0; w < PFD_WARPS; w++) mx = fmaxf(mx, sh_wmax[w][threadIdx.x]);
        __stcg(p.att_max + ((size_t)g * grp + rank) * PFD_MAXREP + threadIdx.x, mx);
    }
    counter_barrier(p.att_counters + g, (unsigned)((2 * epoch + 1) * grp));
    if ((int)threadIdx.x < rep) {
        float mx = -INFINITY;
        for (int c = 0; c < grp; c++) mx = fmaxf(mx, __ldcg(p.att_max + ((size_t)g * grp + c) * PFD_MAXREP + threadIdx.x));
        sh_gmax[threadIdx.x] = mx;
    }
    __syncthreads();
    {
        // e = expf(s - max), row sums in double; warps r, r+rep, ... serve head r
        double part = 0.0;
        const int r = warp % rep, sub = warp / rep;
        const int nsub = (PFD_WARPS - r + rep - 1) / rep;
        const float mx = sh_gmax[r];
        for (int j = sub * 32 + lane; j < cnt; j += nsub * 32) {
            const float e = expf_det(__fsub_rn(sc[(size_t)r * per_cap + j], mx));
            sc[(size_t)r * per_cap + j] = e;
            part += (double)e;
        }
        part = warp_sum_f64(part);
        if (lane == 0) sh_wsum[warp] = part;
    }
    __syncthreads();
    if ((int)threadIdx.x < rep) {
        double s = 0.0;
        for (int w = threadIdx.x; w < PFD_WARPS; w += rep) s += sh_wsum[w];
        __stcg(p.att_sum + ((size_t)g * grp + rank) * PFD_MAXREP + threadIdx.x, s);
    }
    counter_barrier(p.att_counters + g, (unsigned)((2 * epoch + 2) * grp));
    if ((int)threadIdx.x < rep) {
        double s = 0.0;
        for (int c = 0; c < grp; c++) s += __ldcg(p.att_sum + ((size_t)g * grp + c) * PFD_MAXREP + threadIdx.x);
        sh_ginv[threadIdx.x] = (float)(1.0 / s);
    }
    __syncthreads();

    // ---- o += f16(p) * V over my positions (exact products, f64 sums)
    double o[PFD_MAXREP][EPL];
#pragma unroll
    for (int r = 0; r < PFD_MAXREP; r++)
#pragma unroll
        for (int e = 0; e < EPL; e++) o[r][e] = 0.0;
    for (int jb = j0 + warp * PFD_ATT_BATCH; jb < j1; jb += PFD_WARPS * PFD_ATT_BATCH) {
        float vf[PFD_ATT_BATCH][EPL];
#pragma unroll
        for (int t = 0; t < PFD_ATT_BATCH; t++) {
            const int j = min(jb + t, j1 - 1);
            att_load_row<EPL>(P.v_cache + (size_t)j * kvd + (size_t)g * HD + lane * EPL, vf[t]);
        }
#pragma unroll
        for (int t = 0; t < PFD_ATT_BATCH; t++) {
            if (jb + t < j1) {
#pragma unroll
                for (int r = 0; r < PFD_MAXREP; r++) {
                    if (r < rep) {
                        const float pj = __half2float(__float2half_rn(
                            __fmul_rn(sc[(size_t)r * per_cap + (jb + t - j0)], sh_ginv[r])));
#pragma unroll
                        for (int e = 0; e < EPL; e++) o[r][e] += (double)__fmul_rn(pj, vf[t][e]);
                    }
                }
            }
        }
    }
    // CTA-level sum of the warps' partials, then one f64 atomic per output and CTA
    double *part_o = sm.att_part;
#pragma unroll
    for (int r = 0; r < PFD_MAXREP; r++) {
        if (r < rep) {
#pragma unroll
            for (int e = 0; e < EPL; e++) part_o[((size_t)warp * rep + r) * HD + lane * EPL + e] = o[r][e];
        }
    }
    __syncthreads();
    if (cnt > 0) {
        for (int idx = threadIdx.x; idx < rep * HD; idx += PFD_THREADS) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < PFD_WARPS; w++) s += part_o[(size_t)w * rep * HD + idx];
            atomicAdd(p.att_acc + (size_t)g * rep * HD + idx, s);
        }
    }
}

// ---------------------------------------------------------------- the kernel
__device__ __forceinline__ int key_to_index(unsigned long long k) { return 0x7fffffff - (int)(uint32_t)(k & 0xffffffffu); }

// 8 warps = 2 per SM sub-partition (16,384 registers each): up to 255 registers per thread, which holds two
// weight images (current + next item) without spilling.  12 warps would cap at 168 and spill.
__global__ void __launch_bounds__(PFD_THREADS, 1) pfd_kernel(const PfdParams p) {
    extern __shared__ __align__(16) uint8_t smem[];
    PfdSmem sm;
    {
        const size_t img = (act_image_bytes(p.k_max) + 15) / 16 * 16;
        const int rep = p.n_head / p.n_kv;
        sm.act_a = smem;
        sm.act_b = smem + img;
        sm.att_part = reinterpret_cast<double *>(smem + 2 * img);
        sm.red = sm.att_part + (size_t)PFD_WARPS * rep * p.hd;
        sm.scores = reinterpret_cast<float *>(sm.red + PFD_WARPS);
    }
    const int lane = threadIdx.x & 31;
    unsigned bar = 0;
    int att_count = 0;
    int pos = p.pos0;

    Item cur = first_item_from(p, 0, 0);
    AnyRegs A, B;
    bool in_b = false;
    if (cur.valid) load_item(p, cur, A);  // weights of my first item: in flight before anything else happens

    for (int tok = 0; tok < p.n_tokens; tok++) {
        // the token to feed = arg-max of the logits kept from the previous pass
        const unsigned long long key = *reinterpret_cast<volatile unsigned long long *>(p.argmax_key);
        const int token = key_to_index(key);
        pos += 1;
        if (blockIdx.x == 0 && threadIdx.x == 0) p.out_tokens[p.step0 + tok] = token;

        for (int ph = 0; ph < p.n_phases; ph++) {
            const PfdPhase &P = p.phases[ph];
            unsigned long long *tm = (p.timing && blockIdx.x == 0 && threadIdx.x == 0 && tok == p.n_tokens - 1)
                                         ? p.timing + (size_t)ph * 4 : nullptr;
            if (tm) tm[0] = globaltimer_ns();
            if (P.kind == PH_GEMV) {
                if (P.zero_att) {
                    const int n = p.n_head * p.hd;
                    for (int i = blockIdx.x * PFD_THREADS + threadIdx.x; i < n; i += gridDim.x * PFD_THREADS) p.att_acc[i] = 0.0;
                }
                pfd_prologue(p, P, sm, token);
                if (tm) tm[1] = globaltimer_ns();
                double acc[GEMV_R] = {0.0, 0.0};
                unsigned long long best = 0ull;
                while (cur.valid && cur.tok == tok && cur.ph == ph) {
                    const Item nxt = next_item(p, cur);
                    if (!in_b) {
                        if (nxt.valid) load_item(p, nxt, B);
                        compute_item(P, cur, A, sm, acc);
                    } else {
                        if (nxt.valid) load_item(p, nxt, A);
                        compute_item(P, cur, B, sm, acc);
                    }
                    if (cur.chunk + 1 == cur.nchunks) finish_unit(p, P, cur, acc, best);
                    cur = nxt;
                    in_b = !in_b;
                }
                if (P.argmax) {
                    const unsigned long long other = __shfl_sync(0xffffffffu, best, 16);
                    best = other > best ? other : best;
                    if (lane == 0 && best) atomicMax(p.argmax_key, best);
                }
            } else {
                const int epoch = att_count++;  // ordinal of this attention phase within the launch
                if (p.hd == 128) pfd_attention<4>(p, P, sm, pos, epoch);
                else pfd_attention<2>(p, P, sm, pos, epoch);
            }
            if (tm) tm[2] = globaltimer_ns();
            bar++;
            counter_barrier(p.bar_counter, bar * gridDim.x);
            if (tm) tm[3] = globaltimer_ns();
            // every CTA has read this token's key before arriving at the first barrier: clear it for the next lm_head
            if (ph == 0 && blockIdx.x == 0 && threadIdx.x == 0) *p.argmax_key = 0ull;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        p.state[0] = pos;                      // SeqState.pos: position of the last token fed
        p.state[1] = p.step0 + p.n_tokens;     // SeqState.step
    }
}

size_t pfd_smem_bytes(const PfdParams &p) {
    const size_t img = (act_image_bytes(p.k_max) + 15) / 16 * 16;
    const int rep = p.n_head / p.n_kv;
    return 2 * img + ((size_t)PFD_WARPS * rep * p.hd + PFD_WARPS) * sizeof(double) +
           (size_t)rep * p.att_per_cap * sizeof(float) + 64;
}

int pfd_max_grid(int n_sms, const PfdParams &p) {
    const size_t smem = pfd_smem_bytes(p);
    if (smem > 200 * 1024) return 0;
    CUDA_CHECK(cudaFuncSetAttribute(pfd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pfd_kernel, PFD_THREADS, smem));
    if (getenv("AIOS_DEBUG")) {
        cudaFuncAttributes fa{};
        cudaFuncGetAttributes(&fa, pfd_kernel);
        fprintf(stderr, "[aios] pfd_kernel: %d regs, %zu B static smem, %zu B dynamic smem, max %d threads/block, %d CTA/SM\n",
                fa.numRegs, fa.sharedSizeBytes, smem, fa.maxThreadsPerBlock, per_sm);
    }
    return per_sm >= 1 ? n_sms : 0;
}

void launch_pfd(const PfdParams &p, int grid, cudaStream_t st) {
    const size_t smem = pfd_smem_bytes(p);
    CUDA_CHECK(cudaFuncSetAttribute(pfd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(PFD_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;  // all CTAs co-resident: the counter barriers cannot deadlock
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CUDA_CHECK(cudaLaunchKernelEx(&cfg, pfd_kernel, p));
}

}  // namespace aios
