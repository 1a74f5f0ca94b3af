// pfd.cu — the persistent fused decode kernel: ONE cooperative launch generates n tokens.
//
// Why: at batch 1 a Mistral-7B token is 177 dependent kernels of ~4 us of HBM time each; launched one
// by one, every kernel pays a drain + launch + activation prologue + one exposed memory latency, and
// HBM idles for more than half of the step (profiles/r01_v1_summary.md).  Here one CTA per SM stays
// resident for the whole decode loop and walks a host-built table of phases
//   [q][k][v] [attention] [o + residual] [gate/up + SwiGLU] [down + residual] x L, [lm_head]
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

// ---------------------------------------------------------------- the weight loader of one warp
// Walks this warp's items (work unit x K-pass) in order across phases and tokens.  Inside a phase it only
// adds strides; the phase table is read once per phase.
struct AnyRegs {  // one register image for every weight type (fields a type does not use stay dead)
    static constexpr int ITS = 2;
    uint4 a[GEMV_R][2], b[GEMV_R][2], c[GEMV_R][2];
    uint32_t e[GEMV_R][2], f[GEMV_R][2];
};

struct Loader {
    int tok, ph;
    bool valid;
    int type, nu, nchunks, pair, n_rows;
    int unit, chunk, ue;
    RowPtrs rp[GEMV_R];  // plane pointers of the unit's two rows
};

struct Light {  // what the compute side needs to know about an item
    int tok, ph, unit, chunk, nchunks;
    bool valid;
};

__device__ __forceinline__ Light light(const Loader &L) {
    Light l;
    l.tok = L.tok; l.ph = L.ph; l.unit = L.unit; l.chunk = L.chunk; l.nchunks = L.nchunks; l.valid = L.valid;
    return l;
}

__device__ __forceinline__ int unit_elems(int type) { return type == T_Q8_0 ? 32 : 64; }

// bytes of one row inside plane i
__device__ __forceinline__ int plane_row_bytes(int type, int i, int nu) {
    switch (type) {
        case T_Q4_K: return i < 2 ? nu * 16 : (i == 2 ? nu * 4 : 0);
        case T_Q5_K: return i < 2 ? nu * 16 : (i == 2 ? nu * 4 : (i == 3 ? nu * 8 : 0));
        case T_Q6_K: return i < 3 ? nu * 16 : (i == 3 ? nu * 4 : nu / 2);
        default: return i < 2 ? nu * 16 : (i == 2 ? nu * 2 : 0);
    }
}

// this CTA's contiguous share of a phase's work units (32-bit: units_total * gridDim.x < 2^31)
__device__ __forceinline__ void cta_range(int units_total, int &ub, int &ue) {
    const unsigned tot = (unsigned)units_total, g = gridDim.x, b = blockIdx.x;
    ub = (int)(tot * b / g);
    ue = (int)(tot * (b + 1) / g);
}

__device__ __forceinline__ void loader_set_rows(Loader &L, const PfdPhase &P) {
    const int row0 = L.pair ? L.unit : 2 * L.unit;
#pragma unroll
    for (int i = 0; i < 5; i++) {
        const int rb = plane_row_bytes(L.type, i, L.nu);
        L.rp[0].p[i] = P.w0.p[i] ? P.w0.p[i] + (size_t)row0 * rb : nullptr;
        if (L.pair) L.rp[1].p[i] = P.w1.p[i] ? P.w1.p[i] + (size_t)row0 * rb : nullptr;
        else L.rp[1].p[i] = (row0 + 1 < L.n_rows) ? L.rp[0].p[i] + rb : L.rp[0].p[i];
    }
}

// position the loader on this warp's first item at or after (tok, ph)
__device__ __forceinline__ void loader_seek(const PfdParams &p, Loader &L, int tok, int ph) {
    const int warp = threadIdx.x >> 5;
    L.valid = false;
    while (tok < p.n_tokens) {
        for (; ph < p.n_phases; ph++) {
            const PfdPhase &P = p.phases[ph];
            if (P.kind != PH_GEMV) continue;
            int ub, ue;
            cta_range(P.units_total, ub, ue);
            if (ub + warp < ue) {
                L.valid = true; L.tok = tok; L.ph = ph;
                L.type = P.type; L.nu = P.k / unit_elems(P.type); L.nchunks = (L.nu + 63) / 64;
                L.pair = P.epi == EPI_SWIGLU; L.n_rows = P.w0.n_rows;
                L.unit = ub + warp; L.chunk = 0; L.ue = ue;
                loader_set_rows(L, P);
                return;
            }
        }
        tok++; ph = 0;
    }
    L.tok = tok; L.ph = 0;
}

__device__ __forceinline__ void loader_advance(const PfdParams &p, Loader &L) {
    if (L.chunk + 1 < L.nchunks) { L.chunk++; return; }
    if (L.unit + PFD_WARPS < L.ue) {
        L.unit += PFD_WARPS; L.chunk = 0;
        const int rows = L.pair ? PFD_WARPS : 2 * PFD_WARPS;
        const bool last_odd = !L.pair && (2 * L.unit + 1 >= L.n_rows);
#pragma unroll
        for (int i = 0; i < 5; i++) {
            const size_t step = (size_t)rows * plane_row_bytes(L.type, i, L.nu);
            if (L.rp[0].p[i]) {
                L.rp[0].p[i] += step;
                L.rp[1].p[i] = last_odd ? L.rp[0].p[i] : L.rp[1].p[i] + step;
            }
        }
        return;
    }
    loader_seek(p, L, L.tok, L.ph + 1);
}

__device__ __forceinline__ void issue_load(const Loader &L, AnyRegs &g) {
    const int u0 = L.chunk * 64;
    const uint64_t pol = l2_policy(false);
    switch (L.type) {
        case T_Q4_K: load_pass<T_Q4_K, false>(g, L.rp, u0, L.nu, pol); break;
        case T_Q5_K: load_pass<T_Q5_K, false>(g, L.rp, u0, L.nu, pol); break;
        case T_Q6_K: load_pass<T_Q6_K, false>(g, L.rp, u0, L.nu, pol); break;
        default: load_pass<T_Q8_0, false>(g, L.rp, u0, L.nu, pol); break;
    }
}

// Ask L2 to fetch this CTA's slice of the NEXT GEMV phase's weights (cp.async.bulk.prefetch.L2): HBM keeps
// streaming one phase ahead of the compute, through barriers and prologues, and the demand loads hit L2.
__device__ __forceinline__ void l2_prefetch_range(const uint8_t *base, size_t bytes) {
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(base) & ~(uintptr_t)15;
    const uintptr_t a1 = (reinterpret_cast<uintptr_t>(base) + bytes + 15) & ~(uintptr_t)15;
    constexpr size_t PIECE = 8192;
    for (uintptr_t a = a0 + (size_t)threadIdx.x * PIECE; a < a1; a += (size_t)PFD_THREADS * PIECE) {
        const unsigned sz = (unsigned)(a1 - a < PIECE ? a1 - a : PIECE);
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a), "r"(sz) : "memory");
    }
}

__device__ __forceinline__ void l2_prefetch_phase(const PfdPhase &P) {
    int ub, ue;
    cta_range(P.units_total, ub, ue);
    if (ub >= ue) return;
    const int pair = P.epi == EPI_SWIGLU;
    const int nu = P.k / unit_elems(P.type);
    const int r0 = pair ? ub : 2 * ub, r1 = pair ? ue : min(2 * ue, P.w0.n_rows);
#pragma unroll
    for (int i = 0; i < 5; i++) {
        const int rb = plane_row_bytes(P.type, i, nu);
        if (rb == 0 || !P.w0.p[i]) continue;
        l2_prefetch_range(P.w0.p[i] + (size_t)r0 * rb, (size_t)(r1 - r0) * rb);
        if (pair) l2_prefetch_range(P.w1.p[i] + (size_t)r0 * rb, (size_t)(r1 - r0) * rb);
    }
}

struct PfdSmem {
    uint8_t *act_a;    // activation image in Q4_K/Q5_K lane order, or the Q8_0 image
    uint8_t *act_b;    // activation image in Q6_K lane order
    float *scores;     // attention: [rep][per_cap]
    double *att_part;  // attention: [PFD_WARPS][rep*hd]
    double *red;       // [PFD_WARPS]
};

// reduction + fused epilogue of a finished work unit (same arithmetic as gemv_kernel's)
__device__ __forceinline__ void finish_unit(const PfdPhase &P, int unit, double (&acc)[GEMV_R], unsigned long long &best) {
    const int lane = threadIdx.x & 31;
    const double tot = warp_sum2_f64(acc[0], acc[1]);  // lanes 0-15: row 0, lanes 16-31: row 1
    float *out = P.out;
    if (P.epi == EPI_SWIGLU) {
        const float gte = (float)__shfl_sync(0xffffffffu, tot, 0), up = (float)__shfl_sync(0xffffffffu, tot, 16);
        if (lane == 0) out[unit] = __fmul_rn(__fdiv_rn(gte, __fadd_rn(1.0f, expf_det(-gte))), up);
    } else if ((lane & 15) == 0) {
        const int row = 2 * unit + (lane >> 4);
        if (row < P.w0.n_rows) {
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

__device__ __forceinline__ void compute_any(int type, const AnyRegs &g, int u0, int nu, int k, const PfdSmem &sm,
                                            double (&acc)[GEMV_R]) {
    switch (type) {
        case T_Q4_K: compute_pass<T_Q4_K, false>(g, u0, nu, k, carve(sm.act_a, k), acc); break;
        case T_Q5_K: compute_pass<T_Q5_K, false>(g, u0, nu, k, carve(sm.act_a, k), acc); break;
        case T_Q6_K: compute_pass<T_Q6_K, false>(g, u0, nu, k, carve(sm.act_b, k), acc); break;
        default: compute_pass<T_Q8_0, false>(g, u0, nu, k, carve(sm.act_a, k), acc); break;
    }
}

// All of this warp's items of one GEMV phase.  `C` holds the weights of the current item; the next item's
// (possibly of a later phase or token) are fetched into `N` while C is being consumed, then N becomes C.
__device__ __forceinline__ void gemv_phase(const PfdParams &p, const PfdPhase &P, int tok, int ph, Loader &L, Light &cur,
                                           AnyRegs &C, AnyRegs &N, const PfdSmem &sm, unsigned long long &best) {
    const int type = P.type;
    const int nu = P.k / unit_elems(type);
    double acc[GEMV_R] = {0.0, 0.0};
    while (cur.valid && cur.tok == tok && cur.ph == ph) {
        loader_advance(p, L);
        const Light nxt = light(L);
        if (nxt.valid) issue_load(L, N);
        compute_any(type, C, cur.chunk * 64, nu, P.k, sm, acc);
        if (cur.chunk + 1 == cur.nchunks) finish_unit(P, cur.unit, acc, best);
        cur = nxt;
        C = N;
    }
}

__device__ __noinline__ float dequant_elem_nl(int type, const uint8_t *row, int64_t i) { return dequant_elem(type, row, i); }

// ---------------------------------------------------------------- activation prologue of a GEMV phase
__device__ __forceinline__ void src_load8(const PfdParams &p, const PfdPhase &P, int token, int idx, float (&v)[8]) {
    if (P.x64) {
        // (the built-in-scratch variant keeps the f64 accumulators somewhere else than the phase table says)
        const double *src = p.x64_override ? p.x64_override : P.x64;
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = (float)__ldcg(src + idx + i);
    } else {
        const float4 a = __ldcg(reinterpret_cast<const float4 *>(P.x + idx));
        const float4 b = __ldcg(reinterpret_cast<const float4 *>(P.x + idx) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    }
}

// Q8_K-quantise the lane's 8 values of slab b once and store them in the lane order(s) the phase group needs
__device__ __forceinline__ void quant_slab_k(const float (&v)[8], int b, int k, int mask, const ActSmem &sa, const ActSmem &sb) {
    const int lane = threadIdx.x & 31;
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
    int q[8];
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
    int s8 = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s8 += q[i];
    const int s16 = s8 + __shfl_xor_sync(0xffffffffu, s8, 1);
    const int s32 = s16 + __shfl_xor_sync(0xffffffffu, s16, 2);
    const int U = k / 64;
    uint2 pk;
    pk.x = (q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((q[3] & 0xff) << 24);
    pk.y = (q[4] & 0xff) | ((q[5] & 0xff) << 8) | ((q[6] & 0xff) << 16) | ((q[7] & 0xff) << 24);
    if (mask & 1) {
        const int unit = b * 4 + (lane >> 3), kk = (lane & 7) >> 1;
        *reinterpret_cast<uint2 *>(reinterpret_cast<uint8_t *>(sa.act + kk * U + unit) + (lane & 1) * 8) = pk;
        if ((lane & 3) == 0) sa.bs32[b * 8 + (lane >> 2)] = s32;
        if (lane == 0) sa.d8[b] = d;
    }
    if (mask & 2) {
        const int h = lane >> 4, kk = (lane & 15) >> 2, t = (lane >> 1) & 1;
        const int unit = b * 4 + h * 2 + t;
        *reinterpret_cast<uint2 *>(reinterpret_cast<uint8_t *>(sb.act + kk * U + unit) + (lane & 1) * 8) = pk;
        if ((lane & 1) == 0) sb.bs16[unit * 4 + kk] = (int16_t)s16;
        if (lane == 0) sb.d8[b] = d;
    }
}

__device__ void pfd_prologue(const PfdParams &p, const PfdPhase &P, const PfdSmem &sm, int token) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int k = P.k, nslab = k / 256;
    const ActSmem sa = carve(sm.act_a, k), sb = carve(sm.act_b, k);
    const int mask = P.perm_mask;
    auto emit = [&](float (&v)[8], const float (&w)[8], int b, float scale) {
        if (P.norm_w) {
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = __fmul_rn(__fmul_rn(v[i], scale), w[i]);
        }
        if (mask & 4) quant_slab<PERM_Q80>(v, b, k, sa);
        else quant_slab_k(v, b, k, mask, sa, sb);
    };
    constexpr int NB = 4;  // slabs fetched per trip to L2 on the long-vector path
    if (nslab <= 2 * PFD_WARPS) {
        // k <= 4096: every lane keeps its (up to) 16 values in registers: one trip to L2 for x and, concurrently,
        // one for the norm weight
        float v[2][8], w[2][8];
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const int b = warp + j * PFD_WARPS;
#pragma unroll
            for (int i = 0; i < 8; i++) { v[j][i] = 0.f; w[j][i] = 0.f; }
            if (b < nslab) {
                src_load8(p, P, token, b * 256 + lane * 8, v[j]);
                if (P.norm_w) load8(P.norm_w + b * 256 + lane * 8, w[j]);
            }
        }
        float scale = 1.f;
        if (P.norm_w) {
            const double ss = block_sum(sumsq8(v[0]) + sumsq8(v[1]), sm.red);
            const float mean = (float)(ss / (double)k);
            scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, P.eps)));
        }
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const int b = warp + j * PFD_WARPS;
            if (b < nslab) emit(v[j], w[j], b, scale);
        }
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
        for (int b0 = warp; b0 < nslab; b0 += NB * PFD_WARPS) {
            float v[NB][8];
#pragma unroll
            for (int j = 0; j < NB; j++) {
                const int b = b0 + j * PFD_WARPS;
                if (b < nslab) src_load8(p, P, token, b * 256 + lane * 8, v[j]);
            }
#pragma unroll
            for (int j = 0; j < NB; j++) {
                const int b = b0 + j * PFD_WARPS;
                if (b < nslab) {
                    float w[8];
#pragma unroll
                    for (int i = 0; i < 8; i++) w[i] = 0.f;
                    if (P.norm_w) load8(P.norm_w + b * 256 + lane * 8, w);
                    emit(v[j], w, b, scale);
                }
            }
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
    const int kvs = p.kv_stride;  // cache row stride
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
        uint16_t *kc = P.k_cache + (size_t)pos * kvs + (size_t)g * HD;
        uint16_t *vc = P.v_cache + (size_t)pos * kvs + (size_t)g * HD;
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
            att_load_row<EPL>(P.k_cache + (size_t)j * kvs + (size_t)g * HD + lane * EPL, kf[t]);
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
            att_load_row<EPL>(P.v_cache + (size_t)j * kvs + (size_t)g * HD + lane * EPL, vf[t]);
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

    Loader L;
    loader_seek(p, L, 0, 0);
    Light cur = light(L);
    AnyRegs C, N;
    if (cur.valid) issue_load(L, C);  // weights of my first item: in flight before anything else happens

    for (int tok = 0; tok < p.n_tokens; tok++) {
        // the token to feed = arg-max of the logits kept from the previous pass
        const unsigned long long key = *reinterpret_cast<volatile unsigned long long *>(p.argmax_key);
        const int token = key_to_index(key);
        pos += 1;
        if (blockIdx.x == 0 && threadIdx.x == 0) p.out_tokens[p.step0 + tok] = token;
        bool key_cleared = false;
        {
            // residual stream <- embedding row of the token (bit-exact dequantisation), spread over the grid
            const uint8_t *row = p.tok_embd + (int64_t)token * p.tok_embd_row_bytes;
            for (int i = blockIdx.x * PFD_THREADS + threadIdx.x; i < p.n_embd; i += gridDim.x * PFD_THREADS)
                p.x[i] = dequant_elem_nl(p.tok_embd_type, row, i);
            bar++;
            counter_barrier(p.bar_counter, bar * gridDim.x);
            if (blockIdx.x == 0 && threadIdx.x == 0) *p.argmax_key = 0ull;  // every CTA has read the key by now
            key_cleared = true;
        }

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
                {
                    // weights of the next GEMV phase (of this or the next token) start moving HBM -> L2 now
                    int nph = ph + 1;
                    for (int s = 0; s < p.n_phases; s++, nph++) {
                        if (nph >= p.n_phases) nph = 0;
                        if (p.phases[nph].kind == PH_GEMV) break;
                    }
                    if (nph > ph || tok + 1 < p.n_tokens) l2_prefetch_phase(p.phases[nph]);
                }
                if (P.prologue) pfd_prologue(p, P, sm, token);
                if (tm) tm[1] = globaltimer_ns();
                unsigned long long best = 0ull;
                gemv_phase(p, P, tok, ph, L, cur, C, N, sm, best);
                if (P.argmax) {
                    const unsigned long long other = __shfl_sync(0xffffffffu, best, 16);
                    best = other > best ? other : best;
                    if (lane == 0 && best) atomicMax(p.argmax_key, best);
                }
            } else {
                const int epoch = att_count++;  // ordinal of this attention phase within the launch
                if (p.hd == 128) pfd_attention<4>(p, P, sm, pos, epoch);
                else pfd_attention<2>(p, P, sm, pos, epoch);
                if (tm) tm[1] = tm[0];
            }
            if (tm) tm[2] = globaltimer_ns();
            if (P.kind == PH_ATTN || P.barrier_after) {
                bar++;
                counter_barrier(p.bar_counter, bar * gridDim.x);
                // every CTA has read this token's key before arriving at the token's first barrier:
                // clear it for the next lm_head
                if (!key_cleared && blockIdx.x == 0 && threadIdx.x == 0) *p.argmax_key = 0ull;
                key_cleared = true;
            }
            if (tm) tm[3] = globaltimer_ns();
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        p.state[0] = pos;                      // SeqState.pos: position of the last token fed
        p.state[1] = p.step0 + p.n_tokens;     // SeqState.step
    }
}

// ---------------------------------------------------------------- built-in scratch (aios_seq_decode_static)
// The fused kernel's cross-CTA exchange area (barrier counters, per-CTA row maxima and sums, f64 output
// accumulators) normally lives in buffers each sequence cudaMalloc's when it is created.  This array is the
// same area baked into the library image, so a decode call that uses it touches the allocator not at all.
// Every GPU that loads the library carries these 48 KB until the process exits.
static __device__ __align__(256) unsigned char g_pfd_static_scratch[PFD_STATIC_SCRATCH_BYTES];

void *pfd_static_scratch(int device) {
    // In host code the array's name is only a handle for the runtime, not a device address: resolve it once per
    // device, with that device current (the caller has made it current), and hand the result to the kernel.
    static void *addr[64] = {nullptr};
    if (device < 0 || device >= 64) throw CudaError("static scratch: device ordinal out of range");
    if (!addr[device]) {
        void *ptr = nullptr;
        CUDA_CHECK(cudaGetSymbolAddress((void **)&ptr, g_pfd_static_scratch));
        addr[device] = ptr;
    }
    return addr[device];
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
