// attention_split.cu — decode attention as ONE WAVE of independent CTAs: (query head, slice of the head dim).
//
// attention_head.cu runs one 1024-thread CTA per query head: 32 of the 148 SMs work, each walking its head's K and
// then V in sequence (15 us per launch at 512 cached positions, profiles/r01_v3_summary.md).  Here every query head
// is served by SPLIT CTAs.  Each of them computes the head's complete score row (the K rows come from L2; the
// arithmetic is ~65k exact products, nothing next to the latency it replaces), so each CTA knows the row maximum
// and the row sum by itself — and then accumulates P.V for its own HD/SPLIT output dims only.  No CTA ever waits
// for another: no cluster launch, no counters, no atomics, and 128 CTAs (Mistral: 32 heads x 4) fit in one wave.
//
// Instruction diet (the first version of this kernel took as long as the per-head one: IPC 0.2, long chains of
// f32 multiply -> convert -> double add -> 4 levels of 64-bit shuffles per row, profiles/r01_v4_attention.md):
//   * f16 -> f64 conversion in one instruction (F2F.F64.F16 reads either half of a register) and DFMA on the
//     converted values: 2 instructions per product instead of 4.  f16 x f16 products have 22 significant bits, so
//     DFMA(q, k, acc) adds the exact product exactly like the oracle's (double)(float product) + acc.
//   * the lanes that share a row do not reduce row by row: eight rows' partial sums are folded together with a
//     halving butterfly (7 double shuffles per 8 rows instead of 32).
//   * K rows are double buffered in registers; the first V batch is requested before the soft-max.
//
// The kernel is launched with programmatic dependent launch: it lets its successor (the attn_output GEMV) start
// prefetching weights at once, and waits for its predecessor (the q/k/v GEMV) only before touching qkv.
//
// Same arithmetic as attention.cu / attention_head.cu / oracle/aios_oracle.c: rope from the host table, q, K and V
// rounded to f16, exact f16 x f16 products summed in double (order-free), deterministic expf, P rounded to f16.
#include <algorithm>

#include "common.cuh"
#include "kernels.h"

namespace aios {

constexpr int AS_THREADS = 512;
constexpr int AS_WARPS = AS_THREADS / 32;
constexpr int AS_BATCH = 8;  // row-groups in flight per lane (and the width of the butterfly)

__device__ __forceinline__ double h2d_lo(uint32_t w) {
    double d;
    asm("cvt.f64.f16 %0, %1;" : "=d"(d) : "h"((uint16_t)(w & 0xffff)));
    return d;
}
__device__ __forceinline__ double h2d_hi(uint32_t w) {
    double d;
    asm("cvt.f64.f16 %0, %1;" : "=d"(d) : "h"((uint16_t)(w >> 16)));
    return d;
}

// sum_e q[e] * half(v, e): exact products, two independent chains
__device__ __forceinline__ double dot8(const uint4 &v, const double (&q)[8]) {
    double a = h2d_lo(v.x) * q[0], b = h2d_hi(v.x) * q[1];
    a = fma(h2d_lo(v.y), q[2], a); b = fma(h2d_hi(v.y), q[3], b);
    a = fma(h2d_lo(v.z), q[4], a); b = fma(h2d_hi(v.z), q[5], b);
    a = fma(h2d_lo(v.w), q[6], a); b = fma(h2d_hi(v.w), q[7], b);
    return a + b;
}

// o[e] += p * half(v, e)
__device__ __forceinline__ void axpy8(const uint4 &v, double p, double (&o)[8]) {
    o[0] = fma(h2d_lo(v.x), p, o[0]); o[1] = fma(h2d_hi(v.x), p, o[1]);
    o[2] = fma(h2d_lo(v.y), p, o[2]); o[3] = fma(h2d_hi(v.y), p, o[3]);
    o[4] = fma(h2d_lo(v.z), p, o[4]); o[5] = fma(h2d_hi(v.z), p, o[5]);
    o[6] = fma(h2d_lo(v.w), p, o[6]); o[7] = fma(h2d_hi(v.w), p, o[7]);
}

// Fold v[0..8) over the 8 lanes {lane ^ (a*M4 | b*M2 | c*M1)}: afterwards each lane holds the total of ONE
// element, index = 4*bit(M4) + 2*bit(M2) + bit(M1) of its lane id.  7 double shuffles, 7 adds.
template <int M4, int M2, int M1>
__device__ __forceinline__ double butterfly8(const double (&v)[8], int lane) {
    const bool b4 = lane & M4, b2 = lane & M2, b1 = lane & M1;
    double w[4], u[2];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const double send = b4 ? v[i] : v[i + 4], keep = b4 ? v[i + 4] : v[i];
        w[i] = keep + __shfl_xor_sync(0xffffffffu, send, M4);
    }
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const double send = b2 ? w[i] : w[i + 2], keep = b2 ? w[i + 2] : w[i];
        u[i] = keep + __shfl_xor_sync(0xffffffffu, send, M2);
    }
    const double send = b1 ? u[0] : u[1], keep = b1 ? u[1] : u[0];
    return keep + __shfl_xor_sync(0xffffffffu, send, M1);
}
template <int M4, int M2, int M1>
__device__ __forceinline__ int butterfly_index(int lane) {
    return ((lane & M4) ? 4 : 0) + ((lane & M2) ? 2 : 0) + ((lane & M1) ? 1 : 0);
}

// dynamic smem: sc[n_ctx] f32 | part[AS_WARPS][DPC] f64 | knew[HD] f16 | vnew[HD] f16
template <int HD, int SPLIT>
__global__ void __launch_bounds__(AS_THREADS, 1) attn_split_kernel(const AttnParams p_in) {
    constexpr int LPR = HD / 8, RPW = 32 / LPR;       // K: lanes per row, rows per warp-load
    constexpr int DPC = HD / SPLIT;                   // output dims of this CTA
    constexpr int VL = DPC / 8, RPV = 32 / VL;        // V: lanes per row slice, rows per warp-load
    static_assert(DPC % 8 == 0 && VL >= 1 && VL <= 4 && LPR >= 8, "unsupported head-dim split");
    const int h = blockIdx.x / SPLIT, part_id = blockIdx.x % SPLIT;
    // batched decode: blockIdx.y = lane; every lane has its own q/k/v slice, output slice, position and KV range
    AttnParams p = p_in;
    if (p_in.n_lanes > 0) {
        const int l = blockIdx.y;
        p.qkv = p_in.qkv + (size_t)l * p_in.lane_qkv_stride;
        p.out = p_in.out + (size_t)l * p_in.lane_out_stride;
        p.n_past_ptr = p_in.lane_pos + l;
        uint16_t *base = p_in.lane_kv[l];
        p.k_cache = base + (reinterpret_cast<uintptr_t>(p_in.k_cache) >> 1);
        p.v_cache = base + (reinterpret_cast<uintptr_t>(p_in.v_cache) >> 1);
    }
    const int rep = p.n_head / p.n_kv;
    const int g = h / rep;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int kvd = p.n_kv * HD;      // width of a raw k / v projection
    const int kvs = p.kv_stride;      // cache row stride

    extern __shared__ __align__(16) uint8_t dyn[];
    const int sc_cap = (p.n_ctx + 3) & ~3;
    float *sc = reinterpret_cast<float *>(dyn);
    double *part = reinterpret_cast<double *>(dyn + (size_t)sc_cap * 4);
    uint16_t *knew = reinterpret_cast<uint16_t *>(part + (size_t)AS_WARPS * DPC);
    uint16_t *vnew = knew + HD;
    // staged rows (kv_ready launches): K[stage_cap][HD] f16 | V[stage_cap][DPC] f16
    uint16_t *kst = vnew + HD;
    uint16_t *vst = kst + (size_t)p.stage_cap * HD;
    __shared__ float sh_f[AS_WARPS];
    __shared__ double sh_d[AS_WARPS];

    unsigned long long *tr = (blockIdx.x == 0 && threadIdx.x == 0) ? p.trace : nullptr;
    trace_mark(tr, 0);
    if (tr) tr[3] = 2ull;  // launch kind: attention
    // the successor only prefetches weights until this grid has completed: let it start now
    asm volatile("griddepcontrol.launch_dependents;");

    // Before the dependency wait: the cache rows of earlier tokens are final — they were last touched a whole token
    // pass (4 GB of weight traffic) ago, so they come from DRAM.  Copy them into shared memory NOW, all at once
    // (cp.async, 16 B per thread-instruction), while the previous kernel is still finishing: one DRAM round trip
    // that overlaps the dependency wait instead of two to four dependent ones after it.  The position counter read
    // here may in principle be stale; rows at or above the true position are (re)copied after the wait.
    const bool kv_ready = p.kv_ready != 0;
    const size_t head_off = (size_t)g * HD;
    auto stage_rows = [&](int lo, int hi) {  // rows [lo, hi) of K (whole head) and V (my DPC dims) -> shared memory
        for (int c = lo * LPR + threadIdx.x; c < hi * LPR; c += AS_THREADS) {
            const int j = c / LPR, ch = c % LPR;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(kst + (size_t)j * HD + ch * 8);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(p.k_cache + (size_t)j * kvs + head_off + ch * 8) : "memory");
        }
        for (int c = lo * VL + threadIdx.x; c < hi * VL; c += AS_THREADS) {
            const int j = c / VL, ch = c % VL;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(vst + (size_t)j * DPC + ch * 8);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(p.v_cache + (size_t)j * kvs + head_off + part_id * DPC + ch * 8) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int n_stage0 = 0;
    if (kv_ready && p.stage_cap > 0) {
        const int n_hint = *reinterpret_cast<const volatile int32_t *>(p.n_past_ptr);
        n_stage0 = max(0, min(n_hint, p.stage_cap));
        stage_rows(0, n_stage0);
    }
    // qkv, the position counter and the cache rows of earlier tokens are all final once the predecessor is done
    asm volatile("griddepcontrol.wait;" ::: "memory");
    trace_mark(tr, 1);

    const int n_past = __ldcg(p.n_past_ptr);
    const int n_pos = n_past + 1;
    // kv_ready: the q/k/v GEMV's epilogue already roped q, and appended this token's K/V rows to the cache
    const int n_cached = kv_ready ? n_pos : n_past;  // rows that come from the cache
    const int n_stage = kv_ready ? min(n_pos, p.stage_cap) : 0;  // rows served from shared memory
    if (n_stage > 0) stage_rows(min(n_stage0, n_past), n_stage);  // this token's row (and anything a stale hint missed)
    const float *cs = p.rope_cos + (size_t)n_past * (p.n_rot / 2);
    const float *sn = p.rope_sin + (size_t)n_past * (p.n_rot / 2);

    // ---- issue the first K batch right away: rows < n_past do not depend on this token
    const int ksub = lane / LPR, kdl = (lane % LPR) * 8;
    const uint16_t *kbase = p.k_cache + (size_t)g * HD + kdl;
    const int kstep = AS_WARPS * RPW;
    uint4 kk[2][AS_BATCH];
#pragma unroll
    for (int t = 0; t < AS_BATCH; t++) {
        const int j = warp * RPW + t * kstep + ksub;
        kk[0][t] = make_uint4(0, 0, 0, 0);
        if (j >= n_stage && j < n_cached) kk[0][t] = __ldcg(reinterpret_cast<const uint4 *>(kbase + (size_t)j * kvs));
    }

    // ---- new token's k (roped) and v, rounded to f16; one CTA of the kv group appends the cache rows
    if (!kv_ready) {
        const float *kraw = p.qkv + (size_t)p.n_head * HD + (size_t)g * HD;
        const float *vraw = kraw + kvd;
        const bool writer = (h % rep == 0) && part_id == 0;
        for (int i = threadIdx.x; i < HD / 2; i += AS_THREADS) {
            float x0 = kraw[2 * i], x1 = kraw[2 * i + 1];
            if (2 * i < p.n_rot) {
                const float c = cs[i], s = sn[i];
                const float r0 = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, s));
                const float r1 = __fadd_rn(__fmul_rn(x0, s), __fmul_rn(x1, c));
                x0 = r0; x1 = r1;
            }
            const uint16_t k0 = f2h_bits(x0), k1 = f2h_bits(x1);
            const uint16_t v0 = f2h_bits(vraw[2 * i]), v1 = f2h_bits(vraw[2 * i + 1]);
            knew[2 * i] = k0; knew[2 * i + 1] = k1;
            vnew[2 * i] = v0; vnew[2 * i + 1] = v1;
            if (writer) {
                uint16_t *kc = p.k_cache + (size_t)n_past * kvs + (size_t)g * HD;
                uint16_t *vc = p.v_cache + (size_t)n_past * kvs + (size_t)g * HD;
                kc[2 * i] = k0; kc[2 * i + 1] = k1;
                vc[2 * i] = v0; vc[2 * i + 1] = v1;
            }
        }
    }
    // my 8 dims of q: rope, round to f16, keep as doubles
    double q[8];
    {
        const float *qraw = p.qkv + (size_t)h * HD + kdl;
#pragma unroll
        for (int e = 0; e < 8; e += 2) {
            float x0 = qraw[e], x1 = qraw[e + 1];
            const int i = (kdl + e) / 2;
            if (!kv_ready && 2 * i < p.n_rot) {
                const float c = cs[i], s = sn[i];
                const float r0 = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, s));
                const float r1 = __fadd_rn(__fmul_rn(x0, s), __fmul_rn(x1, c));
                x0 = r0; x1 = r1;
            }
            q[e] = (double)__half2float(__float2half_rn(x0));
            q[e + 1] = (double)__half2float(__float2half_rn(x1));
        }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();  // staged rows landed; knew / vnew visible

    // ---- scores of ALL positions (every CTA of the head repeats them): staged rows from shared memory, the rest
    // of the cache from L2/DRAM, (without kv_ready) row n_past from knew
    float wmax = -INFINITY;
    const int kbstride = kstep * AS_BATCH;
    auto kload = [&](uint4 (&dst)[AS_BATCH], int jb) {
#pragma unroll
        for (int t = 0; t < AS_BATCH; t++) {
            const int j = jb + t * kstep + ksub;
            dst[t] = make_uint4(0, 0, 0, 0);
            if (j < n_stage) dst[t] = *reinterpret_cast<const uint4 *>(kst + (size_t)j * HD + kdl);
            else if (j < n_cached) dst[t] = __ldcg(reinterpret_cast<const uint4 *>(kbase + (size_t)j * kvs));
        }
    };
    if (n_stage > 0) {  // the first batch was requested before the staged rows had landed
#pragma unroll
        for (int t = 0; t < AS_BATCH; t++) {
            const int j = warp * RPW + t * kstep + ksub;
            if (j < n_stage) kk[0][t] = *reinterpret_cast<const uint4 *>(kst + (size_t)j * HD + kdl);
        }
    }
    auto kscore = [&](const uint4 (&src)[AS_BATCH], int jb) {
        double sd[AS_BATCH];
#pragma unroll
        for (int t = 0; t < AS_BATCH; t++) {
            const int j = jb + t * kstep + ksub;
            uint4 row = src[t];
            if (!kv_ready && j == n_past) row = *reinterpret_cast<const uint4 *>(knew + kdl);
            sd[t] = dot8(row, q);
        }
        // fold the LPR lanes of each row; lane ends up with the total of row-group `e`
        double tot;
        int e;
        if (LPR == 16) {
            tot = butterfly8<8, 4, 2>(sd, lane);
            tot += __shfl_xor_sync(0xffffffffu, tot, 1);
            e = butterfly_index<8, 4, 2>(lane);
        } else {
            tot = butterfly8<4, 2, 1>(sd, lane);
            e = butterfly_index<4, 2, 1>(lane);
        }
        const int j = jb + e * kstep + ksub;
        if (j < n_pos) {
            const float s = __fmul_rn((float)tot, p.scale);
            sc[j] = s;  // (LPR 16: two lanes store the same value)
            wmax = fmaxf(wmax, s);
        }
    };
    for (int jb = warp * RPW; jb < n_pos; jb += 2 * kbstride) {
        const bool more = jb + kbstride < n_pos;
        if (more) kload(kk[1], jb + kbstride);
        kscore(kk[0], jb);
        if (more) {
            if (jb + 2 * kbstride < n_pos) kload(kk[0], jb + 2 * kbstride);
            kscore(kk[1], jb + kbstride);
        }
    }

    trace_mark(tr, 2);
    // ---- request the first V batch now: it does not depend on the scores
    const int vsub = lane / VL, vdl = part_id * DPC + (lane % VL) * 8;
    const uint16_t *vbase = p.v_cache + (size_t)g * HD + vdl;
    const int vstep = AS_WARPS * RPV;
    uint4 vv[AS_BATCH];
    auto vload = [&](int jb) {
#pragma unroll
        for (int t = 0; t < AS_BATCH; t++) {
            const int j = jb + t * vstep + vsub;
            vv[t] = make_uint4(0, 0, 0, 0);
            if (j < n_stage) vv[t] = *reinterpret_cast<const uint4 *>(vst + (size_t)j * DPC + (lane % VL) * 8);
            else if (j < n_cached) vv[t] = __ldcg(reinterpret_cast<const uint4 *>(vbase + (size_t)j * kvs));
        }
    };
    vload(warp * RPV);

    wmax = warp_max(wmax);
    if (lane == 0) sh_f[warp] = wmax;
    __syncthreads();  // also: every sc[j] written
    float mx = sh_f[0];
#pragma unroll
    for (int w = 1; w < AS_WARPS; w++) mx = fmaxf(mx, sh_f[w]);

    // ---- e = expf(s - max); row sum in double
    {
        double part_sum = 0.0;
        for (int j = threadIdx.x; j < n_pos; j += AS_THREADS) {
            const float e = expf_det(__fsub_rn(sc[j], mx));
            sc[j] = e;
            part_sum += (double)e;
        }
        part_sum = warp_sum_f64(part_sum);
        if (lane == 0) sh_d[warp] = part_sum;
    }
    __syncthreads();
    double tot = 0.0;
#pragma unroll
    for (int w = 0; w < AS_WARPS; w++) tot += sh_d[w];
    const float inv = (float)(1.0 / tot);

    // ---- o[my dims] = sum_j f16(p_j) * v_j  (exact products, double sums)
    double o[8];
#pragma unroll
    for (int e = 0; e < 8; e++) o[e] = 0.0;
    for (int jb = warp * RPV; jb < n_pos; jb += vstep * AS_BATCH) {
        if (jb != warp * RPV) vload(jb);
#pragma unroll
        for (int t = 0; t < AS_BATCH; t++) {
            const int j = jb + t * vstep + vsub;
            if (j < n_pos) {
                uint4 row = vv[t];
                if (!kv_ready && j == n_past) row = *reinterpret_cast<const uint4 *>(vnew + vdl);
                const double pj = (double)__half2float(__float2half_rn(__fmul_rn(sc[j], inv)));
                axpy8(row, pj, o);
            }
        }
    }
    // the RPV row groups of a warp hold the same dims: fold them (halving butterfly), then the warps through smem
    double mine = butterfly8<16, 8, 4>(o, lane);
#pragma unroll
    for (int m = 2; m >= VL; m >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, m);
    part[(size_t)warp * DPC + (lane % VL) * 8 + butterfly_index<16, 8, 4>(lane)] = mine;
    __syncthreads();
    for (int d = threadIdx.x; d < DPC; d += AS_THREADS) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < AS_WARPS; w++) s += part[(size_t)w * DPC + d];
        p.out[(size_t)h * HD + part_id * DPC + d] = (float)s;
    }
    trace_mark(tr, 4);
    if (p.trace && threadIdx.x == 0) {  // [7]: when the last CTA of the launch finished
        unsigned long long now;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
        atomicMax(p.trace + 7, now);
    }
}

static size_t as_smem(int n_ctx, int hd, int split) {
    const size_t sc_cap = ((size_t)n_ctx + 3) & ~(size_t)3;
    return sc_cap * 4 + (size_t)AS_WARPS * (hd / split) * 8 + 2 * (size_t)hd * 2;
}

constexpr int AS_SPLIT = 4;

void attn_split_init() {
    static PerDeviceOnce once;
    once.run([] {
        // same shared-memory carve-out as the GEMVs it sits between (an SM cannot host two carve-outs at once)
        CUDA_CHECK(cudaFuncSetAttribute(attn_split_kernel<128, AS_SPLIT>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        CUDA_CHECK(cudaFuncSetAttribute(attn_split_kernel<64, AS_SPLIT>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        CUDA_CHECK(cudaFuncSetAttribute(attn_split_kernel<128, AS_SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024));
        CUDA_CHECK(cudaFuncSetAttribute(attn_split_kernel<64, AS_SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024));
    });
}

bool attn_split_usable(int n_ctx, int hd) {
    return (hd == 64 || hd == 128) && n_ctx <= 4096 && as_smem(n_ctx, hd, AS_SPLIT) <= 40 * 1024;
}

constexpr int AS_STAGE_ROWS = 512;  // cache rows copied into shared memory before the dependency wait (kv_ready)

void launch_attn_decode_split(const AttnParams &p_in, cudaStream_t st, bool pdl) {
    AttnParams p = p_in;
    p.stage_cap = p.kv_ready ? std::min(p.n_ctx, AS_STAGE_ROWS) : 0;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(p.n_head * AS_SPLIT, p.n_lanes > 0 ? p.n_lanes : 1);
    cfg.blockDim = dim3(AS_THREADS);
    cfg.dynamicSmemBytes = as_smem(p.n_ctx, p.hd, AS_SPLIT) + (size_t)p.stage_cap * (p.hd + p.hd / AS_SPLIT) * 2;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    if (p.hd == 128) CUDA_CHECK(cudaLaunchKernelEx(&cfg, attn_split_kernel<128, AS_SPLIT>, p));
    else CUDA_CHECK(cudaLaunchKernelEx(&cfg, attn_split_kernel<64, AS_SPLIT>, p));
}

}  // namespace aios
