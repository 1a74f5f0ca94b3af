// attention_long.cu — decode attention for ONE token over a LONG context: split over KV positions, the query heads of
// a GQA group sharing every K / V row.
//
// attention_split.cu is a latency design for contexts of a few hundred rows: every CTA of a query head recomputes the
// head's whole score row, and the heads of a GQA group each read K and V for themselves (16x the K traffic and 4x the
// score arithmetic per kv head on Mistral-7B).  That is the right trade at ctx 512 (6.9 us per layer) and the wrong one
// at ctx 4096 (45 us per layer measured, half of the token: profiles/r02_summary.md).  Here the work is cut the other
// way — CTA = (kv head, slice of the positions) — in ONE launch of three phases:
//   A. scores: the CTA reads the K rows of its slice ONCE, converts them once, and scores them against all the query
//      heads of the group (scores stay in shared memory); the slice's per-head maxima go to global memory;
//   B. e = expf(s - row maximum) for the slice, the slice's share of the row sum (double) goes to global memory;
//   C. p = f16(e / row sum), P.V over the slice with the V rows shared by the group; the last CTA of a kv head to
//      finish adds the slices' partial sums (slice order) and writes the output.
// The exact soft-max needs the row maximum before any e and the row sum before any p (p is ROUNDED to f16, so it cannot
// be rescaled afterwards like a flash-decode partial): between the phases the CTAs of a kv head meet at a barrier on a
// device counter.  All CTAs of the grid are resident together (one small CTA per SM at most 148 of them), and nothing
// they wait for needs their SM slots, so the barriers cannot deadlock; a spin that lasts longer than ~1 s gives up
// (wrong output instead of a hung GPU).  A first version with two launches and the row sum recomputed by every CTA
// (4 x 4096 deterministic expf per CTA) only broke even with the latency kernel at ctx 4096.
//
// Arithmetic: exactly attention_split.cu's (and oracle/aios_oracle.c's): q / K / V are f16 values, scores and P.V are
// exact f16 x f16 products summed in double, score = (float)sum * scale, e = expf_det(s - max) with the GLOBAL row
// maximum, row sum in double, p = f16(e * (float)(1 / sum)), out = (float) of the double sum.  Sums of exact products
// do not depend on their order, so the output is the single-row kernels' bit for bit.
// Needs kv_ready launches (the q/k/v GEMV's epilogue has roped q and appended this token's K / V rows).
#include <cuda_fp16.h>

#include "common.cuh"
#include "kernels.h"

namespace aios {

constexpr int AL_THREADS = 256, AL_WARPS = AL_THREADS / 32;
constexpr int AL_HP = 4;     // query heads per pass
constexpr int AL_BATCH = 4;  // row slots in flight per warp (the score fold is written for exactly 4)

__device__ __forceinline__ double al_h2d_lo(uint32_t w) {
    double d;
    asm("cvt.f64.f16 %0, %1;" : "=d"(d) : "h"((uint16_t)(w & 0xffff)));
    return d;
}
__device__ __forceinline__ double al_h2d_hi(uint32_t w) {
    double d;
    asm("cvt.f64.f16 %0, %1;" : "=d"(d) : "h"((uint16_t)(w >> 16)));
    return d;
}
__device__ __forceinline__ void al_row_to_f64(const uint4 &v, double (&x)[8]) {
    x[0] = al_h2d_lo(v.x); x[1] = al_h2d_hi(v.x); x[2] = al_h2d_lo(v.y); x[3] = al_h2d_hi(v.y);
    x[4] = al_h2d_lo(v.z); x[5] = al_h2d_hi(v.z); x[6] = al_h2d_lo(v.w); x[7] = al_h2d_hi(v.w);
}

// positions [lo, hi) of slice s when n_pos rows are cut into n_slices contiguous pieces (multiples of 32 rows)
__device__ __forceinline__ void al_slice(int n_pos, int n_slices, int s, int &lo, int &hi) {
    const int chunk = ((n_pos + n_slices - 1) / n_slices + 31) & ~31;
    lo = min(n_pos, s * chunk);
    hi = min(n_pos, lo + chunk);
}

// ---------------------------------------------------------------- the kernel
// all CTAs of kv head g have arrived `round` times: counter counts arrivals and is never reset inside a sequence
__device__ __forceinline__ void al_group_barrier(unsigned *ctr, int n_ctas) {
    __syncthreads();  // the CTA's global writes are ordered before thread 0's release (cumulativity through the barrier)
    if (threadIdx.x == 0) {
        unsigned old;
        asm volatile("atom.add.release.gpu.global.u32 %0, [%1], 1;" : "=r"(old) : "l"(ctr) : "memory");
        const unsigned target = (old / (unsigned)n_ctas + 1u) * (unsigned)n_ctas;
        const long long t0 = clock64();
        for (;;) {
            unsigned v;
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
            if ((int)(v - target) >= 0) break;
            if (clock64() - t0 > 2000000000ll) break;  // never in a healthy run: do not hang the GPU
        }
    }
    __syncthreads();
}

// grid (n_kv, n_slices).  dynamic smem: sv[rep][chunk_cap] f32 (scores, then e, then p of the slice) |
// part[AL_WARPS][AL_HP][HD] f64
template <int HD>
__global__ void __launch_bounds__(AL_THREADS) attn_long_kernel(const AttnLongParams p) {
    constexpr int LPR = HD / 8, RPW = 32 / LPR;
    unsigned long long *tr = (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) ? p.trace : nullptr;
    trace_mark(tr, 0);
    if (tr) tr[3] = 2ull;  // launch kind: attention
    pdl_launch_dependents();
    pdl_wait();
    trace_mark(tr, 1);
    const int g = blockIdx.x, s = blockIdx.y;
    const int rep = p.n_head / p.n_kv;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n_pos = __ldcg(p.n_past_ptr) + 1;
    int lo, hi;
    al_slice(n_pos, p.n_slices, s, lo, hi);
    const uint16_t *kc = p.k_cache + (size_t)g * HD, *vc = p.v_cache + (size_t)g * HD;
    const size_t kvs = (size_t)p.kv_stride;
    extern __shared__ __align__(16) uint8_t al_smem[];
    float *sv = reinterpret_cast<float *>(al_smem);                                      // [rep][chunk_cap]
    double *part = reinterpret_cast<double *>(al_smem + (size_t)rep * p.chunk_cap * 4);  // [AL_WARPS][AL_HP][HD]
    __shared__ double qd[8 * 128];  // [rep][HD]
    __shared__ float sh_max[AL_WARPS][8];
    __shared__ double sh_sum[AL_WARPS][8];
    __shared__ float s_mx[8], s_inv[8];
    __shared__ int s_last;
    unsigned *bar = p.counters + 3 * g;  // [0] barrier A->B, [1] barrier B->C, [2] finished slices
    for (int i = threadIdx.x; i < rep * HD; i += AL_THREADS)
        qd[i] = (double)__half2float(__float2half_rn(p.qkv[(size_t)g * rep * HD + i]));
    __syncthreads();
    const int sub = lane / LPR, dl = (lane % LPR) * 8;
    const int rows_per_it = AL_WARPS * RPW;

    // ---- phase A: scores of the slice for every head of the group
    for (int h0 = 0; h0 < rep; h0 += AL_HP) {
        const int nh = min(AL_HP, rep - h0);
        double q[AL_HP][8];
#pragma unroll
        for (int h = 0; h < AL_HP; h++)
#pragma unroll
            for (int e = 0; e < 8; e++) q[h][e] = h < nh ? qd[(size_t)(h0 + h) * HD + dl + e] : 0.0;
        float wmax[AL_HP];
#pragma unroll
        for (int h = 0; h < AL_HP; h++) wmax[h] = -INFINITY;
        for (int jb = lo + warp * RPW; jb < hi; jb += rows_per_it * AL_BATCH) {
            uint4 kk[AL_BATCH];
#pragma unroll
            for (int t = 0; t < AL_BATCH; t++) {
                const int j = jb + t * rows_per_it + sub;
                kk[t] = make_uint4(0, 0, 0, 0);
                if (j < hi) kk[t] = __ldcg(reinterpret_cast<const uint4 *>(kc + (size_t)j * kvs + dl));
            }
            double part_s[AL_HP][AL_BATCH];
#pragma unroll
            for (int t = 0; t < AL_BATCH; t++) {
                double kd[8];
                al_row_to_f64(kk[t], kd);
#pragma unroll
                for (int h = 0; h < AL_HP; h++) {
                    double a = kd[0] * q[h][0], b = kd[1] * q[h][1];
                    a = fma(kd[2], q[h][2], a); b = fma(kd[3], q[h][3], b);
                    a = fma(kd[4], q[h][4], a); b = fma(kd[5], q[h][5], b);
                    a = fma(kd[6], q[h][6], a); b = fma(kd[7], q[h][7], b);
                    part_s[h][t] = a + b;
                }
            }
            // fold the LPR lanes of a row; the four row slots are folded together (halving butterfly)
            constexpr int MA = LPR / 2, MB = LPR / 4;
            const bool ba = lane & MA, bb = lane & MB;
            const int my_t = (ba ? 2 : 0) + (bb ? 1 : 0);
            const int j = jb + my_t * rows_per_it + sub;
#pragma unroll
            for (int h = 0; h < AL_HP; h++) {
                const double k0 = ba ? part_s[h][2] : part_s[h][0], k1 = ba ? part_s[h][3] : part_s[h][1];
                const double g0 = ba ? part_s[h][0] : part_s[h][2], g1 = ba ? part_s[h][1] : part_s[h][3];
                const double w0 = k0 + __shfl_xor_sync(0xffffffffu, g0, MA), w1 = k1 + __shfl_xor_sync(0xffffffffu, g1, MA);
                double tot = (bb ? w1 : w0) + __shfl_xor_sync(0xffffffffu, bb ? w0 : w1, MB);
#pragma unroll
                for (int o = MB / 2; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
                if (h < nh && j < hi && (lane & (MB - 1)) == 0) {
                    const float x = __fmul_rn((float)tot, p.scale);
                    sv[(size_t)(h0 + h) * p.chunk_cap + (j - lo)] = x;
                    wmax[h] = fmaxf(wmax[h], x);
                }
            }
        }
#pragma unroll
        for (int h = 0; h < AL_HP; h++) {
            const float m = warp_max(wmax[h]);
            if (lane == 0) sh_max[warp][h] = m;
        }
        __syncthreads();
        if ((int)threadIdx.x < nh) {
            float m = sh_max[0][threadIdx.x];
            for (int w = 1; w < AL_WARPS; w++) m = fmaxf(m, sh_max[w][threadIdx.x]);
            p.smax[(size_t)(g * rep + h0 + threadIdx.x) * p.n_slices + s] = m;
        }
        __syncthreads();
    }
    al_group_barrier(bar + 0, p.n_slices);
    trace_mark(tr, 2);

    // ---- phase B: e = expf(s - row maximum) for the slice; the slice's share of the row sums
    if ((int)threadIdx.x < rep) {
        const float *m = p.smax + (size_t)(g * rep + threadIdx.x) * p.n_slices;
        float mx = -INFINITY;
        for (int i = 0; i < p.n_slices; i++) mx = fmaxf(mx, __ldcg(m + i));
        s_mx[threadIdx.x] = mx;
    }
    __syncthreads();
    for (int h = 0; h < rep; h++) {
        const float mx = s_mx[h];
        float *row = sv + (size_t)h * p.chunk_cap;
        double ps = 0.0;
        for (int jj = threadIdx.x; jj < hi - lo; jj += AL_THREADS) {
            const float e = expf_det(__fsub_rn(row[jj], mx));
            row[jj] = e;
            ps += (double)e;
        }
        ps = warp_sum_f64(ps);
        if (lane == 0) sh_sum[warp][h] = ps;
    }
    __syncthreads();
    if ((int)threadIdx.x < rep) {
        double t = 0.0;
        for (int w = 0; w < AL_WARPS; w++) t += sh_sum[w][threadIdx.x];
        p.ssum[(size_t)(g * rep + threadIdx.x) * p.n_slices + s] = t;
    }
    al_group_barrier(bar + 1, p.n_slices);

    // ---- phase C: p = f16(e * (float)(1 / row sum)); P.V over the slice; combine
    if ((int)threadIdx.x < rep) {
        const double *q2 = p.ssum + (size_t)(g * rep + threadIdx.x) * p.n_slices;
        double t = 0.0;
        for (int i = 0; i < p.n_slices; i++) t += __ldcg(q2 + i);
        s_inv[threadIdx.x] = (float)(1.0 / t);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < rep * (hi - lo); i += AL_THREADS) {
        const int h = i / (hi - lo), jj = i % (hi - lo);
        float *e = sv + (size_t)h * p.chunk_cap + jj;
        *e = __half2float(__float2half_rn(__fmul_rn(*e, s_inv[h])));
    }
    __syncthreads();
    for (int h0 = 0; h0 < rep; h0 += AL_HP) {
        const int nh = min(AL_HP, rep - h0);
        double o[AL_HP][8];
#pragma unroll
        for (int h = 0; h < AL_HP; h++)
#pragma unroll
            for (int e = 0; e < 8; e++) o[h][e] = 0.0;
        for (int jb = lo + warp * RPW; jb < hi; jb += rows_per_it * AL_BATCH) {
            uint4 vv[AL_BATCH];
#pragma unroll
            for (int t = 0; t < AL_BATCH; t++) {
                const int j = jb + t * rows_per_it + sub;
                vv[t] = make_uint4(0, 0, 0, 0);
                if (j < hi) vv[t] = __ldcg(reinterpret_cast<const uint4 *>(vc + (size_t)j * kvs + dl));
            }
#pragma unroll
            for (int t = 0; t < AL_BATCH; t++) {
                const int j = jb + t * rows_per_it + sub;
                if (j < hi) {
                    double vd[8];
                    al_row_to_f64(vv[t], vd);
#pragma unroll
                    for (int h = 0; h < AL_HP; h++) {
                        if (h < nh) {
                            const double pv = (double)sv[(size_t)(h0 + h) * p.chunk_cap + (j - lo)];
#pragma unroll
                            for (int e = 0; e < 8; e++) o[h][e] = fma(vd[e], pv, o[h][e]);
                        }
                    }
                }
            }
        }
        // fold the RPW row groups of the warp, then the warps through shared memory
#pragma unroll
        for (int h = 0; h < AL_HP; h++)
#pragma unroll
            for (int e = 0; e < 8; e++) {
                double v = o[h][e];
#pragma unroll
                for (int s2 = LPR; s2 < 32; s2 <<= 1) v += __shfl_xor_sync(0xffffffffu, v, s2);
                o[h][e] = v;
            }
        if (sub == 0) {
#pragma unroll
            for (int h = 0; h < AL_HP; h++)
#pragma unroll
                for (int e = 0; e < 8; e++) part[((size_t)warp * AL_HP + h) * HD + dl + e] = o[h][e];
        }
        __syncthreads();
        for (int i = threadIdx.x; i < nh * HD; i += AL_THREADS) {
            const int h = i / HD, d = i % HD;
            double t = 0.0;
#pragma unroll
            for (int w = 0; w < AL_WARPS; w++) t += part[((size_t)w * AL_HP + h) * HD + d];
            p.opart[((size_t)s * p.n_head + (g * rep + h0 + h)) * HD + d] = t;
        }
        __syncthreads();
    }
    // the last slice of this kv head to arrive adds the slices' partial sums (slice order) and writes the output
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned old;
        asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(old) : "l"(bar + 2) : "memory");
        s_last = (old + 1u) % (unsigned)p.n_slices == 0u;
    }
    __syncthreads();
    if (s_last) {
        for (int i = threadIdx.x; i < rep * HD; i += AL_THREADS) {
            const size_t hd_idx = (size_t)(g * rep) * HD + i;
            double t = 0.0;
            for (int sl = 0; sl < p.n_slices; sl++) t += __ldcg(p.opart + (size_t)sl * p.n_head * HD + hd_idx);
            p.out[hd_idx] = (float)t;
        }
    }
    trace_mark(tr, 4);
    if (p.trace && threadIdx.x == 0) {  // [7]: when the last CTA of the launch finished
        unsigned long long now;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
        atomicMax(p.trace + 7, now);
    }
}

// ---------------------------------------------------------------- host side
static int al_chunk_cap(int n_ctx, int n_slices) { return (((n_ctx + n_slices - 1) / n_slices + 31) & ~31) + 32; }
static size_t al_smem_bytes(int rep, int n_ctx, int n_slices, int hd) {
    return (size_t)rep * al_chunk_cap(n_ctx, n_slices) * 4 + (size_t)AL_WARPS * AL_HP * hd * 8;
}
// one CTA per SM at most: every CTA of the grid must be resident for the barriers
int attn_long_slices(int n_kv, int n_sms) { return std::max(2, std::min(32, n_sms / std::max(1, n_kv))); }
bool attn_long_usable(int n_head, int n_kv, int n_ctx, int hd) {
    const int rep = n_kv > 0 ? n_head / n_kv : 0;
    return (hd == 64 || hd == 128) && rep >= 1 && rep <= 8 && n_head == rep * n_kv && n_ctx >= 64 && n_kv <= 64;
}
size_t attn_long_scratch_bytes(int n_head, int n_kv, int n_ctx, int hd, int n_slices) {
    return (size_t)n_slices * n_head * hd * 8 + (size_t)n_head * n_slices * 8 + (size_t)n_head * n_slices * 4 + (size_t)n_kv * 3 * 4 + 64;
}

void launch_attn_long(const AttnParams &a, void *scratch, int n_slices, cudaStream_t st, bool pdl) {
    if (!a.kv_ready) throw CudaError("attn_long: needs the q/k/v epilogue's roped q and appended K / V rows");
    AttnLongParams p{};
    p.qkv = a.qkv; p.out = a.out; p.k_cache = a.k_cache; p.v_cache = a.v_cache; p.kv_stride = a.kv_stride;
    p.n_past_ptr = a.n_past_ptr; p.scale = a.scale;
    p.n_head = a.n_head; p.n_kv = a.n_kv; p.hd = a.hd; p.n_ctx = a.n_ctx; p.n_slices = n_slices;
    p.chunk_cap = al_chunk_cap(a.n_ctx, n_slices);
    uint8_t *b = static_cast<uint8_t *>(scratch);
    p.opart = reinterpret_cast<double *>(b); b += (size_t)n_slices * a.n_head * a.hd * 8;
    p.ssum = reinterpret_cast<double *>(b); b += (size_t)a.n_head * n_slices * 8;
    p.smax = reinterpret_cast<float *>(b); b += (size_t)a.n_head * n_slices * 4;
    p.counters = reinterpret_cast<unsigned *>(b);
    p.trace = a.trace;
    const size_t smem = al_smem_bytes(a.n_head / a.n_kv, a.n_ctx, n_slices, a.hd);
    static PerDeviceOnce once;
    once.run([] {
        CUDA_CHECK(cudaFuncSetAttribute(attn_long_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        CUDA_CHECK(cudaFuncSetAttribute(attn_long_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    });
    if (smem > 96 * 1024) throw CudaError("attn_long: slice too long for shared memory");
    PdlScope scope(pdl);
    const dim3 grid((unsigned)a.n_kv, (unsigned)n_slices);
    if (a.hd == 128) launch_k(attn_long_kernel<128>, grid, dim3(AL_THREADS), smem, st, p);
    else launch_k(attn_long_kernel<64>, grid, dim3(AL_THREADS), smem, st, p);
}

}  // namespace aios
