// attention_batch.cu — decode attention for MANY (sequence, token) pairs at once: the continuous batcher's step (one new
// token per live sequence, every lane over its own KV range) and the bit-exact prompt chunks (T tokens of one sequence,
// lane t at position pos0 + t).  attention_split.cu is a latency design for ONE token — 4 CTAs per query head, each
// recomputing the whole score row — and costs 64 x that at 64 lanes (329 us per layer measured).  Here throughput is
// the point: ONE CTA per (lane, kv head) serves all the query heads of the GQA group, so every K / V row is read and
// converted once per group instead of 4 x rep times, and the grid (lanes x kv heads) fills the machine by itself.
//
// Arithmetic: exactly the decode path's (attention_split.cu, oracle/aios_oracle.c) — RoPE from the host table, q / k / v
// rounded to f16, scores and P.V as exact f16 x f16 products summed in double (order-free), the deterministic expf,
// P rounded to f16 — so a lane's output is bit-identical to what the single-token kernels give it.
//
// Two input modes: kv_ready = 0, q|k|v raw f32 projections: the kernel ropes q and k, appends this token's k / v rows to
// the lane's cache (batched decode: lanes are different sequences, no lane reads another's rows); kv_ready = 1, q already
// roped and rounded (f16, attention_prefill.cu rope_kv_rows_kernel) and every row up to the lane's position already in
// the cache (prompt chunks: lane t must see the rows of lanes < t, so the append runs as its own launch first).
#include <cuda_fp16.h>

#include "common.cuh"
#include "kernels.h"

namespace aios {

constexpr int AB_THREADS = 256, AB_WARPS = AB_THREADS / 32;
constexpr int AB_HP = 4;      // query heads per pass (register budget: AB_HP x 8 f64 accumulators per lane)
constexpr int AB_BATCH = 4;   // row slots in flight per warp (the score fold below is written for exactly 4)

__device__ __forceinline__ double ab_h2d_lo(uint32_t w) {
    double d;
    asm("cvt.f64.f16 %0, %1;" : "=d"(d) : "h"((uint16_t)(w & 0xffff)));
    return d;
}
__device__ __forceinline__ double ab_h2d_hi(uint32_t w) {
    double d;
    asm("cvt.f64.f16 %0, %1;" : "=d"(d) : "h"((uint16_t)(w >> 16)));
    return d;
}
__device__ __forceinline__ void ab_row_to_f64(const uint4 &v, double (&x)[8]) {
    x[0] = ab_h2d_lo(v.x); x[1] = ab_h2d_hi(v.x); x[2] = ab_h2d_lo(v.y); x[3] = ab_h2d_hi(v.y);
    x[4] = ab_h2d_lo(v.z); x[5] = ab_h2d_hi(v.z); x[6] = ab_h2d_lo(v.w); x[7] = ab_h2d_hi(v.w);
}

// dynamic smem: sc[min(rep, AB_HP)][n_cap] f32 (one pass of heads at a time) | qd[rep][HD] f64 | part[AB_WARPS][2][AB_HP][8 * LPR ... ] (see below)
template <int HD>
__global__ void __launch_bounds__(AB_THREADS, 2) attn_batch_kernel(const AttnBatchParams p) {
    constexpr int LPR = HD / 8;          // lanes per row (8 dims = one 16-byte load per lane)
    constexpr int RPW = 32 / LPR;        // rows per warp-load: 2 (hd 128) or 4 (hd 64)
    pdl_wait();  // the q/k/v rows (and, inside a decode step, the lane table) come from the launches before this one
    pdl_launch_dependents();
    const int g = blockIdx.x, l = blockIdx.y;
    const int rep = p.n_head / p.n_kv;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int kvd = p.n_kv * HD, qd_all = p.n_head * HD;
    const int n_past = p.lane_pos[l];
    const int n_pos = n_past + 1;
    uint16_t *kc = (p.lane_kv ? p.lane_kv[l] : p.kv_base) + p.k_off + (size_t)g * HD;
    uint16_t *vc = (p.lane_kv ? p.lane_kv[l] : p.kv_base) + p.v_off + (size_t)g * HD;
    const size_t kvs = (size_t)p.kv_stride;

    extern __shared__ __align__(16) uint8_t ab_smem[];
    const int n_cap = (p.n_ctx + 3) & ~3;
    const int sc_heads = min(rep, AB_HP);                                         // score rows of ONE pass of heads
    float *sc = reinterpret_cast<float *>(ab_smem);                               // [sc_heads][n_cap]
    double *qd = reinterpret_cast<double *>(ab_smem + (size_t)sc_heads * n_cap * 4);  // [rep][HD]
    double *part = qd + (size_t)rep * HD;                                         // [AB_WARPS][AB_HP][HD]
    __shared__ float sh_max[AB_WARPS][8];
    __shared__ double sh_sum[AB_WARPS][8];
    __shared__ float s_mx[8], s_inv[8];

    // ---- q of the rep heads as doubles (roped + rounded to f16); raw mode: this token's k / v rows into the cache
    if (p.kv_ready) {
        const __half *qh = p.qh + (size_t)l * qd_all + (size_t)g * rep * HD;
        for (int i = threadIdx.x; i < rep * HD; i += AB_THREADS) qd[i] = (double)__half2float(qh[i]);
    } else {
        const float *row = p.qkv + (size_t)l * p.lane_qkv_stride;
        const float *cs = p.rope_cos + (size_t)n_past * (p.n_rot / 2), *sn = p.rope_sin + (size_t)n_past * (p.n_rot / 2);
        for (int i = threadIdx.x; i < rep * HD / 2; i += AB_THREADS) {
            const int h = (2 * i) / HD, d = (2 * i) % HD;
            const float *qraw = row + (size_t)(g * rep + h) * HD + d;
            float x0 = qraw[0], x1 = qraw[1];
            if (d < p.n_rot) {
                const float c = cs[d >> 1], s = sn[d >> 1];
                const float r0 = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, s));
                const float r1 = __fadd_rn(__fmul_rn(x0, s), __fmul_rn(x1, c));
                x0 = r0; x1 = r1;
            }
            qd[2 * i] = (double)__half2float(__float2half_rn(x0));
            qd[2 * i + 1] = (double)__half2float(__float2half_rn(x1));
        }
        const float *kraw = row + qd_all + (size_t)g * HD, *vraw = kraw + kvd;
        for (int i = threadIdx.x; i < HD / 2; i += AB_THREADS) {
            float x0 = kraw[2 * i], x1 = kraw[2 * i + 1];
            if (2 * i < p.n_rot) {
                const float c = cs[i], s = sn[i];
                const float r0 = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, s));
                const float r1 = __fadd_rn(__fmul_rn(x0, s), __fmul_rn(x1, c));
                x0 = r0; x1 = r1;
            }
            const __half2 kh = __floats2half2_rn(x0, x1), vh = __floats2half2_rn(vraw[2 * i], vraw[2 * i + 1]);
            *reinterpret_cast<__half2 *>(kc + (size_t)n_past * kvs + 2 * i) = kh;
            *reinterpret_cast<__half2 *>(vc + (size_t)n_past * kvs + 2 * i) = vh;
        }
    }
    __syncthreads();  // qd visible; (raw mode) this CTA's own cache row visible to its own later loads

    const int sub = lane / LPR, dl = (lane % LPR) * 8;   // my row inside a warp-load, my first dim
    const int rows_per_it = AB_WARPS * RPW;               // rows the CTA covers per step of one batch slot

    for (int h0 = 0; h0 < rep; h0 += AB_HP) {
        const int nh = min(AB_HP, rep - h0);
        // ---- scores of heads h0 .. h0+nh-1 over every position
        double q[AB_HP][8];
#pragma unroll
        for (int h = 0; h < AB_HP; h++)
#pragma unroll
            for (int e = 0; e < 8; e++) q[h][e] = h < nh ? qd[(size_t)(h0 + h) * HD + dl + e] : 0.0;
        float wmax[AB_HP];
#pragma unroll
        for (int h = 0; h < AB_HP; h++) wmax[h] = -INFINITY;
        for (int jb = warp * RPW; jb < n_pos; jb += rows_per_it * AB_BATCH) {
            uint4 kk[AB_BATCH];
#pragma unroll
            for (int t = 0; t < AB_BATCH; t++) {
                const int j = jb + t * rows_per_it + sub;
                kk[t] = make_uint4(0, 0, 0, 0);
                if (j < n_pos) kk[t] = __ldcg(reinterpret_cast<const uint4 *>(kc + (size_t)j * kvs + dl));
            }
            double part_s[AB_HP][AB_BATCH];  // my lane's share of score (head, row slot)
#pragma unroll
            for (int t = 0; t < AB_BATCH; t++) {
                double kd[8];
                ab_row_to_f64(kk[t], kd);
#pragma unroll
                for (int h = 0; h < AB_HP; h++) {
                    double a = kd[0] * q[h][0], b = kd[1] * q[h][1];
                    a = fma(kd[2], q[h][2], a); b = fma(kd[3], q[h][3], b);
                    a = fma(kd[4], q[h][4], a); b = fma(kd[5], q[h][5], b);
                    a = fma(kd[6], q[h][6], a); b = fma(kd[7], q[h][7], b);
                    part_s[h][t] = a + b;
                }
            }
            // fold the LPR lanes of a row: the four row slots are folded together (halving butterfly: 2 + 1 shuffles
            // for the first two levels instead of 4 + 4), after which a lane holds ONE slot's sum
            constexpr int MA = LPR / 2, MB = LPR / 4;
            const bool ba = lane & MA, bb = lane & MB;
            const int my_t = (ba ? 2 : 0) + (bb ? 1 : 0);
            const int j = jb + my_t * rows_per_it + sub;
#pragma unroll
            for (int h = 0; h < AB_HP; h++) {
                const double k0 = ba ? part_s[h][2] : part_s[h][0], k1 = ba ? part_s[h][3] : part_s[h][1];
                const double g0 = ba ? part_s[h][0] : part_s[h][2], g1 = ba ? part_s[h][1] : part_s[h][3];
                const double w0 = k0 + __shfl_xor_sync(0xffffffffu, g0, MA), w1 = k1 + __shfl_xor_sync(0xffffffffu, g1, MA);
                double tot = (bb ? w1 : w0) + __shfl_xor_sync(0xffffffffu, bb ? w0 : w1, MB);
#pragma unroll
                for (int o = MB / 2; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
                if (h < nh && j < n_pos && (lane & (MB - 1)) == 0) {
                    const float s = __fmul_rn((float)tot, p.scale);
                    sc[(size_t)h * n_cap + j] = s;
                    wmax[h] = fmaxf(wmax[h], s);
                }
            }
        }
#pragma unroll
        for (int h = 0; h < AB_HP; h++) {
            const float m = warp_max(wmax[h]);
            if (lane == 0) sh_max[warp][h] = m;
        }
        __syncthreads();
        if (threadIdx.x < AB_HP) {
            float m = sh_max[0][threadIdx.x];
            for (int w = 1; w < AB_WARPS; w++) m = fmaxf(m, sh_max[w][threadIdx.x]);
            s_mx[threadIdx.x] = m;
        }
        __syncthreads();
        // ---- e = expf(s - max), row sums in double
        {
            double ps[AB_HP];
#pragma unroll
            for (int h = 0; h < AB_HP; h++) {
                ps[h] = 0.0;
                if (h < nh) {
                    const float mx = s_mx[h];
                    float *row = sc + (size_t)h * n_cap;
                    for (int j = threadIdx.x; j < n_pos; j += AB_THREADS) {
                        const float e = expf_det(__fsub_rn(row[j], mx));
                        row[j] = e;
                        ps[h] += (double)e;
                    }
                }
                ps[h] = warp_sum_f64(ps[h]);
                if (lane == 0) sh_sum[warp][h] = ps[h];
            }
        }
        __syncthreads();
        if (threadIdx.x < AB_HP) {
            double t = 0.0;
            for (int w = 0; w < AB_WARPS; w++) t += sh_sum[w][threadIdx.x];
            s_inv[threadIdx.x] = (float)(1.0 / t);
        }
        __syncthreads();
        // ---- o[h][my 8 dims] = sum_j f16(p_hj) * v_j   (exact products, double sums)
        double o[AB_HP][8];
#pragma unroll
        for (int h = 0; h < AB_HP; h++)
#pragma unroll
            for (int e = 0; e < 8; e++) o[h][e] = 0.0;
        float inv[AB_HP];
#pragma unroll
        for (int h = 0; h < AB_HP; h++) inv[h] = s_inv[h];
        for (int jb = warp * RPW; jb < n_pos; jb += rows_per_it * AB_BATCH) {
            uint4 vv[AB_BATCH];
#pragma unroll
            for (int t = 0; t < AB_BATCH; t++) {
                const int j = jb + t * rows_per_it + sub;
                vv[t] = make_uint4(0, 0, 0, 0);
                if (j < n_pos) vv[t] = __ldcg(reinterpret_cast<const uint4 *>(vc + (size_t)j * kvs + dl));
            }
#pragma unroll
            for (int t = 0; t < AB_BATCH; t++) {
                const int j = jb + t * rows_per_it + sub;
                if (j < n_pos) {
                    double vd[8];
                    ab_row_to_f64(vv[t], vd);
#pragma unroll
                    for (int h = 0; h < AB_HP; h++) {
                        if (h < nh) {
                            const double pj = (double)__half2float(__float2half_rn(__fmul_rn(sc[(size_t)h * n_cap + j], inv[h])));
#pragma unroll
                            for (int e = 0; e < 8; e++) o[h][e] = fma(vd[e], pj, o[h][e]);
                        }
                    }
                }
            }
        }
        // fold the RPW row groups of the warp (lanes sub = 0..RPW-1 hold the same dims), then the warps through smem
#pragma unroll
        for (int h = 0; h < AB_HP; h++)
#pragma unroll
            for (int e = 0; e < 8; e++) {
                double v = o[h][e];
#pragma unroll
                for (int s2 = LPR; s2 < 32; s2 <<= 1) v += __shfl_xor_sync(0xffffffffu, v, s2);
                o[h][e] = v;
            }
        if (sub == 0) {
#pragma unroll
            for (int h = 0; h < AB_HP; h++)
#pragma unroll
                for (int e = 0; e < 8; e++) part[((size_t)warp * AB_HP + h) * HD + dl + e] = o[h][e];
        }
        __syncthreads();
        for (int i = threadIdx.x; i < nh * HD; i += AB_THREADS) {
            const int h = i / HD, d = i % HD;
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < AB_WARPS; w++) s += part[((size_t)w * AB_HP + h) * HD + d];
            p.out[(size_t)l * p.lane_out_stride + (size_t)(g * rep + h0 + h) * HD + d] = (float)s;
        }
        __syncthreads();  // part / s_mx / s_inv are reused by the next head group
    }
}

static size_t ab_smem_bytes(int rep, int n_ctx, int hd) {
    const size_t n_cap = ((size_t)n_ctx + 3) & ~(size_t)3;
    return (size_t)std::min(rep, AB_HP) * n_cap * 4 + (size_t)rep * hd * 8 + (size_t)AB_WARPS * AB_HP * hd * 8;
}

bool attn_batch_usable(int n_head, int n_kv, int n_ctx, int hd) {
    const int rep = n_head / n_kv;
    return (hd == 64 || hd == 128) && rep <= 8 && ab_smem_bytes(rep, n_ctx, hd) <= 200 * 1024;
}

void launch_attn_batch(const AttnBatchParams &p, int n_lanes, cudaStream_t st) {
    const int rep = p.n_head / p.n_kv;
    const size_t smem = ab_smem_bytes(rep, p.n_ctx, p.hd);
    if (!attn_batch_usable(p.n_head, p.n_kv, p.n_ctx, p.hd)) throw CudaError("batched attention: unsupported shape or context");
    static PerDeviceOnce once;
    once.run([] {
        CUDA_CHECK(cudaFuncSetAttribute(attn_batch_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CUDA_CHECK(cudaFuncSetAttribute(attn_batch_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    });
    const dim3 grid(p.n_kv, n_lanes);
    if (p.hd == 128) launch_k(attn_batch_kernel<128>, dim3(grid), dim3(AB_THREADS), smem, st, p);
    else launch_k(attn_batch_kernel<64>, dim3(grid), dim3(AB_THREADS), smem, st, p);
    CUDA_CHECK(cudaGetLastError());
}

}  // namespace aios
