// attention.cu — decode attention for one new token (SURVEY.md §2.3 K6 + K9):
// RoPE(q), RoPE(k) + f16 KV-cache write, and softmax(q.K^T/sqrt(d)).V over the cached positions
// with GQA sharing (the q-heads of one kv-head share a single K/V stream).
//
// B200 mapping: one thread-block CLUSTER per kv-head; the cluster's CTAs split the cached
// positions, and the three cross-split exchanges (row max, row sum, partial outputs) go through
// distributed shared memory + cluster barriers — no global scratch, no second launch.
//
// Numerics follow oracle/aios_oracle.c exactly in structure (ggml's
// non-flash path): NORM-style RoPE on adjacent pairs with the host-computed cos/sin table
// (bit-identical to the oracle's glibc values), K/V rounded to f16 in the cache, q rounded to f16
// before the dot, scores in f32, softmax = expf(s - max) / sum with the sum in double, and
// P ROUNDED TO F16 before the V product (ggml converts the f32 operand of an f16 mat-mul to
// f16).  Keeping P in f32 instead (flash-decode style) perturbs the output by ~2e-4, and the
// next layer's Q8_K re-quantisation amplifies that to the quantisation-noise floor (~4e-2 of
// the logit std on tiny models) — so the rounding is mirrored, not skipped.  Only f32
// summation orders differ from the oracle.
#include <cooperative_groups.h>

#include "common.cuh"
#include "kernels.h"

namespace cg = cooperative_groups;

namespace aios {

constexpr int ATT_THREADS = 256;
constexpr int ATT_WARPS = ATT_THREADS / 32;
constexpr int ATT_MAX_REP = 8;  // q-heads per kv-head
constexpr int ATT_BATCH = 4;    // cached positions fetched per warp iteration
constexpr int ATT_CLUSTER = 8;  // CTAs per kv-head (portable cluster size)

size_t attn_scratch_floats(int, int, int) { return 0; }

struct AttnShared {
    float wmax[ATT_WARPS][ATT_MAX_REP];
    double wsum[ATT_WARPS];
    float cta_max[ATT_MAX_REP];
    double cta_sum[ATT_MAX_REP];
    float gmax[ATT_MAX_REP];
    float ginv[ATT_MAX_REP];
};

template <int EPL>
__device__ __forceinline__ void load_row(const uint16_t *p, float (&f)[EPL]) {
    if constexpr (EPL == 4) {
        const uint2 v = __ldcg(reinterpret_cast<const uint2 *>(p));
        f[0] = h2f_bits((uint16_t)(v.x & 0xffff)); f[1] = h2f_bits((uint16_t)(v.x >> 16));
        f[2] = h2f_bits((uint16_t)(v.y & 0xffff)); f[3] = h2f_bits((uint16_t)(v.y >> 16));
    } else {
        const uint32_t v = __ldcg(reinterpret_cast<const uint32_t *>(p));
        f[0] = h2f_bits((uint16_t)(v & 0xffff)); f[1] = h2f_bits((uint16_t)(v >> 16));
    }
}

// EPL = elements per lane (hd / 32): 4 for hd 128, 2 for hd 64.
// dynamic smem: sc[rep][per_cap] scores, then part_o[ATT_WARPS][rep][HD] partial outputs
template <int EPL>
__global__ void __launch_bounds__(ATT_THREADS) attn_decode_kernel(const AttnParams p, int per_cap) {
    constexpr int HD = EPL * 32;
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int g = blockIdx.x / ATT_CLUSTER;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rep = p.n_head / p.n_kv;
    const int kvd = p.n_kv * HD;
    const int kvs = p.kv_stride;  // cache row stride
    const int n_past = *p.n_past_ptr;
    const int n_kv_pos = n_past + 1;  // including the new token
    const int per = (n_kv_pos + ATT_CLUSTER - 1) / ATT_CLUSTER;
    const int j0 = min(n_kv_pos, rank * per), j1 = min(n_kv_pos, j0 + per);
    const int cnt = j1 - j0;

    extern __shared__ __align__(16) float dyn[];
    float *sc = dyn;                                  // [rep][per_cap]
    double *part_o = reinterpret_cast<double *>(dyn + (size_t)rep * per_cap);  // [ATT_WARPS][rep][HD], later [rep][HD] in slot 0
    __shared__ AttnShared sh;

    const float *cs = p.rope_cos + (size_t)n_past * (p.n_rot / 2);
    const float *sn = p.rope_sin + (size_t)n_past * (p.n_rot / 2);

    // the CTA that owns the new position ropes k, rounds k/v to f16 and writes the cache rows
    if (n_past >= j0 && n_past < j1) {
        const float *kraw = p.qkv + (size_t)p.n_head * HD + (size_t)g * HD;
        const float *vraw = p.qkv + (size_t)p.n_head * HD + (size_t)kvd + (size_t)g * HD;
        uint16_t *kc = p.k_cache + (size_t)n_past * kvs + (size_t)g * HD;
        uint16_t *vc = p.v_cache + (size_t)n_past * kvs + (size_t)g * HD;
        for (int i = threadIdx.x; i < HD / 2; i += ATT_THREADS) {
            float x0 = kraw[2 * i], x1 = kraw[2 * i + 1];
            if (2 * i < p.n_rot) {
                const float c = cs[i], s = sn[i];
                const float r0 = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, s));
                const float r1 = __fadd_rn(__fmul_rn(x0, s), __fmul_rn(x1, c));
                x0 = r0; x1 = r1;
            }
            kc[2 * i] = f2h_bits(x0);
            kc[2 * i + 1] = f2h_bits(x1);
            vc[2 * i] = f2h_bits(vraw[2 * i]);
            vc[2 * i + 1] = f2h_bits(vraw[2 * i + 1]);
        }
        __syncthreads();  // this CTA reads the rows back below through the same (L2-coherent) path
    }

    // q of the rep heads sharing kv-head g: rope, round to f16 (vec_dot_type of an f16 K), keep my EPL dims
    float q[ATT_MAX_REP][EPL];
#pragma unroll
    for (int r = 0; r < ATT_MAX_REP; r++) {
        if (r < rep) {
            const float *qraw = p.qkv + (size_t)(g * rep + r) * HD + lane * EPL;
#pragma unroll
            for (int e = 0; e < EPL; e += 2) {
                float x0 = qraw[e], x1 = qraw[e + 1];
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

    // ---- phase 1: scores of my positions, local row max
    float wmax[ATT_MAX_REP];
#pragma unroll
    for (int r = 0; r < ATT_MAX_REP; r++) wmax[r] = -INFINITY;
    for (int jb = j0 + warp * ATT_BATCH; jb < j1; jb += ATT_WARPS * ATT_BATCH) {
        float kf[ATT_BATCH][EPL];
#pragma unroll
        for (int t = 0; t < ATT_BATCH; t++) {
            const int j = min(jb + t, j1 - 1);
            load_row<EPL>(p.k_cache + (size_t)j * kvs + (size_t)g * HD + lane * EPL, kf[t]);
        }
#pragma unroll
        for (int t = 0; t < ATT_BATCH; t++) {
            if (jb + t < j1) {
#pragma unroll
                for (int r = 0; r < ATT_MAX_REP; r++) {
                    if (r < rep) {
                        double sd = 0.0;  // f16 x f16 products are exact in f32; summed order-free in double
#pragma unroll
                        for (int e = 0; e < EPL; e++) sd += (double)__fmul_rn(q[r][e], kf[t][e]);
                        const float sdot = __fmul_rn((float)warp_sum_f64(sd), p.scale);
                        if (lane == 0) sc[(size_t)r * per_cap + (jb + t - j0)] = sdot;
                        wmax[r] = fmaxf(wmax[r], sdot);
                    }
                }
            }
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int r = 0; r < ATT_MAX_REP; r++) sh.wmax[warp][r] = wmax[r];
    }
    __syncthreads();
    if (threadIdx.x < rep) {
        float mx = -INFINITY;
        for (int w = 0; w < ATT_WARPS; w++) mx = fmaxf(mx, sh.wmax[w][threadIdx.x]);
        sh.cta_max[threadIdx.x] = mx;
    }
    cluster.sync();  // #1: every CTA's row max is published
    if (threadIdx.x < rep) {
        float mx = -INFINITY;
        for (int c = 0; c < ATT_CLUSTER; c++) mx = fmaxf(mx, cluster.map_shared_rank(&sh, c)->cta_max[threadIdx.x]);
        sh.gmax[threadIdx.x] = mx;
    }
    __syncthreads();

    // ---- phase 2: e = expf(s - max); row sum in double (ggml_float)
    {
        const int r = warp % rep;               // warps {r, r+rep, ...} serve head r
        const int sub = warp / rep, nsub = ATT_WARPS / rep;
        double part = 0.0;
        if (sub < nsub) {
            const float mx = sh.gmax[r];
            for (int j = sub * 32 + lane; j < cnt; j += nsub * 32) {
                const float e = expf_det(__fsub_rn(sc[(size_t)r * per_cap + j], mx));
                sc[(size_t)r * per_cap + j] = e;
                part += (double)e;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
        if (lane == 0) sh.wsum[warp] = part;
    }
    __syncthreads();
    if (threadIdx.x < rep) {
        double s = 0.0;
        for (int w = threadIdx.x; w < ATT_WARPS; w += rep) s += sh.wsum[w];
        sh.cta_sum[threadIdx.x] = s;
    }
    cluster.sync();  // #2: every CTA's row sum is published
    if (threadIdx.x < rep) {
        double s = 0.0;
        for (int c = 0; c < ATT_CLUSTER; c++) s += cluster.map_shared_rank(&sh, c)->cta_sum[threadIdx.x];
        sh.ginv[threadIdx.x] = (float)(1.0 / s);
    }
    __syncthreads();

    // ---- phase 3: o += f16(p) * V over my positions
    double o[ATT_MAX_REP][EPL];  // exact f16 x f16 products summed in double: order-free
#pragma unroll
    for (int r = 0; r < ATT_MAX_REP; r++)
#pragma unroll
        for (int e = 0; e < EPL; e++) o[r][e] = 0.0;
    for (int jb = j0 + warp * ATT_BATCH; jb < j1; jb += ATT_WARPS * ATT_BATCH) {
        float vf[ATT_BATCH][EPL];
#pragma unroll
        for (int t = 0; t < ATT_BATCH; t++) {
            const int j = min(jb + t, j1 - 1);
            load_row<EPL>(p.v_cache + (size_t)j * kvs + (size_t)g * HD + lane * EPL, vf[t]);
        }
#pragma unroll
        for (int t = 0; t < ATT_BATCH; t++) {
            if (jb + t < j1) {
#pragma unroll
                for (int r = 0; r < ATT_MAX_REP; r++) {
                    if (r < rep) {
                        const float pj = __half2float(__float2half_rn(
                            __fmul_rn(sc[(size_t)r * per_cap + (jb + t - j0)], sh.ginv[r])));
#pragma unroll
                        for (int e = 0; e < EPL; e++) o[r][e] += (double)__fmul_rn(pj, vf[t][e]);
                    }
                }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < ATT_MAX_REP; r++) {
        if (r < rep) {
#pragma unroll
            for (int e = 0; e < EPL; e++) part_o[((size_t)warp * rep + r) * HD + lane * EPL + e] = o[r][e];
        }
    }
    __syncthreads();
    // CTA partial -> slot 0 (fixed warp order)
    for (int idx = threadIdx.x; idx < rep * HD; idx += ATT_THREADS) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < ATT_WARPS; w++) s += part_o[(size_t)w * rep * HD + idx];
        part_o[idx] = s;  // element idx of every warp slot is touched by this thread only
    }
    cluster.sync();  // #3: every CTA's partial output is published
    {
        // each rank finishes a slice of the rep*HD outputs, summing the ranks in fixed order
        const int total = rep * HD;
        const int chunk = (total + ATT_CLUSTER - 1) / ATT_CLUSTER;
        const int i0 = rank * chunk, i1 = min(total, i0 + chunk);
        for (int idx = i0 + threadIdx.x; idx < i1; idx += ATT_THREADS) {
            double s = 0.0;
            for (int c = 0; c < ATT_CLUSTER; c++) s += cluster.map_shared_rank(part_o, c)[idx];
            p.out[(size_t)g * rep * HD + idx] = (float)s;
        }
    }
    cluster.sync();  // #4: nobody leaves while its shared memory may still be read
}

static size_t attn_dyn_smem(int rep, int hd, int per_cap) {
    // scores (f32; rep*per_cap is kept even so the f64 partials stay 8-byte aligned) + per-warp partials (f64)
    return ((size_t)rep * per_cap + 2 * (size_t)ATT_WARPS * rep * hd) * sizeof(float);
}

void attn_init() {
    static PerDeviceOnce once;
    once.run([] {
        CUDA_CHECK(cudaFuncSetAttribute(attn_decode_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        CUDA_CHECK(cudaFuncSetAttribute(attn_decode_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    });
}

void launch_attn_decode(const AttnParams &p, cudaStream_t st, bool pdl, int mode) {
    if (mode == 0 && attn_split_usable(p.n_ctx, p.hd)) {  // one wave of independent CTAs, PDL
        launch_attn_decode_split(p, st, pdl);
        return;
    }
    const int rep = p.n_head / p.n_kv;
    const int per_cap = ((p.n_ctx + ATT_CLUSTER - 1) / ATT_CLUSTER + 1) & ~1;  // even: keeps the f64 area aligned
    const size_t smem = attn_dyn_smem(rep, p.hd, per_cap);
    if (smem > 160 * 1024) throw CudaError("attention: context too long for the cluster's shared memory");
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(p.n_kv * ATT_CLUSTER);
    cfg.blockDim = dim3(ATT_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = ATT_CLUSTER;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (p.hd == 128) CUDA_CHECK(cudaLaunchKernelEx(&cfg, attn_decode_kernel<4>, p, per_cap));
    else if (p.hd == 64) CUDA_CHECK(cudaLaunchKernelEx(&cfg, attn_decode_kernel<2>, p, per_cap));
    else throw CudaError("attention: unsupported head_dim (64 and 128 are built)");
}

}  // namespace aios
