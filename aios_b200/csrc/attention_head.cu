// attention_head.cu — decode attention for short contexts: ONE CTA PER QUERY HEAD, no cross-CTA exchange.
//
// The cluster kernel (attention.cu) splits the cached positions over 8 CTAs per kv-head and pays ~9 us of
// cluster-launch latency plus four cluster barriers (profiles/r01_v2_attention.md: 13 us at 10 cached positions,
// of which the SMs are active for 4).  Up to ~1k cached positions a single CTA can stream one head's K and V
// (256 KB at 512 positions) in less time than that, and then the soft-max row max / row sum stay in shared memory.
// GQA heads that share a kv-head read the same rows; L2 serves the repeats.
//
// Same arithmetic as attention.cu and oracle/aios_oracle.c: rope from the host table, K/V and q rounded to f16,
// exact f16 x f16 products summed in double, deterministic expf, P rounded to f16 before P.V.
#include "common.cuh"
#include "kernels.h"

namespace aios {

constexpr int AH_THREADS = 1024;
constexpr int AH_WARPS = AH_THREADS / 32;
constexpr int AH_BATCH = 4;  // row-groups fetched per warp iteration

__device__ __forceinline__ void unpack8(const uint4 &v, float (&f)[8]) {
    f[0] = h2f_bits((uint16_t)(v.x & 0xffff)); f[1] = h2f_bits((uint16_t)(v.x >> 16));
    f[2] = h2f_bits((uint16_t)(v.y & 0xffff)); f[3] = h2f_bits((uint16_t)(v.y >> 16));
    f[4] = h2f_bits((uint16_t)(v.z & 0xffff)); f[5] = h2f_bits((uint16_t)(v.z >> 16));
    f[6] = h2f_bits((uint16_t)(v.w & 0xffff)); f[7] = h2f_bits((uint16_t)(v.w >> 16));
}

// HD = head_dim; a row of HD halves is read by LPR = HD/8 lanes (16 bytes each), so a warp covers
// RPW = 32/LPR rows per load instruction.
// dynamic smem: sc[n_ctx] f32 | part[AH_WARPS][HD] f64 | knew[HD] f32 | vnew[HD] f32
template <int HD>
__global__ void __launch_bounds__(AH_THREADS, 1) attn_head_kernel(const AttnParams p) {
    constexpr int LPR = HD / 8, RPW = 32 / LPR;
    const int h = blockIdx.x;
    const int rep = p.n_head / p.n_kv;
    const int g = h / rep;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane / LPR, dl = (lane % LPR) * 8;  // my row inside the warp's row group; my first dim
    const int kvd = p.n_kv * HD;
    const int n_past = *p.n_past_ptr;
    const int n_pos = n_past + 1;

    extern __shared__ __align__(16) uint8_t dyn[];
    const int sc_cap = (p.n_ctx + 3) & ~3;
    float *sc = reinterpret_cast<float *>(dyn);
    double *part = reinterpret_cast<double *>(dyn + (size_t)sc_cap * 4);
    float *knew = reinterpret_cast<float *>(part + (size_t)AH_WARPS * HD);
    float *vnew = knew + HD;
    __shared__ float sh_f[AH_WARPS];
    __shared__ double sh_d[AH_WARPS];
    __shared__ float sh_max, sh_inv;

    const float *cs = p.rope_cos + (size_t)n_past * (p.n_rot / 2);
    const float *sn = p.rope_sin + (size_t)n_past * (p.n_rot / 2);

    // new token's k (roped) and v, rounded to f16: kept in smem for this CTA; the first head of the group
    // also writes the cache rows for later tokens
    {
        const float *kraw = p.qkv + (size_t)p.n_head * HD + (size_t)g * HD;
        const float *vraw = kraw + kvd;
        for (int i = threadIdx.x; i < HD / 2; i += AH_THREADS) {
            float x0 = kraw[2 * i], x1 = kraw[2 * i + 1];
            if (2 * i < p.n_rot) {
                const float c = cs[i], s = sn[i];
                const float r0 = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, s));
                const float r1 = __fadd_rn(__fmul_rn(x0, s), __fmul_rn(x1, c));
                x0 = r0; x1 = r1;
            }
            const uint16_t k0 = f2h_bits(x0), k1 = f2h_bits(x1);
            const uint16_t v0 = f2h_bits(vraw[2 * i]), v1 = f2h_bits(vraw[2 * i + 1]);
            knew[2 * i] = h2f_bits(k0); knew[2 * i + 1] = h2f_bits(k1);
            vnew[2 * i] = h2f_bits(v0); vnew[2 * i + 1] = h2f_bits(v1);
            if (h % rep == 0) {
                uint16_t *kc = p.k_cache + (size_t)n_past * kvd + (size_t)g * HD;
                uint16_t *vc = p.v_cache + (size_t)n_past * kvd + (size_t)g * HD;
                kc[2 * i] = k0; kc[2 * i + 1] = k1;
                vc[2 * i] = v0; vc[2 * i + 1] = v1;
            }
        }
    }
    // my 8 dims of q: rope, round to f16
    float q[8];
    {
        const float *qraw = p.qkv + (size_t)h * HD + dl;
#pragma unroll
        for (int e = 0; e < 8; e += 2) {
            float x0 = qraw[e], x1 = qraw[e + 1];
            const int i = (dl + e) / 2;
            if (2 * i < p.n_rot) {
                const float c = cs[i], s = sn[i];
                const float r0 = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, s));
                const float r1 = __fadd_rn(__fmul_rn(x0, s), __fmul_rn(x1, c));
                x0 = r0; x1 = r1;
            }
            q[e] = __half2float(__float2half_rn(x0));
            q[e + 1] = __half2float(__float2half_rn(x1));
        }
    }
    __syncthreads();  // knew / vnew visible

    // ---- scores: cached positions [0, n_past) from HBM/L2, position n_past from smem
    const uint16_t *kbase = p.k_cache + (size_t)g * HD + dl;
    float wmax = -INFINITY;
    const int step = AH_WARPS * RPW;
    for (int jb = warp * RPW; jb < n_pos; jb += step * AH_BATCH) {
        uint4 kk[AH_BATCH];
#pragma unroll
        for (int t = 0; t < AH_BATCH; t++) {
            const int j = jb + t * step + sub;
            kk[t] = make_uint4(0, 0, 0, 0);
            if (j < n_past) kk[t] = __ldcg(reinterpret_cast<const uint4 *>(kbase + (size_t)j * kvd));
        }
#pragma unroll
        for (int t = 0; t < AH_BATCH; t++) {
            const int j = jb + t * step + sub;
            float kf[8];
            unpack8(kk[t], kf);
            if (j == n_past) {
#pragma unroll
                for (int e = 0; e < 8; e++) kf[e] = knew[dl + e];
            }
            double sd = 0.0;  // exact f16 x f16 products, order-free double sum
#pragma unroll
            for (int e = 0; e < 8; e++) sd += (double)__fmul_rn(q[e], kf[e]);
#pragma unroll
            for (int o = LPR / 2; o > 0; o >>= 1) sd += __shfl_xor_sync(0xffffffffu, sd, o);
            if (j < n_pos) {
                const float s = __fmul_rn((float)sd, p.scale);
                if ((lane % LPR) == 0) sc[j] = s;
                wmax = fmaxf(wmax, s);
            }
        }
    }
    wmax = warp_max(wmax);
    if (lane == 0) sh_f[warp] = wmax;
    __syncthreads();
    if (warp == 0) {
        float m = warp_max(sh_f[lane]);
        if (lane == 0) sh_max = m;
    }
    __syncthreads();
    // ---- e = expf(s - max); row sum in double
    {
        const float mx = sh_max;
        double part_sum = 0.0;
        for (int j = threadIdx.x; j < n_pos; j += AH_THREADS) {
            const float e = expf_det(__fsub_rn(sc[j], mx));
            sc[j] = e;
            part_sum += (double)e;
        }
        part_sum = warp_sum_f64(part_sum);
        if (lane == 0) sh_d[warp] = part_sum;
    }
    __syncthreads();
    if (warp == 0) {
        double s = warp_sum_f64(sh_d[lane]);
        if (lane == 0) sh_inv = (float)(1.0 / s);
    }
    __syncthreads();
    // ---- o = sum_j f16(p_j) * v_j  (exact products, double sums)
    const float inv = sh_inv;
    const uint16_t *vbase = p.v_cache + (size_t)g * HD + dl;
    double o[8];
#pragma unroll
    for (int e = 0; e < 8; e++) o[e] = 0.0;
    for (int jb = warp * RPW; jb < n_pos; jb += step * AH_BATCH) {
        uint4 vv[AH_BATCH];
#pragma unroll
        for (int t = 0; t < AH_BATCH; t++) {
            const int j = jb + t * step + sub;
            vv[t] = make_uint4(0, 0, 0, 0);
            if (j < n_past) vv[t] = __ldcg(reinterpret_cast<const uint4 *>(vbase + (size_t)j * kvd));
        }
#pragma unroll
        for (int t = 0; t < AH_BATCH; t++) {
            const int j = jb + t * step + sub;
            if (j < n_pos) {
                float vf[8];
                unpack8(vv[t], vf);
                if (j == n_past) {
#pragma unroll
                    for (int e = 0; e < 8; e++) vf[e] = vnew[dl + e];
                }
                const float pj = __half2float(__float2half_rn(__fmul_rn(sc[j], inv)));
#pragma unroll
                for (int e = 0; e < 8; e++) o[e] += (double)__fmul_rn(pj, vf[e]);
            }
        }
    }
    // the RPW row groups of a warp hold the same dims: fold them, then fold the warps through smem
#pragma unroll
    for (int off = LPR; off < 32; off <<= 1) {
#pragma unroll
        for (int e = 0; e < 8; e++) o[e] += __shfl_xor_sync(0xffffffffu, o[e], off);
    }
    if (sub == 0) {
#pragma unroll
        for (int e = 0; e < 8; e++) part[(size_t)warp * HD + dl + e] = o[e];
    }
    __syncthreads();
    for (int d = threadIdx.x; d < HD; d += AH_THREADS) {
        double s = 0.0;
#pragma unroll 8
        for (int w = 0; w < AH_WARPS; w++) s += part[(size_t)w * HD + d];
        p.out[(size_t)h * HD + d] = (float)s;
    }
}

static size_t ah_smem(int n_ctx, int hd) {
    const size_t sc_cap = ((size_t)n_ctx + 3) & ~(size_t)3;
    return sc_cap * 4 + (size_t)AH_WARPS * hd * 8 + 2 * (size_t)hd * 4;
}

void attn_head_init() {
    static bool done = false;
    if (done) return;
    CUDA_CHECK(cudaFuncSetAttribute(attn_head_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    CUDA_CHECK(cudaFuncSetAttribute(attn_head_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    done = true;
}

bool attn_head_usable(int n_ctx, int hd) { return (hd == 64 || hd == 128) && n_ctx <= 1024 && ah_smem(n_ctx, hd) <= 96 * 1024; }

void launch_attn_decode_head(const AttnParams &p, cudaStream_t st) {
    const size_t smem = ah_smem(p.n_ctx, p.hd);
    if (p.hd == 128) attn_head_kernel<128><<<p.n_head, AH_THREADS, smem, st>>>(p);
    else attn_head_kernel<64><<<p.n_head, AH_THREADS, smem, st>>>(p);
}

}  // namespace aios
