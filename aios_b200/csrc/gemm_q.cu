// gemm_q.cu — multi-token dequant GEMM on the tensor cores: Y[T][N] = X[T][K] . dq(W[N][K])^T for T = 8..32 tokens
// of one prompt chunk (prefill), weights streamed ONCE for all T tokens.
//
// Same numbers as T calls of the GEMV (gemv.cu) — bit for bit.  The GEMV's arithmetic per 64-weight unit is
//   term = fma(d * d8, sc_lo * dot(lo32, a) + sc_hi * dot(hi32, a'), -(dmin * d8) * (m_lo * bs + m_hi * bs'))
// with INTEGER dots of 32 int8 pairs, and the terms are summed in double (order-free).  An integer dot does not care
// how it is computed, so here the 32-wide dots come from `mma.sync.m16n8k32.s8` — one instruction = 16 weight rows x
// 8 tokens x one 32-weight sub-block — and the scale arithmetic around them is the GEMV's, per (row, token) pair.
// (Why not tcgen05 with TMEM accumulators: K-quant scales change every 32 weights, so the accumulator has to be read
// out after every k = 32 step; a register accumulator is the natural home for that.  The integer MMA is what keeps
// the result exact; a bf16/fp8 tensor-core path would not be bit-compatible with the decode path.)
//
// Activations: `quant_rows_kernel` builds, per token, the same Q8_K / Q8_0 image the GEMV prologue builds in shared
// memory (gemv_core.cuh quant_slab, optional RMSNorm first) — in global memory, once per chunk, instead of once per
// CTA.  The GEMM stages the images of all T tokens block by block (256 weights) through shared memory.
//
// Q4_K and Q8_0 weights (sub-blocks of 32).  Q6_K (sub-blocks of 16) is not built yet: the engine runs those
// matrices token by token through the GEMV.
#define GEMV_CORE_WARPS 8
#include "gemv_core.cuh"

namespace aios {

// ---------------------------------------------------------------- per-token activation images in global memory
// grid = T, block = 256.  x row stride ldx (floats); image stride = act_image_bytes(k) rounded to 256.
__host__ __device__ inline size_t gemm_image_stride(int k) { return (act_image_bytes(k) + 255) & ~(size_t)255; }

template <int PERM>
__global__ void __launch_bounds__(GEMV_THREADS) quant_rows_kernel(const float *x, int64_t ldx, const float *norm_w, float eps,
                                                                   int k, uint8_t *images) {
    __shared__ double red[GEMV_WARPS];
    const int t = blockIdx.x;
    const ActSmem s = carve(images + (size_t)t * gemm_image_stride(k), k);
    prep_activations<PERM>(x + (size_t)t * ldx, norm_w, eps, k, s, red);
}

void launch_quant_rows(int weight_type, const float *x, int64_t ldx, const float *norm_w, float eps, int k, int T,
                       uint8_t *images, cudaStream_t st) {
    if (weight_type == T_Q6_K) quant_rows_kernel<PERM_Q6><<<T, GEMV_THREADS, 0, st>>>(x, ldx, norm_w, eps, k, images);
    else if (weight_type == T_Q8_0) quant_rows_kernel<PERM_Q80><<<T, GEMV_THREADS, 0, st>>>(x, ldx, norm_w, eps, k, images);
    else quant_rows_kernel<PERM_Q4><<<T, GEMV_THREADS, 0, st>>>(x, ldx, norm_w, eps, k, images);
    CUDA_CHECK(cudaGetLastError());
}
size_t gemm_images_bytes(int k, int T) { return gemm_image_stride(k) * (size_t)T; }

// ---------------------------------------------------------------- the GEMM
constexpr int GEMM_WARPS = 2, GEMM_THREADS = GEMM_WARPS * 32, GEMM_ROWS = GEMM_WARPS * 16;
constexpr int GEMM_TOK_STRIDE = 17;  // uint4 per token in the staged chunk: 16 used + 1 pad (bank spread over tokens)

__device__ __forceinline__ void mma_s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
        : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0));
}
__device__ __forceinline__ void cp16(void *smem_dst, const void *gsrc) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}

struct GemmSmem {  // one staged 256-weight block of all T token images
    uint4 *act;    // [T][GEMM_TOK_STRIDE]
    int32_t *bs;   // Q4_K: [T][8] 32-wide q8 sums
    float *d8;     // Q4_K: [T] ; Q8_0: [T][8]
};
__host__ __device__ inline size_t gemm_chunk_bytes(int T) { return (size_t)T * (GEMM_TOK_STRIDE * 16 + 32 + 32); }
__device__ __forceinline__ GemmSmem gemm_carve(uint8_t *base, int T) {
    GemmSmem s;
    s.act = reinterpret_cast<uint4 *>(base);
    s.bs = reinterpret_cast<int32_t *>(base + (size_t)T * GEMM_TOK_STRIDE * 16);
    s.d8 = reinterpret_cast<float *>(base + (size_t)T * (GEMM_TOK_STRIDE * 16 + 32));
    return s;
}

// copy block b of every token's image into a staging buffer (cp.async, all threads)
template <int TYPE>
__device__ __forceinline__ void gemm_stage(const GemmSmem &s, const uint8_t *images, size_t img_stride, int k, int T, int b) {
    const int U = TYPE == T_Q8_0 ? k / 32 : k / 64;
    for (int idx = threadIdx.x; idx < T * 16; idx += GEMM_THREADS) {
        const int t = idx >> 4, e = idx & 15;
        const uint4 *act = reinterpret_cast<const uint4 *>(images + (size_t)t * img_stride);
        const uint4 *src = TYPE == T_Q8_0 ? act + (size_t)(e & 1) * U + 8 * b + (e >> 1)   // [blk][kk]
                                          : act + (size_t)(e & 3) * U + 4 * b + (e >> 2);  // [unit][kk]
        cp16(s.act + t * GEMM_TOK_STRIDE + e, src);
    }
    for (int idx = threadIdx.x; idx < T * 2; idx += GEMM_THREADS) {
        const int t = idx >> 1, h = idx & 1;
        const uint8_t *aux = images + (size_t)t * img_stride + k;
        if (TYPE == T_Q8_0) cp16(s.d8 + t * 8 + h * 4, aux + k / 8 + (size_t)(8 * b + h * 4) * 4);
        else cp16(s.bs + t * 8 + h * 4, aux + (size_t)(8 * b + h * 4) * 4);
    }
    if (TYPE != T_Q8_0) {
        for (int t = threadIdx.x; t < T; t += GEMM_THREADS) {
            const uint8_t *aux = images + (size_t)t * img_stride + k;
            const uint32_t d = (uint32_t)__cvta_generic_to_shared(s.d8 + t);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(aux + k / 8 + (size_t)b * 4) : "memory");
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}

struct GemmParams {
    WPlanes w;
    const uint8_t *images;  // T activation images (quant_rows_kernel), stride gemm_image_stride(k)
    float *out;             // [T][ldo]
    int64_t ldo;
    int32_t T, epi;         // epi: EPI_STORE or EPI_ADD
};

// Weight side of a stage: the 256-weight block b of GEMM_ROWS rows, copied plane by plane with 16-byte cp.async
// (a row's bytes of one block are contiguous inside a plane), padded per row so that the fragment reads of the eight
// row groups of a warp fall into different banks.
//   Q4_K: qa[rows][5] uint4 (4 units + pad) | qb[rows][5] | meta[rows] uint4
//   Q8_0: qa[rows][9] uint4 (8 blocks + pad) | qb[rows][9] | d[rows][8] f16
template <int TYPE> struct GemmA;
template <> struct GemmA<T_Q4_K> {
    static constexpr int RS = 5;
    static constexpr size_t bytes = (size_t)GEMM_ROWS * (2 * RS + 1) * 16;
};
template <> struct GemmA<T_Q8_0> {
    static constexpr int RS = 9;
    static constexpr size_t bytes = (size_t)GEMM_ROWS * (2 * RS + 1) * 16;
};

template <int TYPE>
__device__ __forceinline__ void gemm_stage_a(uint8_t *sa, const WPlanes &w, int row0, int b) {
    constexpr int RS = GemmA<TYPE>::RS, UPB = TYPE == T_Q8_0 ? 8 : 4;  // units (blocks of 32) per 256-weight block
    const size_t U = (size_t)w.k / (TYPE == T_Q8_0 ? 32 : 64), nb = (size_t)w.k / 256;
    uint4 *qa = reinterpret_cast<uint4 *>(sa), *qb = qa + GEMM_ROWS * RS, *ex = qb + GEMM_ROWS * RS;
    for (int idx = threadIdx.x; idx < GEMM_ROWS * UPB; idx += GEMM_THREADS) {
        const int r = idx / UPB, u = idx % UPB;
        const size_t row = (size_t)min(row0 + r, w.n_rows - 1);
        cp16(qa + r * RS + u, w.p[0] + (row * U + (size_t)UPB * b + u) * 16);
        cp16(qb + r * RS + u, w.p[1] + (row * U + (size_t)UPB * b + u) * 16);
    }
    for (int r = threadIdx.x; r < GEMM_ROWS; r += GEMM_THREADS) {
        const size_t row = (size_t)min(row0 + r, w.n_rows - 1);
        if (TYPE == T_Q8_0) cp16(ex + r, w.p[2] + (row * U + (size_t)8 * b) * 2);   // eight f16 scales = 16 bytes
        else cp16(ex + r, w.p[2] + (row * nb + b) * 16);                            // META of the block
    }
}

// grid = ceil(N / GEMM_ROWS); warp w of a CTA owns rows [GEMM_ROWS*cta + 16*w, +16) for the whole K and all T = 8*NT
// tokens.  Thread (g = lane/4, q = lane%4) ends up with the sums of rows g and g+8 for tokens 8j+2q, 8j+2q+1, j < NT.
// Three-stage cp.async pipeline over the 256-weight blocks of K: weights and activation images of block b+2 are in
// flight while block b is multiplied.
template <int TYPE, int NT>
__global__ void __launch_bounds__(GEMM_THREADS) gemm_q_kernel(const GemmParams p) {
    extern __shared__ __align__(16) uint8_t smem[];
    constexpr int T = 8 * NT, RS = GemmA<TYPE>::RS, STAGES = 3;
    const size_t stage_bytes = GemmA<TYPE>::bytes + gemm_chunk_bytes(T);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, q = lane & 3;
    const int k = p.w.k, N = p.w.n_rows;
    const int nb = k / 256;
    const size_t img_stride = gemm_image_stride(k);
    const int row0 = blockIdx.x * GEMM_ROWS;
    const int ra = warp * 16 + g, rb = ra + 8;  // my two rows inside the CTA's tile

    double acc[NT][4];
#pragma unroll
    for (int j = 0; j < NT; j++)
#pragma unroll
        for (int e = 0; e < 4; e++) acc[j][e] = 0.0;

    auto stage = [&](int b) {
        uint8_t *base = smem + (size_t)(b % STAGES) * stage_bytes;
        gemm_stage_a<TYPE>(base, p.w, row0, b);
        gemm_stage<TYPE>(gemm_carve(base + GemmA<TYPE>::bytes, T), p.images, img_stride, k, T, b);  // commits the group
    };
    stage(0);
    if (nb > 1) stage(1);
    for (int b = 0; b < nb; b++) {
        if (b + 2 < nb) {
            stage(b + 2);
            asm volatile("cp.async.wait_group 2;" ::: "memory");
        } else if (b + 1 < nb) {
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        const uint8_t *base = smem + (size_t)(b % STAGES) * stage_bytes;
        const uint4 *qa = reinterpret_cast<const uint4 *>(base), *qb = qa + GEMM_ROWS * RS, *ex = qb + GEMM_ROWS * RS;
        const GemmSmem s = gemm_carve(const_cast<uint8_t *>(base) + GemmA<TYPE>::bytes, T);
        if (TYPE == T_Q8_0) {
            const __half *dha = reinterpret_cast<const __half *>(ex + ra), *dhb = reinterpret_cast<const __half *>(ex + rb);
#pragma unroll 2
            for (int u = 0; u < 8; u++) {
                const uint32_t a0 = reinterpret_cast<const uint32_t *>(qa + ra * RS + u)[q];
                const uint32_t a1 = reinterpret_cast<const uint32_t *>(qa + rb * RS + u)[q];
                const uint32_t a2 = reinterpret_cast<const uint32_t *>(qb + ra * RS + u)[q];
                const uint32_t a3 = reinterpret_cast<const uint32_t *>(qb + rb * RS + u)[q];
                const float dwa = __half2float(dha[u]), dwb = __half2float(dhb[u]);
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    const uint32_t *bt = reinterpret_cast<const uint32_t *>(s.act + (8 * j + g) * GEMM_TOK_STRIDE + u * 2);
                    int c[4];
                    mma_s8(c, a0, a1, a2, a3, bt[q], bt[4 + q]);
                    const float d0 = s.d8[(8 * j + 2 * q) * 8 + u], d1 = s.d8[(8 * j + 2 * q + 1) * 8 + u];
                    acc[j][0] += (double)__fmul_rn((float)c[0], __fmul_rn(dwa, d0));
                    acc[j][1] += (double)__fmul_rn((float)c[1], __fmul_rn(dwa, d1));
                    acc[j][2] += (double)__fmul_rn((float)c[2], __fmul_rn(dwb, d0));
                    acc[j][3] += (double)__fmul_rn((float)c[3], __fmul_rn(dwb, d1));
                }
            }
        } else {  // Q4_K
            const uint4 mta = ex[ra], mtb = ex[rb];
            const uint32_t dda = __byte_perm(__byte_perm(mta.x, mta.y, 0x0073), __byte_perm(mta.z, mta.w, 0x0073), 0x5410);
            const uint32_t ddb = __byte_perm(__byte_perm(mtb.x, mtb.y, 0x0073), __byte_perm(mtb.z, mtb.w, 0x0073), 0x5410);
            const float2 da = __half22float2(*reinterpret_cast<const __half2 *>(&dda));  // (d, dmin) of row a
            const float2 db = __half22float2(*reinterpret_cast<const __half2 *>(&ddb));
            const uint32_t mwa[4] = {mta.x, mta.y, mta.z, mta.w}, mwb[4] = {mtb.x, mtb.y, mtb.z, mtb.w};
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const uint32_t qa0 = reinterpret_cast<const uint32_t *>(qa + ra * RS + u)[q];
                const uint32_t qa1 = reinterpret_cast<const uint32_t *>(qa + rb * RS + u)[q];
                const uint32_t qb0 = reinterpret_cast<const uint32_t *>(qb + ra * RS + u)[q];
                const uint32_t qb1 = reinterpret_cast<const uint32_t *>(qb + rb * RS + u)[q];
                const int sa_lo = mwa[u] & 63, sa_hi = (mwa[u] >> 6) & 63, ma_lo = (mwa[u] >> 12) & 63, ma_hi = (mwa[u] >> 18) & 63;
                const int sb_lo = mwb[u] & 63, sb_hi = (mwb[u] >> 6) & 63, mb_lo = (mwb[u] >> 12) & 63, mb_hi = (mwb[u] >> 18) & 63;
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    const uint32_t *bt = reinterpret_cast<const uint32_t *>(s.act + (8 * j + g) * GEMM_TOK_STRIDE + u * 4);
                    int lo[4], hi[4];
                    mma_s8(lo, qa0 & 0x0F0F0F0Fu, qa1 & 0x0F0F0F0Fu, qb0 & 0x0F0F0F0Fu, qb1 & 0x0F0F0F0Fu, bt[q], bt[4 + q]);
                    mma_s8(hi, (qa0 >> 4) & 0x0F0F0F0Fu, (qa1 >> 4) & 0x0F0F0F0Fu, (qb0 >> 4) & 0x0F0F0F0Fu,
                           (qb1 >> 4) & 0x0F0F0F0Fu, bt[8 + q], bt[12 + q]);
#pragma unroll
                    for (int e = 0; e < 2; e++) {  // my two tokens of this n-block
                        const int t = 8 * j + 2 * q + e;
                        const int2 bs = *reinterpret_cast<const int2 *>(s.bs + t * 8 + 2 * u);
                        const float d8 = s.d8[t];
                        const int isa = sa_lo * lo[e] + sa_hi * hi[e], msa = ma_lo * bs.x + ma_hi * bs.y;
                        const int isb = sb_lo * lo[2 + e] + sb_hi * hi[2 + e], msb = mb_lo * bs.x + mb_hi * bs.y;
                        acc[j][e] += (double)__fmaf_rn(__fmul_rn(da.x, d8), (float)isa, -__fmul_rn(__fmul_rn(da.y, d8), (float)msa));
                        acc[j][2 + e] += (double)__fmaf_rn(__fmul_rn(db.x, d8), (float)isb, -__fmul_rn(__fmul_rn(db.y, d8), (float)msb));
                    }
                }
            }
        }
        __syncthreads();  // everyone is done with this stage before it is refilled
    }
    const int row_a = row0 + ra, row_b = row0 + rb;
#pragma unroll
    for (int j = 0; j < NT; j++)
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const int row = (e & 2) ? row_b : row_a, t = 8 * j + 2 * q + (e & 1);
            if (row < N && t < p.T) {
                float *o = p.out + (size_t)t * p.ldo + row;
                const float v = (float)acc[j][e];
                *o = p.epi == EPI_ADD ? __fadd_rn(*o, v) : v;
            }
        }
}

template <int TYPE>
static size_t gemm_smem_bytes(int T) { return 3 * (GemmA<TYPE>::bytes + gemm_chunk_bytes(T)); }

// h[t][i] = silu(gate[t][i]) * up[t][i] for a chunk: gu[t] = gate[ff] | up[ff].  Same expression, operation by
// operation, as the GEMV's fused SwiGLU epilogue.
__global__ void swiglu_rows_kernel(const float *gu, int64_t ldgu, int ff, int T, float *h, int64_t ldh) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)T * ff) return;
    const int t = (int)(i / ff), c = (int)(i % ff);
    const float gte = gu[(size_t)t * ldgu + c], up = gu[(size_t)t * ldgu + ff + c];
    h[(size_t)t * ldh + c] = __fmul_rn(__fdiv_rn(gte, __fadd_rn(1.0f, expf_det(-gte))), up);
}
void launch_swiglu_rows(const float *gu, int64_t ldgu, int ff, int T, float *h, int64_t ldh, cudaStream_t st) {
    const int64_t n = (int64_t)T * ff;
    swiglu_rows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(gu, ldgu, ff, T, h, ldh);
    CUDA_CHECK(cudaGetLastError());
}

bool gemm_q_usable(int type, int k) { return (type == T_Q4_K || type == T_Q8_0) && k % 256 == 0; }

// T in 1..32: the images of tokens T..8*NT-1 (padding) must exist (any finite contents); their outputs are not stored.
void launch_gemm_q(const WPlanes &w, const uint8_t *images, int T, float *out, int64_t ldo, int epi, cudaStream_t st) {
    if (!gemm_q_usable(w.type, w.k)) throw CudaError("launch_gemm_q: only Q4_K / Q8_0 weights");
    if (T < 1 || T > 32) throw CudaError("launch_gemm_q: 1..32 tokens per call");
    GemmParams p{};
    p.w = w; p.images = images; p.out = out; p.ldo = ldo; p.T = T; p.epi = epi;
    const int nt = (T + 7) / 8;
    const dim3 grid((w.n_rows + GEMM_ROWS - 1) / GEMM_ROWS);
    const size_t smem = w.type == T_Q4_K ? gemm_smem_bytes<T_Q4_K>(8 * nt) : gemm_smem_bytes<T_Q8_0>(8 * nt);
#define GEMM_LAUNCH(TY, NTV)                                                                                          \
    do {                                                                                                              \
        static PerDeviceOnce once;                                                                                    \
        once.run([] { CUDA_CHECK(cudaFuncSetAttribute(gemm_q_kernel<TY, NTV>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024)); }); \
        gemm_q_kernel<TY, NTV><<<grid, GEMM_THREADS, smem, st>>>(p);                                                  \
    } while (0)
    if (w.type == T_Q4_K) {
        if (nt == 1) GEMM_LAUNCH(T_Q4_K, 1); else if (nt == 2) GEMM_LAUNCH(T_Q4_K, 2); else if (nt == 3) GEMM_LAUNCH(T_Q4_K, 3); else GEMM_LAUNCH(T_Q4_K, 4);
    } else {
        if (nt == 1) GEMM_LAUNCH(T_Q8_0, 1); else if (nt == 2) GEMM_LAUNCH(T_Q8_0, 2); else if (nt == 3) GEMM_LAUNCH(T_Q8_0, 3); else GEMM_LAUNCH(T_Q8_0, 4);
    }
#undef GEMM_LAUNCH
    CUDA_CHECK(cudaGetLastError());
}

}  // namespace aios
