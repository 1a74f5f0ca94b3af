// gemm_q.cu — multi-token dequant GEMM on the int8 tensor cores: Y[T][N] = X[T][K] . dq(W[N][K])^T for T = 1..64
// tokens, weights streamed ONCE for all T tokens.  Two users:
//   * batched decode — T = the live sequences of the continuous batcher, one token each (BASELINE.json configs[3]);
//   * bit-exact prompt chunks — T consecutive tokens of one prompt.
//
// Same numbers as T calls of the GEMV (gemv.cu) — bit for bit.  The GEMV's arithmetic per 64-weight unit is
//   Q4_K  term = fma(d*d8, sc_lo*dot(lo32, a) + sc_hi*dot(hi32, a'), -(dmin*d8) * (m_lo*bs + m_hi*bs'))
//   Q6_K  term = (d*d8) * sum_k sc_k * (dot(q_k16, a_k) - 32*bs_k)
//   Q8_0  term = dot(q32, a) * (d*d8)                                    (unit = one 32-weight block)
// with INTEGER dots, and the f32 terms are summed in double (order-free, rounded once).  An integer dot does not care
// how it is computed, so here the sub-block dots come from `mma.sync` s8 — m16n8k32 for the 32-wide dots of Q4_K and
// Q8_0, m16n8k16 for the 16-wide dots of Q6_K: one instruction = 16 weight rows x 8 tokens x one sub-block — and the
// scale arithmetic around them is the GEMV's, per (row, token) pair.  This is why the path is exact and also why it is
// issue-bound rather than tensor-bound at T = 64: every (row, token, unit) costs ~12 scalar instructions after the MMA
// (17-20 with the fragment preparation; ncu: 60-70 % of issue slots used, tensor pipe 14 %).
// The tensor-bound path for long prompts is prefill_tc.cu (tcgen05, float arithmetic, stated tolerance).
//
// Activations: `quant_rows_kernel` builds, per token, the same Q8_K / Q8_0 image the GEMV prologue builds in shared
// memory (gemv_core.cuh quant_slab, optional RMSNorm first) — in global memory, once per launch group, instead of once
// per CTA — and then re-files it block by block (256 weights) in exactly the layout a GEMM stage has in shared memory:
// staged[b] = act[T][17] uint4 | aux[T][32] | d8[T], one contiguous record per block for all T tokens.  The GEMM brings
// a block's record in with ONE bulk copy (`cp.async.bulk`, mbarrier completion) issued by one thread, where 1,200
// 16-byte `cp.async` per stage with gathered sources used to keep the LSU busy 40 % of the time.
//
// A CTA owns 32 weight rows (two m16 row groups) of ONE segment and all T tokens: warp (rg, tg) takes row group rg and
// token group tg (8*NT tokens).  Matrices that share an input ride in one launch as segments (q/k/v — also when
// attn_v is Q6_K next to Q4_K q/k — and gate/up): the type is CTA-uniform.
#define GEMV_CORE_WARPS 8
#include "gemv_core.cuh"

namespace aios {

// ---------------------------------------------------------------- per-token activation images in global memory
// grid = T, block = 256.  x row stride ldx (floats); image stride = act_image_bytes(k) rounded to 256.
__host__ __device__ inline size_t gemm_image_stride(int k) { return (act_image_bytes(k) + 255) & ~(size_t)255; }

constexpr int GEMM_TOK_STRIDE = 17;         // uint4 per token in a staged block: 16 used + 1 pad (bank spread over tokens)
// one staged 256-weight block of all T token images
//   act [T][GEMM_TOK_STRIDE] uint4 : Q4_K / Q6_K: entry 4u+kk = activation bytes a_kk of unit u; Q8_0: entry 2u+kk
//   aux [T][32] bytes              : Q4_K: 8 x i32 32-wide q8 sums; Q6_K: 16 x i16 16-wide sums; Q8_0: 8 x f32 d8
//   d8  [T] f32                    : Q4_K / Q6_K block scale
__host__ __device__ inline size_t gemm_img_bytes(int T) { return (size_t)T * (GEMM_TOK_STRIDE * 16 + 32 + 4); }
// the staged records follow the T per-token images
__host__ __device__ inline size_t gemm_staged_offset(int k, int T) { return (gemm_image_stride(k) * (size_t)T + 127) & ~(size_t)127; }

template <int PERM>
__global__ void __launch_bounds__(GEMV_THREADS) quant_rows_kernel(const float *x, int64_t ldx, const float *norm_w, float eps,
                                                                   int k, uint8_t *images) {
    __shared__ double red[GEMV_WARPS];
    pdl_wait();
    pdl_launch_dependents();
    const int t = blockIdx.x, T = gridDim.x;
    uint8_t *img = images + (size_t)t * gemm_image_stride(k);
    const ActSmem s = carve(img, k);
    prep_activations<PERM>(x + (size_t)t * ldx, norm_w, eps, k, s, red);
    // re-file this token's image into the per-block records (this CTA's own global writes, visible after the barrier)
    __threadfence_block();
    __syncthreads();
    constexpr bool Q80 = PERM == PERM_Q80;
    const int U = Q80 ? k / 32 : k / 64, nb = k / 256;
    uint8_t *staged = images + gemm_staged_offset(k, T);
    const size_t rec = gemm_img_bytes(T);
    const uint4 *act = reinterpret_cast<const uint4 *>(img);
    const uint8_t *aux = img + k;
    for (int idx = threadIdx.x; idx < nb * 16; idx += GEMV_THREADS) {
        const int b = idx >> 4, e = idx & 15;
        const uint4 *src = Q80 ? act + (size_t)(e & 1) * U + 8 * b + (e >> 1) : act + (size_t)(e & 3) * U + 4 * b + (e >> 2);
        reinterpret_cast<uint4 *>(staged + (size_t)b * rec)[t * GEMM_TOK_STRIDE + e] = *src;
    }
    for (int idx = threadIdx.x; idx < nb * 2; idx += GEMV_THREADS) {
        const int b = idx >> 1, h = idx & 1;
        const uint4 v = Q80 ? *reinterpret_cast<const uint4 *>(aux + k / 8 + (size_t)(8 * b + h * 4) * 4)
                            : *reinterpret_cast<const uint4 *>(aux + (size_t)b * 32 + h * 16);
        *reinterpret_cast<uint4 *>(staged + (size_t)b * rec + (size_t)T * GEMM_TOK_STRIDE * 16 + (size_t)t * 32 + h * 16) = v;
    }
    if (!Q80)
        for (int b = threadIdx.x; b < nb; b += GEMV_THREADS)
            *reinterpret_cast<float *>(staged + (size_t)b * rec + (size_t)T * (GEMM_TOK_STRIDE * 16 + 32) + (size_t)t * 4) =
                *reinterpret_cast<const float *>(aux + k / 8 + (size_t)b * 4);
}

void launch_quant_rows(int weight_type, const float *x, int64_t ldx, const float *norm_w, float eps, int k, int T,
                       uint8_t *images, cudaStream_t st) {
    const dim3 grid((unsigned)T), block(GEMV_THREADS);
    if (weight_type == T_Q6_K) launch_k(quant_rows_kernel<PERM_Q6>, grid, block, 0, st, x, ldx, norm_w, eps, k, images);
    else if (weight_type == T_Q8_0) launch_k(quant_rows_kernel<PERM_Q80>, grid, block, 0, st, x, ldx, norm_w, eps, k, images);
    else launch_k(quant_rows_kernel<PERM_Q4>, grid, block, 0, st, x, ldx, norm_w, eps, k, images);
    CUDA_CHECK(cudaGetLastError());
}
size_t gemm_images_bytes(int k, int T) { return gemm_staged_offset(k, T) + (size_t)(k / 256) * gemm_img_bytes(T); }

// ---------------------------------------------------------------- the GEMM
constexpr int GEMM_MAX_ROWS = 32;           // weight rows per CTA: 16 * RG, RG = 1 or 2 m16 row groups (template parameter)
constexpr int GEMM_STAGES = 2;

__device__ __forceinline__ void mma_s8_k32(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
        : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0));
}
__device__ __forceinline__ void mma_s8_k16(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%7,%7,%7,%7};"
        : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
        : "r"(a0), "r"(a1), "r"(b0), "r"(0));
}
__device__ __forceinline__ void cp16(void *smem_dst, const void *gsrc) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp4(void *smem_dst, const void *gsrc) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gsrc) : "memory");
}

struct ImgSmem {
    uint4 *act;
    uint8_t *aux;
    float *d8;
};
__device__ __forceinline__ ImgSmem img_carve(uint8_t *base, int T) {
    ImgSmem s;
    s.act = reinterpret_cast<uint4 *>(base);
    s.aux = base + (size_t)T * GEMM_TOK_STRIDE * 16;
    s.d8 = reinterpret_cast<float *>(base + (size_t)T * (GEMM_TOK_STRIDE * 16 + 32));
    return s;
}

// ---- bulk copy of one staged record (one thread) + mbarrier completion
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "GQ_WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra GQ_WAIT_DONE;\n"
        "bra GQ_WAIT_LOOP;\n"
        "GQ_WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
// this GPU's own memory -> this CTA's shared memory (no multicast)
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Weight side of a stage: the 256-weight block b of GEMM_ROWS rows, copied plane by plane with 16-byte cp.async
// (a row's bytes of one block are contiguous inside a plane), padded per row so that the fragment reads of the eight
// row groups of a warp fall into different banks.
//   Q4_K: qa[rows][5] uint4 (4 units + pad) | qb[rows][5] | meta[rows] uint4
//   Q8_0: qa[rows][9] uint4 (8 blocks + pad) | qb[rows][9] | d[rows] uint4 (8 x f16)
//   Q6_K: la[rows][5] | lb[rows][5] | hh[rows][5] | sc[rows] uint4 (4 units x 4 i8) | d[rows] u32 (the aligned word
//         holding the f16 block scale)
template <int TYPE> struct TileA;
template <> struct TileA<T_Q4_K> { static constexpr int RS = 5; static constexpr size_t bytes = (size_t)GEMM_MAX_ROWS * (2 * RS + 1) * 16; };
// Q5_K: the Q4_K tile, then qh[rows][2] uint4 (4 units x (low word, high word) of fifth bits)
template <> struct TileA<T_Q5_K> { static constexpr int RS = 5; static constexpr size_t bytes = (size_t)GEMM_MAX_ROWS * (2 * RS + 3) * 16; };
template <> struct TileA<T_Q8_0> { static constexpr int RS = 9; static constexpr size_t bytes = (size_t)GEMM_MAX_ROWS * (2 * RS + 1) * 16; };
template <> struct TileA<T_Q6_K> { static constexpr int RS = 5; static constexpr size_t bytes = (size_t)GEMM_MAX_ROWS * (3 * RS + 1) * 16 + GEMM_MAX_ROWS * 4; };
constexpr size_t GEMM_A_BYTES = TileA<T_Q8_0>::bytes;  // the largest: one stage layout for every type
static_assert(TileA<T_Q6_K>::bytes <= GEMM_A_BYTES && TileA<T_Q4_K>::bytes <= GEMM_A_BYTES && TileA<T_Q5_K>::bytes <= GEMM_A_BYTES, "stage layout");

template <int TYPE, int GEMM_ROWS>
__device__ __forceinline__ void stage_weights(uint8_t *sa, const WPlanes &w, int row0, int b, int nthreads) {
    constexpr int RS = TileA<TYPE>::RS, UPB = TYPE == T_Q8_0 ? 8 : 4;  // units per 256-weight block
    const size_t U = (size_t)w.k / (TYPE == T_Q8_0 ? 32 : 64), nb = (size_t)w.k / 256;
    uint4 *p0 = reinterpret_cast<uint4 *>(sa), *p1 = p0 + GEMM_ROWS * RS;
    for (int idx = threadIdx.x; idx < GEMM_ROWS * UPB; idx += nthreads) {
        const int r = idx / UPB, u = idx % UPB;
        const size_t row = (size_t)min(row0 + r, w.n_rows - 1);
        cp16(p0 + r * RS + u, w.p[0] + (row * U + (size_t)UPB * b + u) * 16);
        cp16(p1 + r * RS + u, w.p[1] + (row * U + (size_t)UPB * b + u) * 16);
        if (TYPE == T_Q6_K) cp16(p1 + GEMM_ROWS * RS + r * RS + u, w.p[2] + (row * U + (size_t)UPB * b + u) * 16);
    }
    if (TYPE == T_Q6_K) {
        uint4 *sc = p1 + 2 * GEMM_ROWS * RS;
        uint32_t *dd = reinterpret_cast<uint32_t *>(sc + GEMM_ROWS);
        for (int r = threadIdx.x; r < GEMM_ROWS; r += nthreads) {
            const size_t row = (size_t)min(row0 + r, w.n_rows - 1);
            cp16(sc + r, w.p[3] + (row * U + (size_t)4 * b) * 4);   // 4 units x 4 scales
            // the f16 block scale: 2 bytes at a 2-byte-aligned address -> copy the aligned 4-byte word that holds it
            const uintptr_t dsrc = reinterpret_cast<uintptr_t>(w.p[4] + (row * nb + b) * 2);
            cp4(dd + r, reinterpret_cast<const void *>(dsrc & ~(uintptr_t)3));
        }
    } else {
        uint4 *ex = p1 + GEMM_ROWS * RS;
        for (int r = threadIdx.x; r < GEMM_ROWS; r += nthreads) {
            const size_t row = (size_t)min(row0 + r, w.n_rows - 1);
            if (TYPE == T_Q8_0) cp16(ex + r, w.p[2] + (row * U + (size_t)8 * b) * 2);   // eight f16 scales = 16 bytes
            else cp16(ex + r, w.p[2] + (row * nb + b) * 16);                            // META of the block
            if (TYPE == T_Q5_K) {  // fifth bits of the block's four units: 4 x 8 bytes
                cp16(ex + GEMM_ROWS + 2 * r, w.p[3] + (row * U + (size_t)4 * b) * 8);
                cp16(ex + GEMM_ROWS + 2 * r + 1, w.p[3] + (row * U + (size_t)4 * b) * 8 + 16);
            }
        }
    }
}

struct GemmParams {
    int32_t n_seg;           // 1..3 matrices sharing the same input tokens
    int32_t T, epi;          // epi: EPI_STORE or EPI_ADD
    int32_t tile_begin[4];   // segment i owns CTAs [tile_begin[i], tile_begin[i+1])
    WPlanes seg[3];
    float *out[3];           // [T][ldo]
    const uint8_t *images[3];  // T activation images in the permutation of the segment's type, stride gemm_image_stride(k)
    int64_t ldo;
    double *part;            // non-null (one segment): part[t * n_rows + row] = the unrounded double sum, out untouched
};

// the k-loop of one CTA for weight type TYPE
template <int TYPE, int NT, int TW, int GEMM_RG>
__device__ __forceinline__ void gemm_body(const GemmParams &p, int seg, uint8_t *smem) {
    constexpr int GEMM_ROWS = GEMM_RG * 16;
    constexpr int T = 8 * NT * TW, RS = TileA<TYPE>::RS, NTHREADS = GEMM_RG * TW * 32;
    const size_t stage_bytes = GEMM_A_BYTES + gemm_img_bytes(T);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rg = warp % GEMM_RG, tg = warp / GEMM_RG;
    const int g = lane >> 2, q = lane & 3;
    const WPlanes &W = p.seg[seg];
    const int k = W.k, N = W.n_rows;
    const int nb = k / 256;
    const int row0 = ((int)blockIdx.x - p.tile_begin[seg]) * GEMM_ROWS;
    const int ra = rg * 16 + g, rb = ra + 8;  // my two rows inside the CTA's tile
    const int tok0 = tg * 8 * NT;             // first token of my token group

    double acc[NT][4];
#pragma unroll
    for (int j = 0; j < NT; j++)
#pragma unroll
        for (int e = 0; e < 4; e++) acc[j][e] = 0.0;

    // the images' staged records (quant_rows_kernel) sit behind the T per-token images
    const uint8_t *staged = p.images[seg] + gemm_staged_offset(k, T);
    const uint32_t rec_bytes = (uint32_t)gemm_img_bytes(T);
    __shared__ __align__(8) uint64_t full[GEMM_STAGES];
    if (threadIdx.x == 0) {
        for (int i = 0; i < GEMM_STAGES; i++) mbar_init(&full[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto stage_w = [&](int b) {
        stage_weights<TYPE, GEMM_ROWS>(smem + (size_t)(b % GEMM_STAGES) * stage_bytes, W, row0, b, NTHREADS);
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    auto stage_x = [&](int b) {
        if (threadIdx.x == 0) {  // the block's activation record: one bulk copy
            uint8_t *base = smem + (size_t)(b % GEMM_STAGES) * stage_bytes;
            mbar_arrive_expect_tx(&full[b % GEMM_STAGES], rec_bytes);
            bulk_g2s(base + GEMM_A_BYTES, staged + (size_t)b * rec_bytes, rec_bytes, &full[b % GEMM_STAGES]);
        }
    };
    auto stage = [&](int b) { stage_w(b); stage_x(b); };
    // the first weights do not depend on the launch before this one (programmatic dependent launch): they are on
    // their way while that launch finishes; the activation records it writes are asked for after the wait
    stage_w(0);
    pdl_wait();
    pdl_launch_dependents();
    stage_x(0);
    for (int b = 0; b < nb; b++) {
        if (b + 1 < nb) {  // the next block's copies fly while this one is multiplied
            stage(b + 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        mbar_wait(&full[b % GEMM_STAGES], (uint32_t)(b / GEMM_STAGES) & 1u);
        __syncthreads();
        const uint8_t *base = smem + (size_t)(b % GEMM_STAGES) * stage_bytes;
        const uint4 *p0 = reinterpret_cast<const uint4 *>(base), *p1 = p0 + GEMM_ROWS * RS;
        const ImgSmem s = img_carve(const_cast<uint8_t *>(base) + GEMM_A_BYTES, T);
        if (TYPE == T_Q8_0) {
            const uint4 *ex = p1 + GEMM_ROWS * RS;
            const __half *dha = reinterpret_cast<const __half *>(ex + ra), *dhb = reinterpret_cast<const __half *>(ex + rb);
#pragma unroll 2
            for (int u = 0; u < 8; u++) {
                const uint32_t a0 = reinterpret_cast<const uint32_t *>(p0 + ra * RS + u)[q];
                const uint32_t a1 = reinterpret_cast<const uint32_t *>(p0 + rb * RS + u)[q];
                const uint32_t a2 = reinterpret_cast<const uint32_t *>(p1 + ra * RS + u)[q];
                const uint32_t a3 = reinterpret_cast<const uint32_t *>(p1 + rb * RS + u)[q];
                const float dwa = __half2float(dha[u]), dwb = __half2float(dhb[u]);
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    const int tb = tok0 + 8 * j;
                    const uint32_t *bt = reinterpret_cast<const uint32_t *>(s.act + (tb + g) * GEMM_TOK_STRIDE + u * 2);
                    int c[4];
                    mma_s8_k32(c, a0, a1, a2, a3, bt[q], bt[4 + q]);
                    const float d0 = reinterpret_cast<const float *>(s.aux + (tb + 2 * q) * 32)[u];
                    const float d1 = reinterpret_cast<const float *>(s.aux + (tb + 2 * q + 1) * 32)[u];
                    acc[j][0] += (double)__fmul_rn((float)c[0], __fmul_rn(dwa, d0));
                    acc[j][1] += (double)__fmul_rn((float)c[1], __fmul_rn(dwa, d1));
                    acc[j][2] += (double)__fmul_rn((float)c[2], __fmul_rn(dwb, d0));
                    acc[j][3] += (double)__fmul_rn((float)c[3], __fmul_rn(dwb, d1));
                }
            }
        } else if (TYPE == T_Q4_K || TYPE == T_Q5_K) {
            const uint4 *ex = p1 + GEMM_ROWS * RS;
            const uint32_t *qha = reinterpret_cast<const uint32_t *>(ex + GEMM_ROWS + 2 * ra), *qhb = reinterpret_cast<const uint32_t *>(ex + GEMM_ROWS + 2 * rb);
            const uint4 mta = ex[ra], mtb = ex[rb];
            const uint32_t dda = __byte_perm(__byte_perm(mta.x, mta.y, 0x0073), __byte_perm(mta.z, mta.w, 0x0073), 0x5410);
            const uint32_t ddb = __byte_perm(__byte_perm(mtb.x, mtb.y, 0x0073), __byte_perm(mtb.z, mtb.w, 0x0073), 0x5410);
            const float2 da = __half22float2(*reinterpret_cast<const __half2 *>(&dda));  // (d, dmin) of row a
            const float2 db = __half22float2(*reinterpret_cast<const __half2 *>(&ddb));
            const uint32_t mwa[4] = {mta.x, mta.y, mta.z, mta.w}, mwb[4] = {mtb.x, mtb.y, mtb.z, mtb.w};
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const uint32_t qa0 = reinterpret_cast<const uint32_t *>(p0 + ra * RS + u)[q];
                const uint32_t qa1 = reinterpret_cast<const uint32_t *>(p0 + rb * RS + u)[q];
                const uint32_t qb0 = reinterpret_cast<const uint32_t *>(p1 + ra * RS + u)[q];
                const uint32_t qb1 = reinterpret_cast<const uint32_t *>(p1 + rb * RS + u)[q];
                const int sa_lo = mwa[u] & 63, sa_hi = (mwa[u] >> 6) & 63, ma_lo = (mwa[u] >> 12) & 63, ma_hi = (mwa[u] >> 18) & 63;
                const int sb_lo = mwb[u] & 63, sb_hi = (mwb[u] >> 6) & 63, mb_lo = (mwb[u] >> 12) & 63, mb_hi = (mwb[u] >> 18) & 63;
                // A fragments: low nibbles = sub-block 2u (k 0-31 of the unit), high nibbles = sub-block 2u+1
                uint32_t al0 = qa0 & 0x0F0F0F0Fu, al1 = qa1 & 0x0F0F0F0Fu, al2 = qb0 & 0x0F0F0F0Fu, al3 = qb1 & 0x0F0F0F0Fu;
                uint32_t ah0 = (qa0 >> 4) & 0x0F0F0F0Fu, ah1 = (qa1 >> 4) & 0x0F0F0F0Fu, ah2 = (qb0 >> 4) & 0x0F0F0F0Fu, ah3 = (qb1 >> 4) & 0x0F0F0F0Fu;
                if (TYPE == T_Q5_K) {  // the fifth bit of element l of a sub-block is bit l of its QH5 word: my bytes are l = 4q.. and 16+4q..
                    auto sp = [](uint32_t x4) { return ((x4 * 0x00204081u) & 0x01010101u) << 4; };  // bit j -> byte j, value 16
                    const uint32_t ea = qha[2 * u], fa = qha[2 * u + 1], eb = qhb[2 * u], fb = qhb[2 * u + 1];
                    al0 |= sp((ea >> (4 * q)) & 0xFu); al1 |= sp((eb >> (4 * q)) & 0xFu);
                    al2 |= sp((ea >> (16 + 4 * q)) & 0xFu); al3 |= sp((eb >> (16 + 4 * q)) & 0xFu);
                    ah0 |= sp((fa >> (4 * q)) & 0xFu); ah1 |= sp((fb >> (4 * q)) & 0xFu);
                    ah2 |= sp((fa >> (16 + 4 * q)) & 0xFu); ah3 |= sp((fb >> (16 + 4 * q)) & 0xFu);
                }
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    const int tb = tok0 + 8 * j;
                    const uint32_t *bt = reinterpret_cast<const uint32_t *>(s.act + (tb + g) * GEMM_TOK_STRIDE + u * 4);
                    int lo[4], hi[4];
                    mma_s8_k32(lo, al0, al1, al2, al3, bt[q], bt[4 + q]);
                    mma_s8_k32(hi, ah0, ah1, ah2, ah3, bt[8 + q], bt[12 + q]);
#pragma unroll
                    for (int e = 0; e < 2; e++) {  // my two tokens of this n-block
                        const int t = tb + 2 * q + e;
                        const int2 bs = *reinterpret_cast<const int2 *>(s.aux + t * 32 + u * 8);
                        const float d8 = s.d8[t];
                        const int isa = sa_lo * lo[e] + sa_hi * hi[e], msa = ma_lo * bs.x + ma_hi * bs.y;
                        const int isb = sb_lo * lo[2 + e] + sb_hi * hi[2 + e], msb = mb_lo * bs.x + mb_hi * bs.y;
                        acc[j][e] += (double)__fmaf_rn(__fmul_rn(da.x, d8), (float)isa, -__fmul_rn(__fmul_rn(da.y, d8), (float)msa));
                        acc[j][2 + e] += (double)__fmaf_rn(__fmul_rn(db.x, d8), (float)isb, -__fmul_rn(__fmul_rn(db.y, d8), (float)msb));
                    }
                }
            }
        } else {  // Q6_K
            const uint4 *p2 = p1 + GEMM_ROWS * RS;
            const uint4 *sc = p2 + GEMM_ROWS * RS;
            const uint32_t *dd = reinterpret_cast<const uint32_t *>(sc + GEMM_ROWS);
            // the block scale sits in the low or the high half of its aligned word
            const uintptr_t da_addr = reinterpret_cast<uintptr_t>(W.p[4]) + ((size_t)min(row0 + ra, N - 1) * nb + b) * 2;
            const uintptr_t db_addr = reinterpret_cast<uintptr_t>(W.p[4]) + ((size_t)min(row0 + rb, N - 1) * nb + b) * 2;
            const float dwa = h2f_bits((uint16_t)(dd[ra] >> (8 * (da_addr & 2))));
            const float dwb = h2f_bits((uint16_t)(dd[rb] >> (8 * (db_addr & 2))));
            const uint32_t *sca = reinterpret_cast<const uint32_t *>(sc + ra), *scb = reinterpret_cast<const uint32_t *>(sc + rb);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const uint32_t la_a = reinterpret_cast<const uint32_t *>(p0 + ra * RS + u)[q], la_b = reinterpret_cast<const uint32_t *>(p0 + rb * RS + u)[q];
                const uint32_t lb_a = reinterpret_cast<const uint32_t *>(p1 + ra * RS + u)[q], lb_b = reinterpret_cast<const uint32_t *>(p1 + rb * RS + u)[q];
                const uint32_t hh_a = reinterpret_cast<const uint32_t *>(p2 + ra * RS + u)[q], hh_b = reinterpret_cast<const uint32_t *>(p2 + rb * RS + u)[q];
                uint32_t qa[4], qb[4];  // the four 16-weight sub-blocks of the unit (6-bit values, 0..63)
                qa[0] = (la_a & 0x0F0F0F0Fu) | ((hh_a << 4) & 0x30303030u);
                qa[1] = (lb_a & 0x0F0F0F0Fu) | ((hh_a << 2) & 0x30303030u);
                qa[2] = ((la_a >> 4) & 0x0F0F0F0Fu) | (hh_a & 0x30303030u);
                qa[3] = ((lb_a >> 4) & 0x0F0F0F0Fu) | ((hh_a >> 2) & 0x30303030u);
                qb[0] = (la_b & 0x0F0F0F0Fu) | ((hh_b << 4) & 0x30303030u);
                qb[1] = (lb_b & 0x0F0F0F0Fu) | ((hh_b << 2) & 0x30303030u);
                qb[2] = ((la_b >> 4) & 0x0F0F0F0Fu) | (hh_b & 0x30303030u);
                qb[3] = ((lb_b >> 4) & 0x0F0F0F0Fu) | ((hh_b >> 2) & 0x30303030u);
                const uint32_t swa = sca[u], swb = scb[u];
                int sa[4], sb[4];
#pragma unroll
                for (int kk = 0; kk < 4; kk++) {
                    sa[kk] = (int)(int8_t)((swa >> (8 * kk)) & 0xff);
                    sb[kk] = (int)(int8_t)((swb >> (8 * kk)) & 0xff);
                }
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    const int tb = tok0 + 8 * j;
                    const uint32_t *bt = reinterpret_cast<const uint32_t *>(s.act + (tb + g) * GEMM_TOK_STRIDE + u * 4);
                    int dk[4][4];
#pragma unroll
                    for (int kk = 0; kk < 4; kk++) mma_s8_k16(dk[kk], qa[kk], qb[kk], bt[4 * kk + q]);
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        const int t = tb + 2 * q + e;
                        const uint2 bsp = *reinterpret_cast<const uint2 *>(s.aux + t * 32 + u * 8);  // four i16 16-wide sums
                        const int b0 = (int)(int16_t)(bsp.x & 0xffff), b1 = (int)(int16_t)(bsp.x >> 16);
                        const int b2 = (int)(int16_t)(bsp.y & 0xffff), b3 = (int)(int16_t)(bsp.y >> 16);
                        const float d8 = s.d8[t];
                        const int isa = sa[0] * (dk[0][e] - 32 * b0) + sa[1] * (dk[1][e] - 32 * b1) + sa[2] * (dk[2][e] - 32 * b2) + sa[3] * (dk[3][e] - 32 * b3);
                        const int isb = sb[0] * (dk[0][2 + e] - 32 * b0) + sb[1] * (dk[1][2 + e] - 32 * b1) + sb[2] * (dk[2][2 + e] - 32 * b2) + sb[3] * (dk[3][2 + e] - 32 * b3);
                        acc[j][e] += (double)__fmul_rn(__fmul_rn(dwa, d8), (float)isa);
                        acc[j][2 + e] += (double)__fmul_rn(__fmul_rn(dwb, d8), (float)isb);
                    }
                }
            }
        }
        __syncthreads();  // everyone is done with this stage before it is refilled
    }
    const int row_a = row0 + ra, row_b = row0 + rb;
    float *out = p.out[seg];
#pragma unroll
    for (int j = 0; j < NT; j++)
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const int row = (e & 2) ? row_b : row_a, t = tok0 + 8 * j + 2 * q + (e & 1);
            if (row < N && t < p.T) {
                if (p.part) {
                    p.part[(size_t)t * N + row] = acc[j][e];
                    continue;
                }
                float *o = out + (size_t)t * p.ldo + row;
                const float v = (float)acc[j][e];
                *o = p.epi == EPI_ADD ? __fadd_rn(*o, v) : v;
            }
        }
}

template <int NT, int TW, int GEMM_RG>
__global__ void __launch_bounds__(GEMM_RG * TW * 32) gemm_q_kernel(const GemmParams p) {
    extern __shared__ __align__(16) uint8_t smem[];
    int seg = 0;
    while (seg + 1 < p.n_seg && (int)blockIdx.x >= p.tile_begin[seg + 1]) seg++;
    const int type = p.seg[seg].type;  // CTA-uniform
    if (type == T_Q4_K) gemm_body<T_Q4_K, NT, TW, GEMM_RG>(p, seg, smem);
    else if (type == T_Q5_K) gemm_body<T_Q5_K, NT, TW, GEMM_RG>(p, seg, smem);
    else if (type == T_Q6_K) gemm_body<T_Q6_K, NT, TW, GEMM_RG>(p, seg, smem);
    else gemm_body<T_Q8_0, NT, TW, GEMM_RG>(p, seg, smem);
}

static size_t gemm_smem_bytes(int T) { return GEMM_STAGES * (GEMM_A_BYTES + gemm_img_bytes(T)); }

// h[t][i] = silu(gate[t][i]) * up[t][i] for a chunk: gu[t] = gate[ff] | up[ff].  Same expression, operation by
// operation, as the GEMV's fused SwiGLU epilogue.
__global__ void swiglu_rows_kernel(const float *gu, int64_t ldgu, int ff, int T, float *h, int64_t ldh) {
    pdl_wait();
    pdl_launch_dependents();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)T * ff) return;
    const int t = (int)(i / ff), c = (int)(i % ff);
    const float gte = gu[(size_t)t * ldgu + c], up = gu[(size_t)t * ldgu + ff + c];
    h[(size_t)t * ldh + c] = __fmul_rn(__fdiv_rn(gte, __fadd_rn(1.0f, expf_det(-gte))), up);
}
void launch_swiglu_rows(const float *gu, int64_t ldgu, int ff, int T, float *h, int64_t ldh, cudaStream_t st) {
    const int64_t n = (int64_t)T * ff;
    launch_k(swiglu_rows_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st, gu, ldgu, ff, T, h, ldh);
    CUDA_CHECK(cudaGetLastError());
}

__global__ void add_part_rows_kernel(float *x, int64_t ldx, const double *part, int n) {
    const int t = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        float *o = x + (size_t)t * ldx + i;
        *o = __fadd_rn(*o, (float)part[(size_t)t * n + i]);
    }
}
void launch_add_part_rows(float *x, int64_t ldx, const double *part, int n, int T, cudaStream_t st) {
    add_part_rows_kernel<<<dim3((unsigned)((n + 255) / 256), (unsigned)T), 256, 0, st>>>(x, ldx, part, n);
    CUDA_CHECK(cudaGetLastError());
}

bool gemm_q_usable(int type, int k) { return (type == T_Q4_K || type == T_Q5_K || type == T_Q6_K || type == T_Q8_0) && k % 256 == 0; }

// images of tokens T..pad(T)-1 (padding) must exist (any finite contents); their outputs are not stored.
int gemm_pad_tokens(int T) { return T <= 32 ? (T + 7) / 8 * 8 : GEMM_MAX_TOKENS; }

static int n_sms_hint() {
    int dev = 0, n = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    return n;
}

void launch_gemm_multi(const GemmMulti &m, cudaStream_t st) {
    if (m.T < 1 || m.T > GEMM_MAX_TOKENS) throw CudaError("launch_gemm: 1..64 tokens per call");
    if (m.n_seg < 1 || m.n_seg > 3) throw CudaError("launch_gemm: 1..3 segments");
    GemmParams p{};
    p.n_seg = m.n_seg; p.T = m.T; p.epi = m.epi; p.ldo = m.ldo;
    if (m.part && m.n_seg != 1) throw CudaError("launch_gemm: partial sums need a single segment");
    p.part = m.part;
    int64_t total_rows = 0;
    for (int i = 0; i < m.n_seg; i++) total_rows += m.seg[i].n_rows;
    // 32-row CTAs when that still gives the machine >= 2 CTAs per SM, else 16-row CTAs (twice as many)
    const int rg = (total_rows / 32 >= 2 * n_sms_hint()) ? 2 : 1;
    const int rows_per_cta = 16 * rg;
    int tiles = 0;
    for (int i = 0; i < m.n_seg; i++) {
        if (!gemm_q_usable(m.seg[i].type, m.seg[i].k)) throw CudaError("launch_gemm: Q4_K / Q6_K / Q8_0 weights only");
        if (m.seg[i].k != m.seg[0].k) throw CudaError("launch_gemm: segments must share K");
        p.seg[i] = m.seg[i]; p.out[i] = m.out[i]; p.images[i] = m.images[i];
        p.tile_begin[i] = tiles;
        tiles += (m.seg[i].n_rows + rows_per_cta - 1) / rows_per_cta;
    }
    for (int i = m.n_seg; i < 4; i++) p.tile_begin[i] = tiles;
    const int Tp = gemm_pad_tokens(m.T);
    const size_t smem = gemm_smem_bytes(Tp);
#define GEMM_LAUNCH(NTV, TWV, RGV)                                                                                    \
    do {                                                                                                              \
        static PerDeviceOnce once;                                                                                    \
        once.run([] { CUDA_CHECK(cudaFuncSetAttribute(gemm_q_kernel<NTV, TWV, RGV>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024)); }); \
        launch_k(gemm_q_kernel<NTV, TWV, RGV>, dim3((unsigned)tiles), dim3(RGV * TWV * 32), smem, st, p);              \
    } while (0)
#define GEMM_LAUNCH_RG(NTV, TWV) do { if (rg == 2) GEMM_LAUNCH(NTV, TWV, 2); else GEMM_LAUNCH(NTV, TWV, 1); } while (0)
    // tokens per warp = 8 * NT; token groups per CTA = TW: more, smaller warps keep more of them in flight per SM
    switch (Tp) {
        case 8: GEMM_LAUNCH_RG(1, 1); break;
        case 16: GEMM_LAUNCH_RG(1, 2); break;
        case 24: GEMM_LAUNCH_RG(1, 3); break;
        case 32: GEMM_LAUNCH_RG(2, 2); break;
        default: GEMM_LAUNCH_RG(2, 4); break;
    }
#undef GEMM_LAUNCH_RG
#undef GEMM_LAUNCH
    CUDA_CHECK(cudaGetLastError());
}

void launch_gemm_q(const WPlanes &w, const uint8_t *images, int T, float *out, int64_t ldo, int epi, cudaStream_t st) {
    GemmMulti m{};
    m.n_seg = 1; m.T = T; m.epi = epi; m.ldo = ldo;
    m.seg[0] = w; m.out[0] = out; m.images[0] = images;
    launch_gemm_multi(m, st);
}

}  // namespace aios
