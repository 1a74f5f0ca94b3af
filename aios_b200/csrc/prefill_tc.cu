// prefill_tc.cu — the prefill GEMM on the 5th-generation tensor cores (BASELINE.json configs[2]):
//   Y[T][N] (=|+=) X[T][K] . dq(W[N][K])^T     for a whole prompt (T up to thousands of tokens)
// `tcgen05.mma` (kind::f16, f16 operands, f32 accumulation) with the accumulators in TENSOR MEMORY, the activation
// tiles staged by TMA (`cp.async.bulk.tensor`, a tensor map over this GPU's own activation buffer), and the quantised
// weights DEQUANTISED INTO SHARED MEMORY by four producer warps — GGUF bytes stay compressed all the way from HBM to
// the SM; only the 128 x 64 tile that the tensor core is about to read exists as f16.
//
// Numerics (stated, not bit-exact): weights are rounded to f16 after dequantisation (relative error 2^-12), activations
// to f16, products are exact, accumulation is f32 over the whole K in TMEM.  This is NOT the llama.cpp CPU arithmetic
// the decode path reproduces bit for bit (activations there are Q8-quantised, which is ~10x noisier than f16): the
// tolerance against an exact f64 product of the dequantised weights is 4e-3 (max) / 8e-4 (rms) of the output's rms
// (tests/test_gpu_prefill_tc.py).  The engine takes this path for long prompts; the bit-exact chunked path
// (gemm_q.cu) stays available (AIOS_ENGINE_EXACT_PREFILL) and is what short prompts use.
//
// Kernel shape: persistent, one CTA per SM, warp-specialised —
//   warp 0      TMA producer: activation tile B[256 tokens][64 k] f16 per stage (128-byte swizzle)
//   warp 1      MMA issuer: one elected thread, 4 x tcgen05.mma (M 128, N 256, K 16) per stage
//   warp 2      TMEM allocation (512 columns = two 128 x 256 f32 accumulators: epilogue of tile i overlaps tile i+1)
//   warps 4-7   weight producers: thread r dequantises row r's 64 weights of the stage into the swizzled A tile
//   warps 8-11  epilogue: tcgen05.ld accumulator -> registers -> out[t][row] (store or residual add), coalesced over rows
// Stages: 4 x (16 KB A + 32 KB B) = 192 KB of shared memory, mbarrier full/empty pairs; tcgen05.commit frees a stage.
// Tiles are walked token-tile fastest, so CTAs that run together share their weight rows (read from HBM once, then L2).
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "kernels.h"
#include "layout.cuh"

namespace aios {

constexpr int TC_BM = 128, TC_BN = 256, TC_BK = 64, TC_STAGES = 4;
constexpr int TC_A_BYTES = TC_BM * TC_BK * 2, TC_B_BYTES = TC_BN * TC_BK * 2, TC_STAGE_BYTES = TC_A_BYTES + TC_B_BYTES;
constexpr int TC_THREADS = 384;
constexpr int TC_SMEM_BYTES = TC_STAGES * TC_STAGE_BYTES + 1024 /* alignment slack */ + 256 /* barriers */;

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// TMA: 2-D tile of the activation matrix (tensor map over THIS GPU's memory; no multicast) -> shared memory
__device__ __forceinline__ void tma_load_2d(void *smem_dst, const CUtensorMap *map, int x, int y, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y)
        : "memory");
}

// shared-memory matrix descriptor: K-major tile of 128-byte rows, 128-byte swizzle, 8-row atoms 1024 bytes apart
__device__ __forceinline__ uint64_t umma_desc_sw128(const void *smem_tile) {
    const uint32_t addr = smem_u32(smem_tile);
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3FFF);        // start address
    d |= (uint64_t)1 << 16;                       // leading byte offset (unused for swizzled K-major; 1 like CUTLASS)
    d |= (uint64_t)(1024 >> 4) << 32;             // stride byte offset: next 8-row atom
    d |= (uint64_t)1 << 46;                       // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;                       // SWIZZLE_128B
    return d;
}
// instruction descriptor, kind::f16: D f32, A and B f16, both K-major, dense
__device__ __forceinline__ constexpr uint32_t umma_idesc_f16(int M, int N) {
    return (1u << 4) | (0u << 7) | (0u << 10) | (0u << 15) | (0u << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"((uint32_t)accumulate)
        : "memory");
}
// arrive on an mbarrier when every MMA issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---------------------------------------------------------------- dequantisation into the swizzled A tile
// Row r of the stage's A tile is 128 bytes = 8 chunks of 8 halves; logical chunk c lives at physical chunk c ^ (r & 7)
// inside the row (the 128-byte swizzle tcgen05 and TMA use: address bits [4,7) ^= bits [7,10)).
__device__ __forceinline__ void store_chunk(uint8_t *a_tile, int r, int c, uint4 v) {
    *reinterpret_cast<uint4 *>(a_tile + r * 128 + ((c ^ (r & 7)) << 4)) = v;
}
// two bytes of `w` (byte indices 2i, 2i+1) -> half2 (1024 + b_lo, 1024 + b_hi)
__device__ __forceinline__ uint32_t bytes_to_h2_1024(uint32_t w, int pair) {
    return __byte_perm(w, 0x64646464u, pair ? 0x4342 : 0x4140);
}
__device__ __forceinline__ uint32_t h2_sub(uint32_t a, uint32_t b) {
    const __half2 r = __hsub2(*reinterpret_cast<const __half2 *>(&a), *reinterpret_cast<const __half2 *>(&b));
    return *reinterpret_cast<const uint32_t *>(&r);
}
__device__ __forceinline__ uint32_t h2_mul(uint32_t a, __half2 s) {
    const __half2 r = __hmul2(*reinterpret_cast<const __half2 *>(&a), s);
    return *reinterpret_cast<const uint32_t *>(&r);
}
__device__ __forceinline__ uint32_t h2_fma(uint32_t a, __half2 s, __half2 c) {
    const __half2 r = __hfma2(*reinterpret_cast<const __half2 *>(&a), s, c);
    return *reinterpret_cast<const uint32_t *>(&r);
}

// 16 unsigned bytes (values < 1024) -> two chunks of 8 halves: h = (b - offset) * scale [+ add]
template <bool FMA>
__device__ __forceinline__ void bytes16_to_chunks(const uint4 &q, uint32_t offset_h2, __half2 scale, __half2 add, uint4 &c0, uint4 &c1) {
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
    uint32_t o[8];
#pragma unroll
    for (int i = 0; i < 4; i++) {
#pragma unroll
        for (int pr = 0; pr < 2; pr++) {
            const uint32_t h = h2_sub(bytes_to_h2_1024(w[i], pr), offset_h2);  // exact small integers
            o[2 * i + pr] = FMA ? h2_fma(h, scale, add) : h2_mul(h, scale);
        }
    }
    c0 = make_uint4(o[0], o[1], o[2], o[3]);
    c1 = make_uint4(o[4], o[5], o[6], o[7]);
}

// The quantised bytes of one (row, 64-weight stage): requested one stage AHEAD of their use, so that the L2 / HBM
// latency of stage kb+1 hides behind the conversion of stage kb (the loads do not depend on the stage's smem slot).
struct RawStage {
    uint4 v[6];
    uint32_t s[2];
    unsigned short d;
};
template <int TYPE>
__device__ __forceinline__ void load_raw_stage(const WPlanes &w, int64_t row, int kb, RawStage &R) {
    if (TYPE == T_Q8_0) {
        const int64_t u = row * (w.k / 32) + 2 * kb;
        R.v[0] = __ldg(reinterpret_cast<const uint4 *>(w.p[0] + u * 16));
        R.v[1] = __ldg(reinterpret_cast<const uint4 *>(w.p[1] + u * 16));
        R.v[2] = __ldg(reinterpret_cast<const uint4 *>(w.p[0] + (u + 1) * 16));
        R.v[3] = __ldg(reinterpret_cast<const uint4 *>(w.p[1] + (u + 1) * 16));
        R.s[0] = __ldg(reinterpret_cast<const uint32_t *>(w.p[2] + u * 2));  // two f16 scales (u is even: 4-byte aligned)
    } else if (TYPE == T_Q4_K || TYPE == T_Q5_K) {
        const int64_t u = row * (w.k / 64) + kb;
        R.v[0] = __ldg(reinterpret_cast<const uint4 *>(w.p[0] + u * 16));
        R.v[1] = __ldg(reinterpret_cast<const uint4 *>(w.p[1] + u * 16));
        R.v[2] = __ldg(reinterpret_cast<const uint4 *>(w.p[2] + (row * (w.k / 256) + (kb >> 2)) * 16));
        if (TYPE == T_Q5_K) {  // the fifth bits: 32 of the low sub-block, 32 of the high one
            const uint2 h = __ldg(reinterpret_cast<const uint2 *>(w.p[3] + u * 8));
            R.s[0] = h.x; R.s[1] = h.y;
        }
    } else {
        const int b = kb >> 2, hh = (kb >> 1) & 1;
        const int64_t u = row * (w.k / 64) + 4 * b + 2 * hh;
#pragma unroll
        for (int t = 0; t < 2; t++) {
            R.v[3 * t + 0] = __ldg(reinterpret_cast<const uint4 *>(w.p[0] + (u + t) * 16));
            R.v[3 * t + 1] = __ldg(reinterpret_cast<const uint4 *>(w.p[1] + (u + t) * 16));
            R.v[3 * t + 2] = __ldg(reinterpret_cast<const uint4 *>(w.p[2] + (u + t) * 16));
            R.s[t] = __ldg(reinterpret_cast<const uint32_t *>(w.p[3] + (u + t) * 4));
        }
        R.d = __ldg(reinterpret_cast<const unsigned short *>(w.p[4] + (row * (w.k / 256) + b) * 2));
    }
}

// one thread = one weight row: dequantise the 64 weights [64*kb, 64*kb+64) of `row` into row r of the A tile
template <int TYPE>
__device__ __forceinline__ void dequant_row_stage(const RawStage &R, int kb, uint8_t *a_tile, int r) {
    const uint32_t H1024 = 0x64006400u;  // half2(1024, 1024)
    if (TYPE == T_Q8_0) {
#pragma unroll
        for (int blk = 0; blk < 2; blk++) {
            uint4 qa = R.v[2 * blk], qb = R.v[2 * blk + 1];
            const __half d = __ushort_as_half((unsigned short)(R.s[0] >> (16 * blk)));
            const __half2 s = __half2half2(d), z = __half2half2(__ushort_as_half((unsigned short)0));
            // signed bytes: b ^ 0x80 is the value + 128 as an unsigned byte -> subtract 1024 + 128
            qa.x ^= 0x80808080u; qa.y ^= 0x80808080u; qa.z ^= 0x80808080u; qa.w ^= 0x80808080u;
            qb.x ^= 0x80808080u; qb.y ^= 0x80808080u; qb.z ^= 0x80808080u; qb.w ^= 0x80808080u;
            uint4 c0, c1, c2, c3;
            bytes16_to_chunks<false>(qa, 0x64806480u, s, z, c0, c1);  // half2(1152, 1152)
            bytes16_to_chunks<false>(qb, 0x64806480u, s, z, c2, c3);
            store_chunk(a_tile, r, 4 * blk + 0, c0); store_chunk(a_tile, r, 4 * blk + 1, c1);
            store_chunk(a_tile, r, 4 * blk + 2, c2); store_chunk(a_tile, r, 4 * blk + 3, c3);
        }
    } else if (TYPE == T_Q4_K || TYPE == T_Q5_K) {
        // unit kb of the row: k 0-15 QSa low nibbles, 16-31 QSb low, 32-47 QSa high, 48-63 QSb high (layout.cuh);
        // Q5_K: a fifth bit per weight from the QH5 words (bit l = element l of the sub-block)
        const uint4 qa = R.v[0], qb = R.v[1], mt = R.v[2];
        const int pq = kb & 3;
        const uint32_t mw = (pq & 2) ? ((pq & 1) ? mt.w : mt.z) : ((pq & 1) ? mt.y : mt.x);
        const uint32_t ddm = __byte_perm(__byte_perm(mt.x, mt.y, 0x0073), __byte_perm(mt.z, mt.w, 0x0073), 0x5410);
        const float2 dd = __half22float2(*reinterpret_cast<const __half2 *>(&ddm));  // (d, dmin)
        // w = (d*sc)*q - dmin*m per 32-weight sub-block: scale and offset stay f32 (q - dmin*m cancels, so an f16 scale
        // would cost ~10x the error), one rounding to f16 at the end
        const float s_lo = __fmul_rn(dd.x, (float)(mw & 63)), s_hi = __fmul_rn(dd.x, (float)((mw >> 6) & 63));
        const float m_lo = -__fmul_rn(dd.y, (float)((mw >> 12) & 63)), m_hi = -__fmul_rn(dd.y, (float)((mw >> 18) & 63));
        auto deq16 = [&](const uint4 &q, int hi, float sc, float mn, int c, uint32_t bits16) {  // 16 nibbles -> chunks c, c+1
            const uint32_t wds[4] = {q.x, q.y, q.z, q.w};
            uint32_t o[8];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                uint32_t n4 = hi ? (wds[i] >> 4) & 0x0F0F0F0Fu : wds[i] & 0x0F0F0F0Fu;
                if (TYPE == T_Q5_K) n4 |= ((((bits16 >> (4 * i)) & 0xFu) * 0x00204081u) & 0x01010101u) << 4;  // bit j -> byte j, value 16
#pragma unroll
                for (int pr = 0; pr < 2; pr++) {
                    const uint32_t h = h2_sub(bytes_to_h2_1024(n4, pr), H1024);  // exact 0..15
                    const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&h));
                    const __half2 r2 = __floats2half2_rn(__fmaf_rn(f.x, sc, mn), __fmaf_rn(f.y, sc, mn));
                    o[2 * i + pr] = *reinterpret_cast<const uint32_t *>(&r2);
                }
            }
            store_chunk(a_tile, r, c, make_uint4(o[0], o[1], o[2], o[3]));
            store_chunk(a_tile, r, c + 1, make_uint4(o[4], o[5], o[6], o[7]));
        };
        deq16(qa, 0, s_lo, m_lo, 0, R.s[0] & 0xFFFFu);
        deq16(qb, 0, s_lo, m_lo, 2, R.s[0] >> 16);
        deq16(qa, 1, s_hi, m_hi, 4, R.s[1] & 0xFFFFu);
        deq16(qb, 1, s_hi, m_hi, 6, R.s[1] >> 16);
    } else {  // Q6_K
        // 64 weights [64*kb, +64) = half hh = (kb >> 1) & 1 of block b = kb >> 2, sub-blocks k0 = 2*(kb & 1) and k0 + 1 of
        // BOTH units t = 0, 1: order in K is (t0,k0) (t1,k0) (t0,k0+1) (t1,k0+1), 16 weights each
        const int k0 = 2 * (kb & 1);
        const __half d = __ushort_as_half(R.d);
        const float df = __half2float(d);
#pragma unroll
        for (int t = 0; t < 2; t++) {
            const uint4 la = R.v[3 * t], lb = R.v[3 * t + 1], qh = R.v[3 * t + 2];
            const uint32_t scw = R.s[t];
            const uint32_t LA[4] = {la.x, la.y, la.z, la.w}, LB[4] = {lb.x, lb.y, lb.z, lb.w}, QH[4] = {qh.x, qh.y, qh.z, qh.w};
            uint4 v0, v1;  // 6-bit values of sub-blocks k0 and k0 + 1
            uint32_t *p0 = &v0.x, *p1 = &v1.x;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                if (k0 == 0) {
                    p0[i] = (LA[i] & 0x0F0F0F0Fu) | ((QH[i] << 4) & 0x30303030u);
                    p1[i] = (LB[i] & 0x0F0F0F0Fu) | ((QH[i] << 2) & 0x30303030u);
                } else {
                    p0[i] = ((LA[i] >> 4) & 0x0F0F0F0Fu) | (QH[i] & 0x30303030u);
                    p1[i] = ((LB[i] >> 4) & 0x0F0F0F0Fu) | ((QH[i] >> 2) & 0x30303030u);
                }
            }
            const __half2 s0 = __float2half2_rn(__fmul_rn(df, (float)(int)(int8_t)((scw >> (8 * k0)) & 0xff)));
            const __half2 s1 = __float2half2_rn(__fmul_rn(df, (float)(int)(int8_t)((scw >> (8 * (k0 + 1))) & 0xff)));
            const __half2 z = __half2half2(__ushort_as_half((unsigned short)0));
            uint4 c0, c1;
            bytes16_to_chunks<false>(v0, 0x64206420u, s0, z, c0, c1);  // half2(1056, 1056): q - 32
            store_chunk(a_tile, r, 2 * t + 0, c0); store_chunk(a_tile, r, 2 * t + 1, c1);
            bytes16_to_chunks<false>(v1, 0x64206420u, s1, z, c0, c1);
            store_chunk(a_tile, r, 4 + 2 * t + 0, c0); store_chunk(a_tile, r, 4 + 2 * t + 1, c1);
        }
    }
}

struct TcParams {
    WPlanes w;
    float *out;       // [T][ldo]
    int64_t ldo;
    int32_t T, epi;   // EPI_STORE / EPI_ADD
    int32_t m_tiles, n_tiles;
};

template <int TYPE>
__global__ void __launch_bounds__(TC_THREADS, 1) prefill_tc_kernel(const __grid_constant__ CUtensorMap tmap_x, const TcParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + TC_STAGES * TC_STAGE_BYTES);
    uint64_t *full = bars, *empty = bars + TC_STAGES, *tmem_full = bars + 2 * TC_STAGES, *tmem_empty = bars + 2 * TC_STAGES + 2;
    uint32_t *tmem_base_slot = reinterpret_cast<uint32_t *>(bars + 2 * TC_STAGES + 4);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_kb = p.w.k / TC_BK;
    const int n_tiles_total = p.m_tiles * p.n_tiles;

    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_STAGES; s++) {
            mbar_init(&full[s], 1 + 4);   // the TMA producer's arrive (+ transaction bytes) and one per weight-producer warp
            mbar_init(&empty[s], 1);      // tcgen05.commit
        }
        for (int a = 0; a < 2; a++) {
            mbar_init(&tmem_full[a], 1);   // tcgen05.commit after the tile's last MMA
            mbar_init(&tmem_empty[a], 4);  // one per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {  // 512 TMEM columns: two 128-lane x 256-column f32 accumulators
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;

    if (warp == 0) {
        // ===== TMA producer (activations) =====
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_x) : "memory");
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x) {
                const int n_blk = tile % p.n_tiles;
                for (int kb = 0; kb < n_kb; kb++, it++) {
                    const int s = it % TC_STAGES;
                    mbar_wait(&empty[s], ((it / TC_STAGES) & 1) ^ 1);
                    mbar_arrive_expect_tx(&full[s], TC_B_BYTES);
                    tma_load_2d(smem + s * TC_STAGE_BYTES + TC_A_BYTES, &tmap_x, kb * TC_BK, n_blk * TC_BN, &full[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_f16(TC_BM, TC_BN);
            uint32_t it = 0, tile_it = 0;
            for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x, tile_it++) {
                const uint32_t acc = tile_it & 1;
                mbar_wait(&tmem_empty[acc], ((tile_it >> 1) & 1) ^ 1);  // the epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * TC_BN;
                for (int kb = 0; kb < n_kb; kb++, it++) {
                    const int s = it % TC_STAGES;
                    mbar_wait(&full[s], (it / TC_STAGES) & 1);
                    tc_fence_after();
                    const uint64_t a_desc = umma_desc_sw128(smem + s * TC_STAGE_BYTES);
                    const uint64_t b_desc = umma_desc_sw128(smem + s * TC_STAGE_BYTES + TC_A_BYTES);
#pragma unroll
                    for (int k = 0; k < TC_BK / 16; k++)  // 16 halves = 32 bytes further along the swizzled row
                        umma_f16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, kb > 0 || k > 0);
                    umma_commit(&empty[s]);  // the stage is free once these MMAs have read it
                }
                umma_commit(&tmem_full[acc]);  // the accumulator is complete once every MMA of the tile has finished
            }
        }
    } else if (warp >= 4 && warp < 8) {
        // ===== weight producers: thread r owns row r of the tile =====
        const int r = threadIdx.x - 128;
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x) {
            const int m_blk = tile / p.n_tiles;
            const int64_t row = min(m_blk * TC_BM + r, p.w.n_rows - 1);  // rows beyond N repeat the last one (never stored)
            RawStage ra, rb;  // ping-pong: the stage being converted and the one in flight
            load_raw_stage<TYPE>(p.w, row, 0, ra);
            auto do_stage = [&](const RawStage &cur, RawStage &nxt, int kb) {
                const int s = it % TC_STAGES;
                if (kb + 1 < n_kb) load_raw_stage<TYPE>(p.w, row, kb + 1, nxt);  // one stage ahead
                mbar_wait(&empty[s], ((it / TC_STAGES) & 1) ^ 1);
                dequant_row_stage<TYPE>(cur, kb, smem + s * TC_STAGE_BYTES, r);
                fence_proxy_async();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
                __syncwarp();
                if (lane == 0) mbar_arrive(&full[s]);
                it++;
            };
            for (int kb = 0; kb < n_kb; kb += 2) {  // K is a multiple of 256: n_kb is even
                do_stage(ra, rb, kb);
                do_stage(rb, ra, kb + 1);
            }
        }
    } else if (warp >= 8) {
        // ===== epilogue: warp q = warp - 8 reads TMEM lanes [32q, 32q+32) = rows 32q + lane of the tile =====
        const int q = warp - 8;
        uint32_t tile_it = 0;
        for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x, tile_it++) {
            const int m_blk = tile / p.n_tiles, n_blk = tile % p.n_tiles;
            const uint32_t acc = tile_it & 1;
            mbar_wait(&tmem_full[acc], (tile_it >> 1) & 1);
            tc_fence_after();
            const int row = m_blk * TC_BM + 32 * q + lane;
            const int t0 = n_blk * TC_BN;
            const bool row_ok = row < p.w.n_rows;
#pragma unroll 1
            for (int c0 = 0; c0 < TC_BN; c0 += 32) {
                uint32_t v[32];
                tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(32 * q) << 16) + acc * TC_BN + c0, v);
                tmem_ld_wait();
                if (row_ok) {
#pragma unroll
                    for (int j = 0; j < 32; j++) {
                        const int t = t0 + c0 + j;
                        if (t < p.T) {
                            float *o = p.out + (size_t)t * p.ldo + row;  // lanes = consecutive rows: coalesced
                            const float val = __uint_as_float(v[j]);
                            *o = p.epi == EPI_ADD ? *o + val : val;
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

// ---------------------------------------------------------------- host side
namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr) != cudaSuccess || !f) {
            cudaGetLastError();
            throw CudaError("driver entry point cuTensorMapEncodeTiled is not available");
        }
        fn = reinterpret_cast<EncodeTiledFn>(f);
    });
    return fn;
}
}  // namespace

bool prefill_tc_usable(int type, int k) { return (type == T_Q4_K || type == T_Q5_K || type == T_Q6_K || type == T_Q8_0) && k % 256 == 0; }

// xh: activations as f16, [T][k] row-major (k contiguous), in THIS GPU's memory
void launch_prefill_tc(const WPlanes &w, const __half *xh, int T, float *out, int64_t ldo, int epi, int n_sms, cudaStream_t st) {
    if (!prefill_tc_usable(w.type, w.k)) throw CudaError("prefill_tc: Q4_K / Q5_K / Q6_K / Q8_0 weights, K a multiple of 256");
    if (T < 1) return;
    CUtensorMap tmap;
    const cuuint64_t gdim[2] = {(cuuint64_t)w.k, (cuuint64_t)T};
    const cuuint64_t gstride[1] = {(cuuint64_t)w.k * 2};
    const cuuint32_t box[2] = {TC_BK, TC_BN};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = encode_tiled()(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<__half *>(xh), gdim, gstride, box, estr,
                                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) throw CudaError("cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
    TcParams p{};
    p.w = w; p.out = out; p.ldo = ldo; p.T = T; p.epi = epi;
    p.m_tiles = (w.n_rows + TC_BM - 1) / TC_BM;
    p.n_tiles = (T + TC_BN - 1) / TC_BN;
    const int grid = std::min(n_sms, p.m_tiles * p.n_tiles);
#define TC_LAUNCH(TY)                                                                                                  \
    do {                                                                                                               \
        static PerDeviceOnce once;                                                                                     \
        once.run([] { CUDA_CHECK(cudaFuncSetAttribute(prefill_tc_kernel<TY>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES)); }); \
        prefill_tc_kernel<TY><<<grid, TC_THREADS, TC_SMEM_BYTES, st>>>(tmap, p);                                       \
    } while (0)
    if (w.type == T_Q4_K) TC_LAUNCH(T_Q4_K);
    else if (w.type == T_Q5_K) TC_LAUNCH(T_Q5_K);
    else if (w.type == T_Q6_K) TC_LAUNCH(T_Q6_K);
    else TC_LAUNCH(T_Q8_0);
#undef TC_LAUNCH
    CUDA_CHECK(cudaGetLastError());
}

// ---------------------------------------------------------------- activation preparation (f32 -> f16 rows)
// xh[t][0..k) = f16(rmsnorm(x[t]) * norm_w)  (norm_w == nullptr: plain conversion).  One CTA per token.
__global__ void __launch_bounds__(256) rows_to_f16_kernel(const float *x, int64_t ldx, const float *norm_w, float eps, int k, __half *xh) {
    const int t = blockIdx.x;
    const float *xr = x + (size_t)t * ldx;
    __half *o = xh + (size_t)t * k;
    float scale = 1.f;
    if (norm_w) {
        __shared__ double red[8];
        double ss = 0.0;
        for (int i = threadIdx.x; i < k; i += blockDim.x) ss += (double)__fmul_rn(xr[i], xr[i]);
        ss = warp_sum_f64(ss);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
        __syncthreads();
        double tot = 0.0;
        for (int i = 0; i < (int)blockDim.x / 32; i++) tot += red[i];
        scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn((float)(tot / (double)k), eps)));
    }
    for (int i = threadIdx.x * 2; i < k; i += blockDim.x * 2) {
        float a = xr[i], b = xr[i + 1];
        if (norm_w) { a = __fmul_rn(__fmul_rn(a, scale), norm_w[i]); b = __fmul_rn(__fmul_rn(b, scale), norm_w[i + 1]); }
        *reinterpret_cast<__half2 *>(o + i) = __floats2half2_rn(a, b);
    }
}
void launch_rows_to_f16(const float *x, int64_t ldx, const float *norm_w, float eps, int k, int T, __half *xh, cudaStream_t st) {
    rows_to_f16_kernel<<<T, 256, 0, st>>>(x, ldx, norm_w, eps, k, xh);
    CUDA_CHECK(cudaGetLastError());
}

// hh[t][i] = f16(silu(gu[t][i]) * gu[t][ff + i])
__global__ void swiglu_rows_f16_kernel(const float *gu, int64_t ldgu, int ff, int T, __half *hh) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)T * ff) return;
    const int t = (int)(i / ff), c = (int)(i % ff);
    const float gte = gu[(size_t)t * ldgu + c], up = gu[(size_t)t * ldgu + ff + c];
    hh[(size_t)t * ff + c] = __float2half_rn(__fmul_rn(__fdiv_rn(gte, __fadd_rn(1.0f, expf_det(-gte))), up));
}
void launch_swiglu_rows_f16(const float *gu, int64_t ldgu, int ff, int T, __half *hh, cudaStream_t st) {
    const int64_t n = (int64_t)T * ff;
    swiglu_rows_f16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(gu, ldgu, ff, T, hh);
    CUDA_CHECK(cudaGetLastError());
}

}  // namespace aios
