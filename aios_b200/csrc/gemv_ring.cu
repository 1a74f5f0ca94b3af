// gemv_ring.cu — the decode GEMV fed by a TMA ring: y[N] = dq(W[N,K]) . x[K], weights streamed HBM -> shared memory
// by bulk copies that run AHEAD of everything else in the token pass.
//
// Why (profiles/r01_v3_summary.md): the register-pipelined kernel (gemv.cu) streams its big matrices at 55-68 % of
// the HBM peak but pays ~6 us per launch in which HBM idles — the dependency wait on the previous kernel, the
// activation prologue (RMSNorm + Q8 quantisation), the tail.  145 launches per token make that 0.9 of 1.4 ms.  The
// weights do not depend on any of it.  So here:
//   * one CTA per SM, sized to HALF an SM (8 warps x <=128 registers, <=100 KB shared memory), so that the CTA of
//     the NEXT launch is co-resident with the current one (programmatic dependent launch, triggered at kernel entry);
//   * the CTA's list of weight tiles is static (layout.cuh "stream tiles": one contiguous blob per 16 rows x 2048
//     weights).  Thread 0 issues one cp.async.bulk per ring slot at kernel entry — before the dependency wait — and
//     from then on the LAST warp to copy its rows out of a slot re-arms it with the tile S ahead (shared-memory
//     counter + full mbarrier per slot): no producer warp, no waiting.  While the previous kernel finishes, while
//     this kernel waits for it and quantises its activations, 80-92 KB per SM (12-13 MB per GPU) are already on chip;
//   * the 8 warps wait for the dependency, build the activation image (gemv_core.cuh, same arithmetic as gemv.cu),
//     then take the slots two at a time: warp w owns tile rows w and w+8 for the whole K, so the only reduction is
//     the warp-level one; the unit arithmetic is gemv_core.cuh's compute_pass on registers loaded from the slot.
// Numerics are those of gemv.cu (bit-identical results: one f32 term per 64-weight unit, summed in double).
// Status: correct, but slower than gemv.cu (2.41 vs 1.92 ms/token) — the consumers, not the ring, are the limit
// (profiles/r01_v4_ring.md) — so the engine uses it only with AIOS_ENGINE_USE_RING.
//
// Q4_K and Q6_K streams (the Q4_K_M mix); a launch may mix both types (attn_q/k Q4_K + attn_v Q6_K).
#define GEMV_CORE_WARPS 8
#define GEMV_CORE_SYNC() asm volatile("bar.sync 1, 256;" ::: "memory")
#include "gemv_core.cuh"

namespace aios {

// 8 warps; thread 0 doubles as the producer.  (A ninth, dedicated producer warp would land on SMSP 0 in every CTA:
// with the register file split per SMSP, two co-resident CTAs would then have to fit 6 warps into 16K registers.)
constexpr int RING_THREADS = GEMV_THREADS;
constexpr int RING_MAX_SLOTS = 8;
constexpr int RING_SMEM_BUDGET = 100 * 1024;     // dynamic shared memory per CTA: two such CTAs fit one SM

// ---------------------------------------------------------------- tile builder (model load time)
// one thread per (tile, tile row, unit): copies that unit's bytes of every plane, or zeros outside the matrix
__global__ void build_tiles_kernel(WPlanes A, WPlanes B, int pair, int n_rt, int ct, uint8_t *dst) {
    const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t total = (int64_t)n_rt * ct * TILE_ROWS * TILE_UNITS;
    if (gid >= total) return;
    const int ul = (int)(gid % TILE_UNITS);
    const int r = (int)((gid / TILE_UNITS) % TILE_ROWS);
    const int64_t tile = gid / (TILE_UNITS * TILE_ROWS);
    const int c = (int)(tile % ct);
    const int64_t rt = tile / ct;
    const WPlanes &W = (pair && r >= 8) ? B : A;
    const int64_t row = pair ? rt * 8 + (r & 7) : rt * 16 + r;
    const int64_t U = W.k / 64, nb = W.k / 256;
    const int64_t u = (int64_t)c * TILE_UNITS + ul;
    const bool ok = row < W.n_rows && u < U;
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    const size_t tb = W.type == T_Q4_K ? 18432 : 26880;
    uint8_t *t = dst + (size_t)tile * tb;
    const size_t e = (size_t)r * TILE_UNITS + ul;  // (row, unit) index inside the tile
    auto ld16 = [&](int plane, int64_t idx) { return ok ? *reinterpret_cast<const uint4 *>(W.p[plane] + (size_t)idx * 16) : z4; };
    if (W.type == T_Q4_K) {
        *reinterpret_cast<uint4 *>(t + e * 16) = ld16(0, row * U + u);
        *reinterpret_cast<uint4 *>(t + 8192 + e * 16) = ld16(1, row * U + u);
        if ((ul & 3) == 0)
            *reinterpret_cast<uint4 *>(t + 16384 + ((size_t)r * 8 + (ul >> 2)) * 16) = ld16(2, row * nb + (u >> 2));
    } else {
        *reinterpret_cast<uint4 *>(t + e * 16) = ld16(0, row * U + u);
        *reinterpret_cast<uint4 *>(t + 8192 + e * 16) = ld16(1, row * U + u);
        *reinterpret_cast<uint4 *>(t + 16384 + e * 16) = ld16(2, row * U + u);
        *reinterpret_cast<uint32_t *>(t + 24576 + e * 4) =
            ok ? *reinterpret_cast<const uint32_t *>(W.p[3] + (size_t)(row * U + u) * 4) : 0u;
        if ((ul & 3) == 0)
            *reinterpret_cast<uint16_t *>(t + 26624 + ((size_t)r * 8 + (ul >> 2)) * 2) =
                ok ? *reinterpret_cast<const uint16_t *>(W.p[4] + (size_t)(row * nb + (u >> 2)) * 2) : (uint16_t)0;
    }
}

static void tile_geometry(int n_rows, int k, bool pair, int *n_rt, int *ct) {
    *n_rt = pair ? (n_rows + 7) / 8 : (n_rows + TILE_ROWS - 1) / TILE_ROWS;
    *ct = (k / 64 + TILE_UNITS - 1) / TILE_UNITS;
}

size_t tiles_total_bytes(int type, int n_rows, int k, bool pair) {
    int n_rt, ct;
    tile_geometry(n_rows, k, pair, &n_rt, &ct);
    return (size_t)n_rt * ct * tile_bytes(type);
}

void launch_build_tiles(const WPlanes &a, const WPlanes *b, uint8_t *dst, WTiles *out, cudaStream_t st) {
    const bool pair = b != nullptr;
    if (tile_bytes(a.type) == 0) throw CudaError("build_tiles: only Q4_K / Q6_K have stream tiles");
    if (pair && (b->type != a.type || b->n_rows != a.n_rows || b->k != a.k))
        throw CudaError("build_tiles: a PAIR needs two matrices of one type and shape");
    int n_rt, ct;
    tile_geometry(a.n_rows, a.k, pair, &n_rt, &ct);
    const int64_t total = (int64_t)n_rt * ct * TILE_ROWS * TILE_UNITS;
    const int threads = 256;
    build_tiles_kernel<<<(unsigned)((total + threads - 1) / threads), threads, 0, st>>>(a, pair ? *b : a, pair ? 1 : 0, n_rt, ct, dst);
    CUDA_CHECK(cudaGetLastError());
    out->base = dst; out->type = a.type; out->n_out = a.n_rows; out->k = a.k; out->n_rt = n_rt; out->ct = ct; out->pair = pair ? 1 : 0;
}

// ---------------------------------------------------------------- mbarrier / bulk-copy primitives
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {  // non-blocking
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// one contiguous blob HBM -> shared memory; completion is counted in bytes on `bar`; streamed data leaves L2 first
__device__ __forceinline__ void bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
        ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(policy)
        : "memory");
}

// ---------------------------------------------------------------- one tile from a ring slot into registers
template <int TYPE>
struct TileRegs {
    static constexpr int ITS = 1;
    uint4 a[GEMV_R][1], b[GEMV_R][1], c[GEMV_R][1];
    uint32_t e[GEMV_R][1], f[GEMV_R][1];
};

template <int TYPE>
__device__ __forceinline__ void load_tile(TileRegs<TYPE> &g, const uint8_t *slot, int warp, int lane) {
#pragma unroll
    for (int r = 0; r < GEMV_R; r++) {
        const int tr = warp + 8 * r;                    // my tile rows: w and w+8
        const size_t e = (size_t)tr * TILE_UNITS + lane;
        g.a[r][0] = *reinterpret_cast<const uint4 *>(slot + e * 16);
        g.b[r][0] = *reinterpret_cast<const uint4 *>(slot + 8192 + e * 16);
        if (TYPE == T_Q4_K) {
            g.c[r][0] = *reinterpret_cast<const uint4 *>(slot + 16384 + ((size_t)tr * 8 + (lane >> 2)) * 16);
        } else {
            g.c[r][0] = *reinterpret_cast<const uint4 *>(slot + 16384 + e * 16);
            g.e[r][0] = *reinterpret_cast<const uint32_t *>(slot + 24576 + e * 4);
            g.f[r][0] = *reinterpret_cast<const uint16_t *>(slot + 26624 + ((size_t)tr * 8 + (lane >> 2)) * 2);
        }
    }
}

__device__ __forceinline__ void ring_locate(const RingParams &p, int rt, int &seg, int &rt_local) {
    seg = 0;
    while (seg + 1 < p.n_seg && rt >= p.rt_begin[seg + 1]) seg++;
    rt_local = rt - p.rt_begin[seg];
}

__host__ __device__ inline size_t ring_img_bytes(int k) { return (act_image_bytes(k) + 127) & ~(size_t)127; }

// Request job j of CTA c: K tile (j % ct) of its (j / ct)-th row tile, into slot j % S.  Out of line on purpose:
// one lane runs it once per tile, and inlined copies (with their integer division) bloated the hot loop past the
// instruction cache (profiles/r01_v4_ring.md: 18 % of the stall samples were instruction fetches).
__device__ __noinline__ void ring_issue_job(const RingParams &p, int j, int c, int G, uint8_t *ring, uint64_t *full_bar,
                                            uint64_t policy) {
    const int m = j / p.ct, t = j - m * p.ct;
    int seg, rtl;
    ring_locate(p, c + m * G, seg, rtl);
    const uint32_t tb = p.seg[seg].type == T_Q4_K ? 18432u : 26880u;
    const uint8_t *src = p.seg[seg].base + ((size_t)rtl * p.ct + t) * tb;
    const int s = j % p.n_slots;
    const uint32_t fb = smem_u32(&full_bar[s]);
    mbar_expect_tx(fb, tb);
    bulk_load(smem_u32(ring + (size_t)s * p.slot_bytes), src, tb, fb, policy);
}

// ---------------------------------------------------------------- the kernel
// dynamic smem: [activation image, Q4 permutation (if any Q4_K stream)] [image, Q6 permutation (if any Q6_K stream)]
//               [ring: n_slots x slot_bytes]
// MODE: T_Q4_K / T_Q6_K = every stream of the launch has that type (one arithmetic path in the hot loop);
//       0 = mixed (attn_q/k Q4_K with attn_v Q6_K).  FULL: K is a multiple of 2048, no partial K tile.
template <int MODE, bool FULL>
__global__ void __launch_bounds__(RING_THREADS, 2) gemv_ring_kernel(const __grid_constant__ RingParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[RING_MAX_SLOTS];
    __shared__ uint32_t released[RING_MAX_SLOTS];  // per slot: warps that have copied their rows out (running count)
    __shared__ double red[GEMV_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int G = gridDim.x, c = blockIdx.x;
    const int T = p.rt_begin[p.n_seg];
    const int n_my = (T - c + G - 1) / G;        // my row tiles: c, c+G, ...
    const int S = p.n_slots;

    unsigned long long *tr = (c == 0 && lane == 0) ? p.trace : nullptr;
    if (warp == 0) trace_mark(tr, 0);
    constexpr bool any4 = MODE != T_Q6_K, any6 = MODE != T_Q4_K;
    const size_t img = ring_img_bytes(p.k);
    uint8_t *ring = smem + (any4 ? img : 0) + (any6 ? img : 0);

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; s++) {
            mbar_init(smem_u32(&full_bar[s]), 1);
            released[s] = 0;
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    // The next launch of the token pass may start now: until this grid has completed it only streams weights.
    asm volatile("griddepcontrol.launch_dependents;");

    // ---------------- producing: the tile list is static, nothing in it depends on the previous kernel.
    // Job j = K tile (j % ct) of my (j / ct)-th row tile, goes to slot j % S.  Thread 0 puts the first S jobs in
    // flight right here; after that, whichever warp is the LAST to copy its rows out of a slot re-arms that slot
    // with the job S ahead — no producer warp, no waiting, the refill leaves the moment the slot is free.
    const int n_jobs = n_my * p.ct;
    uint64_t policy;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    // lane 0 of a warp, after the warp's shared-memory loads of the job in slot s: count the release; the eighth
    // re-arms the slot
    auto release_slot = [&](int j, int s) {
        uint32_t old;
        asm volatile("atom.acq_rel.cta.shared::cta.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(smem_u32(&released[s])) : "memory");
        if ((old & (GEMV_WARPS - 1)) == GEMV_WARPS - 1 && j + S < n_jobs) ring_issue_job(p, j + S, c, G, ring, full_bar, policy);
    };
    if (threadIdx.x == 0) {
        for (int j = 0; j < min(S, n_jobs); j++) {
            ring_issue_job(p, j, c, G, ring, full_bar, policy);
            if (j == 0) trace_mark(tr, 5);
        }
        trace_mark(tr, 6);
    }
    // reconverge warp 0 here: griddepcontrol.wait below blocks the whole WARP, so lanes 1-31 running ahead into it
    // would freeze thread 0 in the middle of its copies until the previous kernel has finished
    __syncwarp();

    // ---------------- consumers
    ActSmem s4{}, s6{};
    if (any4) s4 = carve(smem, p.k);
    if (any6) s6 = carve(smem + (any4 ? img : 0), p.k);
    const int nslab = p.k / 256;
    const bool small_k = nslab <= 2 * GEMV_WARPS;  // K <= 4096: a lane's share of x (16 values) stays in registers
    // the norm weights are weights: fetch them before the dependency wait
    float nw0[8], nw1[8];
    if (small_k && p.norm_w) {
        if (warp < nslab) load8(p.norm_w + warp * 256 + lane * 8, nw0);
        if (warp + GEMV_WARPS < nslab) load8(p.norm_w + (warp + GEMV_WARPS) * 256 + lane * 8, nw1);
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (warp == 0) trace_mark(tr, 1);
    if (small_k) {
        // one pass over x: load, (norm), quantise once per image this launch needs
        const int b0 = warp, b1 = warp + GEMV_WARPS;
        float v0[8], v1[8];
#pragma unroll
        for (int e = 0; e < 8; e++) v0[e] = v1[e] = 0.f;
        if (b0 < nslab) load8(p.x + b0 * 256 + lane * 8, v0);
        if (b1 < nslab) load8(p.x + b1 * 256 + lane * 8, v1);
        if (p.norm_w) {
            const double ss = block_sum(sumsq8(v0) + sumsq8(v1), red);
            const float mean = (float)(ss / (double)p.k);
            const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, p.eps)));
#pragma unroll
            for (int e = 0; e < 8; e++) {
                v0[e] = __fmul_rn(__fmul_rn(v0[e], scale), nw0[e]);
                v1[e] = __fmul_rn(__fmul_rn(v1[e], scale), nw1[e]);
            }
        }
        if (any4) {
            if (b0 < nslab) quant_slab<PERM_Q4>(v0, b0, p.k, s4);
            if (b1 < nslab) quant_slab<PERM_Q4>(v1, b1, p.k, s4);
        }
        if (any6) {
            if (b0 < nslab) quant_slab<PERM_Q6>(v0, b0, p.k, s6);
            if (b1 < nslab) quant_slab<PERM_Q6>(v1, b1, p.k, s6);
        }
        GEMV_CORE_SYNC();
    } else if (p.norm_w) {
        if (any4) prep_activations<PERM_Q4>(p.x, p.norm_w, p.eps, p.k, s4, red);
        if (any6) prep_activations<PERM_Q6>(p.x, p.norm_w, p.eps, p.k, s6, red);
    } else {
        // big K without a norm (ffn_down): four slabs of x in flight per lane instead of two
        for (int b = warp; b < nslab; b += 4 * GEMV_WARPS) {
            float v[4][8];
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (b + j * GEMV_WARPS < nslab) load8(p.x + (b + j * GEMV_WARPS) * 256 + lane * 8, v[j]);
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (b + j * GEMV_WARPS < nslab) {
                    if (any4) quant_slab<PERM_Q4>(v[j], b + j * GEMV_WARPS, p.k, s4);
                    if (any6) quant_slab<PERM_Q6>(v[j], b + j * GEMV_WARPS, p.k, s6);
                }
            }
        }
        GEMV_CORE_SYNC();
    }

    if (warp == 0) trace_mark(tr, 2);
    const int nu = p.k / 64;
    unsigned long long best = 0ull;
    int i = 0, slot = 0, parity = 0;
    for (int m = 0; m < n_my; m++) {
        int seg, rtl;
        ring_locate(p, c + m * G, seg, rtl);
        const int type = p.seg[seg].type;
        float *out = p.out[seg];
        const int n_out = p.seg[seg].n_out;
        const bool pair = p.seg[seg].pair != 0;
        // the residual this row tile adds to: requested now, needed after the last tile
        const int my_row = rtl * TILE_ROWS + warp + (lane >> 4) * 8;
        float resid = 0.f;
        if (p.epi == EPI_ADD && (lane & 15) == 0 && my_row < n_out) resid = __ldcg(out + my_row);
        double acc[GEMV_R] = {0.0, 0.0};
        // two tiles per step: two independent chains per lane hide the LDS / dp4a latencies
        for (int t = 0; t < p.ct; t += 2) {
            const bool two = t + 1 < p.ct;
            // (slot, phase parity) of tile i and of tile i+1: a running pair instead of i % S, i / S
            const int sa = slot, pa = parity;
            int sb = slot + 1, pb = parity;
            if (sb == S) { sb = 0; pb ^= 1; }
            const uint8_t *slot_a = ring + (size_t)sa * p.slot_bytes, *slot_b = ring + (size_t)sb * p.slot_bytes;
            mbar_wait(smem_u32(&full_bar[sa]), (uint32_t)pa);
            if (two) mbar_wait(smem_u32(&full_bar[sb]), (uint32_t)pb);
            if (i == 0 && warp == 0) trace_mark(tr, 3);
            if (MODE == T_Q4_K || (MODE == 0 && type == T_Q4_K)) {
                TileRegs<T_Q4_K> ga, gb;
                load_tile<T_Q4_K>(ga, slot_a, warp, lane);
                if (two) load_tile<T_Q4_K>(gb, slot_b, warp, lane);
                __syncwarp();
                if (lane == 0) {  // my rows are in registers; the last warp to say so re-arms the slot
                    release_slot(i, sa);
                    if (two) release_slot(i + 1, sb);
                }
                __syncwarp();
                compute_pass<T_Q4_K, FULL>(ga, t * TILE_UNITS, nu, p.k, s4, acc);
                if (two) compute_pass<T_Q4_K, FULL>(gb, (t + 1) * TILE_UNITS, nu, p.k, s4, acc);
            } else {
                TileRegs<T_Q6_K> ga, gb;
                load_tile<T_Q6_K>(ga, slot_a, warp, lane);
                if (two) load_tile<T_Q6_K>(gb, slot_b, warp, lane);
                __syncwarp();
                if (lane == 0) {
                    release_slot(i, sa);
                    if (two) release_slot(i + 1, sb);
                }
                __syncwarp();
                compute_pass<T_Q6_K, FULL>(ga, t * TILE_UNITS, nu, p.k, s6, acc);
                if (two) compute_pass<T_Q6_K, FULL>(gb, (t + 1) * TILE_UNITS, nu, p.k, s6, acc);
            }
            if (two) { i += 2; slot = sb + 1; parity = pb; }
            else { i += 1; slot = sa + 1; }
            if (slot == S) { slot = 0; parity ^= 1; }
        }
        const double tot = warp_sum2_f64(acc[0], acc[1]);  // lanes 0-15: tile row w, lanes 16-31: tile row w+8
        if (pair) {
            const float gte = (float)__shfl_sync(0xffffffffu, tot, 0), up = (float)__shfl_sync(0xffffffffu, tot, 16);
            const int row = rtl * 8 + warp;
            if (lane == 0 && row < n_out) out[row] = __fmul_rn(__fdiv_rn(gte, __fadd_rn(1.0f, expf_det(-gte))), up);
        } else if ((lane & 15) == 0) {
            const int row = my_row;
            if (row < n_out) {
                const float v = (float)tot;
                if (p.epi == EPI_ADD) out[row] = __fadd_rn(resid, v);
                else out[row] = v;
                if (p.argmax_key) {
                    const unsigned long long key = argmax_key(v, row);
                    best = key > best ? key : best;
                }
            }
        }
    }
    if (warp == 0) trace_mark(tr, 4);
    if (p.trace && threadIdx.x == 0) {  // [7] = when the LAST CTA's consumers were done
        unsigned long long now;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
        atomicMax(p.trace + 7, now);
    }
    if (p.argmax_key) {
        const unsigned long long other = __shfl_sync(0xffffffffu, best, 16);
        best = other > best ? other : best;
        if (lane == 0 && best) atomicMax(p.argmax_key, best);
    }
}

// ---------------------------------------------------------------- host side
void gemv_ring_init() {
    static bool done = false;
    if (done) return;
    // two ring CTAs (of consecutive launches) must fit one SM: ask for the largest shared-memory carve-out.  Every
    // kernel of the token pass asks for the same one — an SM cannot host kernels with different carve-outs at once.
    auto prep = [](auto kern) {
        CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, RING_SMEM_BUDGET));
        CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    };
    prep(gemv_ring_kernel<T_Q4_K, true>); prep(gemv_ring_kernel<T_Q4_K, false>);
    prep(gemv_ring_kernel<T_Q6_K, true>); prep(gemv_ring_kernel<T_Q6_K, false>);
    prep(gemv_ring_kernel<0, true>); prep(gemv_ring_kernel<0, false>);
    done = true;
}

static bool ring_plan(RingParams &p) {
    if (p.n_seg < 1 || p.n_seg > 3) return false;
    bool any4 = false, any6 = false;
    int slot = 0, begin = 0;
    for (int i = 0; i < p.n_seg; i++) {
        const WTiles &w = p.seg[i];
        if (!w.base || (w.type != T_Q4_K && w.type != T_Q6_K) || w.k != p.k) return false;
        if ((p.epi == EPI_SWIGLU) != (w.pair != 0)) return false;
        any4 |= w.type == T_Q4_K; any6 |= w.type == T_Q6_K;
        slot = std::max(slot, (int)tile_bytes(w.type));
        p.rt_begin[i] = begin;
        begin += w.n_rt;
        p.ct = w.ct;
    }
    for (int i = p.n_seg; i < 4; i++) p.rt_begin[i] = begin;
    const size_t img = ring_img_bytes(p.k) * ((any4 ? 1 : 0) + (any6 ? 1 : 0));
    if (img + 2 * (size_t)slot > (size_t)RING_SMEM_BUDGET) return false;
    p.slot_bytes = slot;
    p.n_slots = std::min<int>(RING_MAX_SLOTS, (int)((RING_SMEM_BUDGET - img) / slot));
    return p.n_slots >= 2;
}

bool gemv_ring_usable(const RingParams &p) {
    RingParams q = p;
    return ring_plan(q);
}

void launch_gemv_ring(RingParams &p, int n_sms, cudaStream_t st, bool pdl) {
    if (!ring_plan(p)) throw CudaError("launch_gemv_ring: parameters not usable (check gemv_ring_usable first)");
    bool any4 = false, any6 = false;
    for (int i = 0; i < p.n_seg; i++) { any4 |= p.seg[i].type == T_Q4_K; any6 |= p.seg[i].type == T_Q6_K; }
    const int T = p.rt_begin[p.n_seg];
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(std::min(n_sms, T));
    cfg.blockDim = dim3(RING_THREADS);
    cfg.dynamicSmemBytes = ring_img_bytes(p.k) * ((any4 ? 1 : 0) + (any6 ? 1 : 0)) + (size_t)p.n_slots * p.slot_bytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    const bool full = (p.k / 64) % TILE_UNITS == 0;
    const int mode = any4 && any6 ? 0 : (any4 ? (int)T_Q4_K : (int)T_Q6_K);
    if (mode == T_Q4_K) {
        if (full) CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_ring_kernel<T_Q4_K, true>, p));
        else CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_ring_kernel<T_Q4_K, false>, p));
    } else if (mode == T_Q6_K) {
        if (full) CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_ring_kernel<T_Q6_K, true>, p));
        else CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_ring_kernel<T_Q6_K, false>, p));
    } else {
        if (full) CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_ring_kernel<0, true>, p));
        else CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_ring_kernel<0, false>, p));
    }
}

}  // namespace aios
