// gemv.cu — the decode hot loop's kernel: y[N] = dq(W[N,K]) . x[K] for Q4_K / Q5_K / Q6_K / Q8_0
// weights at batch 1, HBM-bandwidth bound.
//
// Numerics mirror llama.cpp's CPU path (oracle/aios_oracle.c): the input
// vector is quantised to Q8_K (K-quants) or Q8_0 (Q8_0 weights); dot products are integer inside a
// 64-weight unit (dp4a); every unit contributes one f32 term, and the terms are summed in double —
// an order-free sum, so the result is bit-identical to the oracle's for any work split.
//
// One launch = [optional RMSNorm] + activation quantisation (done by every CTA into its own shared
// memory: K bytes, so no separate launch and no HBM round trip) + the weight stream (128-bit coalesced
// no-allocate loads from the planes of layout.cuh; the first passes are issued BEFORE the activation
// prologue, and before the programmatic-dependent-launch wait, because weights do not depend on the
// previous kernel) + a double-precision warp-shuffle reduction + a fused epilogue (store / residual
// add / SwiGLU / running arg-max / RoPE + f16 KV append for q,k,v).  Matrices that share the input
// (attn_q/k/v, also when attn_v is Q6_K next to Q4_K q/k; ffn_gate/up) ride in one launch; CTAs are
// split between them by byte share.
//
// Two kernels: gemv_db_kernel (default; K a multiple of 2048) keeps two half-size weight images per
// warp so loads stay in flight through the warp's own arithmetic; gemv_kernel is the single-image
// original and the fallback for other K.  One 16-warp CTA per SM: the activation image is built once
// per SM.  Measured steps, Mistral-7B Q4_K_M ms/token: 1.92 single image -> 1.79 double buffered ->
// 1.72 q/k/v in one launch -> 1.68 RoPE epilogue + staged attention -> 1.66 one CTA per SM
// (profiles/r01_v5_summary.md).
#define GEMV_CORE_WARPS 16
#include "gemv_core.cuh"

namespace aios {

// ---- EPI_PUSH_F64 (tensor parallelism, tp.cu): the row's double sum goes to every rank's receive slab; when the whole
// grid has stored, its last CTA raises this rank's arrival flag on every rank.  Plain stores and a release store over
// NVLink peer mappings; the consumer (tp_wait_add_kernel) acquires the flags and reads the slabs past L1.
__device__ __forceinline__ unsigned tp_push_epoch(const GemvParams &p) {
    return __ldcg(p.push.tick) * (unsigned)p.push.per_token + (unsigned)p.push.index;
}
__device__ __forceinline__ void tp_push_store(const GemvParams &p, int row, int n_rows, double tot, unsigned ep) {
    const size_t off = (size_t)(ep & 1u) * (size_t)n_rows + (size_t)row;
    for (int r = 0; r < p.push.n; r++) p.push.dst[r][off] = tot;
}
__device__ __forceinline__ void tp_push_store_f32(const GemvParams &p, int row, float v) {
    for (int r = 0; r < p.push.n; r++) reinterpret_cast<float *>(p.push.dst[r])[row] = v;
}
__device__ __forceinline__ void tp_push_signal(const GemvParams &p, unsigned ep) {
    __threadfence_system();  // this thread's slab stores are visible system-wide before the CTA counts itself done
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned old = atomicAdd(p.push.done, 1u);
        if (old == gridDim.x - 1) {
            *p.push.done = 0u;  // the next launch on this GPU starts counting from zero
            __threadfence_system();
            for (int r = 0; r < p.push.n; r++) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.push.flag[r]), "r"(ep) : "memory");
        }
    }
}

template <int TYPE>
__global__ void __launch_bounds__(GEMV_THREADS, 1) gemv_kernel(const GemvParams p) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ double red[GEMV_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int UNIT = TYPE == T_Q8_0 ? 32 : 64;
    constexpr int PASS_UNITS = PassCfg<TYPE>::ITS * 32;
    // which segment does this CTA serve?
    int seg = 0;
    if (p.epi != EPI_SWIGLU) {
        while (seg + 1 < p.n_seg && (int)blockIdx.x >= p.cta_begin[seg + 1]) seg++;
    }
    const WPlanes &W = p.seg[seg];
    const WPlanes &W1 = p.seg[p.epi == EPI_SWIGLU ? 1 : seg];
    const int cta_local = (int)blockIdx.x - p.cta_begin[seg];
    const int cta_count = p.cta_begin[seg + 1] - p.cta_begin[seg];
    const ActSmem s = carve(smem, p.k);
    const int N = W.n_rows;
    const int nu = p.k / UNIT;
    const bool full = (nu % PASS_UNITS) == 0;
    const bool pair = p.epi == EPI_SWIGLU;  // rows come as (gate row, up row) pairs of two matrices
    const int rows_per_pass = pair ? 1 : GEMV_R;
    const int row_first = (cta_local * GEMV_WARPS + warp) * rows_per_pass;
    const int row_step = cta_count * GEMV_WARPS * rows_per_pass;

    auto ptrs = [&](int row0, RowPtrs (&rp)[GEMV_R]) {
        if (pair) {
            rp[0] = row_ptrs<TYPE>(W, row0);
            rp[1] = row_ptrs<TYPE>(W1, row0);
        } else {
            rp[0] = row_ptrs<TYPE>(W, row0);
            rp[1] = row_ptrs<TYPE>(W, min(row0 + 1, N - 1));
        }
    };

    // The first pass of weights does not depend on anything the previous kernel wrote: issue it now,
    // so that HBM is busy during the dependency wait and the activation prologue.
    PassRegs<TYPE> g;
    RowPtrs rp[GEMV_R];
    const uint64_t pol = l2_policy(p.evict_first != 0);
    const bool have_work = row_first < N;
    if (have_work) {
        ptrs(row_first, rp);
        if (full) load_pass<TYPE, true>(g, rp, 0, nu, pol);
        else load_pass<TYPE, false>(g, rp, 0, nu, pol);
    }
    // programmatic dependent launch: everything above overlapped the previous kernel's tail
    asm volatile("griddepcontrol.wait;" ::: "memory");

    // (fetching the norm weights before the wait was tried: it costs 16 registers across the held first pass, spills,
    // and gains nothing measurable — 1.915 vs 1.912 ms/token)
    if (p.x_image) copy_activation_image(p.x_image, p.k, smem);
    else if (TYPE == T_Q6_K) prep_activations<PERM_Q6>(p.x, p.norm_w, p.eps, p.k, s, red);
    else if (TYPE == T_Q8_0) prep_activations<PERM_Q80>(p.x, p.norm_w, p.eps, p.k, s, red);
    else prep_activations<PERM_Q4>(p.x, p.norm_w, p.eps, p.k, s, red);
    // let the next kernel's CTAs be scheduled as soon as SM resources free up: until this grid has
    // fully finished they only prefetch their first weights (they block in griddepcontrol.wait)
    asm volatile("griddepcontrol.launch_dependents;");

    float *out = p.out[seg];
    unsigned long long best = 0ull;
    const int rope_pos = p.epi == EPI_QKV_ROPE ? __ldcg(p.rope.n_past_ptr) : 0;
    const unsigned push_ep = p.epi >= EPI_PUSH_F64 ? tp_push_epoch(p) : 0u;
    for (int row0 = row_first; row0 < N; row0 += row_step) {
        double acc[GEMV_R] = {0.0, 0.0};
        if (row0 != row_first) ptrs(row0, rp);
        for (int u0 = 0; u0 < nu; u0 += PASS_UNITS) {
            if (row0 != row_first || u0 != 0) {
                if (full) load_pass<TYPE, true>(g, rp, u0, nu, pol);
                else load_pass<TYPE, false>(g, rp, u0, nu, pol);
            }
            if (full) compute_pass<TYPE, true>(g, u0, nu, p.k, s, acc);
            else compute_pass<TYPE, false>(g, u0, nu, p.k, s, acc);
        }
        const double tot = warp_sum2_f64(acc[0], acc[1]);  // lanes 0-15: row 0, lanes 16-31: row 1
        if (p.epi == EPI_QKV_ROPE) {
            qkv_rope_store(p, seg, row0, tot, rope_pos);
        } else if (pair) {
            const float gte = (float)__shfl_sync(0xffffffffu, tot, 0), up = (float)__shfl_sync(0xffffffffu, tot, 16);
            if (lane == 0) out[row0] = __fmul_rn(__fdiv_rn(gte, __fadd_rn(1.0f, expf_det(-gte))), up);
        } else if ((lane & 15) == 0) {
            const int row = row0 + (lane >> 4);
            if (row < N) {
                const float v = (float)tot;
                if (p.epi == EPI_PUSH_F64) tp_push_store(p, row, N, tot, push_ep);
                else if (p.epi == EPI_PUSH_F32) tp_push_store_f32(p, row, v);
                else if (p.epi == EPI_STORE_F64) reinterpret_cast<double *>(out)[row] = tot;
                else if (p.epi == EPI_ADD) out[row] = __fadd_rn(out[row], v);
                else out[row] = v;
                if (p.argmax_key) {
                    const unsigned long long key = argmax_key(v, row);
                    best = key > best ? key : best;
                }
            }
        }
    }
    if (p.argmax_key) {
        const unsigned long long other = __shfl_sync(0xffffffffu, best, 16);
        best = other > best ? other : best;
        if (lane == 0 && best) atomicMax(p.argmax_key, best);
    }
    if (p.epi >= EPI_PUSH_F64) tp_push_signal(p, push_ep);
}

// ---------------------------------------------------------------- double-buffered variant
// Same work split, prologue, arithmetic and epilogues as gemv_kernel; the difference is the weight pipeline inside a
// warp.  gemv_kernel loads a pass (2 rows x 2 units per lane, 12 loads), waits for it, computes it, loads the next:
// while a warp computes, none of its loads are in flight.  Here a pass is half as big (2 rows x 1 unit, 6 loads) and
// there are two register images: compute(A), refill A, compute(B), refill B — so every warp always has 6-12 loads in
// flight, through its own arithmetic as well.  Same register budget (2 x 24 instead of 1 x 48 weight registers).
// K must be a whole number of passes (K % 2048 == 0); launch_gemv falls back to gemv_kernel otherwise.
template <int TYPE>
struct PassRegs1 {
    static constexpr int ITS = TYPE == T_Q8_0 ? 2 : 1;
    uint4 a[GEMV_R][ITS], b[GEMV_R][ITS], c[GEMV_R][ITS];
    uint32_t e[GEMV_R][ITS], f[GEMV_R][ITS];
};

template <int TYPE, int NB>
__device__ __forceinline__ void gemv_db_body(const GemvParams &p, int seg, uint8_t *smem, double *red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int UNIT = TYPE == T_Q8_0 ? 32 : 64;
    using Regs = PassRegs1<TYPE>;
    constexpr int PU = Regs::ITS * 32;
    unsigned long long *tr = (blockIdx.x == 0 && threadIdx.x == 0) ? p.trace : nullptr;
    trace_mark(tr, 0);
    if (tr) tr[3] = 1ull;  // launch kind: GEMV
    const WPlanes &W = p.seg[seg];
    const WPlanes &W1 = p.seg[p.epi == EPI_SWIGLU ? 1 : seg];
    const int cta_local = (int)blockIdx.x - p.cta_begin[seg];
    const int cta_count = p.cta_begin[seg + 1] - p.cta_begin[seg];
    const ActSmem s = carve(smem, p.k);
    const int N = W.n_rows;
    const int nu = p.k / UNIT;
    const bool pair = p.epi == EPI_SWIGLU;
    const int rows_per_pass = pair ? 1 : GEMV_R;
    const int row_first = (cta_local * GEMV_WARPS + warp) * rows_per_pass;
    const int row_step = cta_count * GEMV_WARPS * rows_per_pass;

    auto ptrs = [&](int row0, RowPtrs (&rp)[GEMV_R]) {
        if (pair) {
            rp[0] = row_ptrs<TYPE>(W, row0);
            rp[1] = row_ptrs<TYPE>(W1, row0);
        } else {
            rp[0] = row_ptrs<TYPE>(W, row0);
            rp[1] = row_ptrs<TYPE>(W, min(row0 + 1, N - 1));
        }
    };

    // load cursor: the pass that will be requested next
    int lrow = row_first, lu = 0;
    RowPtrs lrp[GEMV_R];
    if (lrow < N) ptrs(lrow, lrp);
    const uint64_t pol = l2_policy(p.evict_first != 0);
    auto load_next = [&](Regs &g) {
        if (lrow < N) {
            load_pass<TYPE, true>(g, lrp, lu, nu, pol);
            lu += PU;
            if (lu >= nu) {
                lu = 0;
                lrow += row_step;
                if (lrow < N) ptrs(lrow, lrp);
            }
        }
    };
    Regs g[NB];
    // (Tried instead: hold nothing through the prologue, only pull the first passes into L2, and use the freed
    // registers for a prologue with more loads in flight — 1.81 vs 1.72 ms/token: the held images win.)
#pragma unroll
    for (int b = 0; b < NB; b++) load_next(g[b]);  // all images are requested before the dependency wait
    asm volatile("griddepcontrol.wait;" ::: "memory");
    trace_mark(tr, 1);
    if (p.x_image) copy_activation_image(p.x_image, p.k, smem);
    else if (TYPE == T_Q6_K) prep_activations<PERM_Q6>(p.x, p.norm_w, p.eps, p.k, s, red);
    else if (TYPE == T_Q8_0) prep_activations<PERM_Q80>(p.x, p.norm_w, p.eps, p.k, s, red);
    else prep_activations<PERM_Q4>(p.x, p.norm_w, p.eps, p.k, s, red);
    asm volatile("griddepcontrol.launch_dependents;");
    trace_mark(tr, 2);

    float *out = p.out[seg];
    unsigned long long best = 0ull;
    const int rope_pos = p.epi == EPI_QKV_ROPE ? __ldcg(p.rope.n_past_ptr) : 0;
    const unsigned push_ep = p.epi >= EPI_PUSH_F64 ? tp_push_epoch(p) : 0u;
    int crow = row_first, cu = 0;  // compute cursor
    double acc[GEMV_R] = {0.0, 0.0};
    auto step = [&](Regs &g) {
        compute_pass<TYPE, true>(g, cu, nu, p.k, s, acc);
        load_next(g);  // refill this image; the loads fly while the other image is computed
        cu += PU;
        if (cu >= nu) {
            const double tot = warp_sum2_f64(acc[0], acc[1]);  // lanes 0-15: row 0, lanes 16-31: row 1
            if (p.epi == EPI_QKV_ROPE) {
                qkv_rope_store(p, seg, crow, tot, rope_pos);
            } else if (pair) {
                const float gte = (float)__shfl_sync(0xffffffffu, tot, 0), up = (float)__shfl_sync(0xffffffffu, tot, 16);
                if (lane == 0) out[crow] = __fmul_rn(__fdiv_rn(gte, __fadd_rn(1.0f, expf_det(-gte))), up);
            } else if ((lane & 15) == 0) {
                const int row = crow + (lane >> 4);
                if (row < N) {
                    const float v = (float)tot;
                    if (p.epi == EPI_PUSH_F64) tp_push_store(p, row, N, tot, push_ep);
                    else if (p.epi == EPI_PUSH_F32) tp_push_store_f32(p, row, v);
                    else if (p.epi == EPI_STORE_F64) reinterpret_cast<double *>(out)[row] = tot;
                    else if (p.epi == EPI_ADD) out[row] = __fadd_rn(out[row], v);
                    else out[row] = v;
                    if (p.argmax_key) {
                        const unsigned long long key = argmax_key(v, row);
                        best = key > best ? key : best;
                    }
                }
            }
            acc[0] = acc[1] = 0.0;
            cu = 0;
            crow += row_step;
        }
    };
    while (crow < N) {
#pragma unroll
        for (int b = 0; b < NB; b++) {
            if (crow < N) step(g[b]);
        }
    }
    trace_mark(tr, 4);
    if (p.trace && threadIdx.x == 0) {
        unsigned long long now;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
        atomicMax(p.trace + 7, now);
    }
    if (p.argmax_key) {
        const unsigned long long other = __shfl_sync(0xffffffffu, best, 16);
        best = other > best ? other : best;
        if (lane == 0 && best) atomicMax(p.argmax_key, best);
    }
    if (p.epi >= EPI_PUSH_F64) tp_push_signal(p, push_ep);
}

__device__ __forceinline__ int gemv_my_segment(const GemvParams &p) {
    int seg = 0;
    if (p.epi != EPI_SWIGLU) {
        while (seg + 1 < p.n_seg && (int)blockIdx.x >= p.cta_begin[seg + 1]) seg++;
    }
    return seg;
}

template <int TYPE, int NB>
__global__ void __launch_bounds__(GEMV_THREADS, 1) gemv_db_kernel(const GemvParams p) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ double red[GEMV_WARPS];
    gemv_db_body<TYPE, NB>(p, gemv_my_segment(p), smem, red);
}

// attn_q / attn_k (Q4_K) and attn_v (Q6_K in half of the layers of a Q4_K_M file) in ONE launch: a CTA serves one
// matrix, so the type is CTA-uniform and each CTA builds only the activation image its type needs.  Saves a launch
// (hand-over + a second RMSNorm/quantisation prologue) in every such layer.
__global__ void __launch_bounds__(GEMV_THREADS, 1) gemv_db_mixed_kernel(const GemvParams p) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ double red[GEMV_WARPS];
    const int seg = gemv_my_segment(p);
    if (p.seg[seg].type == T_Q6_K) gemv_db_body<T_Q6_K, 2>(p, seg, smem, red);
    else if (p.seg[seg].type == T_Q5_K) gemv_db_body<T_Q5_K, 2>(p, seg, smem, red);
    else gemv_db_body<T_Q4_K, 2>(p, seg, smem, red);
}

size_t gemv_smem_bytes(int k) { return act_image_bytes(k); }

static size_t seg_row_bytes(const WPlanes &w) { return tensor_bytes(w.type, 1, w.k); }

int gemv_plan_grid(GemvParams &p, int n_sms) {
    const int target = n_sms;  // one 16-warp CTA per SM
    if (p.epi == EPI_SWIGLU) {
        const int groups = (p.seg[0].n_rows + GEMV_WARPS - 1) / GEMV_WARPS;
        const int g = groups < target ? groups : target;
        p.cta_begin[0] = 0;
        p.cta_begin[1] = g;
        p.cta_begin[2] = p.cta_begin[3] = g;
        return g;
    }
    double total = 0;
    int groups[3] = {0, 0, 0};
    int total_groups = 0;
    for (int i = 0; i < p.n_seg; i++) {
        total += (double)seg_row_bytes(p.seg[i]) * p.seg[i].n_rows;
        groups[i] = (p.seg[i].n_rows + GEMV_WARPS * GEMV_R - 1) / (GEMV_WARPS * GEMV_R);
        total_groups += groups[i];
    }
    const int grid_target = total_groups < target ? total_groups : target;
    int begin = 0;
    for (int i = 0; i < p.n_seg; i++) {
        p.cta_begin[i] = begin;
        double share = (double)seg_row_bytes(p.seg[i]) * p.seg[i].n_rows / total;
        int c = (int)(share * grid_target + 0.5);
        if (c < 1) c = 1;
        if (c > groups[i]) c = groups[i];
        begin += c;
    }
    for (int i = p.n_seg; i < 4; i++) p.cta_begin[i] = begin;
    return begin;
}

// N-buffered register pipeline (gemv_db_kernel): 2 = default (1.92 -> 1.79 ms/token), 3 spills (2.01), 0 = gemv_kernel.
// AIOS_GEMV_DB=0/2/3 overrides for A/B runs.
static int g_use_db = 2;

void gemv_init() {
    static PerDeviceOnce once;
    once.run([] {
        // one shared-memory carve-out for every kernel of the token pass: kernels with different carve-outs cannot
        // share an SM, which would serialise the programmatic-dependent-launch overlap (measured: 1.89 -> 2.07 ms/token)
        auto prep = [](auto kern) {
            CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
            CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        };
        prep(gemv_kernel<T_Q4_K>); prep(gemv_kernel<T_Q5_K>); prep(gemv_kernel<T_Q6_K>); prep(gemv_kernel<T_Q8_0>);
        prep(gemv_db_kernel<T_Q4_K, 2>); prep(gemv_db_kernel<T_Q5_K, 2>); prep(gemv_db_kernel<T_Q6_K, 2>); prep(gemv_db_kernel<T_Q8_0, 2>);
        prep(gemv_db_kernel<T_Q4_K, 3>); prep(gemv_db_kernel<T_Q6_K, 3>);
        prep(gemv_db_mixed_kernel);
        if (const char *e = getenv("AIOS_GEMV_DB")) g_use_db = atoi(e);
    });
}

// Q4_K, Q5_K and Q6_K matrices may share a launch (gemv_db_mixed_kernel) when K is a whole number of 2048-weight passes
// (70B-class Q4_K_M files: attn_q / attn_k Q4_K next to an attn_v of Q5_K or Q6_K)
bool gemv_can_mix(const GemvParams &p) {
    if (g_use_db != 2 || p.epi == EPI_SWIGLU || (p.k / 64) % 32 != 0) return false;
    for (int i = 0; i < p.n_seg; i++)
        if (p.seg[i].type != T_Q4_K && p.seg[i].type != T_Q5_K && p.seg[i].type != T_Q6_K) return false;
    return true;
}

// All segments of a launch must have one type, except for the K-quant mixes gemv_can_mix() allows.
void launch_gemv(const GemvParams &p_in, int grid, cudaStream_t st, bool pdl) {
    const GemvParams &p = p_in;
    const int t = p.seg[0].type;
    bool mixed = false;
    for (int i = 1; i < p.n_seg; i++) mixed |= p.seg[i].type != t;
    if (mixed && !gemv_can_mix(p)) throw CudaError("launch_gemv: segments of one launch must share a weight type");
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(GEMV_THREADS);
    cfg.dynamicSmemBytes = gemv_smem_bytes(p.k);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    if (mixed) {
        CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_db_mixed_kernel, p));
        return;
    }
    const int pass_units = t == T_Q8_0 ? 64 : 32, unit = t == T_Q8_0 ? 32 : 64;
    if (g_use_db && (p.k / unit) % pass_units == 0) {
        if (g_use_db == 3 && t == T_Q4_K) { CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_db_kernel<T_Q4_K, 3>, p)); return; }
        if (g_use_db == 3 && t == T_Q6_K) { CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_db_kernel<T_Q6_K, 3>, p)); return; }
        switch (t) {
            case T_Q4_K: CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_db_kernel<T_Q4_K, 2>, p)); return;
            case T_Q5_K: CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_db_kernel<T_Q5_K, 2>, p)); return;
            case T_Q6_K: CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_db_kernel<T_Q6_K, 2>, p)); return;
            case T_Q8_0: CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_db_kernel<T_Q8_0, 2>, p)); return;
            default: break;
        }
    }
    switch (t) {
        case T_Q4_K: CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_kernel<T_Q4_K>, p)); break;
        case T_Q5_K: CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_kernel<T_Q5_K>, p)); break;
        case T_Q6_K: CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_kernel<T_Q6_K>, p)); break;
        case T_Q8_0: CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_kernel<T_Q8_0>, p)); break;
        default: throw CudaError("launch_gemv: unsupported weight type");
    }
}

}  // namespace aios
