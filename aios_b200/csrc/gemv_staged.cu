// gemv_staged.cu — the dequant+GEMV kernel for LARGE matrices (ffn gate/up, ffn down, output.weight: 80 % of the
// bytes of a token): same arithmetic as gemv.cu, different memory pipeline.
//
// gemv.cu keeps a pass of weights in registers, so a warp has nothing in flight while it computes; with ~1.9 us
// of loaded HBM latency that caps a big matrix at ~55 % of the measured copy bandwidth (profiles/r01_v2_gemv.md).
// Here every warp owns a private 3-stage ring in shared memory filled by cp.async (LDGSTS, L1-bypassing, no
// register cost): two passes (2 x 4.6 KB) are always in flight per warp while a third is being consumed —
// 12 warps x 9.2 KB = 110 KB in flight per SM.  A lane reads back exactly the bytes it copied (except the
// block meta words shared by 4 lanes), so the ring needs no barrier beyond a __syncwarp.
// One CTA of 12 warps per SM; grid = number of SMs.
#define GEMV_CORE_WARPS 12
#include "gemv_core.cuh"

namespace aios {

constexpr int ST_STAGES = 3;
constexpr int ST_STAGE_BYTES = 4608;

// lane-iterations per pass, chosen so that 2 rows fit a 4608-byte stage
template <int TYPE> struct StageCfg { static constexpr int ITS = 2; };           // Q4_K: 2*2*(512+512+128) = 4608
template <> struct StageCfg<T_Q5_K> { static constexpr int ITS = 1; };           // 2*(1024+128+256)       = 2816
template <> struct StageCfg<T_Q6_K> { static constexpr int ITS = 1; };           // 2*(1536+128+16)        = 3360
template <> struct StageCfg<T_Q8_0> { static constexpr int ITS = 2; };           // 2*2*(1024+64)          = 4352

template <int TYPE>
struct StageRegs {
    static constexpr int ITS = StageCfg<TYPE>::ITS;
    uint4 a[GEMV_R][ITS], b[GEMV_R][ITS], c[GEMV_R][ITS];
    uint32_t e[GEMV_R][ITS], f[GEMV_R][ITS];
};

template <int TYPE> __device__ __forceinline__ constexpr int slot_bytes() {  // bytes of one (row, lane-iteration) slot
    return TYPE == T_Q4_K ? 1152 : TYPE == T_Q5_K ? 1408 : TYPE == T_Q6_K ? 1680 : 1088;
}

__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8(uint32_t dst, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// start copying one pass (GEMV_R rows x ITS*32 units from unit u0) into a stage
template <int TYPE>
__device__ __forceinline__ void stage_issue(uint32_t stage, const RowPtrs (&rp)[GEMV_R], int u0, int nu) {
    const int lane = threadIdx.x & 31;
    constexpr int ITS = StageCfg<TYPE>::ITS;
    constexpr int SB = slot_bytes<TYPE>();
#pragma unroll
    for (int r = 0; r < GEMV_R; r++) {
#pragma unroll
        for (int it = 0; it < ITS; it++) {
            const int u = u0 + it * 32 + lane;
            const uint32_t s = stage + (r * ITS + it) * SB;
            if (u < nu) {
                cp_async16(s + lane * 16, rp[r].p[0] + (size_t)u * 16);
                cp_async16(s + 512 + lane * 16, rp[r].p[1] + (size_t)u * 16);
                if (TYPE == T_Q4_K || TYPE == T_Q5_K) {
                    if ((lane & 3) == 0) cp_async16(s + 1024 + (lane >> 2) * 16, rp[r].p[2] + (size_t)(u >> 2) * 16);
                    if (TYPE == T_Q5_K) cp_async8(s + 1152 + lane * 8, rp[r].p[3] + (size_t)u * 8);
                } else if (TYPE == T_Q6_K) {
                    cp_async16(s + 1024 + lane * 16, rp[r].p[2] + (size_t)u * 16);
                    cp_async4(s + 1536 + lane * 4, rp[r].p[3] + (size_t)u * 4);
                    if ((lane & 7) == 0) cp_async4(s + 1664 + (lane >> 3) * 4, rp[r].p[4] + (size_t)(u >> 3) * 4);  // d of 2 blocks
                } else {
                    if ((lane & 1) == 0) cp_async4(s + 1024 + (lane >> 1) * 4, rp[r].p[2] + (size_t)(u >> 1) * 4);  // d of 2 blocks
                }
            }
        }
    }
}

// read a landed pass back into registers (each lane reads the bytes it copied; meta / d words are shared)
template <int TYPE>
__device__ __forceinline__ void stage_read(const uint8_t *stage, StageRegs<TYPE> &g) {
    const int lane = threadIdx.x & 31;
    constexpr int ITS = StageCfg<TYPE>::ITS;
    constexpr int SB = slot_bytes<TYPE>();
#pragma unroll
    for (int r = 0; r < GEMV_R; r++) {
#pragma unroll
        for (int it = 0; it < ITS; it++) {
            const uint8_t *s = stage + (r * ITS + it) * SB;
            g.a[r][it] = *reinterpret_cast<const uint4 *>(s + lane * 16);
            g.b[r][it] = *reinterpret_cast<const uint4 *>(s + 512 + lane * 16);
            if (TYPE == T_Q4_K || TYPE == T_Q5_K) {
                g.c[r][it] = *reinterpret_cast<const uint4 *>(s + 1024 + (lane >> 2) * 16);
                if (TYPE == T_Q5_K) {
                    const uint2 h = *reinterpret_cast<const uint2 *>(s + 1152 + lane * 8);
                    g.e[r][it] = h.x; g.f[r][it] = h.y;
                }
            } else if (TYPE == T_Q6_K) {
                g.c[r][it] = *reinterpret_cast<const uint4 *>(s + 1024 + lane * 16);
                g.e[r][it] = *reinterpret_cast<const uint32_t *>(s + 1536 + lane * 4);
                g.f[r][it] = *reinterpret_cast<const uint16_t *>(s + 1664 + (lane >> 2) * 2);
            } else {
                g.e[r][it] = *reinterpret_cast<const uint16_t *>(s + 1024 + lane * 2);
            }
        }
    }
}

template <int TYPE>
__global__ void __launch_bounds__(GEMV_THREADS, 1) gemv_staged_kernel(const GemvParams p) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ double red[GEMV_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int UNIT = TYPE == T_Q8_0 ? 32 : 64;
    constexpr int PASS_UNITS = StageCfg<TYPE>::ITS * 32;
    int seg = 0;
    if (p.epi != EPI_SWIGLU) {
        while (seg + 1 < p.n_seg && (int)blockIdx.x >= p.cta_begin[seg + 1]) seg++;
    }
    const WPlanes &W = p.seg[seg];
    const WPlanes &W1 = p.seg[p.epi == EPI_SWIGLU ? 1 : seg];
    const int cta_local = (int)blockIdx.x - p.cta_begin[seg];
    const int cta_count = p.cta_begin[seg + 1] - p.cta_begin[seg];
    const size_t img = (act_image_bytes(p.k) + 15) / 16 * 16;
    const ActSmem s = carve(smem, p.k);
    uint8_t *ring = smem + img + (size_t)warp * ST_STAGES * ST_STAGE_BYTES;
    const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(ring);
    const int N = W.n_rows;
    const int nu = p.k / UNIT;
    const int nchunks = (nu + PASS_UNITS - 1) / PASS_UNITS;
    const bool pair = p.epi == EPI_SWIGLU;
    const int rows_per_pass = pair ? 1 : GEMV_R;
    const int row_first = (cta_local * GEMV_WARPS + warp) * rows_per_pass;
    const int row_step = cta_count * GEMV_WARPS * rows_per_pass;
    const int n_units = row_first < N ? (N - row_first + row_step - 1) / row_step : 0;  // row pairs of this warp
    const int n_pass = n_units * nchunks;

    auto ptrs = [&](int row0, RowPtrs (&rp)[GEMV_R]) {
        if (pair) {
            rp[0] = row_ptrs<TYPE>(W, row0);
            rp[1] = row_ptrs<TYPE>(W1, row0);
        } else {
            rp[0] = row_ptrs<TYPE>(W, row0);
            rp[1] = row_ptrs<TYPE>(W, min(row0 + 1, N - 1));
        }
    };
    // producer cursor: pass index -> (row pair, chunk)
    int pi = 0;  // next pass to issue
    auto issue_next = [&]() {
        if (pi < n_pass) {
            const int unit = pi / nchunks, chunk = pi - unit * nchunks;
            RowPtrs rp[GEMV_R];
            ptrs(row_first + unit * row_step, rp);
            stage_issue<TYPE>(ring_s + (uint32_t)(pi % ST_STAGES) * ST_STAGE_BYTES, rp, chunk * PASS_UNITS, nu);
        }
        cp_async_commit();  // (an empty group keeps the group count uniform at the tail)
        pi++;
    };
    // weights do not depend on the previous kernel: fill the ring before the dependency wait and the prologue
#pragma unroll
    for (int i = 0; i < ST_STAGES - 1; i++) issue_next();
    l2_prefetch_next(p);
    asm volatile("griddepcontrol.wait;" ::: "memory");

    if (TYPE == T_Q6_K) prep_activations<PERM_Q6>(p.x, p.norm_w, p.eps, p.k, s, red);
    else if (TYPE == T_Q8_0) prep_activations<PERM_Q80>(p.x, p.norm_w, p.eps, p.k, s, red);
    else prep_activations<PERM_Q4>(p.x, p.norm_w, p.eps, p.k, s, red);
    asm volatile("griddepcontrol.launch_dependents;");

    float *out = p.out[seg];
    unsigned long long best = 0ull;
    double acc[GEMV_R] = {0.0, 0.0};
    for (int ci = 0; ci < n_pass; ci++) {
        cp_async_wait<ST_STAGES - 2>();  // pass ci has landed (for this lane's own copies)
        __syncwarp();                    // ... and for the meta / d words other lanes copied; it also orders every lane's
                                         // reads of the stage consumed last iteration before that stage is refilled below
        StageRegs<TYPE> g;
        stage_read<TYPE>(ring + (size_t)(ci % ST_STAGES) * ST_STAGE_BYTES, g);
        issue_next();                    // refills stage (ci+2)%3 == the one consumed at iteration ci-1
        const int unit = ci / nchunks, chunk = ci - unit * nchunks;
        compute_pass<TYPE, false>(g, chunk * PASS_UNITS, nu, p.k, s, acc);
        if (chunk + 1 == nchunks) {
            const int row0 = row_first + unit * row_step;
            const double tot = warp_sum2_f64(acc[0], acc[1]);  // lanes 0-15: row 0, lanes 16-31: row 1
            if (pair) {
                const float gte = (float)__shfl_sync(0xffffffffu, tot, 0), up = (float)__shfl_sync(0xffffffffu, tot, 16);
                if (lane == 0) out[row0] = __fmul_rn(__fdiv_rn(gte, __fadd_rn(1.0f, expf_det(-gte))), up);
            } else if ((lane & 15) == 0) {
                const int row = row0 + (lane >> 4);
                if (row < N) {
                    const float v = (float)tot;
                    if (p.epi == EPI_ADD) out[row] = __fadd_rn(out[row], v);
                    else out[row] = v;
                    if (p.argmax_key) {
                        const unsigned long long key = argmax_key(v, row);
                        best = key > best ? key : best;
                    }
                }
            }
            acc[0] = 0.0;
            acc[1] = 0.0;
        }
    }
    cp_async_wait<0>();
    if (p.argmax_key) {
        const unsigned long long other = __shfl_sync(0xffffffffu, best, 16);
        best = other > best ? other : best;
        if (lane == 0 && best) atomicMax(p.argmax_key, best);
    }
}

static size_t staged_smem(int k) {
    return (act_image_bytes(k) + 15) / 16 * 16 + (size_t)GEMV_WARPS * ST_STAGES * ST_STAGE_BYTES;
}

void gemv_staged_init() {
    static bool done = false;
    if (done) return;
    CUDA_CHECK(cudaFuncSetAttribute(gemv_staged_kernel<T_Q4_K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    CUDA_CHECK(cudaFuncSetAttribute(gemv_staged_kernel<T_Q5_K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    CUDA_CHECK(cudaFuncSetAttribute(gemv_staged_kernel<T_Q6_K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    CUDA_CHECK(cudaFuncSetAttribute(gemv_staged_kernel<T_Q8_0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    done = true;
}

// Is the launch big enough to keep the ring busy, and are the 4-byte side-band copies aligned?
bool gemv_staged_usable(const GemvParams &p, int n_sms) {
    if (staged_smem(p.k) > 220 * 1024) return false;
    const int t = p.seg[0].type;
    const int nb = p.k / 256;
    if (t == T_Q6_K && (nb & 1)) return false;  // d words are copied two blocks at a time
    const int unit = t == T_Q8_0 ? 32 : 64;
    const int pass_units = (t == T_Q5_K || t == T_Q6_K) ? 32 : 64;
    const long long nchunks = (p.k / unit + pass_units - 1) / pass_units;
    long long units = 0;
    if (p.epi == EPI_SWIGLU) units = p.seg[0].n_rows;
    else for (int i = 0; i < p.n_seg; i++) units += (p.seg[i].n_rows + 1) / 2;
    return units * nchunks >= (long long)n_sms * GEMV_WARPS * 3;  // at least ~3 passes per warp
}

void launch_gemv_staged(const GemvParams &p_in, int n_sms, cudaStream_t st, bool pdl) {
    GemvParams p = p_in;
    // one CTA per SM; segments share the CTAs by bytes
    if (p.epi == EPI_SWIGLU || p.n_seg == 1) {
        p.cta_begin[0] = 0;
        for (int i = 1; i < 4; i++) p.cta_begin[i] = n_sms;
    } else {
        double total = 0;
        for (int i = 0; i < p.n_seg; i++) total += (double)tensor_bytes(p.seg[i].type, p.seg[i].n_rows, p.seg[i].k);
        int begin = 0;
        for (int i = 0; i < p.n_seg; i++) {
            p.cta_begin[i] = begin;
            int c = (int)((double)tensor_bytes(p.seg[i].type, p.seg[i].n_rows, p.seg[i].k) / total * n_sms + 0.5);
            if (c < 1) c = 1;
            if (i == p.n_seg - 1) c = n_sms - begin;
            begin += c;
        }
        for (int i = p.n_seg; i < 4; i++) p.cta_begin[i] = begin;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(n_sms);
    cfg.blockDim = dim3(GEMV_THREADS);
    cfg.dynamicSmemBytes = staged_smem(p.k);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    switch (p.seg[0].type) {
        case T_Q4_K: CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_staged_kernel<T_Q4_K>, p)); break;
        case T_Q5_K: CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_staged_kernel<T_Q5_K>, p)); break;
        case T_Q6_K: CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_staged_kernel<T_Q6_K>, p)); break;
        case T_Q8_0: CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_staged_kernel<T_Q8_0>, p)); break;
        default: throw CudaError("launch_gemv_staged: unsupported weight type");
    }
}

}  // namespace aios
