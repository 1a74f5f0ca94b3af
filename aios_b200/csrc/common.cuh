// common.cuh — small device/host helpers shared by every kernel TU (sm_100a only).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <stdexcept>
#include <string>

#define AIOS_QK_K 256
#define AIOS_QK8_0 32

namespace aios {

struct CudaError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

inline void cuda_check(cudaError_t e, const char *what, const char *file, int line) {
    if (e != cudaSuccess) {
        throw CudaError(std::string(what) + ": " + cudaGetErrorString(e) + " (" + file + ":" + std::to_string(line) + ")");
    }
}
#define CUDA_CHECK(x) ::aios::cuda_check((x), #x, __FILE__, __LINE__)

// Function attributes (dynamic shared-memory limit, carve-out) are per-DEVICE settings: a process that drives several
// GPUs (one engine each) must set them once on each.  run(f) calls f the first time it is reached with a given
// device current; thread-safe.
struct PerDeviceOnce {
    std::mutex mu;
    bool done[64] = {};
    template <typename F>
    void run(F &&f) {
        int dev = 0;
        CUDA_CHECK(cudaGetDevice(&dev));
        std::lock_guard<std::mutex> lk(mu);
        if (dev < 0 || dev >= 64) { f(); return; }
        if (!done[dev]) { f(); done[dev] = true; }
    }
};

// Programmatic dependent launch for a whole sequence of launches (the batched decode step): inside a PdlScope(true)
// launch_k gives every kernel the programmatic-stream-serialization attribute.  EVERY kernel launched that way must
// execute pdl_wait() before it touches anything an earlier kernel of the stream wrote (the wait also carries the
// ordering on to the kernels after it); pdl_launch_dependents() lets the next kernel's CTAs come up early.
inline thread_local bool g_launch_pdl = false;
struct PdlScope {
    bool prev;
    explicit PdlScope(bool on) : prev(g_launch_pdl) { g_launch_pdl = on; }
    ~PdlScope() { g_launch_pdl = prev; }
};
#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
inline void launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args &&...args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = g_launch_pdl ? 1 : 0;
    CUDA_CHECK(cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...));
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }
#endif

#ifdef __CUDACC__
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// 128-bit streaming load of weights: read-only path, do not allocate in L1 (each byte is used once)
__device__ __forceinline__ uint4 ldg_stream(const void *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
// L2 eviction policy for streamed weights: every weight byte is used once per token, so it is marked evict-first and
// does not push the activations, the KV rows and the next launch's first weights out of L2.  0 = no hint.
__device__ __forceinline__ uint64_t l2_policy(bool evict_first) {
    uint64_t pol = 0;
    if (evict_first) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint4 ldg_stream(const void *p, uint64_t pol) {
    if (pol == 0) return ldg_stream(p);
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ uint32_t ldg_stream_u32(const void *p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ uint16_t ldg_stream_u16(const void *p) {
    uint16_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(r) : "l"(p));
    return r;
}

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// expf with a fixed operation sequence, bit-identical to oracle_expf() in oracle/aios_oracle.c
// (CUDA's and glibc's expf differ in the last ulp, which is enough to flip a Q8 rounding
// downstream): n = rint(x*log2e), two-step Cody-Waite reduction, degree-6 Horner, scale by 2^n.
__device__ __forceinline__ float expf_det(float x) {
    if (x < -87.0f) return 0.0f;
    if (x > 88.0f) return __int_as_float(0x7f800000);
    const float t = __fmul_rn(x, 1.44269504088896341f);
    const float n = rintf(t);
    float r = __fmaf_rn(n, -0.693145751953125f, x);
    r = __fmaf_rn(n, -1.42860682030941723212e-6f, r);
    float p = 1.0f / 720.0f;
    p = __fmaf_rn(p, r, 1.0f / 120.0f);
    p = __fmaf_rn(p, r, 1.0f / 24.0f);
    p = __fmaf_rn(p, r, 1.0f / 6.0f);
    p = __fmaf_rn(p, r, 0.5f);
    p = __fmaf_rn(p, r, 1.0f);
    p = __fmaf_rn(p, r, 1.0f);
    const int ni = (int)n;
    return __fmul_rn(p, __int_as_float((ni + 127) << 23));
}

// launch timeline (aios_seq_trace): 8 u64 per launch, nanoseconds of %globaltimer, written by CTA 0 only
__device__ __forceinline__ void trace_mark(unsigned long long *trace, int slot) {
    if (trace) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        trace[slot] = t;
    }
}
__device__ __forceinline__ float h2f_bits(uint16_t h) { return __half2float(__ushort_as_half(h)); }
__device__ __forceinline__ uint16_t f2h_bits(float f) { return __half_as_ushort(__float2half_rn(f)); }

__device__ __forceinline__ int dp4a_su(uint32_t a_signed, uint32_t b_unsigned, int c) {
    // signed x unsigned is not needed: all our "unsigned" operands are < 128, so s8 x s8 is exact
    return __dp4a((int)a_signed, (int)b_unsigned, c);
}
#endif

}  // namespace aios
