// pfd.h — descriptors of the persistent fused decode kernel (pfd.cu), built on the host per sequence.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "kernels.h"
#include "layout.cuh"

namespace aios {

enum : int { PH_GEMV = 0, PH_ATTN = 1 };

// One phase = one weight matrix (or the gate/up pair) streamed against one input vector, or one attention.
// Matrices that share an input (q,k,v) are consecutive phases: the first carries `prologue`, the last
// `barrier_after`.
struct PfdPhase {
    int32_t kind;            // PH_GEMV or PH_ATTN
    // ---- GEMV phase
    int32_t type;            // ggml type of w0 (and w1)
    int32_t epi;             // GemvEpilogue; EPI_SWIGLU: unit u = (row u of w0, row u of w1)
    int32_t k;               // input length
    WPlanes w0, w1;
    float *out;
    const float *x;          // f32 input vector [k] ...
    const double *x64;       // ... or the f64 attention accumulators [k] (attn_output projection)
    const float *norm_w;     // RMSNorm weight or null
    float eps;
    int32_t prologue;        // 1: build the activation image(s) at the start of this phase
    int32_t barrier_after;   // 1: grid barrier at the end of this phase
    int32_t embed;           // 1: the input is the embedding row of the current token (first phase of a token)
    int32_t argmax;          // 1: lm_head — the epilogue updates the running arg-max key
    int32_t zero_att;        // 1: zero the f64 attention accumulators at the start of this phase
    int32_t perm_mask;       // images the prologue builds: 1 Q4_K/Q5_K order, 2 Q6_K order, 4 Q8_0
    int32_t units_total;     // work units: row pairs (rows 2u, 2u+1), or rows for EPI_SWIGLU
    // ---- attention phase
    uint16_t *k_cache, *v_cache;
};

struct PfdParams {
    const PfdPhase *phases;
    int32_t n_phases, n_tokens;
    int32_t pos0;            // position of the last token already fed (SeqState.pos at launch)
    int32_t step0;           // first index of out_tokens to write
    int32_t *out_tokens;
    int32_t *state;          // SeqState: {pos, step, cur_token}
    unsigned long long *argmax_key;
    float *x;                // residual stream [n_embd]
    const float *qkv;        // raw projections of the current token
    const uint8_t *tok_embd; // embedding table, GGUF natural layout
    int64_t tok_embd_row_bytes;
    int32_t tok_embd_type;
    int32_t n_embd;
    const float *rope_cos, *rope_sin;
    int32_t n_head, n_kv, hd, n_rot;
    int32_t kv_stride;       // f16 elements between the cache rows of consecutive positions
    float att_scale;
    int32_t att_group;       // CTAs per kv-head
    int32_t att_per_cap;     // max cached positions per CTA
    float *att_max;          // [n_kv][att_group][8]
    double *att_sum;         // [n_kv][att_group][8]
    double *att_acc;         // [n_head*hd] f64 accumulators, zero between uses
    unsigned *att_counters;  // [n_kv] monotonic group barriers (zeroed by the host before the launch)
    unsigned *bar_counter;   // grid barrier (zeroed by the host before the launch)
    int32_t k_max;           // largest GEMV input length (sizes the activation images)
    const double *x64_override;  // null, or where the f64 attention accumulators really are (built-in scratch variant)
    unsigned long long *timing;  // optional [n_phases][4] ns stamps of CTA 0 for the last token: start, after prologue,
                                 // after compute, after barrier (null = off)
};

// Size of the library's built-in exchange scratch (a static __device__ array in pfd.cu) used by
// aios_seq_decode_static instead of per-sequence cudaMalloc'd buffers.
constexpr size_t PFD_STATIC_SCRATCH_BYTES = 48 * 1024;
constexpr int PFD_STATIC_MAX_KV_HEADS = 8;        // att_max / att_sum rows
constexpr int PFD_STATIC_MAX_QDIM = 4096;         // n_head * head_dim (f64 accumulators)
// device address of that array on `device` (made current by the caller); the first call per device resolves it
void *pfd_static_scratch(int device);

size_t pfd_smem_bytes(const PfdParams &p);
int pfd_max_grid(int n_sms, const PfdParams &p);   // 0 if the kernel cannot be made resident
void launch_pfd(const PfdParams &p, int grid, cudaStream_t st);

}  // namespace aios
