// kernels.h — host-callable launchers of the CUDA kernels (internal; not part of the C ABI).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "layout.cuh"

namespace aios {

// ---- dequant.cu
void launch_dequantize(int type, const void *d_blocks, int64_t n, float *d_out, cudaStream_t st);
void launch_embed_row(int type, const void *d_table, int64_t row_bytes, int k, const int32_t *d_tok, float *d_x,
                      int32_t *d_pos, int pos_value, int32_t *d_cur_token, cudaStream_t st);
// T rows at once: x[t] = embedding row of d_toks[t] (row stride ldx floats)
void launch_embed_rows(int type, const void *d_table, int64_t row_bytes, int k, const int32_t *d_toks, int T, float *d_x,
                       int64_t ldx, cudaStream_t st);
void launch_repack(int type, const void *d_raw, int64_t n_rows, int k, uint8_t *const planes[5], cudaStream_t st);

// ---- gemv.cu : fused [RMSNorm] -> activation quantisation -> dequant+GEMV -> epilogue
enum GemvEpilogue : int {
    EPI_STORE = 0,   // out[n] = dot
    EPI_ADD = 1,     // out[n] += dot            (residual stream update, in place)
    EPI_SWIGLU = 2,  // out[n] = silu(dot(seg0 row n)) * dot(seg1 row n)
    // segments = (attn_q, attn_k, attn_v): RoPE + f16 rounding in the epilogue; q -> out[0] (f32 holding f16 values),
    // k and v -> row *n_past_ptr of the layer's KV cache.  Attention then reads every position from the cache.
    EPI_QKV_ROPE = 3,
    // out[n] (a double array) = the row's double sum, unrounded: the partial of a K-split matrix whose other parts live on
    // other GPUs (tensor parallelism): the f64 partials are all-reduced and rounded to f32 once, like the unsplit sum
    EPI_STORE_F64 = 4,
    // the same partial sums, pushed straight into every rank's receive slab over NVLink (plain peer stores), followed by
    // an arrival flag from the launch's last CTA: the all-reduce's transport rides in the GEMV (TpPush; tp.cu)
    EPI_PUSH_F64 = 5,
    // out[n] = dot on every rank: this rank's slice of the logits stored into every rank's logits buffer (push.dst as
    // float pointers, already offset to this rank's slice; no double buffering: see tp.cu), then the arrival flag
    EPI_PUSH_F32 = 6,
};

// EPI_PUSH_F64: where this rank's partials go.  An all-reduce has an epoch = *tick * per_token + index (tick: the
// device-side count of token passes started, so that a captured launch needs no per-call arguments); slabs are double
// buffered by the epoch's parity, flags carry the epoch of the last complete push.
constexpr int TP_MAX_RANKS = 8;
struct TpPush {
    double *dst[TP_MAX_RANKS];     // rank r's slab for THIS rank's partials: [2][n_rows] doubles (r = own rank included)
    unsigned *flag[TP_MAX_RANKS];  // rank r's arrival flag for this rank
    unsigned *done;                // this GPU's finished-CTA counter (zero between launches)
    const unsigned *tick;
    int32_t n, per_token, index;
};

struct RopeKv {              // EPI_QKV_ROPE only
    const float *cos, *sin;  // [n_ctx][n_rot/2]
    const int32_t *n_past_ptr;
    uint16_t *k_cache, *v_cache;  // this layer, this sequence: row of position j at + j * kv_stride (f16 elements)
    int32_t n_rot, hd, kvd, kv_stride;
};

struct GemvParams {
    int32_t n_seg;          // 1..3 weight matrices sharing the same input vector
    int32_t epi;            // GemvEpilogue
    int32_t k;              // input length (all segments)
    int32_t cta_begin[4];   // CTA ranges per segment: segment i owns CTAs [cta_begin[i], cta_begin[i+1])
    WPlanes seg[3];
    float *out[3];
    const float *x;         // f32 input vector [k]
    const float *norm_w;    // if non-null: x <- rmsnorm(x) * norm_w first (llama attn_norm / ffn_norm / output_norm)
    float eps;
    unsigned long long *argmax_key;  // if non-null (EPI_STORE, 1 segment): running packed (value,index) max
    // timeline of CTA 0 (null = off; gemv_db_kernel): [0] entry [1] dependency wait over [2] activations ready
    // [3] launch kind (1 = GEMV, 2 = attention) [4] CTA 0 done [7] last CTA done
    unsigned long long *trace;
    RopeKv rope;
    // ready-made activation image in global memory (quant_rows_kernel, gemm_q.cu), in this launch's shared-memory
    // image layout: the prologue becomes a copy and x / norm_w are not read
    const uint8_t *x_image;
    int32_t evict_first;    // stream this launch's weights with the L2 evict-first policy
    TpPush push;            // EPI_PUSH_F64 only
};

// fills cta_begin[] and returns the grid size; n_sms from the device
int gemv_plan_grid(GemvParams &p, int n_sms);
size_t gemv_smem_bytes(int k);
void launch_gemv(const GemvParams &p, int grid, cudaStream_t st, bool pdl = false);
bool gemv_can_mix(const GemvParams &p);  // may Q4_K and Q6_K segments of p ride in one launch?
void gemv_init();  // one-time function attributes (dynamic smem)

// ---- gemm_q.cu : multi-token (batched decode / prompt chunk) dequant GEMM on int8 tensor cores, bit-identical to T
// GEMV calls.  Per-token Q8_K / Q8_0 activation images in global memory (optionally RMS-normed first); x row stride
// ldx floats.
constexpr int GEMM_MAX_TOKENS = 64;
void launch_quant_rows(int weight_type, const float *x, int64_t ldx, const float *norm_w, float eps, int k, int T,
                       uint8_t *images, cudaStream_t st);
size_t gemm_images_bytes(int k, int T);  // allocate for gemm_pad_tokens(T) tokens
int gemm_pad_tokens(int T);              // tokens whose images must exist for a T-token launch: T rounded up to 8 (64 above 32)
bool gemm_q_usable(int type, int k);     // Q4_K / Q6_K / Q8_0
// out[t][row] (row stride ldo floats) = or += dot, t < T <= 64; epi = EPI_STORE or EPI_ADD
void launch_gemm_q(const WPlanes &w, const uint8_t *images, int T, float *out, int64_t ldo, int epi, cudaStream_t st);
// 1..3 matrices that share the input tokens in ONE launch (q/k/v, gate/up): images[i] = the image set in the
// permutation of seg[i]'s type (Q4_K/Q5_K, Q6_K and Q8_0 weights read differently ordered images)
struct GemmMulti {
    int32_t n_seg, T, epi;
    WPlanes seg[3];
    float *out[3];
    const uint8_t *images[3];
    int64_t ldo;
    // one segment, tensor-parallel K shards: instead of touching out, store the rows' unrounded double sums to
    // part[t * n_rows + row] (the partials are all-reduced across the ranks and rounded once, like EPI_STORE_F64)
    double *part = nullptr;
};
void launch_gemm_multi(const GemmMulti &m, cudaStream_t st);
// x[t][i] += (float) part[t * n + i]   (t < T): the all-reduced partials of a K-split GEMM, rounded once
void launch_add_part_rows(float *x, int64_t ldx, const double *part, int n, int T, cudaStream_t st);
// h[t][0..ff) = silu(gu[t][0..ff)) * gu[t][ff..2ff)
void launch_swiglu_rows(const float *gu, int64_t ldgu, int ff, int T, float *h, int64_t ldh, cudaStream_t st);

// ---- prefill_tc.cu : the prefill GEMM on tcgen05 tensor cores (TMEM accumulators, TMA-staged f16 activations, weights
// dequantised into shared memory); f16 x f16 products, f32 accumulation — a stated tolerance, not the bit-exact path
bool prefill_tc_usable(int type, int k);  // Q4_K / Q6_K / Q8_0
// out[t][row] (row stride ldo floats) = or += sum_k xh[t][k] * dq(w)[row][k], t < T; xh = f16 [T][k] in this GPU's memory
void launch_prefill_tc(const WPlanes &w, const __half *xh, int T, float *out, int64_t ldo, int epi, int n_sms, cudaStream_t st);
// xh[t] = f16(rmsnorm(x[t]) * norm_w) (norm_w null: plain conversion); x row stride ldx floats
void launch_rows_to_f16(const float *x, int64_t ldx, const float *norm_w, float eps, int k, int T, __half *xh, cudaStream_t st);
// hh[t][i] = f16(silu(gu[t][i]) * gu[t][ff+i])
void launch_swiglu_rows_f16(const float *gu, int64_t ldgu, int ff, int T, __half *hh, cudaStream_t st);

// ---- attention_prefill.cu : RoPE + KV append for T tokens, and causal flash attention over the KV range (tensor-core
// prefill path; f16 operands, f32 soft-max and accumulation)
void launch_rope_kv_rows(const float *qkv, __half *qh, uint16_t *k_cache, uint16_t *v_cache, const float *cos, const float *sin, int T,
                         int pos0, int n_head, int n_kv, int hd, int n_rot, int kv_stride, cudaStream_t st);
void launch_attn_prefill(const __half *qh, __half *out, const uint16_t *k_cache, const uint16_t *v_cache, int T, int pos0, int n_head,
                         int n_kv, int hd, int kv_stride, cudaStream_t st);

// ---- attention.cu : RoPE + KV write + flash-decode over the f16 KV cache (one new token)
struct AttnParams {
    const float *qkv;        // [n_head*hd + 2*n_kv*hd] raw projections of the new token
    float *out;              // [n_head*hd]
    uint16_t *k_cache;       // this layer: row of position j = [n_kv*hd] f16 at + j * kv_stride
    uint16_t *v_cache;
    int32_t kv_stride;       // f16 elements between the rows of consecutive positions (position-major KV pages)
    const float *rope_cos;   // [n_ctx][hd/2]
    const float *rope_sin;
    const int32_t *n_past_ptr;  // device: number of cached positions before this token
    float *scratch;          // [n_head][n_split][hd+2] partial (m, l, o[hd])
    uint32_t *counters;      // [n_kv] arrival counters (zero on entry, zero on exit)
    int32_t n_head, n_kv, hd, n_rot, n_split, n_ctx;
    float scale;
    unsigned long long *trace;  // split kernel, CTA 0 (null = off): [0] entry [1] wait over [2] scores done [4] end
    // split kernel: the q/k/v GEMV already applied RoPE, rounded to f16 and appended this token's K/V rows
    // (EPI_QKV_ROPE): qkv holds the final q only, every position comes from the cache
    int32_t kv_ready;
    int32_t stage_cap;  // set by launch_attn_decode_split: cache rows staged in shared memory (kv_ready launches)
    // Batched decode (split kernel only): n_lanes > 0 -> blockIdx.y = lane, one new token per live sequence.  Lane l
    // reads qkv + l * lane_qkv_stride, writes out + l * lane_out_stride, is at position lane_pos[l] and keeps its KV
    // rows in its own range: k_cache / v_cache above are then OFFSETS (this layer's K and V rows inside a position,
    // i.e. pointers relative to null) added to lane_kv[l].
    int32_t n_lanes;
    int64_t lane_qkv_stride, lane_out_stride;
    const int32_t *lane_pos;
    uint16_t *const *lane_kv;
};
// ---- attention_long.cu : one token over a long context — CTA = (kv head, slice of the positions), the GQA group's
// query heads sharing every K / V row; one launch, the CTAs of a kv head meeting at two device-counter barriers.  Same
// arithmetic, bit for bit.
struct AttnLongParams {
    const float *qkv;
    float *out;
    const uint16_t *k_cache, *v_cache;
    int32_t kv_stride;
    const int32_t *n_past_ptr;
    float scale;
    int32_t n_head, n_kv, hd, n_ctx, n_slices, chunk_cap;
    float *smax;         // [n_head][n_slices] slice maxima
    double *ssum;        // [n_head][n_slices] slice shares of the row sums
    double *opart;       // [n_slices][n_head][hd] partial P.V sums
    unsigned *counters;  // [n_kv][3]: two barrier counters and the finished-slice counter (they only ever count up)
    unsigned long long *trace;  // CTA (0, 0), null = off: [0] entry [1] wait over [2] scores done [3] = 2 [4] CTA done [7] last CTA done
};
bool attn_long_usable(int n_head, int n_kv, int n_ctx, int hd);
int attn_long_slices(int n_kv, int n_sms);
size_t attn_long_scratch_bytes(int n_head, int n_kv, int n_ctx, int hd, int n_slices);  // zero it once
// a = the AttnParams of a kv_ready split launch (qkv, out, caches, n_past_ptr, shapes, scale)
void launch_attn_long(const AttnParams &a, void *scratch, int n_slices, cudaStream_t st, bool pdl);

// mode: 0 = pick (split kernel when usable, else cluster); 2 = cluster kernel
void launch_attn_decode(const AttnParams &p, cudaStream_t st, bool pdl = false, int mode = 0);
// one wave of independent (query head, head-dim slice) CTAs, launched with PDL (attention_split.cu)
bool attn_split_usable(int n_ctx, int hd);
void attn_split_init();
void launch_attn_decode_split(const AttnParams &p, cudaStream_t st, bool pdl);
void attn_init();  // one-time function attributes (dynamic smem)
size_t attn_scratch_floats(int n_head, int n_split, int hd);

// ---- attention_batch.cu : decode attention for many (sequence, token) lanes at once, one CTA per (lane, kv head)
struct AttnBatchParams {
    const float *qkv;          // kv_ready = 0: [lane] raw q|k|v projections, lane stride lane_qkv_stride floats
    const __half *qh;          // kv_ready = 1: [lane][n_head*hd] q, roped and rounded
    float *out;                // [lane] attention output, lane stride lane_out_stride floats
    const float *rope_cos, *rope_sin;
    const int32_t *lane_pos;   // device [lanes]: cached positions before the lane's token (= its position)
    uint16_t *const *lane_kv;  // device [lanes]: base of each lane's KV range, or null: every lane uses kv_base
    uint16_t *kv_base;
    int64_t k_off, v_off;      // this layer's K / V rows inside a position block (f16 elements)
    int64_t lane_qkv_stride, lane_out_stride;
    int32_t n_head, n_kv, hd, n_rot, n_ctx, kv_stride, kv_ready;
    float scale;
};
bool attn_batch_usable(int n_head, int n_kv, int n_ctx, int hd);
void launch_attn_batch(const AttnBatchParams &p, int n_lanes, cudaStream_t st);

// ---- sampler.cu
struct SampleParams {
    const float *logits;
    int32_t n_vocab;
    float temperature;
    int32_t top_k;
    float top_p, min_p;
    uint64_t seed;
    int32_t *pos_ptr;        // device state: position of the token being fed; incremented by one
    int32_t *step_ptr;       // device state: index into out_tokens, incremented
    int32_t *cur_token;      // device: sampled token (input of the next step)
    int32_t *out_tokens;     // device array
    int32_t out_cap;
    const uint8_t *mask;     // optional: bit i of byte i/8 set = token i may be sampled (JSON mode); null = all
    int32_t *cand_out;       // optional (tests): the K candidate ids, best first
    int32_t *counts_out;     // optional (tests): {K, n after top-p, n after min-p}
};
void launch_sample(const SampleParams &p, cudaStream_t st);
// batched decode: one CTA per lane; lane l samples from logits + l * n_vocab with its own parameters
struct SampleLane {
    float temperature;
    int32_t top_k;
    float top_p, min_p;
    uint64_t seed;
    const uint8_t *mask;   // device pointer or null
};
struct SampleBatchParams {
    const float *logits;      // [n_lanes][n_vocab]
    int32_t n_vocab;
    const SampleLane *lanes;  // device array
    int32_t *pos;             // device [n_lanes]: position of the last token fed; incremented
    int32_t *tok;             // device [n_lanes]: sampled token (input of the forward pass that follows)
};
void launch_sample_batch(const SampleBatchParams &p, int n_lanes, cudaStream_t st);

}  // namespace aios
