// engine.h — host-side runtime: engine (device + stream), model (weights in HBM, RoPE table,
// KV-cache slots) and sequence (device-resident decode state + captured CUDA graph).
// This is what replaces a llama-server child process (reference: runtime/src/model_manager.rs
// :187-204 spawns one per model; runtime/src/inference.rs:94-186 proxies requests to it).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/aios_b200.h"
#include "gguf.h"
#include "kernels.h"
#include "kvpool.h"
#include "layout.cuh"
#include "pfd.h"

namespace aios {

struct DevBuf {
    void *p = nullptr;
    size_t bytes = 0;
    DevBuf() = default;
    explicit DevBuf(size_t n);
    ~DevBuf();
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    DevBuf(DevBuf &&o) noexcept : p(o.p), bytes(o.bytes) { o.p = nullptr; o.bytes = 0; }
    DevBuf &operator=(DevBuf &&o) noexcept;
    template <typename T>
    T *as() const { return static_cast<T *>(p); }
};

struct Engine {
    int device = 0;
    int n_sms = 148;
    int max_batch = 1;
    int flags = 0;
    // streamed weights carry the L2 evict-first policy (bytes used once should not push out activations, KV rows or
    // the next launch's first weights); AIOS_EVICT_FIRST=0/1 overrides for A/B runs
    bool evict_first = true;
    // contexts of at least this many positions decode with the position-split attention (AIOS_ATTN_LONG_MIN overrides;
    // measured crossover against the latency kernel: DESIGN.md §4)
    int attn_long_min = 2048;
    cudaStream_t stream = nullptr;
    std::mutex mu;  // serialises use of the stream by API calls
    explicit Engine(const aios_engine_opts &o);
    ~Engine();
    void make_current() const;
};

struct DevMatrix {
    WPlanes w{};      // planes layout (layout.cuh)
    DevBuf storage;   // one allocation holding all planes
    size_t src_bytes = 0;
};

struct Layer {
    DevMatrix wq, wk, wv, wo, w_gate, w_up, w_down;
    DevBuf attn_norm, ffn_norm;  // f32
};

struct HParams {
    int n_layer = 0, n_embd = 0, n_head = 0, n_head_kv = 0, head_dim = 0, n_ff = 0, n_vocab = 0, n_ctx = 0, n_rot = 0;
    int file_type = 0;
    float rms_eps = 1e-5f, rope_freq_base = 10000.f;
};

struct Seq;

// activations of one prompt chunk on the tensor-core prefill path (prefill_tc.cu): f32 residual stream and projections,
// f16 GEMM inputs.  One per model, used under the engine lock, grown to the longest chunk seen.
struct TcScratch {
    int cap = 0;  // tokens
    DevBuf X, Xh, QKV, Qh, ATTh, GU, Hh;
};

struct Model {
    Engine *eng;
    std::string name, arch, path;
    HParams hp;
    std::unique_ptr<GgufFile> file;
    std::vector<Layer> layers;
    DevBuf tok_embd_raw;  // natural GGUF layout: one row is gathered per token
    int tok_embd_type = 0;
    int64_t tok_embd_row_bytes = 0;
    DevMatrix output;     // output.weight (or the tied embedding) in planes layout
    DevBuf output_norm;
    DevBuf rope_cos, rope_sin;  // [n_ctx][n_rot/2], host-computed (glibc cosf/sinf, as the oracle)
    // KV cache: a pool of physical pages; every sequence maps pages into its own virtual range (kvpool.h).
    // The caller may hold max_batch sequences (aios_seq_create, aios_generate); the library's own sequences
    // (aios_forward_logits' scratch, the request scheduler's batch lanes) are not counted against that budget.
    KvPool kv_pool;
    TcScratch tc;
    bool tc_ok = false;   // every matrix has a type the tensor-core prefill GEMM reads (Q4_K / Q6_K / Q8_0)
    int user_seqs = 0;    // live caller-visible sequences (guarded by mu)
    int live_seqs = 0;    // all live sequences (guarded by mu)
    int64_t weight_bytes = 0, weight_bytes_per_token = 0;
    std::mutex mu;
    int attn_splits = 16;

    // tp_size > 1: this is shard tp_rank of a tensor-parallel model (tp.cu): q/k/v, gate/up and the lm_head hold this
    // rank's rows (heads / FF columns / vocabulary slice), attn_output and ffn_down this rank's K slice; hp holds the
    // LOCAL head and FF counts, n_embd and n_vocab stay global
    Model(Engine *e, const std::string &name, const std::string &path, int n_ctx, int tp_rank = 0, int tp_size = 1);
    int tp_rank = 0, tp_size = 1;
    int vocab_local = 0;  // rows of the lm_head held here (= n_vocab without tensor parallelism)
    ~Model();
    // position-major KV layout: [position][layer][K|V][n_kv*hd] f16
    size_t kvd() const { return (size_t)hp.n_head_kv * hp.head_dim; }
    size_t kv_pos_elems() const { return (size_t)hp.n_layer * 2 * kvd(); }   // f16 elements per position = row stride
    size_t kv_pos_bytes() const { return kv_pos_elems() * 2; }

   private:
    // rows [row0, row0 + n_rows) x weights [k0, k0 + k_len) of the tensor (n_rows < 0: all rows, k_len < 0: the whole K)
    void upload_matrix(const GgufTensor &t, DevMatrix &dst, DevBuf &staging, void *pinned, size_t pinned_bytes, int64_t row0 = 0,
                       int64_t n_rows = -1, int64_t k0 = 0, int64_t k_len = -1);
};

struct SeqState {  // lives in device memory
    int32_t pos;        // position of the token being processed (= #cached positions before it); -1 when empty
    int32_t step;       // tokens generated so far in the current decode call
    int32_t cur_token;  // token fed at `pos`
    int32_t pad;
    unsigned long long argmax_key;
};

// Tensor-parallel prompt chunks: what a shard does with the f64 partial sums of a K-split projection — all-reduce them
// over the ranks in place (tp.cu); the chunk then adds them, rounded once, to the residual rows.
struct ChunkReduce {
    virtual void all_reduce(double *part, size_t count, cudaStream_t st) = 0;
    virtual ~ChunkReduce() = default;
};

// one launch of the token pass (Seq::build_token_pass)
struct PassLaunch {
    int kind = 0;  // 0 GEMV, 1 attention, 2 attention over a long context (attention_long.cu: two kernels)
    GemvParams g{};
    AttnParams a{};
    int grid = 0;
    bool pdl = false;
    bool reduce_after = false;  // tensor parallel: g.out[0] holds f64 partial sums to all-reduce and add to x
};

struct Seq {
    Model *m;
    bool internal;   // one of the library's own sequences: KV pages are mapped as the position grows
    KvSpace kv;      // this sequence's virtual KV range (caller-visible sequences map their whole context up front)
    int n_past = 0;  // host mirror: number of positions in the KV cache
    uint16_t *k_cache(int layer) const { return kv.as<uint16_t>() + ((size_t)layer * 2 + 0) * m->kvd(); }
    uint16_t *v_cache(int layer) const { return kv.as<uint16_t>() + ((size_t)layer * 2 + 1) * m->kvd(); }
    void ensure_kv(int n_positions);  // back positions [0, n_positions) with pages
    bool have_logits = false;
    DevBuf state, x, qkv, att, h, logits, out_tokens, prompt, attn_scratch, attn_counters, snap_logits;
    int snap_n_past = -1;
    struct KernelTimer *timer = nullptr;  // non-null only inside profile_decode
    // prompt chunks (prefill_chunk): activations of up to 32 tokens side by side, their quantised images, positions
    DevBuf cx, cqkv, catt, ch, cgu, cimg, cpos, cqh, cpart;
    DevBuf trace;
    DevBuf mask;           // optional logit mask (n_vocab bits), see set_mask
    bool mask_on = false;
    const uint8_t *ext_mask = nullptr;  // a mask owned by someone else (the batcher's lane mask); wins over `mask`
    unsigned long long *trace_base = nullptr;  // non-null only inside trace_decode
    // the captured decode step, in two variants: [0] the latency attention kernel (short contexts), [1] the
    // position-split GQA-sharing one (attention_long.cu) that the host switches to once the context is long
    struct StepGraph {
        cudaGraphExec_t exec = nullptr;
        aios_sampling sampling{};
        const uint8_t *mask_ptr = nullptr;
        bool mask = false;
    } graphs[2];
    DevBuf attn_long_scratch;  // scores / slice maxima / partial sums of attention_long.cu (null: not usable here)
    int attn_long_n_slices = 0;
    bool want_long_attn(int n_pos) const { return attn_long_scratch.p && n_pos >= m->eng->attn_long_min; }
    int kernels_per_step = 0;
    // persistent fused decode (pfd.cu): phase table + global exchange buffers; pfd_grid == 0 -> not usable
    DevBuf pfd_phases, pfd_att_max, pfd_att_sum, pfd_att_acc, pfd_counters, pfd_timing;
    PfdParams pfd{};
    int pfd_grid = 0;

    // internal = true: one of the library's own sequences (takes a hidden slot, never a caller's)
    explicit Seq(Model *m, bool internal = false);
    ~Seq();
    void reset();
    void prefill(const int32_t *tokens, int n);
    int decode(const aios_sampling &sp, int max_tokens, int32_t *out);
    void step(const aios_sampling &sp, int32_t *tok);
    int decode_static(int max_tokens, int32_t *out);  // fused kernel on the library's built-in scratch: no allocation
    void bench_decode(int n_tokens, aios_bench_result *out);
    void profile_decode(int n_tokens, aios_bench_result *out);
    int trace_decode(unsigned long long *out, int cap_launches);  // per-launch timeline of one token (kernels.h marks)
    void snapshot();
    void restore();
    void read_logits(float *out);
    void set_mask(const uint8_t *bits, int n_bytes);  // null: sample from every token again
    // the token pass in pieces, for the tensor-parallel driver (tp.cu), which runs the ranks' lists in lockstep
    DevBuf tp_part64;  // [n_embd] doubles, tensor-parallel shards only
    TpPush tp_push{};  // n > 0: the K-split projections push their partials to the peers (EPI_PUSH_F64) instead
    bool tp_pdl_wait = false;  // the kernel after a pushed all-reduce is a programmatic dependent of the wait kernel
    TpPush tp_push_logits{};  // ... and the lm_head its slice of the logits (EPI_PUSH_F32)
    void launch_embed(bool from_prompt, int prompt_idx, int pos_value, cudaStream_t st);
    void build_token_pass(bool want_logits, std::vector<PassLaunch> &ls, bool long_attn = false);
    void launch_pass_item(PassLaunch &L, cudaStream_t st, int launch_index);
    void launch_sample_kernel(const aios_sampling &sp, cudaStream_t st);
    // tokens prompt[idx0 .. idx0+T) at positions pos0...; on a tensor-parallel shard `reduce` carries the K-split
    // projections' partial sums across the ranks and want_logits must be false (the group gathers the logits itself)
    void prefill_chunk(int idx0, int T, int pos0, bool want_logits, cudaStream_t st, ChunkReduce *reduce = nullptr);

   private:
    void launch_token_pass(bool from_prompt, int prompt_idx, int pos_value, bool want_logits, cudaStream_t st, int *n_launch,
                           bool long_attn = false);
    void prefill_tc(int idx0, int T, int pos0, bool want_logits, cudaStream_t st);     // the same on the tensor-core path
    cudaGraphExec_t ensure_graph(const aios_sampling &sp, bool long_attn);
    void build_pfd();
    bool use_pfd(const aios_sampling &sp) const;
    void launch_pfd_tokens(int n, cudaStream_t st);
};

// ---------------------------------------------------------------- continuous batching (batch.cu)
// Up to GEMM_MAX_TOKENS (64) sequences of one model decode together: one token per live lane per step, every weight
// matrix streamed once per step by the multi-token GEMM (gemm_q.cu), attention per lane over the lane's own KV range,
// one sampler CTA per lane.  The request scheduler (capi.cu) admits and retires lanes between steps.  All arithmetic
// is the single-sequence path's, so a lane's tokens do not depend on what else is in the batch.
struct Batch {
    Model *m;
    int n = 0;                          // live lanes: 0 .. n-1 (kept dense)
    std::vector<std::unique_ptr<Seq>> seqs;  // lane i's sequence (KV range, prompt prefill); pooled, created on demand
    std::vector<SampleLane> cfg;        // host mirror of the per-lane sampler parameters
    // where a lane's current logits live: its own sequence's buffer (after its prefill, and while the batch is small
    // enough that every lane runs the single-sequence token pass) or its row of LOGITS (batched steps)
    std::vector<char> in_seq;
    static constexpr int SMALL_BATCH = 3;  // up to this many lanes, one GEMV token pass per lane beats the multi-token GEMM
    DevBuf X, QKV, ATT, H, GU, LOGITS, IMG_A, IMG_B, POS, TOK, KVB, CFG, MASKS;
    int32_t *h_tok = nullptr;           // pinned: the step's sampled tokens
    int32_t *h_pos = nullptr;           // pinned staging for POS
    void **h_kvb = nullptr;             // pinned staging for KVB
    SampleLane *h_cfg = nullptr;        // pinned staging for CFG
    bool dirty = true;                  // lane set / parameters changed: re-upload POS, KVB, CFG before the next step
    std::vector<cudaGraphExec_t> graphs;  // captured step, indexed by n (lazily)
    size_t mask_bytes = 0;

    explicit Batch(Model *m);
    ~Batch();
    int capacity() const { return GEMM_MAX_TOKENS; }
    // bring a request in: prefill its prompt into a pooled sequence, whose logits become lane n's row.  Returns the lane.
    int admit(const int32_t *prompt, int n_prompt, const aios_sampling &sp);
    void retire(int lane);              // the last lane moves into the hole
    void set_mask(int lane, const uint8_t *bits);  // null: no mask; else n_vocab bits, copied
    // one step: every live lane samples a token from its logits and feeds it.  Returns after the tokens are in h_tok.
    void step();
    int kernels_per_step = 0;

   private:
    void launch_step(cudaStream_t st, int *n_launch);
    void upload_lane_state(cudaStream_t st);
};

}  // namespace aios
