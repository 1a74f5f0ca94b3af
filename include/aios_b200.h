/*
 * aios_b200.h — C ABI of libaios_b200.so, the B200-native (sm_100a) GGUF inference engine
 * that replaces the llama-server child processes behind aiOS's `aios-runtime`.
 *
 * The reference has no FFI for this path: its Rust `runtime` crate spawns llama-server
 * (runtime/src/model_manager.rs:187-204) and proxies HTTP (runtime/src/inference.rs:94-186).
 * Each entry point below names the reference interface it stands in for; INTEGRATION.md
 * shows the Rust `-sys` binding a maintainer would add.
 *
 * Conventions: plain C types only; every call returns 0 (AIOS_OK) or a negative AIOS_E_*
 * code; the message is in the calling thread's aios_last_error().  Nothing throws or
 * aborts across this boundary.  Handles are opaque and owned by the library.  All pointer
 * arguments are HOST pointers unless a comment says otherwise.  There is no CPU fallback:
 * every compute entry point fails with AIOS_E_CUDA when no sm_100 device is usable.
 */
#ifndef AIOS_B200_H
#define AIOS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AIOS_OK 0
#define AIOS_E_INVAL (-1)   /* bad argument                          -> tonic InvalidArgument */
#define AIOS_E_IO (-2)      /* file missing / unreadable / not GGUF  -> "error: ..." status   */
#define AIOS_E_FORMAT (-3)  /* unsupported GGUF contents                                      */
#define AIOS_E_CUDA (-4)    /* CUDA runtime / no device              -> tonic Internal        */
#define AIOS_E_NOMEM (-5)   /* device or host memory exhausted                                */
#define AIOS_E_STATE (-6)   /* wrong state (context full, unloaded)  -> tonic FailedPrecondition */
#define AIOS_E_RANGE (-7)   /* size beyond a stated maximum                                   */
#define AIOS_E_UNSUPPORTED (-8) /* valid input this build does not implement (e.g. BPE vocab) */

/* ggml tensor type ids as stored in GGUF (gguf-py constants.py GGMLQuantizationType) */
#define AIOS_TYPE_F32 0
#define AIOS_TYPE_F16 1
#define AIOS_TYPE_Q8_0 8
#define AIOS_TYPE_Q4_K 12
#define AIOS_TYPE_Q5_K 13
#define AIOS_TYPE_Q6_K 14

typedef struct aios_engine aios_engine;   /* one per GPU: device, stream, scratch            */
typedef struct aios_model aios_model;     /* a loaded GGUF: replaces one llama-server child  */
typedef struct aios_seq aios_seq;         /* one sequence = one KV-cache slot of a model     */
typedef struct aios_vocab aios_vocab;     /* a GGUF's tokenizer (host only, needs no GPU)    */

const char *aios_last_error(void);
/* library / build identification, e.g. "aios_b200 0.1 sm_100a" */
const char *aios_version(void);

/* ---- engine: replaces runtime/src/main.rs:50-51 (ModelManager + InferenceEngine construction) */
typedef struct aios_engine_opts {
    int32_t device;      /* CUDA ordinal                                                     */
    int32_t max_batch;   /* max concurrent sequences per model the CALLER may hold (aios_seq_create,
                            aios_generate); 0 -> 1.  The library's own sequences (aios_forward_logits' scratch,
                            the request scheduler's) live in extra hidden KV slots and never count against it */
    int32_t flags;       /* AIOS_ENGINE_* bits                                               */
    int32_t reserved;
} aios_engine_opts;
#define AIOS_ENGINE_NO_GRAPH 1 /* launch kernels one by one instead of via CUDA graphs (debug) */
#define AIOS_ENGINE_NO_PDL 2   /* no programmatic dependent launch between the GEMV kernels (debug) */
#define AIOS_ENGINE_NO_PFD 4   /* greedy decode through per-op kernels instead of the persistent fused kernel */
#define AIOS_ENGINE_USE_PFD 8  /* greedy decode through the persistent fused decode kernel (pfd.cu) */
#define AIOS_ENGINE_ATTN_CLUSTER 64 /* decode attention: the cluster / DSMEM kernel (attention.cu) at every context length */
#define AIOS_ENGINE_NO_ROPE_EPILOGUE 256/* RoPE / KV append in the attention kernel instead of the q/k/v GEMV's epilogue */
#define AIOS_ENGINE_EXACT_PREFILL 2048 /* prompts of every length through the bit-exact integer path (64-token chunks,
                                           gemm_q.cu); default: prompts of 256+ tokens take the tcgen05 float path (prefill_tc.cu) */
#define AIOS_ENGINE_NO_GEMM 1024 /* prefill token by token through the decode kernels instead of prompt chunks through the tensor-core GEMM */
/* (bits 16, 32, 128 and 512 selected round-1 experiments — L2 prefetch, per-head attention, TMA-ring GEMV, producer-side
 * quantisation — that measured slower than the default path and were removed; their profiles are under profiles/r01_*.
 * The bits are ignored.) */

int aios_engine_create(const aios_engine_opts *opts, aios_engine **out);
void aios_engine_destroy(aios_engine *e);
int aios_device_count(int32_t *n);

/* ---- model: replaces ModelManager::load_model / unload_model
 *      (runtime/src/model_manager.rs:149-277, :285-345).  n_ctx<=0 -> 2048 like the
 *      reference default (model_manager.rs:167-171). */
int aios_model_load(aios_engine *e, const char *name, const char *gguf_path, int32_t n_ctx, aios_model **out);
int aios_model_unload(aios_model *m);

typedef struct aios_model_info {
    char name[128];
    char arch[32];
    int32_t n_layer, n_embd, n_head, n_head_kv, head_dim, n_ff, n_vocab, n_ctx, n_rot;
    int32_t file_type;            /* general.file_type                                      */
    int32_t n_tensors;
    int32_t max_batch;
    float rms_eps, rope_freq_base;
    int64_t file_bytes;           /* GGUF file size                                         */
    int64_t weight_bytes;         /* device bytes held by weights                           */
    int64_t kv_bytes;             /* device bytes held by the KV cache                      */
    int64_t weight_bytes_per_token; /* algorithmic bytes one decode step streams            */
    int64_t kv_bytes_per_position;
} aios_model_info;
/* replaces the ModelStatus / ManagedModel fields (runtime/src/model_manager.rs:42-56, :539-548) */
int aios_model_info_get(const aios_model *m, aios_model_info *out);

/* tensor directory (loader parity tests): i in [0, n_tensors) */
typedef struct aios_tensor_info {
    char name[96];
    int32_t type;
    int32_t n_dims;
    int64_t ne[4];      /* ne[0] = innermost (K) */
    int64_t offset;     /* byte offset inside the GGUF data section */
    int64_t nbytes;
} aios_tensor_info;
int aios_model_tensor_info(const aios_model *m, int32_t i, aios_tensor_info *out);

/* a string-valued GGUF metadata key ("general.name", "tokenizer.chat_template", ...): AIOS_E_INVAL if absent.
 * Output-buffer convention here and below: *len / *n_out always receives the full size; pass buf/out = NULL to
 * query it; AIOS_E_RANGE if cap is too short. */
int aios_model_meta_str(aios_model *m, const char *key, char *buf, int32_t cap, int32_t *len);

/* ---- vocabulary, tokenizer and chat prompt: the stage llama-server runs on the {system,user} messages the
 *      reference posts (runtime/src/inference.rs:363-376 build_messages), read from the
 *      GGUF's tokenizer.ggml.* metadata.  Host-only: a vocabulary can be opened straight from a file without a GPU
 *      (aios_vocab_open / aios_vocab_close), or borrowed from a loaded model (aios_model_vocab: owned by the model,
 *      valid until aios_model_unload, never closed by the caller).  SentencePiece-style ("llama") vocabularies
 *      only; encoding with a "gpt2"/BPE vocabulary answers AIOS_E_UNSUPPORTED. */
int aios_vocab_open(const char *gguf_path, aios_vocab **out);
void aios_vocab_close(aios_vocab *v);
int aios_model_vocab(aios_model *m, aios_vocab **out);
typedef struct aios_vocab_info {
    int32_t n_tokens;    /* length of tokenizer.ggml.tokens (0 if the file carries no vocabulary)          */
    int32_t bos, eos;    /* tokenizer.ggml.{bos,eos}_token_id, -1 if absent                               */
    int32_t add_bos;     /* tokenizer.ggml.add_bos_token (default 1)                                      */
    int32_t chat_format; /* 0 chatml, 1 llama-2, 2 mistral, 3 zephyr, 4 llama-3 (see aios_chat_prompt)    */
    char model[32];      /* tokenizer.ggml.model: "llama" (SPM), "gpt2" (BPE), ...                        */
} aios_vocab_info;
int aios_vocab_info_get(aios_vocab *v, aios_vocab_info *out);
/* piece bytes (not NUL-terminated) of token id; type = token_type (1 normal, 2 unknown, 3 control,
 * 4 user-defined, 6 byte), score = tokenizer.ggml.scores[id] */
int aios_vocab_token_piece(aios_vocab *v, int32_t id, char *buf, int32_t cap, int32_t *len, int32_t *type, float *score);
#define AIOS_TOK_ADD_SPECIAL 1     /* tokenize: prepend BOS when the vocabulary asks for it                  */
#define AIOS_TOK_PARSE_SPECIAL 2   /* tokenize: control-token strings in the text become their single ids    */
#define AIOS_TOK_RENDER_SPECIAL 4  /* detokenize: emit control tokens' strings instead of skipping them      */
#define AIOS_TOK_STRIP_SPACE 8     /* detokenize: drop the leading space of the SentencePiece prefix         */
/* text_len < 0: text is NUL-terminated */
int aios_tokenize(aios_vocab *v, const char *text, int32_t text_len, int32_t flags, int32_t *out, int32_t cap, int32_t *n_out);
int aios_detokenize(aios_vocab *v, const int32_t *tokens, int32_t n, int32_t flags, char *buf, int32_t cap, int32_t *len);
/* [system?, user] (the message list of runtime/src/inference.rs:363-376) rendered for generation, ready for
 * aios_tokenize(ADD_SPECIAL|PARSE_SPECIAL).  First the file's own tokenizer.chat_template is evaluated (a Jinja subset:
 * csrc/jinja_lite.h); if the template uses something outside that subset or raises (Mistral's refuses a system role),
 * the format recognised from its marker strings is used instead (chatml, llama-2, mistral, zephyr, llama-3; chatml when
 * the file has no template).  A BOS string the template itself emits is dropped when the tokenizer adds BOS anyway.
 * system_prompt may be NULL or "". */
int aios_chat_prompt(aios_vocab *v, const char *system_prompt, const char *user_prompt, char *buf, int32_t cap, int32_t *len);
/* any message list through tokenizer.chat_template only (what the OpenAI-compatible endpoint receives,
 * api-gateway/src/openai.rs:148-165): AIOS_E_STATE if the file has no template, AIOS_E_UNSUPPORTED (message in
 * aios_last_error) if the template is outside the supported subset or raises for these messages. */
int aios_chat_apply_template(aios_vocab *v, const char *const *roles, const char *const *contents, int32_t n_messages,
                             int32_t add_generation_prompt, char *buf, int32_t cap, int32_t *len);
/* 1 if id ends generation: tokenizer.ggml.eos_token_id or a turn terminator (<|im_end|>, <|eot_id|>, ...) */
int aios_token_is_eog(aios_vocab *v, int32_t id, int32_t *is_eog);

/* ---- sampling parameters: the request fields InferenceEngine::infer forwards
 *      (runtime/src/inference.rs:103-122: max_tokens, temperature) plus the llama-server
 *      defaults it leaves implicit (top_k 40, top_p 0.95, min_p 0.05). */
typedef struct aios_sampling {
    float temperature;   /* <= 0 -> greedy argmax (lowest index wins ties)                   */
    int32_t top_k;       /* <= 0 -> disabled                                                 */
    float top_p;         /* >= 1 -> disabled                                                 */
    float min_p;         /* <= 0 -> disabled                                                 */
    uint64_t seed;
} aios_sampling;

/* ---- sequences: the decode hot loop llama-server runs per slot */
int aios_seq_create(aios_model *m, aios_seq **out);
void aios_seq_free(aios_seq *s);
int aios_seq_reset(aios_seq *s);                     /* drop the KV contents, position 0     */
int aios_seq_position(const aios_seq *s, int32_t *n_past);
/* feed n prompt tokens at the current position (prefill); keeps the last token's logits */
int aios_seq_prefill(aios_seq *s, const int32_t *tokens, int32_t n);
/* generate up to max_tokens tokens after the fed ones; out_tokens[cap]; stops at the
 * context limit.  One H2D of nothing, one D2H of the token ids. */
int aios_seq_decode(aios_seq *s, const aios_sampling *sp, int32_t max_tokens, int32_t *out_tokens, int32_t *n_out);
/* one streaming step: sample from the kept logits, feed that token, return it.
 * H2D: 0 bytes; D2H: 4 bytes; synchronises. */
int aios_seq_step(aios_seq *s, const aios_sampling *sp, int32_t *token_out);
/* Greedy decode for callers that must not allocate device memory while serving.  (No counterpart in the
 * reference: llama-server hides its allocator behind a process boundary.)
 *
 * Who needs it: a serving thread whose device-memory budget was handed out completely at start-up and is frozen
 * afterwards (an embedding host that owns the rest of HBM for its own pools, or a real-time loop that forbids
 * allocator calls because cudaMalloc/cudaFree synchronise the whole device and take a process-wide lock).  Such a
 * caller creates the model and the sequence during start-up, when allocating is still allowed, and later only calls
 * this function.
 *
 * Same result as aios_seq_decode with greedy sampling (temperature <= 0): up to max_tokens tokens into the caller's
 * HOST buffer out_tokens[max_tokens], *n_out = how many (fewer at the context limit).  It always runs the persistent
 * fused decode kernel, and that kernel's cross-CTA exchange area — barrier counters, per-CTA soft-max row maxima and
 * sums, the f64 attention-output accumulators — is a 48 KB array compiled into the library image instead of buffers
 * from cudaMalloc.  The call issues only memset / kernel launch / memcpy on the engine's stream: no cudaMalloc, no
 * cudaFree, no allocating library routine.
 *
 * Limits: the array fits models with at most 8 kv heads and n_head*head_dim <= 4096 (TinyLlama, Mistral-7B);
 * anything larger is refused with AIOS_E_RANGE.  AIOS_E_STATE if the sequence has no logits yet or the fused kernel
 * cannot be made resident on this GPU.
 *
 * Sharing: there is ONE array per GPU, shared by every sequence and every call in the process, and it keeps what
 * the previous call left in it; each call clears the counters and accumulators it needs at zero before launching.
 * Two calls whose kernels overlap on the same GPU (two engines, two streams) would trample each other's
 * counters — serialise them.  Every GPU that loads the library holds the 48 KB until the process exits.
 *
 * Set-up: the first call on a device looks the array's device address up once (cudaGetSymbolAddress, with that
 * device current) and remembers it; make that first call before starting any stream capture. */
int aios_seq_decode_static(aios_seq *s, int32_t max_tokens, int32_t *out_tokens, int32_t *n_out);
/* Logit mask for this sequence's sampler (greedy or stochastic): bit i of byte i/8 of mask_bits set = token i may be
 * sampled; n_bytes >= ceil(n_vocab / 8).  It applies to every later step until replaced, or cleared with
 * mask_bits = NULL.  This is the device-side half of JSON mode: unary Infer always asks for response_format
 * json_object (runtime/src/inference.rs:119-121), llama-server enforces it with a grammar that masks logits before its
 * sampler chain; here the host advances the grammar and uploads the allowed set (n_vocab / 8 bytes) per step, and the
 * on-device sampler applies it — the logits themselves never leave the GPU.  H2D: n_vocab / 8 bytes; D2H: 0. */
int aios_seq_set_mask(aios_seq *s, const uint8_t *mask_bits, int32_t n_bytes);
/* copy the kept logits (n_vocab floats) to the host */
int aios_seq_logits(aios_seq *s, float *logits_out);
/* remember / return to the current position and its logits (prompt caching, benchmark replays).
 * KV rows past the snapshot position are simply overwritten by later tokens. */
int aios_seq_snapshot(aios_seq *s);
int aios_seq_restore(aios_seq *s);

/* ---- convenience: InferenceEngine::infer at token level (runtime/src/inference.rs:94-186):
 *      prefill prompt, decode max_tokens.  tokens_used = n_prompt + *n_out. */
int aios_generate(aios_model *m, const int32_t *prompt, int32_t n_prompt, const aios_sampling *sp,
                  int32_t max_tokens, int32_t *out_tokens, int32_t *n_out);

/* ---- queued requests: what concurrent tonic handlers need (SURVEY.md §8b "Threading": the reference lets Infer calls
 *      run in parallel against llama-server, runtime/src/main.rs:50-51, grpc_service.rs:86-141).  submit / poll /
 *      wait / cancel / free are thread-safe and submit / poll / cancel never block on the GPU.  One scheduler thread
 *      per engine does CONTINUOUS BATCHING: up to 64 requests of a model decode together, one token per request per
 *      step, every weight matrix streamed once per step for all of them; requests join and leave between steps, and
 *      synchronous calls on the engine interleave between steps as well.  A request's tokens do not depend on what
 *      else is in the batch (same arithmetic as the single-sequence path, bit for bit).  KV memory comes from the
 *      model's page pool as the request grows; the library's batch lanes do not count against max_batch.
 *      The prompt is copied on submit.  max_tokens is clipped to the context. */
typedef struct aios_request aios_request;
#define AIOS_REQ_STOP_AT_EOG 1  /* stop at (and do not return) the first end-of-generation token, see aios_token_is_eog */
int aios_request_submit(aios_model *m, const int32_t *tokens, int32_t n, const aios_sampling *sp, int32_t max_tokens,
                        int32_t flags, aios_request **out);
/* Same, with a per-step logit mask (JSON mode through the queue: unary Infer always asks for json_object,
 * runtime/src/inference.rs:119-121).  Before every decode step of this request the scheduler thread calls
 * cb(user, last_token, mask_out, n_bytes): last_token = the token sampled in the previous step (-1 before the first);
 * the callback advances its grammar by that token and writes the allowed set of the NEXT token into mask_out
 * (n_bytes = ceil(n_vocab / 8), zeroed on entry; bit i of byte i/8 = token i allowed).  Return 0: mask written;
 * 1: no restriction for this step; 2: the output is complete, finish the request without sampling; < 0: fail the
 * request.  The mask goes to the device (n_vocab / 8 bytes H2D) and is applied by the on-device sampler; logits never
 * leave the GPU.  The callback runs on the scheduler thread between steps: keep it short, do not call back into the
 * engine from it. */
typedef int (*aios_mask_cb)(void *user, int32_t last_token, uint8_t *mask_out, int32_t n_bytes);
int aios_request_submit_masked(aios_model *m, const int32_t *tokens, int32_t n, const aios_sampling *sp, int32_t max_tokens,
                               int32_t flags, aios_mask_cb cb, void *user, aios_request **out);
/* JSON mode without any host code per token on the caller's side: a JSON-object grammar (root ::= object, RFC 8259 as
 * llama.cpp's json.gbnf restricts it) over a vocabulary.  One grammar per request: create it, submit with
 * cb = aios_json_grammar_mask_cb and user = the grammar, free it after the request.  The callback advances the grammar
 * by every sampled token, allows exactly the tokens whose bytes keep the text a prefix of one JSON object, and ends
 * the request (return 2) when the object closes or nothing fits. */
typedef struct aios_json_grammar aios_json_grammar;
int aios_json_grammar_create(aios_vocab *v, aios_json_grammar **out);
void aios_json_grammar_free(aios_json_grammar *g);
int aios_json_grammar_complete(aios_json_grammar *g, int32_t *complete);
int aios_json_grammar_mask_cb(void *user, int32_t last_token, uint8_t *mask_out, int32_t n_bytes);
/* take up to cap new tokens; *done = 1 once the request has finished AND everything was taken.  A request that
 * failed returns its error code (message in aios_last_error) from the poll that would have set *done. */
int aios_request_poll(aios_request *r, int32_t *out_tokens, int32_t cap, int32_t *n_out, int32_t *done);
/* block until there is something to poll or the request is finished (timeout_ms < 0: no timeout); *ready = 0 on timeout */
int aios_request_wait(aios_request *r, int32_t timeout_ms, int32_t *ready);
/* ask the scheduler to stop at the next chunk boundary; tokens already generated stay pollable */
int aios_request_cancel(aios_request *r);
/* may be called at any time; a request that is still queued or running is cancelled and reclaimed by the scheduler */
void aios_request_free(aios_request *r);

/* ---- test / parity hooks (SURVEY.md §8b "test hooks") */
/* dequantise n elements (whole blocks) of ggml_type on the GPU: host blocks -> host floats.
 * Bit-exact against ggml's dequantize_row_* as restated by gguf-py. */
int aios_dequantize(aios_engine *e, int32_t ggml_type, const void *blocks, int64_t n, float *out);
/* y[n_rows] = W[n_rows,k] . x[k] with llama.cpp's CPU numerics (x quantised to Q8_K / Q8_0,
 * integer block dots): host in, host out.  If norm_w != NULL, x is RMS-normed first. */
int aios_matvec(aios_engine *e, int32_t ggml_type, const void *w_blocks, int64_t n_rows, int64_t k,
                const float *x, const float *norm_w, float rms_eps, float *y);
/* Y[n_tok][n_rows] = X[n_tok][k] . W[n_rows,k]^T for 1..64 tokens through the multi-token kernel (gemm_q.cu: int8 MMA
 * on the sub-block dots, weights streamed once for all tokens — batched decode and bit-exact prompt chunks).  Same
 * numerics as n_tok calls of aios_matvec, bit for bit.  Q4_K, Q5_K, Q6_K and Q8_0 weights (AIOS_E_UNSUPPORTED otherwise);
 * host in, host out. */
int aios_matmul(aios_engine *e, int32_t ggml_type, const void *w_blocks, int64_t n_rows, int64_t k, const float *x,
                int32_t n_tok, const float *norm_w, float rms_eps, float *y);
/* the same product through the prefill path's tensor-core GEMM (prefill_tc.cu: tcgen05 / TMEM, f16 operands, f32
 * accumulation): any n_tok >= 1; Q4_K, Q5_K, Q6_K and Q8_0 weights.  Float arithmetic with a stated tolerance — max error 4e-3,
 * rms error 8e-4 of the output's rms against an exact product of the dequantised weights — not the bit-exact integer path of aios_matmul. */
int aios_matmul_tc(aios_engine *e, int32_t ggml_type, const void *w_blocks, int64_t n_rows, int64_t k, const float *x,
                   int32_t n_tok, const float *norm_w, float rms_eps, float *y);
/* one run of the on-device sampler chain on host logits (parity hook for oracle_sample): mask_bits may be NULL;
 * `counter` is the draw's counter (a sequence uses the position of the last token fed, n_past - 1).  Outputs: *token,
 * cand[128] (candidate ids, best first; -1 beyond K), counts[3] = {K, n after top-p, n after min-p}; cand / counts may
 * be NULL. */
int aios_sample_logits(aios_engine *e, const float *logits, int32_t n_vocab, const uint8_t *mask_bits, const aios_sampling *sp,
                       uint64_t counter, int32_t *token, int32_t *cand, int32_t *counts);
/* logits of the last of n tokens fed at position n_past into a scratch sequence */
int aios_forward_logits(aios_model *m, const int32_t *tokens, int32_t n, int32_t n_past, float *logits_out);
/* read back one layer's KV rows [pos0, pos0+n) as fp16 bit patterns: k_out/v_out[n][n_head_kv*head_dim] */
int aios_seq_read_kv(aios_seq *s, int32_t layer, int32_t pos0, int32_t n, uint16_t *k_out, uint16_t *v_out);
/* read one of the sequence's f32 activation buffers as the last token pass left it:
 * which = 0 residual x[n_embd], 1 raw q|k|v[(n_head+2*n_head_kv)*head_dim], 2 attention output
 * [n_head*head_dim], 3 SwiGLU output h[n_ff].  n = number of floats to copy. */
int aios_seq_read_activation(aios_seq *s, int32_t which, float *out, int64_t n);

/* ---- measurement hooks (bench.py): device-timed decode with CUDA events on the engine's
 *      own stream.  Runs n_tokens greedy decode steps from the current position and returns
 *      the elapsed milliseconds and the number of kernels launched. */
typedef struct aios_bench_result {
    float ms_total;          /* CUDA-event time of the whole region                          */
    int32_t n_tokens;
    int32_t kernel_launches; /* kernel nodes executed inside the region                      */
    int32_t n_past_begin, n_past_end;
    float ms_gemv;           /* if per-kernel timing was requested, time in GEMV kernels     */
    float reserved[3];
} aios_bench_result;
int aios_seq_bench_decode(aios_seq *s, int32_t n_tokens, aios_bench_result *out);
/* same work launched kernel by kernel with a CUDA-event pair around every launch: fills
 * ms_gemv (sum of the dequant+GEMV launches), reserved[0] = ms in attention launches,
 * reserved[1] = ms in the remaining launches, reserved[2] = number of GEMV launches. */
int aios_seq_profile_decode(aios_seq *s, int32_t n_tokens, aios_bench_result *out);
/* Launch timeline of one greedy token: two tokens are decoded kernel by kernel on the engine's stream (same
 * programmatic-dependent-launch attributes as in the captured graph); during the second one, CTA 0 of every TMA-ring
 * GEMV / split-attention launch records %globaltimer nanoseconds.  out[launch][8]: [0] kernel entry, [1] dependency
 * wait over, [2] activations ready (GEMV) / scores done (attention), [3] first weight tile landed, [4] CTA done,
 * [5] first / [6] last bulk copy issued by the producer warp; zeros for launches of other kernels. */
int aios_seq_trace_decode(aios_seq *s, uint64_t *out, int32_t cap_launches, int32_t *n_launches);
/* per-phase time stamps (ns) that CTA 0 of the persistent fused decode kernel took for the last token of the last
 * fused launch: out[n_phases][4] = phase start, after the activation prologue, after the compute loop, after the grid
 * barrier.  *n_phases returns the phase count; cap = capacity of out in phases. */
int aios_seq_phase_timing(aios_seq *s, uint64_t *out, int32_t cap, int32_t *n_phases);
/* prefill of an n-token prompt from position 0 (bench.py configs[2]): the sequence is reset, one untimed prefill warms
 * the path up, then `iters` prefills are timed with CUDA events on the engine's stream (token ids already on the host
 * side of the call: the 4n-byte H2D is inside the timed region, like in aios_seq_prefill).  The sequence ends prefilled. */
int aios_seq_bench_prefill(aios_seq *s, const int32_t *tokens, int32_t n, int32_t iters, float *ms_per_prefill);
/* batched decode as the continuous batcher runs it (bench.py configs[3]): n_lanes sequences are admitted with their
 * n_prompt-token prompts (prompts[n_lanes][n_prompt]), then n_steps greedy steps run — one token per lane per step, the
 * first step untimed (it captures the step's CUDA graph), the others timed with CUDA events on the engine's stream,
 * each step's sampled tokens copied to the host and synchronised like in the scheduler.  out_tokens[n_lanes][n_steps]
 * (may be NULL) receives every lane's tokens.  ms_prefill includes the first step. */
typedef struct aios_batch_bench {
    float ms_decode;         /* CUDA-event time of the timed steps                            */
    float ms_prefill;        /* admission: the prompts' prefill + the untimed first step      */
    int32_t n_lanes, n_steps_timed;
    int32_t kernel_launches; /* kernel nodes executed inside the timed steps                  */
    int32_t n_past_begin, n_past_end;  /* positions in every lane's cache at the first / after the last timed step */
    int32_t reserved;
} aios_batch_bench;
int aios_bench_batch_decode(aios_model *m, int32_t n_lanes, const int32_t *prompts, int32_t n_prompt, int32_t n_steps,
                            int32_t *out_tokens, aios_batch_bench *out);
/* time ONE GEMV launch shape in isolation (roofline leg): type/n_rows/k on random weights,
 * `iters` launches over `n_copies` distinct weight copies (so the working set exceeds L2) */
int aios_bench_gemv(aios_engine *e, int32_t ggml_type, int64_t n_rows, int64_t k, int32_t n_copies,
                    int32_t iters, float *ms_per_launch);

/* ---- tensor parallelism: one model split over 2..8 GPUs of this node --------------------------------------------
 * (BASELINE.json configs[4], Llama-3-70B at TP 2/4/8.  The reference's hardware table lists
 *  70B-class local models, docs/HARDWARE.md:135-138; llama-server's counterpart is --tensor-split / --split-mode row.)
 * attn_q/k/v, ffn_gate/up and the lm_head are split by rows, attn_output and ffn_down along K; per layer two
 * all-reduces of the projections' f64 partial sums (so the tokens are those of the unsplit model), per token one
 * all-gather of the logits.  With peer access between all the devices the all-reduce rides in the GEMV: its epilogue
 * stores the partials into every GPU's receive slab over NVLink and raises an arrival flag, and one small kernel per
 * GPU waits for the flags and adds the sum to the residual; otherwise (or with AIOS_TP_ALLREDUCE=nccl in the
 * environment) it is ncclAllReduce.  libnccl.so.2 is opened on the first aios_tp_model_load (communicators, the logits
 * all-gather, the fallback); the library does not link to it.
 * One process drives all the shards; calls on one aios_tp_model must not overlap.  The model must split evenly:
 * n_head_kv, n_ff and n_vocab divisible by n_devices, and n_head*head_dim / n_devices and n_ff / n_devices
 * multiples of 256 (AIOS_E_FORMAT otherwise). */
typedef struct aios_tp_model aios_tp_model;
int aios_tp_model_load(const int32_t *devices, int32_t n_devices, const char *name, const char *gguf_path, int32_t n_ctx,
                       int32_t engine_flags, aios_tp_model **out);
int aios_tp_model_unload(aios_tp_model *m);
int aios_tp_model_info(const aios_tp_model *m, aios_model_info *out);   /* weight_bytes: one shard's           */
int aios_tp_reset(aios_tp_model *m);
int aios_tp_prefill(aios_tp_model *m, const int32_t *tokens, int32_t n_tokens);
int aios_tp_decode(aios_tp_model *m, const aios_sampling *sp, int32_t max_tokens, int32_t *out_tokens, int32_t *n_out);
int aios_tp_get_logits(aios_tp_model *m, float *out, int32_t n);
/* n_tokens greedy steps timed with CUDA events on rank 0's stream (every rank runs the same launches and the
 * collectives keep them in step), then 8 more with an event pair around every all-reduce */
typedef struct aios_tp_bench {
    float ms_total;
    int32_t n_tokens, n_ranks;
    int32_t kernel_launches;        /* rank 0: kernels + collectives inside the timed region     */
    int32_t n_past_begin, n_past_end;
    int32_t n_allreduce_per_token;
    float us_per_allreduce;         /* average, n_embd doubles each                              */
    int32_t allreduce_kind;         /* 1: partial sums pushed to the peers from the GEMV epilogue over NVLink
                                     *    (peer stores + arrival flags); 0: ncclAllReduce             */
    int32_t graphs;                 /* 1: each rank replays one CUDA graph per token (push mode)   */
} aios_tp_bench;
int aios_tp_bench_decode(aios_tp_model *m, int32_t n_tokens, aios_tp_bench *out);

#ifdef __cplusplus
}
#endif
#endif /* AIOS_B200_H */
