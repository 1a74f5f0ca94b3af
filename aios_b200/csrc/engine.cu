// engine.cu — see engine.h.
#include "engine.h"

#include <math.h>
#include <string.h>

#include <algorithm>

#include "common.cuh"

namespace aios {

// ---------------------------------------------------------------- DevBuf / Engine
DevBuf::DevBuf(size_t n) : bytes(n) {
    if (n) {
        cudaError_t e = cudaMalloc(&p, n);
        if (e != cudaSuccess) {
            p = nullptr;
            bytes = 0;
            throw CudaError(std::string("cudaMalloc(") + std::to_string(n) + " bytes): " + cudaGetErrorString(e));
        }
    }
}
DevBuf::~DevBuf() {
    if (p) cudaFree(p);
}
DevBuf &DevBuf::operator=(DevBuf &&o) noexcept {
    if (this != &o) {
        if (p) cudaFree(p);
        p = o.p;
        bytes = o.bytes;
        o.p = nullptr;
        o.bytes = 0;
    }
    return *this;
}

Engine::Engine(const aios_engine_opts &o) : device(o.device), max_batch(o.max_batch > 0 ? o.max_batch : 1), flags(o.flags) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) throw CudaError("no CUDA device is usable (this library has no CPU fallback)");
    if (device < 0 || device >= n) throw CudaError("device ordinal out of range");
    CUDA_CHECK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        throw CudaError(std::string("device '") + prop.name + "' is sm_" + std::to_string(prop.major) +
                        std::to_string(prop.minor) + "; this library is built for sm_100a (B200) only");
    }
    n_sms = prop.multiProcessorCount;
    CUDA_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    if (const char *e = getenv("AIOS_EVICT_FIRST")) evict_first = atoi(e) != 0;
    if (const char *e = getenv("AIOS_ATTN_LONG_MIN")) attn_long_min = std::max(1, std::min(4096, atoi(e)));
    // function attributes are per-device settings: every engine sets them for its own device
    gemv_init();
    attn_split_init();
    attn_init();
}
Engine::~Engine() {
    if (stream) cudaStreamDestroy(stream);
}
void Engine::make_current() const { CUDA_CHECK(cudaSetDevice(device)); }

// ---------------------------------------------------------------- Model
static bool is_quant_matrix_type(int t) { return t == T_Q4_K || t == T_Q5_K || t == T_Q6_K || t == T_Q8_0; }

static void alloc_planes(int type, int64_t n_rows, int64_t k, DevMatrix &dst) {
    size_t total = 0, off[5];
    for (int i = 0; i < 5; i++) {
        off[i] = total;
        total += (plane_bytes(type, i, n_rows, k) + 255) / 256 * 256;
    }
    dst.storage = DevBuf(total);
    for (int i = 0; i < 5; i++)
        dst.w.p[i] = plane_bytes(type, i, n_rows, k) ? dst.storage.as<uint8_t>() + off[i] : nullptr;
    dst.w.type = type;
    dst.w.n_rows = (int32_t)n_rows;
    dst.w.k = (int32_t)k;
}

void Model::upload_matrix(const GgufTensor &t, DevMatrix &dst, DevBuf &staging, void *pinned, size_t pinned_bytes, int64_t row0,
                          int64_t n_rows, int64_t k0, int64_t k_len) {
    if (!is_quant_matrix_type(t.type))
        throw GgufError(AIOS_E_FORMAT, "tensor '" + t.name + "': matrix type " + std::to_string(t.type) +
                                           " is not supported (Q4_K, Q5_K, Q6_K, Q8_0 are)");
    if (t.n_dims != 2) throw GgufError(AIOS_E_FORMAT, "tensor '" + t.name + "' is not a matrix");
    const int64_t K = t.ne[0], N = t.ne[1];
    if (K % 256) throw GgufError(AIOS_E_FORMAT, "tensor '" + t.name + "': K must be a multiple of 256");
    const int64_t n = n_rows < 0 ? N : n_rows, k = k_len < 0 ? K : k_len;
    if (row0 < 0 || row0 + n > N || k0 < 0 || k0 + k > K || k % 256 || k0 % 256)
        throw GgufError(AIOS_E_FORMAT, "tensor '" + t.name + "': bad shard (quant blocks run along K: K slices are cut at multiples of 256)");
    // the slice as a contiguous [n][k] tensor of the same type: whole rows are contiguous in the file, a K slice is a
    // contiguous run of blocks inside every row
    const size_t be = (size_t)type_block_elems(t.type), bb = (size_t)type_block_bytes(t.type);
    const size_t row_bytes_full = (size_t)K / be * bb, row_bytes = (size_t)k / be * bb, nbytes = (size_t)n * row_bytes;
    std::vector<uint8_t> gathered;
    const uint8_t *src = t.data + (size_t)row0 * row_bytes_full;
    if (k != K) {
        gathered.resize(nbytes);
        for (int64_t r = 0; r < n; r++)
            memcpy(gathered.data() + (size_t)r * row_bytes, t.data + (size_t)(row0 + r) * row_bytes_full + (size_t)k0 / be * bb, row_bytes);
        src = gathered.data();
    }
    alloc_planes(t.type, n, k, dst);
    dst.src_bytes = nbytes;
    if (staging.bytes < nbytes) staging = DevBuf(nbytes);
    cudaStream_t st = eng->stream;
    for (size_t off = 0; off < nbytes; off += pinned_bytes) {
        const size_t c = std::min(pinned_bytes, nbytes - off);
        memcpy(pinned, src + off, c);
        CUDA_CHECK(cudaMemcpyAsync(staging.as<uint8_t>() + off, pinned, c, cudaMemcpyHostToDevice, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
    }
    uint8_t *planes[5];
    for (int i = 0; i < 5; i++) planes[i] = const_cast<uint8_t *>(dst.w.p[i]);
    launch_repack(t.type, staging.p, n, (int)k, planes, st);
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaStreamSynchronize(st));
    weight_bytes += (int64_t)dst.storage.bytes;
    weight_bytes_per_token += (int64_t)nbytes;
}

static DevBuf upload_f32(const GgufTensor &t, int64_t expect, cudaStream_t st) {
    if (t.type != T_F32 || t.ne[0] != expect)
        throw GgufError(AIOS_E_FORMAT, "tensor '" + t.name + "' must be F32[" + std::to_string(expect) + "]");
    DevBuf b((size_t)expect * 4);
    CUDA_CHECK(cudaMemcpyAsync(b.p, t.data, (size_t)expect * 4, cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    return b;
}

Model::Model(Engine *e, const std::string &name_, const std::string &path_, int n_ctx, int tp_rank_, int tp_size_)
    : eng(e), name(name_), path(path_), tp_rank(tp_rank_), tp_size(tp_size_) {
    e->make_current();
    file.reset(new GgufFile(path));
    const GgufFile &f = *file;
    arch = f.get_s("general.architecture", "");
    if (arch != "llama") throw GgufError(AIOS_E_FORMAT, "unsupported architecture '" + arch + "' (only 'llama' is built)");
    hp.n_layer = (int)f.get_u("llama.block_count", 0);
    hp.n_embd = (int)f.get_u("llama.embedding_length", 0);
    hp.n_head = (int)f.get_u("llama.attention.head_count", 0);
    hp.n_head_kv = (int)f.get_u("llama.attention.head_count_kv", hp.n_head);
    hp.n_ff = (int)f.get_u("llama.feed_forward_length", 0);
    hp.rms_eps = (float)f.get_f("llama.attention.layer_norm_rms_epsilon", 1e-5);
    hp.rope_freq_base = (float)f.get_f("llama.rope.freq_base", 10000.0);
    hp.file_type = (int)f.get_u("general.file_type", 0);
    if (!hp.n_layer || !hp.n_embd || !hp.n_head || !hp.n_ff || hp.n_embd % hp.n_head || hp.n_head % hp.n_head_kv)
        throw GgufError(AIOS_E_FORMAT, "GGUF: missing or inconsistent llama.* hyper-parameters");
    hp.head_dim = hp.n_embd / hp.n_head;
    hp.n_rot = (int)f.get_u("llama.rope.dimension_count", hp.head_dim);
    if (hp.head_dim != 64 && hp.head_dim != 128) throw GgufError(AIOS_E_FORMAT, "head_dim must be 64 or 128");
    {
        const int rep = hp.n_head / hp.n_head_kv;
        if (rep != 1 && rep != 2 && rep != 4 && rep != 8)
            throw GgufError(AIOS_E_FORMAT, "query heads per kv head must be 1, 2, 4 or 8");
    }
    if (hp.n_rot > hp.head_dim || hp.n_rot % 2) throw GgufError(AIOS_E_FORMAT, "bad rope dimension count");
    hp.n_ctx = n_ctx > 0 ? n_ctx : 2048;  // reference default (model_manager.rs:167-171)
    const GgufTensor &te = f.require("token_embd.weight");
    hp.n_vocab = (int)te.ne[1];
    if (te.ne[0] != hp.n_embd) throw GgufError(AIOS_E_FORMAT, "token_embd.weight has the wrong width");
    // full-model sizes (the GGUF tensors' shapes) and this shard's share of them
    const int QD_full = hp.n_head * hp.head_dim, KVD_full = hp.n_head_kv * hp.head_dim, FF_full = hp.n_ff;
    vocab_local = hp.n_vocab;
    if (tp_size > 1) {
        if (tp_rank < 0 || tp_rank >= tp_size) throw GgufError(AIOS_E_INVAL, "bad tensor-parallel rank");
        if (hp.n_head_kv % tp_size || hp.n_ff % tp_size || hp.n_vocab % tp_size || (QD_full / tp_size) % 256 || (FF_full / tp_size) % 256)
            throw GgufError(AIOS_E_FORMAT, "model cannot be split " + std::to_string(tp_size) + " ways: kv heads, FF and vocabulary "
                            "must divide, and the K slices of attn_output / ffn_down must be multiples of 256");
        hp.n_head /= tp_size; hp.n_head_kv /= tp_size; hp.n_ff /= tp_size;
        vocab_local = hp.n_vocab / tp_size;
    }

    cudaStream_t st = e->stream;
    const size_t pinned_bytes = 64u << 20;
    void *pinned = nullptr;
    CUDA_CHECK(cudaMallocHost(&pinned, pinned_bytes));
    DevBuf staging;
    try {
        // embedding table stays in GGUF's natural layout: one row is gathered per token
        tok_embd_type = te.type;
        tok_embd_row_bytes = (int64_t)(te.nbytes / (uint64_t)hp.n_vocab);
        tok_embd_raw = DevBuf(te.nbytes);
        for (size_t off = 0; off < te.nbytes; off += pinned_bytes) {
            const size_t c = std::min(pinned_bytes, (size_t)te.nbytes - off);
            memcpy(pinned, te.data + off, c);
            CUDA_CHECK(cudaMemcpyAsync(tok_embd_raw.as<uint8_t>() + off, pinned, c, cudaMemcpyHostToDevice, st));
            CUDA_CHECK(cudaStreamSynchronize(st));
        }
        weight_bytes += (int64_t)te.nbytes;
        weight_bytes_per_token += tok_embd_row_bytes;

        layers.resize(hp.n_layer);
        for (int i = 0; i < hp.n_layer; i++) {
            const std::string p = "blk." + std::to_string(i) + ".";
            Layer &L = layers[i];
            L.attn_norm = upload_f32(f.require(p + "attn_norm.weight"), hp.n_embd, st);
            L.ffn_norm = upload_f32(f.require(p + "ffn_norm.weight"), hp.n_embd, st);
            weight_bytes_per_token += 2 * (int64_t)hp.n_embd * 4;
            // rows / cols = the tensor's full shape; split: 1 = this rank's rows (column-parallel), 2 = this rank's K slice
            struct { const char *n; DevMatrix *d; int64_t rows, cols; int split; } ms[] = {
                {"attn_q.weight", &L.wq, QD_full, hp.n_embd, 1},      {"attn_k.weight", &L.wk, KVD_full, hp.n_embd, 1},
                {"attn_v.weight", &L.wv, KVD_full, hp.n_embd, 1},     {"attn_output.weight", &L.wo, hp.n_embd, QD_full, 2},
                {"ffn_gate.weight", &L.w_gate, FF_full, hp.n_embd, 1}, {"ffn_up.weight", &L.w_up, FF_full, hp.n_embd, 1},
                {"ffn_down.weight", &L.w_down, hp.n_embd, FF_full, 2}};
            for (auto &m : ms) {
                const GgufTensor &t = f.require(p + m.n);
                if (t.ne[0] != m.cols || t.ne[1] != m.rows)
                    throw GgufError(AIOS_E_FORMAT, "tensor '" + t.name + "' has unexpected shape");
                if (tp_size == 1) upload_matrix(t, *m.d, staging, pinned, pinned_bytes);
                else if (m.split == 1) upload_matrix(t, *m.d, staging, pinned, pinned_bytes, tp_rank * (m.rows / tp_size), m.rows / tp_size);
                else upload_matrix(t, *m.d, staging, pinned, pinned_bytes, 0, -1, tp_rank * (m.cols / tp_size), m.cols / tp_size);
            }
            if (L.w_gate.w.type != L.w_up.w.type)
                throw GgufError(AIOS_E_FORMAT, "ffn_gate and ffn_up of layer " + std::to_string(i) + " differ in type");
        }
        output_norm = upload_f32(f.require("output_norm.weight"), hp.n_embd, st);
        weight_bytes_per_token += (int64_t)hp.n_embd * 4;
        const GgufTensor *out = f.find("output.weight");
        if (tp_size == 1) upload_matrix(out ? *out : te, output, staging, pinned, pinned_bytes);  // tied embeddings when absent
        else upload_matrix(out ? *out : te, output, staging, pinned, pinned_bytes, (int64_t)tp_rank * vocab_local, vocab_local);
        tc_ok = true;
        for (const Layer &L : layers)
            for (const DevMatrix *dm : {&L.wq, &L.wk, &L.wv, &L.wo, &L.w_gate, &L.w_up, &L.w_down})
                tc_ok = tc_ok && prefill_tc_usable(dm->w.type, dm->w.k);
    } catch (...) {
        cudaFreeHost(pinned);
        throw;
    }
    cudaFreeHost(pinned);

    // RoPE table, built like the oracle (repeated multiplication by theta_scale, glibc cosf/sinf)
    {
        const int half = hp.n_rot / 2;
        std::vector<float> c((size_t)hp.n_ctx * half), s((size_t)hp.n_ctx * half);
        const float theta_scale = powf(hp.rope_freq_base, -2.0f / (float)hp.n_rot);
        for (int pos = 0; pos < hp.n_ctx; pos++) {
            float theta = (float)pos;
            for (int i = 0; i < half; i++) {
                c[(size_t)pos * half + i] = cosf(theta);
                s[(size_t)pos * half + i] = sinf(theta);
                theta *= theta_scale;
            }
        }
        rope_cos = DevBuf(c.size() * 4);
        rope_sin = DevBuf(s.size() * 4);
        CUDA_CHECK(cudaMemcpy(rope_cos.p, c.data(), c.size() * 4, cudaMemcpyHostToDevice));
        CUDA_CHECK(cudaMemcpy(rope_sin.p, s.data(), s.size() * 4, cudaMemcpyHostToDevice));
    }
    {   // KV pages of 8 MB (64 positions of Mistral-7B, 372 of TinyLlama); AIOS_KV_POOL_MB caps the pool
        size_t cap = 0;
        if (const char *c = getenv("AIOS_KV_POOL_MB")) cap = (size_t)atoll(c) << 20;
        kv_pool.init(e->device, 8u << 20, cap);
    }
    attn_splits = std::max(1, std::min(32, (2 * e->n_sms) / std::max(1, hp.n_head_kv)));
}

Model::~Model() {}

// ---------------------------------------------------------------- Seq
struct KernelTimer {  // per-launch CUDA events (profile_decode only)
    struct Rec { cudaEvent_t a, b; int kind; };
    std::vector<Rec> recs;
    cudaStream_t st;
    void begin(int kind) {
        Rec r; r.kind = kind;
        CUDA_CHECK(cudaEventCreate(&r.a));
        CUDA_CHECK(cudaEventCreate(&r.b));
        CUDA_CHECK(cudaEventRecord(r.a, st));
        recs.push_back(r);
    }
    void end() { CUDA_CHECK(cudaEventRecord(recs.back().b, st)); }
};
#define TIMED(kind, stmt) do { if (timer) timer->begin(kind); stmt; if (timer) timer->end(); } while (0)

Seq::Seq(Model *m_, bool internal_) : m(m_), internal(internal_) {
    m->eng->make_current();
    {
        std::lock_guard<std::mutex> lk(m->mu);
        if (!internal && m->user_seqs >= m->eng->max_batch) throw std::runtime_error("no free KV slot (engine max_batch reached)");
        if (!internal) m->user_seqs++;
        m->live_seqs++;
    }
    const HParams &hp = m->hp;
    try {
        kv.init(&m->kv_pool, (size_t)hp.n_ctx * m->kv_pos_bytes());
        // a caller's sequence owns its whole context from the start (it may later be driven by the allocation-free
        // aios_seq_decode_static); the library's own sequences grow page by page
        if (!internal) kv.ensure((size_t)hp.n_ctx * m->kv_pos_bytes());
        state = DevBuf(sizeof(SeqState));
        x = DevBuf((size_t)hp.n_embd * 4);
        qkv = DevBuf((size_t)(hp.n_head + 2 * hp.n_head_kv) * hp.head_dim * 4);
        att = DevBuf((size_t)hp.n_head * hp.head_dim * 4);
        h = DevBuf((size_t)hp.n_ff * 4);
        logits = DevBuf((size_t)hp.n_vocab * 4);
        out_tokens = DevBuf((size_t)hp.n_ctx * 4);
        prompt = DevBuf((size_t)hp.n_ctx * 4);
        attn_scratch = DevBuf(attn_scratch_floats(hp.n_head, m->attn_splits, hp.head_dim) * 4);
        attn_counters = DevBuf((size_t)hp.n_head_kv * 4);
        CUDA_CHECK(cudaMemset(attn_counters.p, 0, attn_counters.bytes));
        pfd_att_max = DevBuf((size_t)hp.n_head_kv * 16 * 8 * sizeof(float));
        pfd_att_sum = DevBuf((size_t)hp.n_head_kv * 16 * 8 * sizeof(double));
        pfd_att_acc = DevBuf((size_t)hp.n_head * hp.head_dim * sizeof(double));
        pfd_counters = DevBuf(((size_t)hp.n_head_kv + 16) * sizeof(unsigned));  // [n_kv] group barriers, then the grid barrier
        CUDA_CHECK(cudaMemset(pfd_att_acc.p, 0, pfd_att_acc.bytes));
        CUDA_CHECK(cudaMemset(pfd_counters.p, 0, pfd_counters.bytes));
        if (m->tp_size > 1) tp_part64 = DevBuf((size_t)hp.n_embd * sizeof(double));
        if (attn_long_usable(hp.n_head, hp.n_head_kv, hp.n_ctx, hp.head_dim) &&
            (hp.n_ctx >= m->eng->attn_long_min || !attn_split_usable(hp.n_ctx, hp.head_dim))) {
            attn_long_n_slices = attn_long_slices(hp.n_head_kv, m->eng->n_sms);
            attn_long_scratch = DevBuf(attn_long_scratch_bytes(hp.n_head, hp.n_head_kv, hp.n_ctx, hp.head_dim, attn_long_n_slices));
            CUDA_CHECK(cudaMemset(attn_long_scratch.p, 0, attn_long_scratch.bytes));
        }
        reset();
        if (m->tp_size == 1) build_pfd();
    } catch (...) {
        std::lock_guard<std::mutex> lk(m->mu);
        if (!internal) m->user_seqs--;
        m->live_seqs--;
        throw;
    }
}

Seq::~Seq() {
    for (auto &g : graphs)
        if (g.exec) cudaGraphExecDestroy(g.exec);
    cudaStreamSynchronize(m->eng->stream);  // nothing in flight may still touch the pages that are about to be unmapped
    std::lock_guard<std::mutex> lk(m->mu);
    if (!internal) m->user_seqs--;
    m->live_seqs--;
}

void Seq::ensure_kv(int n_positions) {
    if (!internal) return;  // fully mapped at creation
    const int n = std::min(n_positions, m->hp.n_ctx);
    kv.ensure((size_t)n * m->kv_pos_bytes());
}

void Seq::reset() {
    n_past = 0;
    have_logits = false;
    if (internal && !kv.pages.empty()) {  // give the pages back (all but the first, which the next request will need anyway)
        CUDA_CHECK(cudaStreamSynchronize(m->eng->stream));
        kv.trim(1);
    }
    // the position-split attention's barrier counters only count up: a sequence's worth of launches fits 32 bits many
    // times over, a pooled sequence object that serves requests for days does not — start every sequence at zero
    if (attn_long_scratch.p) CUDA_CHECK(cudaMemsetAsync(attn_long_scratch.p, 0, attn_long_scratch.bytes, m->eng->stream));
    SeqState s{-1, 0, 0, 0, 0ull};
    CUDA_CHECK(cudaMemcpy(state.p, &s, sizeof(s), cudaMemcpyHostToDevice));
}

// ---- the token pass as a list of launches (PassLaunch: engine.h)
void Seq::launch_embed(bool from_prompt, int prompt_idx, int pos_value, cudaStream_t st) {
    const HParams &hp = m->hp;
    SeqState *sd = state.as<SeqState>();
    const int32_t *tok_ptr = from_prompt ? prompt.as<int32_t>() + prompt_idx : &sd->cur_token;
    TIMED(2, launch_embed_row(m->tok_embd_type, m->tok_embd_raw.p, m->tok_embd_row_bytes, hp.n_embd, tok_ptr, x.as<float>(),
                     from_prompt ? &sd->pos : nullptr, pos_value, from_prompt ? &sd->cur_token : nullptr, st));
}

// With a tensor-parallel shard (m->tp_size > 1) hp holds the LOCAL head / FF counts, the K-split projections
// (attn_output, ffn_down) store their unrounded double sums into tp_part64 and are marked reduce_after: the driver
// (tp.cu) all-reduces the partials and adds them to x; the lm_head writes this rank's slice of the logits.
void Seq::build_token_pass(bool want_logits, std::vector<PassLaunch> &ls, bool long_attn) {
    const HParams &hp = m->hp;
    SeqState *sd = state.as<SeqState>();
    const int n_sms = m->eng->n_sms;
    const bool pdl = !(m->eng->flags & AIOS_ENGINE_NO_PDL);
    const bool tp = m->tp_size > 1;
    const int QD = hp.n_head * hp.head_dim, KVD = hp.n_head_kv * hp.head_dim;
    const int amode = (m->eng->flags & AIOS_ENGINE_ATTN_CLUSTER) ? 2 : 0;
    // with the split attention kernel, RoPE + f16 rounding + KV append move into the q/k/v GEMV's epilogue
    // (EPI_QKV_ROPE): the attention kernel then starts on final data, with no dependent table loads of its own
    // (the position-split kernel of attention_long.cu takes such launches too: it is the decode attention of contexts
    // beyond the split kernel's 4096 rows, and of long contexts below that)
    // With the position-split kernel taking over at attn_long_min (<= 4096) rows, the latency kernel never sees more
    // than that: a model loaded with a bigger n_ctx still starts on it, its shared-memory score row sized for 4096.
    const bool long_ok = attn_long_scratch.p != nullptr;
    const int split_ctx = long_ok ? std::min(hp.n_ctx, 4096) : hp.n_ctx;
    const bool split_ok = attn_split_usable(split_ctx, hp.head_dim);
    const bool rope_in_gemv = amode == 0 && (split_ok || long_ok) && !(m->eng->flags & AIOS_ENGINE_NO_ROPE_EPILOGUE);
    ls.clear();
    ls.reserve((size_t)hp.n_layer * 6 + 2);
    int n_reduce = 0;
    auto add_gemv = [&](const GemvParams &p, bool use_pdl, bool reduce_after = false) {
        PassLaunch L;
        L.kind = 0; L.g = p; L.pdl = use_pdl; L.reduce_after = reduce_after;
        if (reduce_after && tp_push.n > 0) {
            L.g.epi = EPI_PUSH_F64;
            L.g.push = tp_push;
            L.g.push.index = n_reduce;
        }
        if (reduce_after) n_reduce++;
        L.g.evict_first = m->eng->evict_first ? 1 : 0;
        L.grid = gemv_plan_grid(L.g, n_sms);
        ls.push_back(L);
    };
    for (int il = 0; il < hp.n_layer; il++) {
        const Layer &L = m->layers[il];
        bool kv_ready = false;
        {   // RMSNorm + Q,K,V projections: one launch when the three matrices share a type (else q,k + v)
            GemvParams p{};
            bool same = L.wv.w.type == L.wq.w.type && L.wk.w.type == L.wq.w.type;
            p.n_seg = 3; p.epi = EPI_STORE; p.k = hp.n_embd;
            p.seg[0] = L.wq.w; p.seg[1] = L.wk.w; p.seg[2] = L.wv.w;
            if (!same && gemv_can_mix(p)) same = true;  // Q4_K + Q6_K: one launch, the type is per CTA
            p.n_seg = same ? 3 : 2;
            p.out[0] = qkv.as<float>(); p.out[1] = qkv.as<float>() + QD; p.out[2] = qkv.as<float>() + QD + KVD;
            p.x = x.as<float>(); p.norm_w = L.attn_norm.as<float>(); p.eps = hp.rms_eps;
            if (same && rope_in_gemv) {  // RoPE, f16 rounding and the KV append ride in this launch's epilogue
                p.epi = EPI_QKV_ROPE;
                p.rope.cos = m->rope_cos.as<float>(); p.rope.sin = m->rope_sin.as<float>();
                p.rope.n_past_ptr = &sd->pos;
                p.rope.k_cache = k_cache(il); p.rope.v_cache = v_cache(il);
                p.rope.n_rot = hp.n_rot; p.rope.hd = hp.head_dim; p.rope.kvd = KVD; p.rope.kv_stride = (int32_t)m->kv_pos_elems();
                kv_ready = true;
            }
            add_gemv(p, pdl);
            if (!same) {
                GemvParams v{};
                v.n_seg = 1; v.epi = EPI_STORE; v.k = hp.n_embd;
                v.seg[0] = L.wv.w; v.out[0] = qkv.as<float>() + QD + KVD;
                v.x = x.as<float>(); v.norm_w = L.attn_norm.as<float>(); v.eps = hp.rms_eps;
                add_gemv(v, pdl);
            }
        }
        {   // RoPE + KV write + flash-decode
            PassLaunch A;
            A.kind = 1; A.pdl = pdl;
            AttnParams &a = A.a;
            a.qkv = qkv.as<float>(); a.out = att.as<float>();
            a.k_cache = k_cache(il); a.v_cache = v_cache(il); a.kv_stride = (int32_t)m->kv_pos_elems();
            a.rope_cos = m->rope_cos.as<float>(); a.rope_sin = m->rope_sin.as<float>();
            a.n_past_ptr = &sd->pos;
            a.scratch = attn_scratch.as<float>(); a.counters = attn_counters.as<uint32_t>();
            a.n_head = hp.n_head; a.n_kv = hp.n_head_kv; a.hd = hp.head_dim; a.n_rot = hp.n_rot;
            a.n_split = m->attn_splits; a.n_ctx = hp.n_ctx;
            a.scale = 1.0f / sqrtf((float)hp.head_dim);
            a.kv_ready = kv_ready ? 1 : 0;
            if (kv_ready && long_ok && (long_attn || !split_ok)) A.kind = 2;  // the position-split kernel
            else if (kv_ready && long_ok) a.n_ctx = split_ctx;
            ls.push_back(A);
        }
        {   // attention output projection + residual add
            GemvParams p{};
            p.n_seg = 1; p.epi = tp ? EPI_STORE_F64 : EPI_ADD; p.k = QD;
            p.seg[0] = L.wo.w; p.out[0] = tp ? tp_part64.as<float>() : x.as<float>(); p.x = att.as<float>();
            add_gemv(p, pdl, tp);
        }
        {   // RMSNorm + gate/up + SwiGLU (after an all-reduce the previous launch on the stream is not a PDL primary)
            GemvParams p{};
            p.n_seg = 2; p.epi = EPI_SWIGLU; p.k = hp.n_embd;
            p.seg[0] = L.w_gate.w; p.seg[1] = L.w_up.w; p.out[0] = h.as<float>();
            p.x = x.as<float>(); p.norm_w = L.ffn_norm.as<float>(); p.eps = hp.rms_eps;
            add_gemv(p, pdl && (!tp || tp_pdl_wait));
        }
        {   // down projection + residual add
            GemvParams p{};
            p.n_seg = 1; p.epi = tp ? EPI_STORE_F64 : EPI_ADD; p.k = hp.n_ff;
            p.seg[0] = L.w_down.w; p.out[0] = tp ? tp_part64.as<float>() : x.as<float>(); p.x = h.as<float>();
            add_gemv(p, pdl, tp);
        }
    }
    if (want_logits) {
        // the lm_head epilogue keeps the packed (value, index) arg-max of the logits it writes
        GemvParams p{};
        p.n_seg = 1; p.epi = EPI_STORE; p.k = hp.n_embd;
        p.seg[0] = m->output.w; p.out[0] = logits.as<float>() + (size_t)m->tp_rank * m->vocab_local;
        p.x = x.as<float>(); p.norm_w = m->output_norm.as<float>(); p.eps = hp.rms_eps;
        p.argmax_key = tp ? nullptr : &sd->argmax_key;
        add_gemv(p, false);
        if (tp && tp_push_logits.n > 0) {
            ls.back().g.epi = EPI_PUSH_F32;
            ls.back().g.push = tp_push_logits;
        }
    }
    if (tp && !tp_pdl_wait)  // the launch after an all-reduce follows NCCL's / a plain add kernel: no PDL primary
        for (size_t i = 1; i < ls.size(); i++)
            if (ls[i - 1].reduce_after) ls[i].pdl = false;
}

void Seq::launch_pass_item(PassLaunch &L, cudaStream_t st, int launch_index) {
    SeqState *sd = state.as<SeqState>();
    unsigned long long *tr = trace_base && launch_index < 4096 ? trace_base + (size_t)launch_index * 8 : nullptr;
    if (L.kind == 0) {
        if (L.g.argmax_key) CUDA_CHECK(cudaMemsetAsync(&sd->argmax_key, 0, sizeof(unsigned long long), st));
        L.g.trace = tr;
        TIMED(0, launch_gemv(L.g, L.grid, st, L.pdl));
    } else if (L.kind == 2) {
        L.a.trace = tr;
        TIMED(1, launch_attn_long(L.a, attn_long_scratch.p, attn_long_n_slices, st, L.pdl));
    } else {
        L.a.trace = tr;
        const int amode = (m->eng->flags & AIOS_ENGINE_ATTN_CLUSTER) ? 2 : 0;
        TIMED(1, launch_attn_decode(L.a, st, L.pdl, amode));
    }
}

void Seq::launch_token_pass(bool from_prompt, int prompt_idx, int pos_value, bool want_logits, cudaStream_t st, int *n_launch,
                            bool long_attn) {
    if (m->tp_size > 1) throw std::logic_error("a tensor-parallel shard is driven by its group (tp.cu)");
    int launches = 0;
    launch_embed(from_prompt, prompt_idx, pos_value, st);
    launches++;
    std::vector<PassLaunch> ls;
    build_token_pass(want_logits, ls, long_attn);
    for (size_t i = 0; i < ls.size(); i++) launch_pass_item(ls[i], st, launches++);
    if (n_launch) *n_launch += launches;
}

void Seq::launch_sample_kernel(const aios_sampling &sp, cudaStream_t st) {
    SeqState *sd = state.as<SeqState>();
    SampleParams s{};
    s.logits = logits.as<float>();
    s.n_vocab = m->hp.n_vocab;
    s.temperature = sp.temperature; s.top_k = sp.top_k; s.top_p = sp.top_p; s.min_p = sp.min_p; s.seed = sp.seed;
    s.pos_ptr = &sd->pos; s.step_ptr = &sd->step; s.cur_token = &sd->cur_token;
    s.out_tokens = out_tokens.as<int32_t>(); s.out_cap = m->hp.n_ctx;
    s.mask = ext_mask ? ext_mask : (mask_on ? mask.as<uint8_t>() : nullptr);
    TIMED(2, launch_sample(s, st));
}

void Seq::prefill(const int32_t *tokens, int n) {
    if (n <= 0) return;
    if (n_past + n > m->hp.n_ctx) throw std::runtime_error("prefill would exceed the context length");
    for (int i = 0; i < n; i++)
        if (tokens[i] < 0 || tokens[i] >= m->hp.n_vocab) throw std::invalid_argument("token id out of range");
    m->eng->make_current();
    ensure_kv(n_past + n);
    cudaStream_t st = m->eng->stream;
    CUDA_CHECK(cudaMemcpyAsync(prompt.p, tokens, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    constexpr int TC_MIN_TOKENS = 256, TC_MAX_TOKENS = 4096;
    if (n >= TC_MIN_TOKENS && m->tc_ok && !(m->eng->flags & (AIOS_ENGINE_EXACT_PREFILL | AIOS_ENGINE_NO_GEMM))) {
        // long prompts: the tcgen05 GEMM + flash attention (float arithmetic, stated tolerance)
        for (int i = 0; i < n; i += TC_MAX_TOKENS) {
            const int T = std::min(TC_MAX_TOKENS, n - i);
            prefill_tc(i, T, n_past + i, i + T == n, st);
        }
    } else if (n >= 4 && !(m->eng->flags & AIOS_ENGINE_NO_GEMM)) {
        // prompt chunks of up to 64 tokens: every matrix is streamed once per chunk (gemm_q.cu)
        for (int i = 0; i < n; i += GEMM_MAX_TOKENS) {
            const int T = std::min((int)GEMM_MAX_TOKENS, n - i);
            prefill_chunk(i, T, n_past + i, i + T == n, st);
        }
    } else {
        for (int i = 0; i < n; i++) launch_token_pass(true, i, n_past + i, i == n - 1, st, nullptr, want_long_attn(n_past + i + 1));
    }
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaStreamSynchronize(st));
    n_past += n;
    have_logits = true;
}

static bool same_sampling(const aios_sampling &a, const aios_sampling &b) {
    return a.temperature == b.temperature && a.top_k == b.top_k && a.top_p == b.top_p && a.min_p == b.min_p && a.seed == b.seed;
}

cudaGraphExec_t Seq::ensure_graph(const aios_sampling &sp, bool long_attn) {
    StepGraph &G = graphs[long_attn ? 1 : 0];
    if (G.exec && same_sampling(sp, G.sampling) && G.mask == mask_on && G.mask_ptr == ext_mask) return G.exec;
    if (G.exec) { cudaGraphExecDestroy(G.exec); G.exec = nullptr; }
    cudaStream_t st = m->eng->stream;
    cudaGraph_t g = nullptr;
    kernels_per_step = 0;
    CUDA_CHECK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    try {
        launch_sample_kernel(sp, st);
        kernels_per_step++;
        launch_token_pass(false, 0, 0, true, st, &kernels_per_step, long_attn);
    } catch (...) {
        cudaStreamEndCapture(st, &g);
        if (g) cudaGraphDestroy(g);
        throw;
    }
    CUDA_CHECK(cudaStreamEndCapture(st, &g));
    cudaError_t e = cudaGraphInstantiate(&G.exec, g, 0);
    cudaGraphDestroy(g);
    CUDA_CHECK(e);
    G.sampling = sp;
    G.mask = mask_on;
    G.mask_ptr = ext_mask;
    return G.exec;
}

// One prompt chunk of T <= 64 tokens at positions pos0 .. pos0+T-1.  Same arithmetic as T token passes (the logits and
// the KV rows come out bit-identical); what changes is the weight traffic: every matrix is streamed once for the whole
// chunk by the multi-token GEMM (gemm_q.cu; q/k/v and gate/up as segments of one launch); RoPE + KV append for all
// rows, then batched attention (one CTA per (token, kv head), each token over the rows before it).  A matrix the GEMM
// cannot take (more than two weight-type classes in one group) falls back to the GEMV, token by token.
void Seq::prefill_chunk(int idx0, int T, int pos0, bool want_logits, cudaStream_t st, ChunkReduce *reduce) {
    const HParams &hp = m->hp;
    constexpr int CH = GEMM_MAX_TOKENS;
    const int E = hp.n_embd, FF = hp.n_ff, QD = hp.n_head * hp.head_dim, KVD = hp.n_head_kv * hp.head_dim;
    const int LQ = QD + 2 * KVD, Tp = gemm_pad_tokens(T), n_sms = m->eng->n_sms;
    SeqState *sd = state.as<SeqState>();
    const bool tp = m->tp_size > 1;
    if (tp && (!reduce || want_logits)) throw std::logic_error("a tensor-parallel shard's chunks are driven by its group (tp.cu)");
    if (!cx.p) {  // first chunk of this sequence: room for CH tokens side by side
        cx = DevBuf((size_t)CH * E * 4); cqkv = DevBuf((size_t)CH * LQ * 4); catt = DevBuf((size_t)CH * QD * 4);
        ch = DevBuf((size_t)CH * FF * 4); cgu = DevBuf((size_t)CH * 2 * FF * 4); cqh = DevBuf((size_t)CH * QD * 2);
        cimg = DevBuf(2 * gemm_images_bytes(std::max(std::max(E, QD), FF), CH)); cpos = DevBuf(CH * 4);
        if (m->tp_size > 1) cpart = DevBuf((size_t)CH * E * sizeof(double));
        CUDA_CHECK(cudaMemsetAsync(cx.p, 0, cx.bytes, st)); CUDA_CHECK(cudaMemsetAsync(catt.p, 0, catt.bytes, st));
        CUDA_CHECK(cudaMemsetAsync(ch.p, 0, ch.bytes, st));  // rows of padding tokens stay zero (finite images)
    }
    float *X = cx.as<float>(), *QKV = cqkv.as<float>(), *ATT = catt.as<float>(), *H = ch.as<float>(), *GU = cgu.as<float>();
    uint8_t *IMG[2] = {cimg.as<uint8_t>(), cimg.as<uint8_t>() + cimg.bytes / 2};
    int32_t *POS = cpos.as<int32_t>();
    const int32_t *toks = prompt.as<int32_t>() + idx0;
    launch_embed_rows(m->tok_embd_type, m->tok_embd_raw.p, m->tok_embd_row_bytes, E, toks, T, X, E, st);
    {   // positions of the chunk's tokens; the last token also leaves its position / id in the sequence state
        std::vector<int32_t> hp_pos((size_t)T);
        for (int t = 0; t < T; t++) hp_pos[(size_t)t] = pos0 + t;
        CUDA_CHECK(cudaMemcpyAsync(POS, hp_pos.data(), (size_t)T * 4, cudaMemcpyHostToDevice, st));
        CUDA_CHECK(cudaStreamSynchronize(st));  // hp_pos is a stack buffer
        CUDA_CHECK(cudaMemcpyAsync(&sd->pos, POS + (T - 1), 4, cudaMemcpyDeviceToDevice, st));
        CUDA_CHECK(cudaMemcpyAsync(&sd->cur_token, toks + (T - 1), 4, cudaMemcpyDeviceToDevice, st));
    }
    auto perm_class = [](int type) { return type == T_Q6_K ? 1 : type == T_Q8_0 ? 2 : 0; };
    auto gemv1 = [&](const WPlanes &w, const float *xin, const float *norm, float *out, int epi) {
        GemvParams p{};
        p.n_seg = 1; p.epi = epi; p.k = w.k; p.seg[0] = w; p.out[0] = out; p.x = xin; p.norm_w = norm; p.eps = hp.rms_eps;
        p.evict_first = m->eng->evict_first ? 1 : 0;
        launch_gemv(p, gemv_plan_grid(p, n_sms), st, false);
    };
    // Y[t] (=|+=) W . act(in[t]) for all tokens of the chunk and every matrix of the group (they share the input)
    auto matmul = [&](std::initializer_list<const WPlanes *> ws, std::initializer_list<float *> outs, const float *in, int64_t ldin,
                      const float *norm, int64_t ldo, int epi) {
        GemmMulti g{};
        g.T = T; g.epi = epi; g.ldo = ldo;
        uint8_t *img_of_class[3] = {nullptr, nullptr, nullptr};
        int n_bufs = 0;
        auto o = outs.begin();
        for (const WPlanes *w : ws) {
            float *out = *o++;
            const int pc = perm_class(w->type);
            if (!gemm_q_usable(w->type, w->k) || (!img_of_class[pc] && n_bufs >= 2)) {
                for (int t = 0; t < T; t++) gemv1(*w, in + (size_t)t * ldin, norm, out + (size_t)t * ldo, epi);
                continue;
            }
            if (!img_of_class[pc]) {
                img_of_class[pc] = IMG[n_bufs++];
                launch_quant_rows(w->type, in, ldin, norm, hp.rms_eps, w->k, Tp, img_of_class[pc], st);
            }
            g.seg[g.n_seg] = *w; g.out[g.n_seg] = out; g.images[g.n_seg] = img_of_class[pc];
            g.n_seg++;
        }
        if (tp && epi == EPI_ADD) {
            // a K shard of attn_output / ffn_down: f64 partial sums, summed over the ranks, then rounded into the rows
            if (g.n_seg != 1 || ws.size() != 1) throw CudaError("tensor-parallel chunk: the K-split matrix must take the GEMM path");
            g.part = cpart.as<double>();
            launch_gemm_multi(g, st);
            reduce->all_reduce(cpart.as<double>(), (size_t)T * E, st);
            launch_add_part_rows(g.out[0], ldo, cpart.as<double>(), E, T, st);
            return;
        }
        if (g.n_seg) launch_gemm_multi(g, st);
    };
    for (int il = 0; il < hp.n_layer; il++) {
        const Layer &L = m->layers[il];
        matmul({&L.wq.w, &L.wk.w, &L.wv.w}, {QKV, QKV + QD, QKV + QD + KVD}, X, E, L.attn_norm.as<float>(), LQ, EPI_STORE);
        if (attn_batch_usable(hp.n_head, hp.n_head_kv, hp.n_ctx, hp.head_dim)) {
            // RoPE + KV append for all T rows first (token t must see the rows of the tokens before it), then attention
            // with one CTA per (token, kv head); same operations as the single-token kernels, bit for bit
            launch_rope_kv_rows(QKV, cqh.as<__half>(), k_cache(il), v_cache(il), m->rope_cos.as<float>(), m->rope_sin.as<float>(), T,
                                pos0, hp.n_head, hp.n_head_kv, hp.head_dim, hp.n_rot, (int)m->kv_pos_elems(), st);
            AttnBatchParams a{};
            a.qh = cqh.as<__half>(); a.out = ATT;
            a.lane_pos = POS; a.lane_kv = nullptr; a.kv_base = kv.as<uint16_t>();
            a.k_off = (int64_t)(((size_t)il * 2 + 0) * m->kvd()); a.v_off = (int64_t)(((size_t)il * 2 + 1) * m->kvd());
            a.lane_out_stride = QD;
            a.n_head = hp.n_head; a.n_kv = hp.n_head_kv; a.hd = hp.head_dim; a.n_rot = hp.n_rot; a.n_ctx = hp.n_ctx;
            a.kv_stride = (int32_t)m->kv_pos_elems(); a.kv_ready = 1;
            a.scale = 1.0f / sqrtf((float)hp.head_dim);
            launch_attn_batch(a, T, st);
        } else
        for (int t = 0; t < T; t++) {  // RoPE + KV append + attention, token by token (each sees the rows before it)
            AttnParams a{};
            a.qkv = QKV + (size_t)t * LQ; a.out = ATT + (size_t)t * QD;
            a.k_cache = k_cache(il); a.v_cache = v_cache(il); a.kv_stride = (int32_t)m->kv_pos_elems();
            a.rope_cos = m->rope_cos.as<float>(); a.rope_sin = m->rope_sin.as<float>();
            a.n_past_ptr = POS + t;
            a.scratch = attn_scratch.as<float>(); a.counters = attn_counters.as<uint32_t>();
            a.n_head = hp.n_head; a.n_kv = hp.n_head_kv; a.hd = hp.head_dim; a.n_rot = hp.n_rot;
            a.n_split = m->attn_splits; a.n_ctx = hp.n_ctx;
            a.scale = 1.0f / sqrtf((float)hp.head_dim);
            launch_attn_decode(a, st, false, (m->eng->flags & AIOS_ENGINE_ATTN_CLUSTER) ? 2 : 0);
        }
        matmul({&L.wo.w}, {X}, ATT, QD, nullptr, E, EPI_ADD);
        matmul({&L.w_gate.w, &L.w_up.w}, {GU, GU + FF}, X, E, L.ffn_norm.as<float>(), 2 * FF, EPI_STORE);
        launch_swiglu_rows(GU, 2 * FF, FF, T, H, FF, st);
        matmul({&L.w_down.w}, {X}, H, FF, nullptr, E, EPI_ADD);
    }
    if (want_logits) {
        CUDA_CHECK(cudaMemsetAsync(&sd->argmax_key, 0, sizeof(unsigned long long), st));
        GemvParams p{};
        p.n_seg = 1; p.epi = EPI_STORE; p.k = E;
        p.seg[0] = m->output.w; p.out[0] = logits.as<float>();
        p.x = X + (size_t)(T - 1) * E; p.norm_w = m->output_norm.as<float>(); p.eps = hp.rms_eps;
        p.argmax_key = &sd->argmax_key;
        p.evict_first = m->eng->evict_first ? 1 : 0;
        launch_gemv(p, gemv_plan_grid(p, n_sms), st, false);
    }
}

// One prompt chunk of T tokens on the tensor-core path: every matrix product through the tcgen05 GEMM (prefill_tc.cu),
// RoPE + KV append for all T rows at once, causal flash attention over the KV range (attention_prefill.cu).  The
// residual stream stays f32; GEMM inputs are f16.  The last token's logits come from the exact lm_head GEMV.
void Seq::prefill_tc(int idx0, int T, int pos0, bool want_logits, cudaStream_t st) {
    const HParams &hp = m->hp;
    const int E = hp.n_embd, FF = hp.n_ff, QD = hp.n_head * hp.head_dim, KVD = hp.n_head_kv * hp.head_dim, LQ = QD + 2 * KVD;
    const int n_sms = m->eng->n_sms;
    TcScratch &S = m->tc;
    if (S.cap < T) {
        CUDA_CHECK(cudaStreamSynchronize(st));
        S = TcScratch{};
        S.X = DevBuf((size_t)T * E * 4); S.Xh = DevBuf((size_t)T * std::max(E, QD) * 2); S.QKV = DevBuf((size_t)T * LQ * 4);
        S.Qh = DevBuf((size_t)T * QD * 2); S.ATTh = DevBuf((size_t)T * QD * 2); S.GU = DevBuf((size_t)T * 2 * FF * 4);
        S.Hh = DevBuf((size_t)T * FF * 2);
        S.cap = T;
    }
    SeqState *sd = state.as<SeqState>();
    float *X = S.X.as<float>(), *QKV = S.QKV.as<float>(), *GU = S.GU.as<float>();
    __half *Xh = S.Xh.as<__half>(), *Qh = S.Qh.as<__half>(), *ATTh = S.ATTh.as<__half>(), *Hh = S.Hh.as<__half>();
    const int32_t *toks = prompt.as<int32_t>() + idx0;
    launch_embed_rows(m->tok_embd_type, m->tok_embd_raw.p, m->tok_embd_row_bytes, E, toks, T, X, E, st);
    const int kvs = (int)m->kv_pos_elems();
    for (int il = 0; il < hp.n_layer; il++) {
        const Layer &L = m->layers[il];
        launch_rows_to_f16(X, E, L.attn_norm.as<float>(), hp.rms_eps, E, T, Xh, st);
        launch_prefill_tc(L.wq.w, Xh, T, QKV, LQ, EPI_STORE, n_sms, st);
        launch_prefill_tc(L.wk.w, Xh, T, QKV + QD, LQ, EPI_STORE, n_sms, st);
        launch_prefill_tc(L.wv.w, Xh, T, QKV + QD + KVD, LQ, EPI_STORE, n_sms, st);
        launch_rope_kv_rows(QKV, Qh, k_cache(il), v_cache(il), m->rope_cos.as<float>(), m->rope_sin.as<float>(), T, pos0, hp.n_head,
                            hp.n_head_kv, hp.head_dim, hp.n_rot, kvs, st);
        launch_attn_prefill(Qh, ATTh, k_cache(il), v_cache(il), T, pos0, hp.n_head, hp.n_head_kv, hp.head_dim, kvs, st);
        launch_prefill_tc(L.wo.w, ATTh, T, X, E, EPI_ADD, n_sms, st);
        launch_rows_to_f16(X, E, L.ffn_norm.as<float>(), hp.rms_eps, E, T, Xh, st);
        launch_prefill_tc(L.w_gate.w, Xh, T, GU, 2 * FF, EPI_STORE, n_sms, st);
        launch_prefill_tc(L.w_up.w, Xh, T, GU + FF, 2 * FF, EPI_STORE, n_sms, st);
        launch_swiglu_rows_f16(GU, 2 * FF, FF, T, Hh, st);
        launch_prefill_tc(L.w_down.w, Hh, T, X, E, EPI_ADD, n_sms, st);
    }
    // the sequence state after the chunk: position and id of its last token
    const int32_t last_pos = pos0 + T - 1;
    CUDA_CHECK(cudaMemcpyAsync(&sd->pos, &last_pos, 4, cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(&sd->cur_token, toks + (T - 1), 4, cudaMemcpyDeviceToDevice, st));
    CUDA_CHECK(cudaStreamSynchronize(st));  // last_pos is a stack variable
    if (want_logits) {
        CUDA_CHECK(cudaMemsetAsync(&sd->argmax_key, 0, sizeof(unsigned long long), st));
        GemvParams p{};
        p.n_seg = 1; p.epi = EPI_STORE; p.k = E;
        p.seg[0] = m->output.w; p.out[0] = logits.as<float>();
        p.x = X + (size_t)(T - 1) * E; p.norm_w = m->output_norm.as<float>(); p.eps = hp.rms_eps;
        p.argmax_key = &sd->argmax_key;
        p.evict_first = m->eng->evict_first ? 1 : 0;
        launch_gemv(p, gemv_plan_grid(p, n_sms), st, false);
    }
}

int Seq::decode(const aios_sampling &sp, int max_tokens, int32_t *out) {
    if (!have_logits) throw std::runtime_error("decode needs logits: prefill at least one token first");
    // the sampled token n is fed at position n_past+n, which must fit; the last sampled token is returned un-fed
    const int steps = std::min(max_tokens, m->hp.n_ctx - n_past);
    if (steps <= 0) return 0;
    m->eng->make_current();
    ensure_kv(n_past + steps);
    cudaStream_t st = m->eng->stream;
    const bool use_graph = !(m->eng->flags & AIOS_ENGINE_NO_GRAPH);
    int32_t hdr[2] = {n_past - 1, 0};  // pos, step
    CUDA_CHECK(cudaMemcpyAsync(state.p, hdr, sizeof(hdr), cudaMemcpyHostToDevice, st));
    if (use_pfd(sp)) {
        launch_pfd_tokens(steps, st);  // one cooperative launch for all the tokens
    } else if (use_graph) {
        // step i feeds the token at position n_past + i: n_past + i + 1 rows to attend to
        for (int i = 0; i < steps; i++) CUDA_CHECK(cudaGraphLaunch(ensure_graph(sp, want_long_attn(n_past + i + 1)), st));
    } else {
        for (int i = 0; i < steps; i++) {
            launch_sample_kernel(sp, st);
            launch_token_pass(false, 0, 0, true, st, nullptr, want_long_attn(n_past + i + 1));
        }
    }
    CUDA_CHECK(cudaMemcpyAsync(out, out_tokens.p, (size_t)steps * 4, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    CUDA_CHECK(cudaGetLastError());
    n_past += steps;
    return steps;
}

void Seq::step(const aios_sampling &sp, int32_t *tok) {
    if (decode(sp, 1, tok) != 1) throw std::runtime_error("context is full");
}

void Seq::bench_decode(int n_tokens, aios_bench_result *out) {
    if (!have_logits) throw std::runtime_error("bench_decode needs a prefilled sequence");
    const int steps = std::min(n_tokens, m->hp.n_ctx - n_past);
    if (steps <= 0) throw std::runtime_error("context is full");
    m->eng->make_current();
    ensure_kv(n_past + steps);
    cudaStream_t st = m->eng->stream;
    aios_sampling sp{};  // greedy
    const bool fused = use_pfd(sp);
    if (!fused)
        for (int i = 0; i < steps; i++) ensure_graph(sp, want_long_attn(n_past + i + 1));  // capture outside the timed region
    int32_t hdr[2] = {n_past - 1, 0};
    CUDA_CHECK(cudaMemcpyAsync(state.p, hdr, sizeof(hdr), cudaMemcpyHostToDevice, st));
    cudaEvent_t e0, e1;
    CUDA_CHECK(cudaEventCreate(&e0));
    CUDA_CHECK(cudaEventCreate(&e1));
    CUDA_CHECK(cudaStreamSynchronize(st));
    CUDA_CHECK(cudaEventRecord(e0, st));
    if (fused) launch_pfd_tokens(steps, st);
    else for (int i = 0; i < steps; i++) CUDA_CHECK(cudaGraphLaunch(ensure_graph(sp, want_long_attn(n_past + i + 1)), st));
    CUDA_CHECK(cudaEventRecord(e1, st));
    CUDA_CHECK(cudaEventSynchronize(e1));
    float ms = 0;
    CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    memset(out, 0, sizeof(*out));
    out->ms_total = ms;
    out->n_tokens = steps;
    out->kernel_launches = fused ? 1 : steps * kernels_per_step;
    out->n_past_begin = n_past;
    out->n_past_end = n_past + steps;
    n_past += steps;
}

void Seq::profile_decode(int n_tokens, aios_bench_result *out) {
    if (!have_logits) throw std::runtime_error("profile_decode needs a prefilled sequence");
    const int steps = std::min(n_tokens, m->hp.n_ctx - n_past);
    if (steps <= 0) throw std::runtime_error("context is full");
    m->eng->make_current();
    ensure_kv(n_past + steps);
    cudaStream_t st = m->eng->stream;
    aios_sampling sp{};
    int32_t hdr[2] = {n_past - 1, 0};
    CUDA_CHECK(cudaMemcpyAsync(state.p, hdr, sizeof(hdr), cudaMemcpyHostToDevice, st));
    KernelTimer kt;
    kt.st = st;
    timer = &kt;
    int launches = 0;
    try {
        for (int i = 0; i < steps; i++) {
            launch_sample_kernel(sp, st);
            launches++;
            launch_token_pass(false, 0, 0, true, st, &launches, want_long_attn(n_past + i + 1));
        }
        CUDA_CHECK(cudaStreamSynchronize(st));
    } catch (...) {
        timer = nullptr;
        throw;
    }
    timer = nullptr;
    float ms_kind[3] = {0, 0, 0};
    int n_gemv = 0;
    for (auto &r : kt.recs) {
        float ms = 0;
        cudaEventElapsedTime(&ms, r.a, r.b);
        ms_kind[r.kind] += ms;
        if (r.kind == 0) n_gemv++;
        cudaEventDestroy(r.a);
        cudaEventDestroy(r.b);
    }
    memset(out, 0, sizeof(*out));
    out->ms_total = ms_kind[0] + ms_kind[1] + ms_kind[2];
    out->n_tokens = steps;
    out->kernel_launches = launches;
    out->n_past_begin = n_past;
    out->n_past_end = n_past + steps;
    out->ms_gemv = ms_kind[0];
    out->reserved[0] = ms_kind[1];
    out->reserved[1] = ms_kind[2];
    out->reserved[2] = (float)n_gemv;
    n_past += steps;
}

// Two greedy tokens launched kernel by kernel on the stream (PDL attributes as in the graph); during the second one
// CTA 0 of every ring-GEMV / split-attention launch writes its %globaltimer marks (kernels.h) into `trace`.
int Seq::trace_decode(unsigned long long *out, int cap_launches) {
    if (!have_logits) throw std::runtime_error("trace_decode needs a prefilled sequence");
    if (m->hp.n_ctx - n_past < 2) throw std::runtime_error("context is full");
    m->eng->make_current();
    ensure_kv(n_past + 2);
    cudaStream_t st = m->eng->stream;
    aios_sampling sp{};
    int32_t hdr[2] = {n_past - 1, 0};
    CUDA_CHECK(cudaMemcpyAsync(state.p, hdr, sizeof(hdr), cudaMemcpyHostToDevice, st));
    const size_t cap = 4096;
    if (trace.bytes < cap * 8 * sizeof(unsigned long long)) trace = DevBuf(cap * 8 * sizeof(unsigned long long));
    CUDA_CHECK(cudaMemsetAsync(trace.p, 0, trace.bytes, st));
    int launches = 0;
    launch_sample_kernel(sp, st);
    launch_token_pass(false, 0, 0, true, st, nullptr, want_long_attn(n_past + 1));
    launch_sample_kernel(sp, st);
    trace_base = trace.as<unsigned long long>();
    try {
        launch_token_pass(false, 0, 0, true, st, &launches, want_long_attn(n_past + 2));
        CUDA_CHECK(cudaStreamSynchronize(st));
    } catch (...) {
        trace_base = nullptr;
        throw;
    }
    trace_base = nullptr;
    n_past += 2;
    const int n = std::min(launches, cap_launches);
    CUDA_CHECK(cudaMemcpy(out, trace.p, (size_t)n * 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    return n;
}

void Seq::snapshot() {
    if (!have_logits) throw std::runtime_error("nothing to snapshot: prefill first");
    m->eng->make_current();
    if (!snap_logits.p) snap_logits = DevBuf(logits.bytes + sizeof(unsigned long long));
    CUDA_CHECK(cudaMemcpyAsync(snap_logits.p, logits.p, logits.bytes, cudaMemcpyDeviceToDevice, m->eng->stream));
    CUDA_CHECK(cudaMemcpyAsync(snap_logits.as<uint8_t>() + logits.bytes, &state.as<SeqState>()->argmax_key,
                               sizeof(unsigned long long), cudaMemcpyDeviceToDevice, m->eng->stream));  // the logits' arg-max key
    CUDA_CHECK(cudaStreamSynchronize(m->eng->stream));
    snap_n_past = n_past;
}

void Seq::restore() {
    if (snap_n_past < 0) throw std::runtime_error("no snapshot");
    m->eng->make_current();
    CUDA_CHECK(cudaMemcpyAsync(logits.p, snap_logits.p, logits.bytes, cudaMemcpyDeviceToDevice, m->eng->stream));
    CUDA_CHECK(cudaMemcpyAsync(&state.as<SeqState>()->argmax_key, snap_logits.as<uint8_t>() + logits.bytes,
                               sizeof(unsigned long long), cudaMemcpyDeviceToDevice, m->eng->stream));
    n_past = snap_n_past;
    have_logits = true;
}

// ---------------------------------------------------------------- persistent fused decode (pfd.cu)
static int perm_bit(int type) { return type == T_Q6_K ? 2 : (type == T_Q8_0 ? 4 : 1); }

void Seq::build_pfd() {
    const HParams &hp = m->hp;
    const int QD = hp.n_head * hp.head_dim, KVD = hp.n_head_kv * hp.head_dim;
    std::vector<PfdPhase> ph;
    // one phase per matrix (or gate/up pair); matrices sharing an input form a group: the first builds the
    // activation image(s) for all of them, only the last is followed by a grid barrier
    auto gemv = [&](const WPlanes &w0, const WPlanes *w1, float *out, int epi, int k, const float *xin, const double *x64,
                    const float *norm_w, bool prologue, bool barrier_after, int group_mask) {
        PfdPhase P{};
        P.kind = PH_GEMV; P.type = w0.type; P.epi = epi; P.k = k;
        P.w0 = w0; P.w1 = w1 ? *w1 : w0; P.out = out;
        P.x = xin; P.x64 = x64; P.norm_w = norm_w; P.eps = hp.rms_eps;
        P.prologue = prologue; P.barrier_after = barrier_after; P.perm_mask = group_mask;
        P.units_total = epi == EPI_SWIGLU ? w0.n_rows : (w0.n_rows + 1) / 2;
        ph.push_back(P);
        return ph.size() - 1;
    };
    float *xq = x.as<float>(), *qk = qkv.as<float>(), *hh = h.as<float>();
    bool mixed80 = false;
    for (int il = 0; il < hp.n_layer; il++) {
        const Layer &L = m->layers[il];
        {
            const int mask = perm_bit(L.wq.w.type) | perm_bit(L.wk.w.type) | perm_bit(L.wv.w.type);
            if ((mask & 4) && (mask & 3)) mixed80 = true;
            const float *nw = L.attn_norm.as<float>();
            const size_t i = gemv(L.wq.w, nullptr, qk, EPI_STORE, hp.n_embd, xq, nullptr, nw, true, false, mask);
            gemv(L.wk.w, nullptr, qk + QD, EPI_STORE, hp.n_embd, xq, nullptr, nw, false, false, mask);
            gemv(L.wv.w, nullptr, qk + QD + KVD, EPI_STORE, hp.n_embd, xq, nullptr, nw, false, true, mask);
            if (il == 0) ph[i].embed = 1;
        }
        {
            PfdPhase P{};
            P.kind = PH_ATTN;
            P.k_cache = k_cache(il); P.v_cache = v_cache(il);
            ph.push_back(P);
        }
        gemv(L.wo.w, nullptr, xq, EPI_ADD, QD, nullptr, pfd_att_acc.as<double>(), nullptr, true, true, perm_bit(L.wo.w.type));
        {
            const size_t i = gemv(L.w_gate.w, &L.w_up.w, hh, EPI_SWIGLU, hp.n_embd, xq, nullptr, L.ffn_norm.as<float>(), true,
                                  true, perm_bit(L.w_gate.w.type));
            ph[i].zero_att = 1;
        }
        gemv(L.w_down.w, nullptr, xq, EPI_ADD, hp.n_ff, hh, nullptr, nullptr, true, true, perm_bit(L.w_down.w.type));
    }
    {
        const size_t i = gemv(m->output.w, nullptr, logits.as<float>(), EPI_STORE, hp.n_embd, xq, nullptr,
                              m->output_norm.as<float>(), true, true, perm_bit(m->output.w.type));
        ph[i].argmax = 1;
    }
    pfd_phases = DevBuf(ph.size() * sizeof(PfdPhase));
    CUDA_CHECK(cudaMemcpy(pfd_phases.p, ph.data(), ph.size() * sizeof(PfdPhase), cudaMemcpyHostToDevice));

    SeqState *sd = state.as<SeqState>();
    PfdParams &p = pfd;
    p = PfdParams{};
    p.phases = pfd_phases.as<PfdPhase>();
    p.n_phases = (int)ph.size();
    p.out_tokens = out_tokens.as<int32_t>();
    p.state = reinterpret_cast<int32_t *>(sd);
    p.argmax_key = &sd->argmax_key;
    p.x = xq; p.qkv = qk;
    p.tok_embd = m->tok_embd_raw.as<uint8_t>();
    p.tok_embd_row_bytes = m->tok_embd_row_bytes; p.tok_embd_type = m->tok_embd_type; p.n_embd = hp.n_embd;
    p.rope_cos = m->rope_cos.as<float>(); p.rope_sin = m->rope_sin.as<float>();
    p.n_head = hp.n_head; p.n_kv = hp.n_head_kv; p.hd = hp.head_dim; p.n_rot = hp.n_rot;
    p.kv_stride = (int32_t)m->kv_pos_elems();
    p.att_scale = 1.0f / sqrtf((float)hp.head_dim);
    const int n_sms = m->eng->n_sms;
    p.att_group = std::max(1, std::min(16, n_sms / hp.n_head_kv));
    p.att_per_cap = ((hp.n_ctx + p.att_group - 1) / p.att_group + 1) & ~1;
    p.att_max = pfd_att_max.as<float>(); p.att_sum = pfd_att_sum.as<double>(); p.att_acc = pfd_att_acc.as<double>();
    p.att_counters = pfd_counters.as<unsigned>(); p.bar_counter = pfd_counters.as<unsigned>() + hp.n_head_kv;
    p.k_max = std::max(std::max(hp.n_embd, hp.n_ff), QD);
    pfd_timing = DevBuf(ph.size() * 4 * sizeof(unsigned long long));
    CUDA_CHECK(cudaMemset(pfd_timing.p, 0, pfd_timing.bytes));
    p.timing = pfd_timing.as<unsigned long long>();
    pfd_grid = (hp.n_head_kv <= n_sms && !mixed80) ? pfd_max_grid(n_sms, p) : 0;
}

bool Seq::use_pfd(const aios_sampling &sp) const {
    // opt-in until it beats the per-op path (AIOS_ENGINE_USE_PFD); AIOS_ENGINE_NO_PFD always wins
    const int f = m->eng->flags;
    return pfd_grid > 0 && !mask_on && !ext_mask && !(sp.temperature > 0.f) && (f & AIOS_ENGINE_USE_PFD) && !(f & AIOS_ENGINE_NO_PFD);
}

// n greedy tokens after the current position; device state (pos, step) must already be set
void Seq::launch_pfd_tokens(int n, cudaStream_t st) {
    PfdParams p = pfd;
    p.n_tokens = n;
    p.pos0 = n_past - 1;
    p.step0 = 0;
    CUDA_CHECK(cudaMemsetAsync(pfd_counters.p, 0, pfd_counters.bytes, st));
    launch_pfd(p, pfd_grid, st);
}

// Greedy decode through the fused kernel with its exchange scratch taken from the library's built-in
// static device array (pfd.cu) instead of this sequence's cudaMalloc'd buffers.  Nothing on this path
// allocates: only memset / launch / memcpy on the engine's stream.
int Seq::decode_static(int max_tokens, int32_t *out) {
    if (!have_logits) throw std::runtime_error("decode needs logits: prefill at least one token first");
    if (pfd_grid <= 0) throw std::runtime_error("the fused decode kernel is not usable for this model on this device");
    const HParams &hp = m->hp;
    if (hp.n_head_kv > PFD_STATIC_MAX_KV_HEADS || hp.n_head * hp.head_dim > PFD_STATIC_MAX_QDIM)
        throw std::length_error("model too large for the built-in scratch: at most " +
                                std::to_string(PFD_STATIC_MAX_KV_HEADS) + " kv heads and n_head*head_dim <= " +
                                std::to_string(PFD_STATIC_MAX_QDIM) + " (" + std::to_string(PFD_STATIC_SCRATCH_BYTES) +
                                " bytes)");
    const int steps = std::min(max_tokens, hp.n_ctx - n_past);
    if (steps <= 0) return 0;
    m->eng->make_current();
    cudaStream_t st = m->eng->stream;
    // first call on a device resolves the array's device address (cudaGetSymbolAddress)
    uint8_t *base = static_cast<uint8_t *>(pfd_static_scratch(m->eng->device));
    constexpr size_t OFF_COUNTERS = 0;                                                // 32 x u32
    constexpr size_t OFF_MAX = 256;                                                   // [8][16][8] f32
    constexpr size_t OFF_SUM = OFF_MAX + PFD_STATIC_MAX_KV_HEADS * 16 * 8 * sizeof(float);   // [8][16][8] f64
    constexpr size_t OFF_ACC = OFF_SUM + PFD_STATIC_MAX_KV_HEADS * 16 * 8 * sizeof(double);  // [4096] f64
    static_assert(OFF_ACC + PFD_STATIC_MAX_QDIM * sizeof(double) <= PFD_STATIC_SCRATCH_BYTES, "scratch layout");
    PfdParams p = pfd;
    p.att_counters = reinterpret_cast<unsigned *>(base + OFF_COUNTERS);
    p.bar_counter = p.att_counters + 16;
    p.att_max = reinterpret_cast<float *>(base + OFF_MAX);
    p.att_sum = reinterpret_cast<double *>(base + OFF_SUM);
    p.att_acc = reinterpret_cast<double *>(base + OFF_ACC);
    p.timing = nullptr;
    // the attn_output phases read the accumulators through the phase table, which points at this sequence's own
    // buffer; the static variant therefore uses a second, static-scratch copy of the table's pointer: see below
    p.x64_override = p.att_acc;
    p.n_tokens = steps;
    p.pos0 = n_past - 1;
    p.step0 = 0;
    // the array keeps whatever the previous call left: counters and accumulators must start at zero
    CUDA_CHECK(cudaMemsetAsync(base + OFF_COUNTERS, 0, 256, st));
    CUDA_CHECK(cudaMemsetAsync(base + OFF_ACC, 0, (size_t)hp.n_head * hp.head_dim * sizeof(double), st));
    int32_t hdr[2] = {n_past - 1, 0};
    CUDA_CHECK(cudaMemcpyAsync(state.p, hdr, sizeof(hdr), cudaMemcpyHostToDevice, st));
    launch_pfd(p, pfd_grid, st);
    CUDA_CHECK(cudaMemcpyAsync(out, out_tokens.p, (size_t)steps * 4, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    CUDA_CHECK(cudaGetLastError());
    n_past += steps;
    return steps;
}

// Logit mask for the sampler: bit i of byte i/8 set = token i may be sampled.  It stays in force (for greedy and
// stochastic sampling alike) until replaced or cleared; JSON mode uploads the grammar's allowed set before every step.
void Seq::set_mask(const uint8_t *bits, int n_bytes) {
    if (!bits) { mask_on = false; return; }
    const int need = (m->hp.n_vocab + 7) / 8;
    if (n_bytes < need) throw std::invalid_argument("mask shorter than n_vocab bits");
    m->eng->make_current();
    if (!mask.p) mask = DevBuf((size_t)need);
    CUDA_CHECK(cudaMemcpyAsync(mask.p, bits, (size_t)need, cudaMemcpyHostToDevice, m->eng->stream));
    CUDA_CHECK(cudaStreamSynchronize(m->eng->stream));  // the caller's buffer is free again on return
    mask_on = true;
}

void Seq::read_logits(float *out) {
    if (!have_logits) throw std::runtime_error("no logits yet");
    m->eng->make_current();
    CUDA_CHECK(cudaMemcpyAsync(out, logits.p, (size_t)m->hp.n_vocab * 4, cudaMemcpyDeviceToHost, m->eng->stream));
    CUDA_CHECK(cudaStreamSynchronize(m->eng->stream));
}

}  // namespace aios
