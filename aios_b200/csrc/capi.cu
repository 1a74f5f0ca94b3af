// capi.cu — the extern "C" boundary declared in include/aios_b200.h.  Every entry point catches
// all C++ exceptions and turns them into an AIOS_E_* code + thread-local message.
#include <string.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include <cuda_fp16.h>
#include <cuda_profiler_api.h>

#include "../../include/aios_b200.h"
#include "common.cuh"
#include "engine.h"
#include "json_mask.h"
#include "tokenizer.h"
#include "tp.h"

using namespace aios;

// One queued generation (aios_request_submit).  The engine's scheduler thread fills `out`; the owner polls it.
struct aios_request {
    aios_model *m = nullptr;
    std::vector<int32_t> prompt;
    aios_sampling sp{};
    int32_t max_tokens = 0, flags = 0;
    aios_mask_cb mask_cb = nullptr;   // optional: the allowed-token set of the next step (JSON mode)
    void *mask_user = nullptr;
    std::mutex mu;
    std::condition_variable cv;
    std::vector<int32_t> out;   // generated tokens so far
    size_t read = 0;            // how many of them the owner has taken
    bool done = false, cancel = false, orphan = false;
    int err = AIOS_OK;
    std::string msg;
    // scheduler-side state (touched by the scheduler thread only)
    int generated = 0;
    int32_t last_token = -1;
};

struct aios_engine {
    Engine impl;
    // scheduler thread: continuous batching over the requests of all models of this engine (one GPU stream)
    std::thread worker;
    std::mutex qmu;
    std::condition_variable qcv;
    std::deque<aios_request *> queue;
    bool stop = false;
    explicit aios_engine(const aios_engine_opts &o) : impl(o) {}
    ~aios_engine();
    void ensure_worker();
    void run();
};
// host-only: the tokenizer of one GGUF file; either owns its file (aios_vocab_open) or borrows the model's
struct aios_vocab {
    std::unique_ptr<GgufFile> own;
    Tokenizer tok;
    std::mutex json_mu;
    std::shared_ptr<JsonTables> json_tables;  // built on first use (aios_json_grammar_create)
    std::shared_ptr<JsonTables> json() {
        std::lock_guard<std::mutex> lk(json_mu);
        if (!json_tables) json_tables = std::make_shared<JsonTables>(tok);
        return json_tables;
    }
    explicit aios_vocab(const char *path) : own(new GgufFile(path)), tok(*own) {}
    explicit aios_vocab(const GgufFile &f) : tok(f) {}
};
struct aios_model {
    Model impl;
    std::unique_ptr<Seq> scratch;  // for aios_forward_logits
    // built from the tokenizer.ggml.* metadata on first use, under tok_mu (not the engine's GPU lock)
    std::mutex tok_mu;
    std::unique_ptr<aios_vocab> vocab_;
    aios_vocab *vocab() {
        std::lock_guard<std::mutex> lk(tok_mu);
        if (!vocab_) vocab_.reset(new aios_vocab(*impl.file));
        return vocab_.get();
    }
    aios_engine *owner = nullptr;
    // the scheduler's batch of this model: up to 64 requests decode together (created on first use)
    std::unique_ptr<Batch> batch;
    std::vector<aios_request *> lane_req;  // lane -> request
    std::atomic<int> inflight{0};        // queued or running requests
    aios_model(Engine *e, const char *n, const char *p, int c) : impl(e, n, p, c) {}
};
struct aios_seq { Seq impl; explicit aios_seq(Model *m) : impl(m) {} };

// ---------------------------------------------------------------- request scheduler (one thread per engine)
static void finish_request(aios_request *r, int err, const std::string &msg) {
    aios_model *m = r->m;
    bool orphan;
    {
        std::lock_guard<std::mutex> lk(r->mu);
        r->err = err;
        r->msg = msg;
        r->done = true;
        orphan = r->orphan;
        r->cv.notify_all();
    }
    // from here on the owner may free r at any moment (it sees done): do not touch it unless it is ours
    m->inflight.fetch_sub(1);
    if (orphan) delete r;  // the owner freed it while it was queued or running
}

static void error_of(const std::exception_ptr &ep, int *err, std::string *msg) {
    try {
        std::rethrow_exception(ep);
    } catch (const GgufError &e) {
        *err = e.code; *msg = e.what();
    } catch (const CudaError &e) {
        cudaGetLastError();
        *err = AIOS_E_CUDA; *msg = e.what();
    } catch (const std::exception &e) {
        *err = AIOS_E_STATE; *msg = e.what();
    } catch (...) {
        *err = AIOS_E_STATE; *msg = "unknown error";
    }
}

// Continuous batching: every pass of the loop (1) admits waiting requests into free lanes of their model's batch —
// the prompt is prefilled into the lane's sequence — and (2) runs ONE decode step for every model that has live lanes:
// each lane samples a token and feeds it, all lanes of a model through the same weight stream.  Finished, cancelled and
// failed requests leave their lane between steps; the next waiting request takes it in the following pass.
void aios_engine::run() {
    std::deque<aios_request *> waiting;   // submitted, not yet in a lane (first come first served per model)
    std::vector<aios_model *> live;       // models with lanes in use
    auto fail_all = [&](const std::string &why) {
        for (aios_request *r : waiting) finish_request(r, AIOS_E_STATE, why);
        waiting.clear();
        for (aios_model *m : live) {
            std::lock_guard<std::mutex> lk(impl.mu);
            while (m->batch && m->batch->n > 0) {
                aios_request *r = m->lane_req[(size_t)m->batch->n - 1];
                try { m->batch->retire(m->batch->n - 1); } catch (...) { m->batch->n--; }
                finish_request(r, AIOS_E_STATE, why);
            }
        }
        live.clear();
    };
    for (;;) {
        {
            std::unique_lock<std::mutex> lk(qmu);
            if (waiting.empty() && live.empty()) qcv.wait(lk, [&] { return stop || !queue.empty(); });
            while (!queue.empty()) { waiting.push_back(queue.front()); queue.pop_front(); }
            if (stop) {
                lk.unlock();
                fail_all("engine destroyed before the request finished");
                return;
            }
        }
        // ---- (1) admission
        for (auto it = waiting.begin(); it != waiting.end();) {
            aios_request *r = *it;
            aios_model *m = r->m;
            bool cancelled;
            {
                std::lock_guard<std::mutex> lk(r->mu);
                cancelled = r->cancel;
            }
            if (cancelled || r->max_tokens <= 0) {
                it = waiting.erase(it);
                finish_request(r, AIOS_OK, "");
                continue;
            }
            int err = AIOS_OK;
            std::string msg;
            bool admitted = false;
            try {
                std::lock_guard<std::mutex> lk(impl.mu);
                impl.make_current();
                if (!m->batch) {
                    m->batch.reset(new Batch(&m->impl));
                    m->lane_req.assign((size_t)m->batch->capacity(), nullptr);
                }
                Batch &b = *m->batch;
                // a lane is free and the KV pool can back prompt + generation (pages are only counted when capped)
                const size_t need = ((size_t)(r->prompt.size() + (size_t)r->max_tokens) * m->impl.kv_pos_bytes() + m->impl.kv_pool.page_bytes - 1) /
                                    m->impl.kv_pool.page_bytes;
                if (b.n < b.capacity() && m->impl.kv_pool.pages_available() >= need) {
                    const int lane = b.admit(r->prompt.data(), (int)r->prompt.size(), r->sp);
                    m->lane_req[(size_t)lane] = r;
                    r->generated = 0;
                    r->last_token = -1;
                    admitted = true;
                    if (std::find(live.begin(), live.end(), m) == live.end()) live.push_back(m);
                }
            } catch (...) {
                error_of(std::current_exception(), &err, &msg);
            }
            if (err != AIOS_OK) {
                it = waiting.erase(it);
                finish_request(r, err, msg);
            } else if (admitted) {
                it = waiting.erase(it);
            } else {
                ++it;
            }
        }
        // ---- (2) one step per model with live lanes
        for (size_t mi = 0; mi < live.size();) {
            aios_model *m = live[mi];
            Batch &b = *m->batch;
            int err = AIOS_OK;
            std::string msg;
            std::vector<uint8_t> mask_buf;
            try {
                std::lock_guard<std::mutex> lk(impl.mu);  // synchronous API calls of this engine interleave between steps
                impl.make_current();
                // cancelled lanes leave before the step; masked lanes get this step's allowed set
                for (int lane = b.n - 1; lane >= 0; lane--) {
                    aios_request *r = m->lane_req[(size_t)lane];
                    bool cancelled;
                    {
                        std::lock_guard<std::mutex> rl(r->mu);
                        cancelled = r->cancel;
                    }
                    int cb_rc = 1;
                    if (!cancelled && r->mask_cb) {
                        mask_buf.assign(b.mask_bytes, 0);
                        cb_rc = r->mask_cb(r->mask_user, r->last_token, mask_buf.data(), (int32_t)((m->impl.hp.n_vocab + 7) / 8));
                    }
                    if (cancelled || cb_rc == 2 || cb_rc < 0) {
                        b.retire(lane);
                        m->lane_req[(size_t)lane] = m->lane_req[(size_t)b.n];  // the former last lane moved here
                        m->lane_req[(size_t)b.n] = nullptr;
                        finish_request(r, cb_rc < 0 ? AIOS_E_STATE : AIOS_OK, cb_rc < 0 ? "the mask callback failed" : "");
                        continue;
                    }
                    if (r->mask_cb) b.set_mask(lane, cb_rc == 0 ? mask_buf.data() : nullptr);
                }
                if (b.n > 0) b.step();
                const Tokenizer *tok = nullptr;
                for (int lane = b.n - 1; lane >= 0; lane--) {
                    aios_request *r = m->lane_req[(size_t)lane];
                    const int32_t t = b.h_tok[lane];
                    r->generated++;
                    r->last_token = t;
                    bool finished = r->generated >= r->max_tokens;
                    bool keep = true;
                    if (r->flags & AIOS_REQ_STOP_AT_EOG) {
                        if (!tok) tok = &m->vocab()->tok;
                        if (tok->is_eog(t)) { keep = false; finished = true; }
                    }
                    if (keep) {
                        std::lock_guard<std::mutex> rl(r->mu);
                        r->out.push_back(t);
                    }
                    r->cv.notify_all();
                    if (finished) {
                        b.retire(lane);
                        m->lane_req[(size_t)lane] = m->lane_req[(size_t)b.n];
                        m->lane_req[(size_t)b.n] = nullptr;
                        finish_request(r, AIOS_OK, "");
                    }
                }
            } catch (...) {
                error_of(std::current_exception(), &err, &msg);
            }
            if (err != AIOS_OK) {  // a failed step fails every request of the batch
                std::lock_guard<std::mutex> lk(impl.mu);
                while (b.n > 0) {
                    aios_request *r = m->lane_req[(size_t)b.n - 1];
                    m->lane_req[(size_t)b.n - 1] = nullptr;
                    try { b.retire(b.n - 1); } catch (...) { b.n--; }
                    finish_request(r, err, msg);
                }
            }
            if (b.n == 0) live.erase(live.begin() + (long)mi);
            else mi++;
        }
    }
}

void aios_engine::ensure_worker() {
    if (!worker.joinable()) worker = std::thread([this] { run(); });
}

aios_engine::~aios_engine() {
    {
        std::lock_guard<std::mutex> lk(qmu);
        stop = true;
    }
    qcv.notify_all();
    if (worker.joinable()) worker.join();
}

static thread_local std::string g_err;

static int fail(int code, const std::string &msg) {
    g_err = msg;
    return code;
}

template <typename F>
static int guarded(F &&f) {
    try {
        g_err.clear();
        return f();
    } catch (const GgufError &e) {
        return fail(e.code, e.what());
    } catch (const CudaError &e) {
        cudaGetLastError();  // clear the sticky flag of non-fatal errors
        return fail(AIOS_E_CUDA, e.what());
    } catch (const std::invalid_argument &e) {
        return fail(AIOS_E_INVAL, e.what());
    } catch (const std::length_error &e) {
        return fail(AIOS_E_RANGE, e.what());
    } catch (const std::bad_alloc &) {
        return fail(AIOS_E_NOMEM, "out of host memory");
    } catch (const std::exception &e) {
        return fail(AIOS_E_STATE, e.what());
    } catch (...) {
        return fail(AIOS_E_STATE, "unknown error");
    }
}

// ---------------------------------------------------------------- tensor parallelism (tp.cu)
struct aios_tp_model {
    std::unique_ptr<aios::TpModel> impl;
    std::mutex mu;
};


extern "C" {

const char *aios_last_error(void) { return g_err.c_str(); }
const char *aios_version(void) { return "aios_b200 0.1 sm_100a"; }

int aios_device_count(int32_t *n) {
    return guarded([&] {
        if (!n) return fail(AIOS_E_INVAL, "n is null");
        int c = 0;
        cudaError_t e = cudaGetDeviceCount(&c);
        if (e != cudaSuccess) { cudaGetLastError(); c = 0; }
        *n = c;
        return AIOS_OK;
    });
}

int aios_engine_create(const aios_engine_opts *opts, aios_engine **out) {
    return guarded([&] {
        if (!out) return fail(AIOS_E_INVAL, "out is null");
        aios_engine_opts o{};
        if (opts) o = *opts;
        *out = new aios_engine(o);
        return AIOS_OK;
    });
}

void aios_engine_destroy(aios_engine *e) { delete e; }

int aios_model_load(aios_engine *e, const char *name, const char *gguf_path, int32_t n_ctx, aios_model **out) {
    return guarded([&] {
        if (!e || !gguf_path || !out) return fail(AIOS_E_INVAL, "null argument");
        std::lock_guard<std::mutex> lk(e->impl.mu);
        *out = new aios_model(&e->impl, name ? name : "", gguf_path, n_ctx);
        (*out)->owner = e;
        return AIOS_OK;
    });
}

int aios_model_unload(aios_model *m) {
    return guarded([&] {
        if (!m) return fail(AIOS_E_INVAL, "null model");
        if (m->inflight.load() > 0) return fail(AIOS_E_STATE, "model still has queued or running requests");
        {
            std::lock_guard<std::mutex> lk(m->impl.eng->mu);
            m->impl.eng->make_current();
            m->scratch.reset();
            m->batch.reset();
        }
        if (m->impl.live_seqs > 0) return fail(AIOS_E_STATE, "model still has live sequences");
        delete m;
        return AIOS_OK;
    });
}

int aios_model_info_get(const aios_model *m, aios_model_info *o) {
    return guarded([&] {
        if (!m || !o) return fail(AIOS_E_INVAL, "null argument");
        memset(o, 0, sizeof(*o));
        const Model &M = m->impl;
        strncpy(o->name, M.name.c_str(), sizeof(o->name) - 1);
        strncpy(o->arch, M.arch.c_str(), sizeof(o->arch) - 1);
        o->n_layer = M.hp.n_layer; o->n_embd = M.hp.n_embd; o->n_head = M.hp.n_head; o->n_head_kv = M.hp.n_head_kv;
        o->head_dim = M.hp.head_dim; o->n_ff = M.hp.n_ff; o->n_vocab = M.hp.n_vocab; o->n_ctx = M.hp.n_ctx;
        o->n_rot = M.hp.n_rot; o->file_type = M.hp.file_type; o->n_tensors = (int32_t)M.file->tensors().size();
        o->max_batch = M.eng->max_batch; o->rms_eps = M.hp.rms_eps; o->rope_freq_base = M.hp.rope_freq_base;
        o->file_bytes = (int64_t)M.file->file_bytes(); o->weight_bytes = M.weight_bytes;
        o->kv_bytes = (int64_t)(const_cast<Model &>(M).kv_pool.pages_in_use() * M.kv_pool.page_bytes);
        o->weight_bytes_per_token = M.weight_bytes_per_token;
        o->kv_bytes_per_position = (int64_t)M.hp.n_layer * 2 * M.hp.n_head_kv * M.hp.head_dim * 2;
        return AIOS_OK;
    });
}

int aios_model_tensor_info(const aios_model *m, int32_t i, aios_tensor_info *o) {
    return guarded([&] {
        if (!m || !o) return fail(AIOS_E_INVAL, "null argument");
        const auto &ts = m->impl.file->tensors();
        if (i < 0 || (size_t)i >= ts.size()) return fail(AIOS_E_INVAL, "tensor index out of range");
        memset(o, 0, sizeof(*o));
        strncpy(o->name, ts[i].name.c_str(), sizeof(o->name) - 1);
        o->type = ts[i].type; o->n_dims = ts[i].n_dims;
        for (int d = 0; d < 4; d++) o->ne[d] = ts[i].ne[d];
        o->offset = (int64_t)ts[i].offset; o->nbytes = (int64_t)ts[i].nbytes;
        return AIOS_OK;
    });
}

// copy a std::string into a caller buffer: *len always gets the full length; AIOS_E_RANGE if it does not fit
static int put_bytes(const std::string &s, char *buf, int32_t cap, int32_t *len) {
    *len = (int32_t)s.size();
    if (buf && cap > 0) {
        if ((size_t)cap < s.size()) return fail(AIOS_E_RANGE, "output buffer too small");
        memcpy(buf, s.data(), s.size());
    } else if (buf == nullptr) {
        return (int)AIOS_OK;  // length query
    } else if (!s.empty()) {
        return fail(AIOS_E_RANGE, "output buffer too small");
    }
    return (int)AIOS_OK;
}

int aios_vocab_open(const char *gguf_path, aios_vocab **out) {
    return guarded([&] {
        if (!gguf_path || !out) return fail(AIOS_E_INVAL, "null argument");
        *out = new aios_vocab(gguf_path);
        return (int)AIOS_OK;
    });
}

void aios_vocab_close(aios_vocab *v) { delete v; }

int aios_model_vocab(aios_model *m, aios_vocab **out) {
    return guarded([&] {
        if (!m || !out) return fail(AIOS_E_INVAL, "null argument");
        *out = m->vocab();
        return (int)AIOS_OK;
    });
}

int aios_vocab_info_get(aios_vocab *v, aios_vocab_info *o) {
    return guarded([&] {
        if (!v || !o) return fail(AIOS_E_INVAL, "null argument");
        const Tokenizer &t = v->tok;
        memset(o, 0, sizeof(*o));
        o->n_tokens = t.n_tokens();
        o->bos = t.bos();
        o->eos = t.eos();
        o->add_bos = t.add_bos() ? 1 : 0;
        o->chat_format = (int32_t)t.chat_format();
        strncpy(o->model, t.model().c_str(), sizeof(o->model) - 1);
        return (int)AIOS_OK;
    });
}

int aios_vocab_token_piece(aios_vocab *v, int32_t id, char *buf, int32_t cap, int32_t *len, int32_t *type, float *score) {
    return guarded([&] {
        if (!v || !len) return fail(AIOS_E_INVAL, "null argument");
        const Tokenizer &t = v->tok;
        if (id < 0 || id >= t.n_tokens()) return fail(AIOS_E_INVAL, "token id out of range");
        if (type) *type = t.type(id);
        if (score) *score = t.score(id);
        return put_bytes(t.piece(id), buf, cap, len);
    });
}

int aios_model_meta_str(aios_model *m, const char *key, char *buf, int32_t cap, int32_t *len) {
    return guarded([&] {
        if (!m || !key || !len) return fail(AIOS_E_INVAL, "null argument");
        const GgufValue *v = m->impl.file->get(key);
        if (!v || v->kind != GgufValue::STRING) return fail(AIOS_E_INVAL, std::string("no string metadata key ") + key);
        return put_bytes(v->s, buf, cap, len);
    });
}

static int need_vocab(const Tokenizer &t) {
    if (t.n_tokens() == 0) return fail(AIOS_E_STATE, "the GGUF file carries no tokenizer.ggml.tokens");
    if (t.model() != "llama" && t.model() != "gpt2")
        return fail(AIOS_E_UNSUPPORTED, "tokenizer.ggml.model '" + t.model() + "' is not supported ('llama' = SentencePiece and 'gpt2' = byte-level BPE are)");
    return (int)AIOS_OK;
}

int aios_tokenize(aios_vocab *v, const char *text, int32_t text_len, int32_t flags, int32_t *out, int32_t cap, int32_t *n_out) {
    return guarded([&] {
        if (!v || (!text && text_len != 0) || !n_out) return fail(AIOS_E_INVAL, "null argument");
        const Tokenizer &t = v->tok;
        if (int rc = need_vocab(t)) return rc;
        std::string s = text_len < 0 ? std::string(text) : std::string(text, (size_t)text_len);
        std::vector<int32_t> ids = t.encode(s, (flags & AIOS_TOK_ADD_SPECIAL) != 0, (flags & AIOS_TOK_PARSE_SPECIAL) != 0);
        *n_out = (int32_t)ids.size();
        if (!out) return (int)AIOS_OK;  // count query
        if ((size_t)cap < ids.size()) return fail(AIOS_E_RANGE, "token buffer too small");
        if (!ids.empty()) memcpy(out, ids.data(), ids.size() * sizeof(int32_t));
        return (int)AIOS_OK;
    });
}

int aios_detokenize(aios_vocab *v, const int32_t *tokens, int32_t n, int32_t flags, char *buf, int32_t cap, int32_t *len) {
    return guarded([&] {
        if (!v || (!tokens && n > 0) || !len || n < 0) return fail(AIOS_E_INVAL, "bad argument");
        const Tokenizer &t = v->tok;
        if (t.n_tokens() == 0) return fail(AIOS_E_STATE, "the GGUF file carries no tokenizer.ggml.tokens");
        for (int32_t i = 0; i < n; i++)
            if (tokens[i] < 0 || tokens[i] >= t.n_tokens()) return fail(AIOS_E_INVAL, "token id out of range");
        return put_bytes(t.decode(tokens, n, (flags & AIOS_TOK_RENDER_SPECIAL) != 0, (flags & AIOS_TOK_STRIP_SPACE) != 0),
                         buf, cap, len);
    });
}

int aios_chat_prompt(aios_vocab *v, const char *system_prompt, const char *user_prompt, char *buf, int32_t cap, int32_t *len) {
    return guarded([&] {
        if (!v || !user_prompt || !len) return fail(AIOS_E_INVAL, "null argument");
        return put_bytes(v->tok.chat_prompt(system_prompt ? system_prompt : "", user_prompt), buf, cap, len);
    });
}

int aios_chat_apply_template(aios_vocab *v, const char *const *roles, const char *const *contents, int32_t n_messages,
                             int32_t add_generation_prompt, char *buf, int32_t cap, int32_t *len) {
    return guarded([&] {
        if (!v || !len || n_messages < 0 || (n_messages > 0 && (!roles || !contents))) return fail(AIOS_E_INVAL, "null argument");
        if (v->tok.chat_template().empty()) return fail(AIOS_E_STATE, "the GGUF file carries no tokenizer.chat_template");
        std::vector<std::pair<std::string, std::string>> msgs;
        for (int32_t i = 0; i < n_messages; i++) {
            if (!roles[i] || !contents[i]) return fail(AIOS_E_INVAL, "null message");
            msgs.emplace_back(roles[i], contents[i]);
        }
        std::string out, err;
        if (!v->tok.apply_template(msgs, add_generation_prompt != 0, &out, &err)) return fail(AIOS_E_UNSUPPORTED, "chat template: " + err);
        return put_bytes(out, buf, cap, len);
    });
}

int aios_token_is_eog(aios_vocab *v, int32_t id, int32_t *is_eog) {
    return guarded([&] {
        if (!v || !is_eog) return fail(AIOS_E_INVAL, "null argument");
        *is_eog = v->tok.is_eog(id) ? 1 : 0;
        return (int)AIOS_OK;
    });
}

struct aios_json_grammar { JsonMasker impl; explicit aios_json_grammar(std::shared_ptr<JsonTables> t) : impl(std::move(t)) {} };

int aios_json_grammar_create(aios_vocab *v, aios_json_grammar **out) {
    return guarded([&] {
        if (!v || !out) return fail(AIOS_E_INVAL, "null argument");
        if (v->tok.n_tokens() == 0) return fail(AIOS_E_STATE, "the GGUF file carries no tokenizer.ggml.tokens");
        *out = new aios_json_grammar(v->json());
        return (int)AIOS_OK;
    });
}
void aios_json_grammar_free(aios_json_grammar *g) { delete g; }
int aios_json_grammar_complete(aios_json_grammar *g, int32_t *complete) {
    if (!g || !complete) return fail(AIOS_E_INVAL, "null argument");
    *complete = g->impl.state.complete() ? 1 : 0;
    return AIOS_OK;
}
// aios_mask_cb: user = the aios_json_grammar of this request
int aios_json_grammar_mask_cb(void *user, int32_t last_token, uint8_t *mask_out, int32_t n_bytes) {
    if (!user || !mask_out) return -1;
    try {
        return static_cast<aios_json_grammar *>(user)->impl.step(last_token, mask_out, n_bytes);
    } catch (...) {
        return -1;
    }
}

int aios_seq_create(aios_model *m, aios_seq **out) {
    return guarded([&] {
        if (!m || !out) return fail(AIOS_E_INVAL, "null argument");
        std::lock_guard<std::mutex> lk(m->impl.eng->mu);
        *out = new aios_seq(&m->impl);
        return AIOS_OK;
    });
}
void aios_seq_free(aios_seq *s) {
    if (!s) return;
    std::lock_guard<std::mutex> lk(s->impl.m->eng->mu);
    delete s;
}
int aios_seq_reset(aios_seq *s) {
    return guarded([&] {
        if (!s) return fail(AIOS_E_INVAL, "null seq");
        std::lock_guard<std::mutex> lk(s->impl.m->eng->mu);
        s->impl.reset();
        return AIOS_OK;
    });
}
int aios_seq_position(const aios_seq *s, int32_t *n_past) {
    if (!s || !n_past) return fail(AIOS_E_INVAL, "null argument");
    *n_past = s->impl.n_past;
    return AIOS_OK;
}
int aios_seq_prefill(aios_seq *s, const int32_t *tokens, int32_t n) {
    return guarded([&] {
        if (!s || (!tokens && n > 0) || n < 0) return fail(AIOS_E_INVAL, "bad argument");
        if (s->impl.n_past + n > s->impl.m->hp.n_ctx) return fail(AIOS_E_STATE, "prompt does not fit in the context");
        std::lock_guard<std::mutex> lk(s->impl.m->eng->mu);
        s->impl.prefill(tokens, n);
        return AIOS_OK;
    });
}
int aios_seq_decode(aios_seq *s, const aios_sampling *sp, int32_t max_tokens, int32_t *out_tokens, int32_t *n_out) {
    return guarded([&] {
        if (!s || !out_tokens || !n_out || max_tokens < 0) return fail(AIOS_E_INVAL, "bad argument");
        aios_sampling d{};
        if (sp) d = *sp;
        std::lock_guard<std::mutex> lk(s->impl.m->eng->mu);
        *n_out = s->impl.decode(d, max_tokens, out_tokens);
        return AIOS_OK;
    });
}
int aios_seq_step(aios_seq *s, const aios_sampling *sp, int32_t *token_out) {
    return guarded([&] {
        if (!s || !token_out) return fail(AIOS_E_INVAL, "bad argument");
        aios_sampling d{};
        if (sp) d = *sp;
        std::lock_guard<std::mutex> lk(s->impl.m->eng->mu);
        s->impl.step(d, token_out);
        return AIOS_OK;
    });
}
// Greedy decode that never touches the device allocator: the fused kernel's exchange scratch is the static
// __device__ array in pfd.cu (resolved with cudaGetSymbolAddress on the first call per device).  See the header.
int aios_seq_decode_static(aios_seq *s, int32_t max_tokens, int32_t *out_tokens, int32_t *n_out) {
    return guarded([&] {
        if (!s || !out_tokens || !n_out || max_tokens < 0) return fail(AIOS_E_INVAL, "bad argument");
        std::lock_guard<std::mutex> lk(s->impl.m->eng->mu);
        *n_out = s->impl.decode_static(max_tokens, out_tokens);
        return (int)AIOS_OK;
    });
}
int aios_seq_set_mask(aios_seq *s, const uint8_t *mask_bits, int32_t n_bytes) {
    return guarded([&] {
        if (!s || (mask_bits && n_bytes <= 0)) return fail(AIOS_E_INVAL, "bad argument");
        std::lock_guard<std::mutex> lk(s->impl.m->eng->mu);
        s->impl.set_mask(mask_bits, n_bytes);
        return (int)AIOS_OK;
    });
}
int aios_seq_logits(aios_seq *s, float *logits_out) {
    return guarded([&] {
        if (!s || !logits_out) return fail(AIOS_E_INVAL, "bad argument");
        std::lock_guard<std::mutex> lk(s->impl.m->eng->mu);
        s->impl.read_logits(logits_out);
        return AIOS_OK;
    });
}

int aios_generate(aios_model *m, const int32_t *prompt, int32_t n_prompt, const aios_sampling *sp, int32_t max_tokens,
                  int32_t *out_tokens, int32_t *n_out) {
    return guarded([&] {
        if (!m || !prompt || n_prompt <= 0 || !out_tokens || !n_out) return fail(AIOS_E_INVAL, "bad argument");
        if (n_prompt > m->impl.hp.n_ctx) return fail(AIOS_E_STATE, "prompt does not fit in the context");
        aios_sampling d{};
        if (sp) d = *sp;
        std::lock_guard<std::mutex> lk(m->impl.eng->mu);
        Seq s(&m->impl);
        s.prefill(prompt, n_prompt);
        *n_out = s.decode(d, max_tokens, out_tokens);
        return AIOS_OK;
    });
}

int aios_request_submit_masked(aios_model *m, const int32_t *tokens, int32_t n, const aios_sampling *sp, int32_t max_tokens,
                               int32_t flags, aios_mask_cb cb, void *user, aios_request **out);

int aios_request_submit(aios_model *m, const int32_t *tokens, int32_t n, const aios_sampling *sp, int32_t max_tokens,
                        int32_t flags, aios_request **out) {
    return aios_request_submit_masked(m, tokens, n, sp, max_tokens, flags, nullptr, nullptr, out);
}

int aios_request_submit_masked(aios_model *m, const int32_t *tokens, int32_t n, const aios_sampling *sp, int32_t max_tokens,
                               int32_t flags, aios_mask_cb cb, void *user, aios_request **out) {
    return guarded([&] {
        if (!m || !tokens || n <= 0 || !out || max_tokens < 0) return fail(AIOS_E_INVAL, "bad argument");
        if (!m->owner) return fail(AIOS_E_STATE, "model has no engine");
        if (n >= m->impl.hp.n_ctx) return fail(AIOS_E_STATE, "prompt does not fit in the context");
        for (int32_t i = 0; i < n; i++)
            if (tokens[i] < 0 || tokens[i] >= m->impl.hp.n_vocab) return fail(AIOS_E_INVAL, "token id out of range");
        aios_request *r = new aios_request();
        r->m = m;
        r->prompt.assign(tokens, tokens + n);  // copied: the caller's buffer is free again when this returns
        if (sp) r->sp = *sp;
        r->max_tokens = std::min(max_tokens, m->impl.hp.n_ctx - n);
        r->flags = flags;
        r->mask_cb = cb;
        r->mask_user = user;
        m->inflight.fetch_add(1);
        aios_engine *e = m->owner;
        {
            std::lock_guard<std::mutex> lk(e->qmu);
            e->ensure_worker();
            e->queue.push_back(r);
        }
        e->qcv.notify_one();
        *out = r;
        return (int)AIOS_OK;
    });
}

int aios_request_poll(aios_request *r, int32_t *out, int32_t cap, int32_t *n_out, int32_t *done) {
    return guarded([&] {
        if (!r || !n_out || !done || cap < 0 || (cap > 0 && !out)) return fail(AIOS_E_INVAL, "bad argument");
        std::lock_guard<std::mutex> lk(r->mu);
        const size_t avail = r->out.size() - r->read;
        const size_t take = std::min(avail, (size_t)cap);
        if (take) memcpy(out, r->out.data() + r->read, take * sizeof(int32_t));
        r->read += take;
        *n_out = (int32_t)take;
        *done = (r->done && r->read == r->out.size()) ? 1 : 0;
        if (*done && r->err != AIOS_OK) return fail(r->err, r->msg);  // a failed request reports its error last
        return (int)AIOS_OK;
    });
}

int aios_request_wait(aios_request *r, int32_t timeout_ms, int32_t *ready) {
    return guarded([&] {
        if (!r || !ready) return fail(AIOS_E_INVAL, "bad argument");
        std::unique_lock<std::mutex> lk(r->mu);
        auto pred = [&] { return r->done || r->read < r->out.size(); };
        if (timeout_ms < 0) r->cv.wait(lk, pred);
        else r->cv.wait_for(lk, std::chrono::milliseconds(timeout_ms), pred);
        *ready = pred() ? 1 : 0;
        return (int)AIOS_OK;
    });
}

int aios_request_cancel(aios_request *r) {
    return guarded([&] {
        if (!r) return fail(AIOS_E_INVAL, "null request");
        std::lock_guard<std::mutex> lk(r->mu);
        r->cancel = true;
        return (int)AIOS_OK;
    });
}

void aios_request_free(aios_request *r) {
    if (!r) return;
    bool del;
    {
        std::lock_guard<std::mutex> lk(r->mu);
        r->cancel = true;
        r->orphan = true;
        del = r->done;  // finished: nobody else holds it.  Otherwise the scheduler deletes it when it is done with it.
    }
    if (del) delete r;
}

int aios_dequantize(aios_engine *e, int32_t type, const void *blocks, int64_t n, float *out) {
    return guarded([&] {
        if (!e || !blocks || !out || n < 0) return fail(AIOS_E_INVAL, "bad argument");
        const int be = type_block_elems(type), bb = type_block_bytes(type);
        if (!bb) return fail(AIOS_E_INVAL, "unsupported ggml type");
        if (n % be) return fail(AIOS_E_INVAL, "n is not a multiple of the block size");
        if (n == 0) return (int)AIOS_OK;
        std::lock_guard<std::mutex> lk(e->impl.mu);
        e->impl.make_current();
        const size_t nbytes = (size_t)(n / be) * bb;
        DevBuf in(nbytes), o((size_t)n * 4);
        cudaStream_t st = e->impl.stream;
        CUDA_CHECK(cudaMemcpyAsync(in.p, blocks, nbytes, cudaMemcpyHostToDevice, st));
        launch_dequantize(type, in.p, n, o.as<float>(), st);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaMemcpyAsync(out, o.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        return (int)AIOS_OK;
    });
}

static void make_dev_matrix(Engine &E, int type, const void *host_blocks, int64_t n_rows, int64_t k, DevMatrix &dm) {
    size_t total = 0, off[5];
    for (int i = 0; i < 5; i++) { off[i] = total; total += (plane_bytes(type, i, n_rows, k) + 255) / 256 * 256; }
    dm.storage = DevBuf(total);
    uint8_t *planes[5];
    for (int i = 0; i < 5; i++) {
        planes[i] = plane_bytes(type, i, n_rows, k) ? dm.storage.as<uint8_t>() + off[i] : nullptr;
        dm.w.p[i] = planes[i];
    }
    dm.w.type = type; dm.w.n_rows = (int32_t)n_rows; dm.w.k = (int32_t)k;
    const size_t nbytes = tensor_bytes(type, n_rows, k);
    DevBuf raw(nbytes);
    CUDA_CHECK(cudaMemcpyAsync(raw.p, host_blocks, nbytes, cudaMemcpyHostToDevice, E.stream));
    launch_repack(type, raw.p, n_rows, (int)k, planes, E.stream);
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaStreamSynchronize(E.stream));
}

int aios_matvec(aios_engine *e, int32_t type, const void *w_blocks, int64_t n_rows, int64_t k, const float *x,
                const float *norm_w, float rms_eps, float *y) {
    return guarded([&] {
        if (!e || !w_blocks || !x || !y || n_rows <= 0 || k <= 0) return fail(AIOS_E_INVAL, "bad argument");
        if (!(type == T_Q4_K || type == T_Q5_K || type == T_Q6_K || type == T_Q8_0)) return fail(AIOS_E_INVAL, "unsupported weight type");
        if (k % 256) return fail(AIOS_E_INVAL, "k must be a multiple of 256");
        if (gemv_smem_bytes((int)k) > 100 * 1024) return fail(AIOS_E_RANGE, "k too large");
        std::lock_guard<std::mutex> lk(e->impl.mu);
        Engine &E = e->impl;
        E.make_current();
        DevMatrix dm;
        make_dev_matrix(E, type, w_blocks, n_rows, k, dm);
        DevBuf dx((size_t)k * 4), dy((size_t)n_rows * 4), dn;
        CUDA_CHECK(cudaMemcpyAsync(dx.p, x, (size_t)k * 4, cudaMemcpyHostToDevice, E.stream));
        if (norm_w) {
            dn = DevBuf((size_t)k * 4);
            CUDA_CHECK(cudaMemcpyAsync(dn.p, norm_w, (size_t)k * 4, cudaMemcpyHostToDevice, E.stream));
        }
        GemvParams p{};
        p.n_seg = 1; p.epi = EPI_STORE; p.k = (int32_t)k; p.seg[0] = dm.w; p.out[0] = dy.as<float>();
        p.x = dx.as<float>(); p.norm_w = norm_w ? dn.as<float>() : nullptr; p.eps = rms_eps;
        launch_gemv(p, gemv_plan_grid(p, E.n_sms), E.stream);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaMemcpyAsync(y, dy.p, (size_t)n_rows * 4, cudaMemcpyDeviceToHost, E.stream));
        CUDA_CHECK(cudaStreamSynchronize(E.stream));
        return (int)AIOS_OK;
    });
}

int aios_matmul(aios_engine *e, int32_t type, const void *w_blocks, int64_t n_rows, int64_t k, const float *x, int32_t n_tok,
                const float *norm_w, float rms_eps, float *y) {
    return guarded([&] {
        if (!e || !w_blocks || !x || !y || n_rows <= 0 || k <= 0 || n_tok <= 0) return fail(AIOS_E_INVAL, "bad argument");
        if (!gemm_q_usable(type, (int)k)) return fail(AIOS_E_UNSUPPORTED, "aios_matmul: Q4_K / Q5_K / Q6_K / Q8_0 weights, k a multiple of 256");
        if (n_tok > GEMM_MAX_TOKENS) return fail(AIOS_E_RANGE, "aios_matmul: at most 64 tokens per call");
        std::lock_guard<std::mutex> lk(e->impl.mu);
        Engine &E = e->impl;
        E.make_current();
        DevMatrix dm;
        make_dev_matrix(E, type, w_blocks, n_rows, k, dm);
        const int T8 = gemm_pad_tokens(n_tok);
        DevBuf dx((size_t)T8 * k * 4), dy((size_t)n_tok * n_rows * 4), dn, img(gemm_images_bytes((int)k, T8));
        CUDA_CHECK(cudaMemsetAsync(dx.p, 0, dx.bytes, E.stream));  // padding tokens: zeros
        CUDA_CHECK(cudaMemcpyAsync(dx.p, x, (size_t)n_tok * k * 4, cudaMemcpyHostToDevice, E.stream));
        if (norm_w) {
            dn = DevBuf((size_t)k * 4);
            CUDA_CHECK(cudaMemcpyAsync(dn.p, norm_w, (size_t)k * 4, cudaMemcpyHostToDevice, E.stream));
        }
        launch_quant_rows(type, dx.as<float>(), k, norm_w ? dn.as<float>() : nullptr, rms_eps, (int)k, T8, img.as<uint8_t>(), E.stream);
        launch_gemm_q(dm.w, img.as<uint8_t>(), n_tok, dy.as<float>(), n_rows, EPI_STORE, E.stream);
        CUDA_CHECK(cudaMemcpyAsync(y, dy.p, dy.bytes, cudaMemcpyDeviceToHost, E.stream));
        CUDA_CHECK(cudaStreamSynchronize(E.stream));
        return (int)AIOS_OK;
    });
}

int aios_sample_logits(aios_engine *e, const float *logits, int32_t n_vocab, const uint8_t *mask_bits, const aios_sampling *sp,
                       uint64_t counter, int32_t *token, int32_t *cand, int32_t *counts) {
    return guarded([&] {
        if (!e || !logits || n_vocab <= 0 || !token) return fail(AIOS_E_INVAL, "bad argument");
        if (counter > 0x7fffffffull) return fail(AIOS_E_RANGE, "counter beyond 2^31-1");
        aios_sampling d{};
        if (sp) d = *sp;
        std::lock_guard<std::mutex> lk(e->impl.mu);
        Engine &E = e->impl;
        E.make_current();
        const size_t mb = ((size_t)n_vocab + 7) / 8;
        DevBuf dl((size_t)n_vocab * 4), dm(mask_bits ? mb : 0), st(64), outs((128 + 3 + 4) * 4);
        int32_t *o = outs.as<int32_t>();  // [0] token slot (out_tokens), [1..129) cand, [129..132) counts
        int32_t hdr[4] = {(int32_t)counter, 0, 0, 0};  // pos (the draw's counter), step, cur_token
        CUDA_CHECK(cudaMemcpyAsync(st.p, hdr, sizeof(hdr), cudaMemcpyHostToDevice, E.stream));
        CUDA_CHECK(cudaMemcpyAsync(dl.p, logits, (size_t)n_vocab * 4, cudaMemcpyHostToDevice, E.stream));
        if (mask_bits) CUDA_CHECK(cudaMemcpyAsync(dm.p, mask_bits, mb, cudaMemcpyHostToDevice, E.stream));
        CUDA_CHECK(cudaMemsetAsync(outs.p, 0xff, outs.bytes, E.stream));
        SampleParams s{};
        s.logits = dl.as<float>(); s.n_vocab = n_vocab;
        s.temperature = d.temperature; s.top_k = d.top_k; s.top_p = d.top_p; s.min_p = d.min_p; s.seed = d.seed;
        s.pos_ptr = st.as<int32_t>(); s.step_ptr = st.as<int32_t>() + 1; s.cur_token = st.as<int32_t>() + 2;
        s.out_tokens = o; s.out_cap = 1;
        s.mask = mask_bits ? dm.as<uint8_t>() : nullptr;
        s.cand_out = o + 1; s.counts_out = o + 129;
        launch_sample(s, E.stream);
        CUDA_CHECK(cudaGetLastError());
        int32_t host[132];
        CUDA_CHECK(cudaMemcpyAsync(host, outs.p, sizeof(host), cudaMemcpyDeviceToHost, E.stream));
        CUDA_CHECK(cudaStreamSynchronize(E.stream));
        *token = host[0];
        if (cand) memcpy(cand, host + 1, 128 * 4);
        if (counts) memcpy(counts, host + 129, 3 * 4);
        return (int)AIOS_OK;
    });
}

int aios_matmul_tc(aios_engine *e, int32_t type, const void *w_blocks, int64_t n_rows, int64_t k, const float *x, int32_t n_tok,
                   const float *norm_w, float rms_eps, float *y) {
    return guarded([&] {
        if (!e || !w_blocks || !x || !y || n_rows <= 0 || k <= 0 || n_tok <= 0) return fail(AIOS_E_INVAL, "bad argument");
        if (!prefill_tc_usable(type, (int)k)) return fail(AIOS_E_UNSUPPORTED, "aios_matmul_tc: Q4_K / Q5_K / Q6_K / Q8_0 weights, k a multiple of 256");
        std::lock_guard<std::mutex> lk(e->impl.mu);
        Engine &E = e->impl;
        E.make_current();
        DevMatrix dm;
        make_dev_matrix(E, type, w_blocks, n_rows, k, dm);
        DevBuf dx((size_t)n_tok * k * 4), dh((size_t)n_tok * k * 2), dy((size_t)n_tok * n_rows * 4), dn;
        CUDA_CHECK(cudaMemcpyAsync(dx.p, x, (size_t)n_tok * k * 4, cudaMemcpyHostToDevice, E.stream));
        if (norm_w) {
            dn = DevBuf((size_t)k * 4);
            CUDA_CHECK(cudaMemcpyAsync(dn.p, norm_w, (size_t)k * 4, cudaMemcpyHostToDevice, E.stream));
        }
        launch_rows_to_f16(dx.as<float>(), k, norm_w ? dn.as<float>() : nullptr, rms_eps, (int)k, n_tok, dh.as<__half>(), E.stream);
        launch_prefill_tc(dm.w, dh.as<__half>(), n_tok, dy.as<float>(), n_rows, EPI_STORE, E.n_sms, E.stream);
        CUDA_CHECK(cudaMemcpyAsync(y, dy.p, dy.bytes, cudaMemcpyDeviceToHost, E.stream));
        CUDA_CHECK(cudaStreamSynchronize(E.stream));
        CUDA_CHECK(cudaGetLastError());
        return (int)AIOS_OK;
    });
}

int aios_forward_logits(aios_model *m, const int32_t *tokens, int32_t n, int32_t n_past, float *logits_out) {
    return guarded([&] {
        if (!m || !tokens || n <= 0 || !logits_out || n_past < 0) return fail(AIOS_E_INVAL, "bad argument");
        std::lock_guard<std::mutex> lk(m->impl.eng->mu);
        if (!m->scratch) m->scratch.reset(new Seq(&m->impl, true));  // a hidden slot, not one of the caller's
        Seq &s = *m->scratch;
        if (n_past == 0) s.reset();
        else if (s.n_past != n_past) return fail(AIOS_E_STATE, "scratch sequence is at a different position");
        if (s.n_past + n > m->impl.hp.n_ctx) return fail(AIOS_E_STATE, "tokens do not fit in the context");
        s.prefill(tokens, n);
        s.read_logits(logits_out);
        return (int)AIOS_OK;
    });
}

int aios_seq_read_kv(aios_seq *s, int32_t layer, int32_t pos0, int32_t n, uint16_t *k_out, uint16_t *v_out) {
    return guarded([&] {
        if (!s || !k_out || !v_out) return fail(AIOS_E_INVAL, "bad argument");
        Model &M = *s->impl.m;
        if (layer < 0 || layer >= M.hp.n_layer || pos0 < 0 || n < 0 || pos0 + n > M.hp.n_ctx) return fail(AIOS_E_INVAL, "range");
        std::lock_guard<std::mutex> lk(M.eng->mu);
        M.eng->make_current();
        if (n == 0) return (int)AIOS_OK;
        if (pos0 + n > s->impl.n_past) return fail(AIOS_E_INVAL, "rows beyond the sequence's position");
        // rows of one layer are kv_pos_bytes apart (position-major pages): strided copy
        const size_t kvd = M.kvd(), pitch = M.kv_pos_bytes();
        CUDA_CHECK(cudaMemcpy2D(k_out, kvd * 2, s->impl.k_cache(layer) + (size_t)pos0 * M.kv_pos_elems(), pitch, kvd * 2, (size_t)n, cudaMemcpyDeviceToHost));
        CUDA_CHECK(cudaMemcpy2D(v_out, kvd * 2, s->impl.v_cache(layer) + (size_t)pos0 * M.kv_pos_elems(), pitch, kvd * 2, (size_t)n, cudaMemcpyDeviceToHost));
        return (int)AIOS_OK;
    });
}

int aios_seq_read_activation(aios_seq *s, int32_t which, float *out, int64_t n) {
    return guarded([&] {
        if (!s || !out || n < 0) return fail(AIOS_E_INVAL, "bad argument");
        Seq &q = s->impl;
        const DevBuf *b = which == 0 ? &q.x : which == 1 ? &q.qkv : which == 2 ? &q.att : which == 3 ? &q.h : nullptr;
        if (!b) return fail(AIOS_E_INVAL, "unknown activation buffer");
        if ((size_t)n * 4 > b->bytes) return fail(AIOS_E_RANGE, "n exceeds the buffer");
        std::lock_guard<std::mutex> lk(q.m->eng->mu);
        q.m->eng->make_current();
        CUDA_CHECK(cudaMemcpy(out, b->p, (size_t)n * 4, cudaMemcpyDeviceToHost));
        return (int)AIOS_OK;
    });
}

int aios_seq_bench_decode(aios_seq *s, int32_t n_tokens, aios_bench_result *out) {
    return guarded([&] {
        if (!s || !out || n_tokens <= 0) return fail(AIOS_E_INVAL, "bad argument");
        std::lock_guard<std::mutex> lk(s->impl.m->eng->mu);
        s->impl.bench_decode(n_tokens, out);
        return (int)AIOS_OK;
    });
}

int aios_seq_profile_decode(aios_seq *s, int32_t n_tokens, aios_bench_result *out) {
    return guarded([&] {
        if (!s || !out || n_tokens <= 0) return fail(AIOS_E_INVAL, "bad argument");
        std::lock_guard<std::mutex> lk(s->impl.m->eng->mu);
        s->impl.profile_decode(n_tokens, out);
        return (int)AIOS_OK;
    });
}
int aios_seq_trace_decode(aios_seq *s, uint64_t *out, int32_t cap_launches, int32_t *n_launches) {
    return guarded([&] {
        if (!s || !out || !n_launches || cap_launches <= 0) return fail(AIOS_E_INVAL, "bad argument");
        std::lock_guard<std::mutex> lk(s->impl.m->eng->mu);
        *n_launches = s->impl.trace_decode(reinterpret_cast<unsigned long long *>(out), cap_launches);
        return (int)AIOS_OK;
    });
}

int aios_seq_phase_timing(aios_seq *s, uint64_t *out, int32_t cap, int32_t *n_phases) {
    return guarded([&] {
        if (!s || !out || !n_phases) return fail(AIOS_E_INVAL, "bad argument");
        Seq &q = s->impl;
        const int n = q.pfd.n_phases;
        *n_phases = n;
        if (cap < n) return fail(AIOS_E_RANGE, "out is too small");
        std::lock_guard<std::mutex> lk(q.m->eng->mu);
        q.m->eng->make_current();
        CUDA_CHECK(cudaMemcpy(out, q.pfd_timing.p, (size_t)n * 4 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
        return (int)AIOS_OK;
    });
}
int aios_seq_snapshot(aios_seq *s) {
    return guarded([&] {
        if (!s) return fail(AIOS_E_INVAL, "null seq");
        std::lock_guard<std::mutex> lk(s->impl.m->eng->mu);
        s->impl.snapshot();
        return (int)AIOS_OK;
    });
}
int aios_seq_restore(aios_seq *s) {
    return guarded([&] {
        if (!s) return fail(AIOS_E_INVAL, "null seq");
        std::lock_guard<std::mutex> lk(s->impl.m->eng->mu);
        s->impl.restore();
        return (int)AIOS_OK;
    });
}

int aios_seq_bench_prefill(aios_seq *s, const int32_t *tokens, int32_t n, int32_t iters, float *ms_per_prefill) {
    return guarded([&] {
        if (!s || !tokens || n <= 0 || iters <= 0 || !ms_per_prefill) return fail(AIOS_E_INVAL, "bad argument");
        Seq &q = s->impl;
        if (n > q.m->hp.n_ctx) return fail(AIOS_E_STATE, "prompt does not fit in the context");
        std::lock_guard<std::mutex> lk(q.m->eng->mu);
        q.m->eng->make_current();
        cudaStream_t st = q.m->eng->stream;
        q.reset();
        q.prefill(tokens, n);  // warm-up: scratch buffers, tensor maps, function attributes
        cudaEvent_t e0, e1;
        CUDA_CHECK(cudaEventCreate(&e0));
        CUDA_CHECK(cudaEventCreate(&e1));
        float total = 0.f;
        const bool prof = getenv("AIOS_PROFILE_STEPS") != nullptr;
        for (int i = 0; i < iters; i++) {
            q.reset();
            CUDA_CHECK(cudaStreamSynchronize(st));
            if (prof && i == 0) cudaProfilerStart();
            CUDA_CHECK(cudaEventRecord(e0, st));
            q.prefill(tokens, n);
            CUDA_CHECK(cudaEventRecord(e1, st));
            CUDA_CHECK(cudaEventSynchronize(e1));
            if (prof && i == 0) cudaProfilerStop();
            float ms = 0.f;
            CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
            total += ms;
        }
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        *ms_per_prefill = total / (float)iters;
        return (int)AIOS_OK;
    });
}

int aios_bench_batch_decode(aios_model *m, int32_t n_lanes, const int32_t *prompts, int32_t n_prompt, int32_t n_steps,
                            int32_t *out_tokens, aios_batch_bench *out) {
    return guarded([&] {
        if (!m || !prompts || !out || n_lanes <= 0 || n_lanes > GEMM_MAX_TOKENS || n_prompt <= 0 || n_steps <= 0)
            return fail(AIOS_E_INVAL, "bad argument");
        if (n_prompt + n_steps > m->impl.hp.n_ctx) return fail(AIOS_E_STATE, "prompt + steps do not fit in the context");
        for (int64_t i = 0; i < (int64_t)n_lanes * n_prompt; i++)
            if (prompts[i] < 0 || prompts[i] >= m->impl.hp.n_vocab) return fail(AIOS_E_INVAL, "token id out of range");
        Engine &E = *m->impl.eng;
        std::lock_guard<std::mutex> lk(E.mu);
        E.make_current();
        Batch b(&m->impl);  // a private batch: the scheduler's own lanes are untouched
        aios_sampling greedy{};
        cudaEvent_t e0, e1, e2;
        CUDA_CHECK(cudaEventCreate(&e0));
        CUDA_CHECK(cudaEventCreate(&e1));
        CUDA_CHECK(cudaEventCreate(&e2));
        CUDA_CHECK(cudaEventRecord(e0, E.stream));
        for (int l = 0; l < n_lanes; l++) b.admit(prompts + (size_t)l * n_prompt, n_prompt, greedy);
        CUDA_CHECK(cudaEventRecord(e1, E.stream));
        b.step();  // untimed: captures the step's graph for this lane count
        if (out_tokens)
            for (int l = 0; l < n_lanes; l++) out_tokens[(size_t)l * n_steps] = b.h_tok[l];
        CUDA_CHECK(cudaEventRecord(e1, E.stream));
        const bool prof = getenv("AIOS_PROFILE_STEPS") != nullptr;  // ncu --profile-from-start off: the timed steps only
        if (prof) cudaProfilerStart();
        for (int s = 1; s < n_steps; s++) {
            b.step();
            if (out_tokens)
                for (int l = 0; l < n_lanes; l++) out_tokens[(size_t)l * n_steps + s] = b.h_tok[l];
        }
        CUDA_CHECK(cudaEventRecord(e2, E.stream));
        CUDA_CHECK(cudaEventSynchronize(e2));
        if (prof) cudaProfilerStop();
        memset(out, 0, sizeof(*out));
        CUDA_CHECK(cudaEventElapsedTime(&out->ms_decode, e1, e2));
        CUDA_CHECK(cudaEventElapsedTime(&out->ms_prefill, e0, e1));
        cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
        out->n_lanes = n_lanes;
        out->n_steps_timed = n_steps - 1;
        out->kernel_launches = b.kernels_per_step * (n_steps - 1);
        out->n_past_begin = n_prompt + 1;
        out->n_past_end = n_prompt + n_steps;
        while (b.n > 0) b.retire(b.n - 1);
        return (int)AIOS_OK;
    });
}

int aios_bench_gemv(aios_engine *e, int32_t type, int64_t n_rows, int64_t k, int32_t n_copies, int32_t iters,
                    float *ms_per_launch) {
    return guarded([&] {
        if (!e || !ms_per_launch || n_rows <= 0 || k <= 0 || n_copies <= 0 || iters <= 0) return fail(AIOS_E_INVAL, "bad argument");
        if (!(type == T_Q4_K || type == T_Q5_K || type == T_Q6_K || type == T_Q8_0)) return fail(AIOS_E_INVAL, "unsupported weight type");
        if (k % 256 || gemv_smem_bytes((int)k) > 100 * 1024) return fail(AIOS_E_RANGE, "bad k");
        std::lock_guard<std::mutex> lk(e->impl.mu);
        Engine &E = e->impl;
        E.make_current();
        // random but valid blocks: fp16 scale fields are overwritten with small finite values
        const size_t nbytes = tensor_bytes(type, n_rows, k);
        std::vector<uint8_t> host(nbytes);
        std::mt19937_64 rng(1234);
        for (size_t i = 0; i + 8 <= nbytes; i += 8) { uint64_t r = rng(); memcpy(&host[i], &r, 8); }
        const int bb = type_block_bytes(type);
        for (size_t b = 0; b < nbytes / bb; b++) {
            uint8_t *blk = &host[b * bb];
            const uint16_t small = 0x1400;  // 2^-10 as fp16
            if (type == T_Q6_K) memcpy(blk + 208, &small, 2);
            else { memcpy(blk, &small, 2); if (type != T_Q8_0) memcpy(blk + 2, &small, 2); }
        }
        std::vector<DevMatrix> ws((size_t)n_copies);
        for (auto &w : ws) make_dev_matrix(E, type, host.data(), n_rows, k, w);
        std::vector<float> hx((size_t)k);
        for (auto &v : hx) v = (float)((int)(rng() % 2001) - 1000) / 1000.f;
        DevBuf dx((size_t)k * 4), dy((size_t)n_rows * 4);
        CUDA_CHECK(cudaMemcpy(dx.p, hx.data(), (size_t)k * 4, cudaMemcpyHostToDevice));
        GemvParams p{};
        p.n_seg = 1; p.epi = EPI_STORE; p.k = (int32_t)k; p.out[0] = dy.as<float>(); p.x = dx.as<float>();
        p.seg[0] = ws[0].w;
        const int grid = gemv_plan_grid(p, E.n_sms);
        for (int i = 0; i < 3 * n_copies; i++) { p.seg[0] = ws[(size_t)i % n_copies].w; launch_gemv(p, grid, E.stream); }
        cudaEvent_t e0, e1;
        CUDA_CHECK(cudaEventCreate(&e0));
        CUDA_CHECK(cudaEventCreate(&e1));
        CUDA_CHECK(cudaStreamSynchronize(E.stream));
        CUDA_CHECK(cudaEventRecord(e0, E.stream));
        for (int i = 0; i < iters; i++) { p.seg[0] = ws[(size_t)i % n_copies].w; launch_gemv(p, grid, E.stream); }
        CUDA_CHECK(cudaEventRecord(e1, E.stream));
        CUDA_CHECK(cudaEventSynchronize(e1));
        CUDA_CHECK(cudaGetLastError());
        float ms = 0;
        CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        *ms_per_launch = ms / (float)iters;
        return (int)AIOS_OK;
    });
}

// ---- tensor parallelism
int aios_tp_model_load(const int32_t *devices, int32_t n_devices, const char *name, const char *gguf_path, int32_t n_ctx,
                       int32_t engine_flags, aios_tp_model **out) {
    return guarded([&] {
        if (!devices || !name || !gguf_path || !out) return fail(AIOS_E_INVAL, "null argument");
        if (n_devices < 2 || n_devices > 8) return fail(AIOS_E_INVAL, "tensor parallelism takes 2 to 8 devices");
        std::vector<int> devs(devices, devices + n_devices);
        std::unique_ptr<aios_tp_model> m(new aios_tp_model());
        m->impl.reset(new aios::TpModel(devs, name, gguf_path, n_ctx, engine_flags));
        *out = m.release();
        return (int)AIOS_OK;
    });
}
int aios_tp_model_unload(aios_tp_model *m) {
    return guarded([&] {
        delete m;
        return (int)AIOS_OK;
    });
}
int aios_tp_model_info(const aios_tp_model *m, aios_model_info *out) {
    return guarded([&] {
        if (!m || !out) return fail(AIOS_E_INVAL, "null argument");
        m->impl->info(out);
        return (int)AIOS_OK;
    });
}
int aios_tp_reset(aios_tp_model *m) {
    return guarded([&] {
        if (!m) return fail(AIOS_E_INVAL, "null argument");
        std::lock_guard<std::mutex> lk(m->mu);
        m->impl->reset();
        return (int)AIOS_OK;
    });
}
int aios_tp_prefill(aios_tp_model *m, const int32_t *tokens, int32_t n_tokens) {
    return guarded([&] {
        if (!m || (!tokens && n_tokens > 0)) return fail(AIOS_E_INVAL, "null argument");
        std::lock_guard<std::mutex> lk(m->mu);
        m->impl->prefill(tokens, n_tokens);
        return (int)AIOS_OK;
    });
}
int aios_tp_decode(aios_tp_model *m, const aios_sampling *sp, int32_t max_tokens, int32_t *out_tokens, int32_t *n_out) {
    return guarded([&] {
        if (!m || !sp || !out_tokens || !n_out) return fail(AIOS_E_INVAL, "null argument");
        std::lock_guard<std::mutex> lk(m->mu);
        *n_out = m->impl->decode(*sp, max_tokens, out_tokens);
        return (int)AIOS_OK;
    });
}
int aios_tp_get_logits(aios_tp_model *m, float *out, int32_t n) {
    return guarded([&] {
        if (!m || !out) return fail(AIOS_E_INVAL, "null argument");
        if (n != m->impl->n_vocab()) return fail(AIOS_E_RANGE, "n must equal n_vocab");
        std::lock_guard<std::mutex> lk(m->mu);
        m->impl->read_logits(out);
        return (int)AIOS_OK;
    });
}
int aios_tp_bench_decode(aios_tp_model *m, int32_t n_tokens, aios_tp_bench *out) {
    return guarded([&] {
        if (!m || !out) return fail(AIOS_E_INVAL, "null argument");
        std::lock_guard<std::mutex> lk(m->mu);
        m->impl->bench_decode(n_tokens, out);
        return (int)AIOS_OK;
    });
}

}  // extern "C"
