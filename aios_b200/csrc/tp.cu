// tp.cu — tensor parallelism for models that need more than one GPU (BASELINE.json configs[4]:
// Llama-3-70B at TP 2/4/8).  The reference's hardware table lists 70B-class local models
// (docs/HARDWARE.md:135-138); north_star asks for 2/4/8-way tensor parallelism with NCCL all-reduce over NVLink.
//
// Megatron-style split inside ONE process (one engine, one stream and one NCCL communicator per GPU, driven by the
// calling thread): attn_q/k/v, ffn_gate/up and the lm_head are split by rows (heads / FF columns / vocabulary slice),
// attn_output and ffn_down by K at multiples of 256 (quant blocks run along K).  Every rank holds the replicated
// residual stream; per layer two all-reduces (after attn_output and after ffn_down) and, per token, one all-gather of
// the logits slices.
//
// How the order-free double sum crosses ranks: a K-split GEMV stores its rows' UNROUNDED double sums (EPI_STORE_F64);
// the f64 partials are all-reduced (ncclDouble, ncclSum) and rounded to f32 once, then added to the residual — the same
// set of f32 terms as the unsplit sum, so the result is the single-GPU result bit for bit (up to the usual ~1e-7
// chance per value that a double sum's association moves the final rounding).  Every rank then samples from the same
// full logits with the same deterministic sampler: no broadcast of the token is needed.
//
// Two transports for those partial sums.  NCCL (the reference behaviour and the fallback without peer access): the GEMV
// stores its partials locally, ncclAllReduce(ncclDouble) + ncclAllGather per token, launched eagerly.  Push (default):
// the GEMV's epilogue stores every row's partial straight into each rank's receive slab over the NVLink peer mapping
// and its last CTA raises an arrival flag (EPI_PUSH_F64 / EPI_PUSH_F32, gemv.cu); a small wait kernel, chained to its
// neighbours by programmatic dependent launch, adds the slabs to the residual.  No host-side collective is left in a
// decode step, which is then one CUDA graph per rank.  Peer traffic is plain stores and flag loads: no TMA involved.
#include <dlfcn.h>
#include <math.h>
#include <nccl.h>
#include <string.h>

#include <exception>
#include <mutex>
#include <thread>

#include <algorithm>
#include <memory>
#include <vector>

#include "common.cuh"
#include "engine.h"
#include "tp.h"

namespace aios {

namespace {
struct Nccl {
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
// libnccl is opened on first use: the library itself has no link-time dependency on it (single-GPU users never load it)
const Nccl &nccl() {
    static Nccl n;
    static std::once_flag once;
    std::call_once(once, [] {
        void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) throw CudaError(std::string("tensor parallelism needs NCCL: ") + dlerror());
        auto get = [&](const char *name, void *slot) {
            void *f = dlsym(h, name);
            if (!f) throw CudaError(std::string("NCCL symbol missing: ") + name);
            memcpy(slot, &f, sizeof(f));
        };
        get("ncclCommInitAll", &n.CommInitAll); get("ncclCommDestroy", &n.CommDestroy); get("ncclAllReduce", &n.AllReduce);
        get("ncclAllGather", &n.AllGather); get("ncclGroupStart", &n.GroupStart); get("ncclGroupEnd", &n.GroupEnd);
        get("ncclGetErrorString", &n.GetErrorString);
    });
    return n;
}
void nccl_check(ncclResult_t r, const char *what) {
    if (r != ncclSuccess) throw CudaError(std::string(what) + ": " + nccl().GetErrorString(r));
}

// x[i] += (float) part[i]   (the all-reduced f64 partial sums of a K-split projection, rounded once)
__global__ void add_f64_kernel(float *x, const double *part, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) x[i] = __fadd_rn(x[i], (float)part[i]);
}

// ---- the push all-reduce's other half (the first half is the GEMV's EPI_PUSH_F64 epilogue, gemv.cu)
__global__ void tp_tick_kernel(unsigned *tick) { *tick = *tick + 1u; }

// wait until every rank's partials of this epoch have landed in this GPU's slabs, then x[i] += (float) sum over ranks
// (rank order, the same on every GPU: every rank computes the same residual stream bit for bit).
// slabs: [n][2][E] doubles; flags: one per source rank, 32 words apart.  A peer that never arrives (a failed launch on
// another GPU) must not hang this one: after ~2 s the kernel gives up and raises *err, which the host checks.
__global__ void tp_wait_add_kernel(float *x, const double *slabs, const unsigned *flags, const unsigned *tick, int per_token,
                                   int index, int n, int E, unsigned *err) {
    // Launched as a programmatic dependent of the GEMV that pushes (its CTAs come up as the GEMV's CTAs leave, and spin
    // on the flags instead of sitting in the launch queue), and the next GEMV is a dependent of this kernel: it may
    // start fetching its first weights right away and blocks in its own griddepcontrol.wait until x is updated.
    asm volatile("griddepcontrol.launch_dependents;");
    const unsigned ep = __ldcg(tick) * (unsigned)per_token + (unsigned)index;
    if ((int)threadIdx.x < n) {
        const unsigned *f = flags + threadIdx.x * 32;
        const long long t0 = clock64();
        for (;;) {
            unsigned v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
            if ((int)(v - ep) >= 0) break;
            if (clock64() - t0 > 4000000000ll) {
                atomicExch(err, 1u);
                break;
            }
        }
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");  // this GPU's own GEMV has completely finished
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < E) {
        const double *s = slabs + (size_t)(ep & 1u) * E + i;
        double tot = __ldcg(s);
        for (int r = 1; r < n; r++) tot += __ldcg(s + (size_t)r * 2 * E);
        x[i] = __fadd_rn(x[i], (float)tot);
    }
}

// every rank's slice of this token's logits has landed in this GPU's logits buffer (EPI_PUSH_F32): one CTA, n waiters
__global__ void tp_wait_logits_kernel(const unsigned *flags, const unsigned *tick, int n, unsigned *err) {
    const unsigned ep = __ldcg(tick);
    if ((int)threadIdx.x < n) {
        const unsigned *f = flags + threadIdx.x * 32;
        const long long t0 = clock64();
        for (;;) {
            unsigned v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
            if ((int)(v - ep) >= 0) break;
            if (clock64() - t0 > 4000000000ll) {
                atomicExch(err, 1u);
                break;
            }
        }
    }
}

void launch_wait_add(float *x, const double *slabs, const unsigned *flags, const unsigned *tick, int per_token, int index, int n,
                     int E, unsigned *err, cudaStream_t st, bool pdl) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)((E + 255) / 256));
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    CUDA_CHECK(cudaLaunchKernelEx(&cfg, tp_wait_add_kernel, x, slabs, flags, tick, per_token, index, n, E, err));
}
}  // namespace

struct TpRank {
    std::unique_ptr<Engine> eng;
    std::unique_ptr<Model> model;
    std::unique_ptr<Seq> seq;
    ncclComm_t comm = nullptr;
    // push all-reduce: this GPU's receive slabs [n][2][E] doubles, then n flags (32 words apart), done, tick, err
    DevBuf xchg;
    double *slabs = nullptr;
    unsigned *flags = nullptr, *lflags = nullptr, *done = nullptr, *tick = nullptr, *err = nullptr;
    // one decode step (sample, tick, embed, the layers, lm_head, logits wait): [0] with the latency attention kernel,
    // [1] with the position-split one (long contexts)
    cudaGraphExec_t graph[2] = {nullptr, nullptr};
};

struct TpModelImpl {
    std::vector<TpRank> ranks;
    int n = 0;
    int n_past = 0;
    bool have_logits = false;
    bool pdl_wait = false;  // programmatic dependent launch around the wait kernel (push mode, unless AIOS_ENGINE_NO_PDL)
    bool push = false;  // the all-reduce rides in the GEMV epilogue over peer mappings (else: ncclAllReduce)
    void setup_push();
    void check_push_errors();
    // push mode: one CUDA graph per rank holds a whole decode step (no host-side collective call is left in it)
    bool use_graphs = false;
    bool graphs_built[2] = {false, false};
    aios_sampling graph_sp[2] = {};
    int graph_launches = 0;  // kernels inside one rank's graph
    void rank_step(int r, const aios_sampling &sp, int *launches, bool long_attn);
    void ensure_graphs(const aios_sampling &sp, bool long_attn);
    bool want_long(int n_pos) const { return ranks[0].seq->want_long_attn(n_pos); }
    void decode_steps(const aios_sampling &sp, int n_steps, int *launches);
    // all-reduce timing (rank 0's stream), filled when timing is on
    bool timing = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ar_events;

    ~TpModelImpl() {
        for (auto &r : ranks) {
            if (r.eng) cudaSetDevice(r.eng->device);
            for (auto &gx : r.graph)
                if (gx) cudaGraphExecDestroy(gx);
            r.seq.reset();
            r.xchg = DevBuf();
            r.model.reset();
            if (r.comm) nccl().CommDestroy(r.comm);
            r.eng.reset();
        }
        for (auto &e : ar_events) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
    }
    const HParams &hp() const { return ranks[0].model->hp; }
    void all_reduce_add_to_x(int reduce_index);
    // one token through every rank's shard in lockstep; launches counts rank 0's kernels (collectives included)
    void token_pass(bool from_prompt, int prompt_idx, int pos_value, bool want_logits, int *launches, bool long_attn = false);
    void gather_logits();
    bool chunks_usable() const;
    void prefill_chunks(int n_tokens);  // tokens already in every rank's prompt buffer; asynchronous on the streams
    void sample_all(const aios_sampling &sp);
    void sync_all();
    void set_pos_step();
    // the token pass as launch lists per rank, by (want_logits, long_attn); built on first use
    std::vector<std::vector<PassLaunch>> list_cache[2][2];
    std::vector<std::vector<PassLaunch>> &lists_for(bool want_logits, bool long_attn) {
        auto &L = list_cache[want_logits ? 1 : 0][long_attn ? 1 : 0];
        if (L.empty()) {
            L.resize((size_t)n);
            for (int r = 0; r < n; r++) ranks[(size_t)r].seq->build_token_pass(want_logits, L[(size_t)r], long_attn);
        }
        return L;
    }
};

void TpModelImpl::sync_all() {
    for (auto &r : ranks) {
        r.eng->make_current();
        CUDA_CHECK(cudaStreamSynchronize(r.eng->stream));
        CUDA_CHECK(cudaGetLastError());
    }
}

// part64 of every rank: all-reduce (sum of doubles), then x += (float) sum on every rank
void TpModelImpl::all_reduce_add_to_x(int reduce_index) {
    const int E = hp().n_embd;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (timing) {
        ranks[0].eng->make_current();
        CUDA_CHECK(cudaEventCreate(&e0));
        CUDA_CHECK(cudaEventCreate(&e1));
        CUDA_CHECK(cudaEventRecord(e0, ranks[0].eng->stream));
    }
    if (push) {
        const int per_token = 2 * hp().n_layer;
        for (auto &r : ranks) {
            r.eng->make_current();
            launch_wait_add(r.seq->x.as<float>(), r.slabs, r.flags, r.tick, per_token, reduce_index, n, E, r.err, r.eng->stream, pdl_wait);
        }
    } else {
        nccl_check(nccl().GroupStart(), "ncclGroupStart");
        for (auto &r : ranks)
            nccl_check(nccl().AllReduce(r.seq->tp_part64.p, r.seq->tp_part64.p, (size_t)E, ncclDouble, ncclSum, r.comm, r.eng->stream), "ncclAllReduce");
        nccl_check(nccl().GroupEnd(), "ncclGroupEnd");
    }
    if (timing) {
        ranks[0].eng->make_current();
        CUDA_CHECK(cudaEventRecord(e1, ranks[0].eng->stream));
        ar_events.emplace_back(e0, e1);
    }
    if (!push)
        for (auto &r : ranks) {
            r.eng->make_current();
            add_f64_kernel<<<(E + 255) / 256, 256, 0, r.eng->stream>>>(r.seq->x.as<float>(), r.seq->tp_part64.as<double>(), E);
        }
}

// Peer mappings between every pair of shards, the receive slabs, and the push descriptors of every rank's sequence.
// Without full peer access (or with AIOS_TP_ALLREDUCE=nccl) the group stays on ncclAllReduce.
void TpModelImpl::setup_push() {
    const char *mode = getenv("AIOS_TP_ALLREDUCE");
    if (mode && !strcmp(mode, "nccl")) return;
    for (int a = 0; a < n; a++)
        for (int b = 0; b < n; b++) {
            if (a == b) continue;
            int can = 0;
            CUDA_CHECK(cudaDeviceCanAccessPeer(&can, ranks[(size_t)a].eng->device, ranks[(size_t)b].eng->device));
            if (!can) return;
        }
    for (int a = 0; a < n; a++) {
        ranks[(size_t)a].eng->make_current();
        for (int b = 0; b < n; b++) {
            if (a == b) continue;
            const cudaError_t e = cudaDeviceEnablePeerAccess(ranks[(size_t)b].eng->device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else CUDA_CHECK(e);
        }
    }
    const int E = hp().n_embd;
    const size_t slab_bytes = (size_t)n * 2 * E * sizeof(double);
    for (auto &r : ranks) {
        r.eng->make_current();
        r.xchg = DevBuf(slab_bytes + ((size_t)n * 64 + 96) * sizeof(unsigned));
        CUDA_CHECK(cudaMemset(r.xchg.p, 0, r.xchg.bytes));
        r.slabs = r.xchg.as<double>();
        r.flags = reinterpret_cast<unsigned *>(r.xchg.as<uint8_t>() + slab_bytes);
        r.lflags = r.flags + (size_t)n * 32;
        r.done = r.lflags + (size_t)n * 32;
        r.tick = r.done + 32;
        r.err = r.tick + 32;
    }
    for (int a = 0; a < n; a++) {
        TpPush &P = ranks[(size_t)a].seq->tp_push;
        P = TpPush{};
        for (int b = 0; b < n; b++) {
            P.dst[b] = ranks[(size_t)b].slabs + (size_t)a * 2 * E;  // rank b's slab for rank a's partials
            P.flag[b] = ranks[(size_t)b].flags + (size_t)a * 32;
        }
        P.done = ranks[(size_t)a].done;
        P.tick = ranks[(size_t)a].tick;
        P.n = n;
        P.per_token = 2 * hp().n_layer;
        // the lm_head's slice of the logits: straight into every rank's logits buffer.  No double buffering needed: a
        // rank samples from its logits before its next token pass, and nobody reaches the next lm_head without that
        // rank's contributions to the token's all-reduces.
        TpPush &L = ranks[(size_t)a].seq->tp_push_logits;
        L = TpPush{};
        const size_t vl = (size_t)ranks[0].model->vocab_local;
        for (int b = 0; b < n; b++) {
            L.dst[b] = reinterpret_cast<double *>(ranks[(size_t)b].seq->logits.as<float>() + (size_t)a * vl);
            L.flag[b] = ranks[(size_t)b].lflags + (size_t)a * 32;
        }
        L.done = ranks[(size_t)a].done;
        L.tick = ranks[(size_t)a].tick;
        L.n = n;
        L.per_token = 1;
        L.index = 0;
    }
    push = true;
    pdl_wait = !(ranks[0].eng->flags & AIOS_ENGINE_NO_PDL) && !getenv("AIOS_TP_NO_PDL_WAIT");
    for (auto &r : ranks) r.seq->tp_pdl_wait = pdl_wait;
    use_graphs = !(ranks[0].eng->flags & AIOS_ENGINE_NO_GRAPH);
}

void TpModelImpl::check_push_errors() {
    if (!push) return;
    for (auto &r : ranks) {
        r.eng->make_current();
        unsigned e = 0;
        CUDA_CHECK(cudaMemcpy(&e, r.err, sizeof(e), cudaMemcpyDeviceToHost));
        if (e) throw CudaError("tensor-parallel all-reduce: a peer's partial sums never arrived");
    }
}

void TpModelImpl::token_pass(bool from_prompt, int prompt_idx, int pos_value, bool want_logits, int *launches, bool long_attn) {
    auto &lists = lists_for(want_logits, long_attn);
    int li = 0, n_reduce = 0;
    for (auto &r : ranks) {
        r.eng->make_current();
        if (push) tp_tick_kernel<<<1, 1, 0, r.eng->stream>>>(r.tick);
        r.seq->launch_embed(from_prompt, prompt_idx, pos_value, r.eng->stream);
    }
    li += push ? 2 : 1;
    const size_t n_items = lists[0].size();
    for (size_t i = 0; i < n_items; i++) {
        for (int r = 0; r < n; r++) {
            TpRank &R = ranks[(size_t)r];
            R.eng->make_current();
            R.seq->launch_pass_item(lists[(size_t)r][i], R.eng->stream, li);
        }
        li++;
        if (lists[0][i].reduce_after) {
            all_reduce_add_to_x(n_reduce++);
            li += push ? 1 : 2;
        }
    }
    if (want_logits) {
        gather_logits();
        li++;
    }
    if (launches) *launches += li;
}

// after the lm_head launches: every rank ends up with every slice of the logits
void TpModelImpl::gather_logits() {
    if (push) {  // the slices were pushed by the lm_head's epilogue: wait for the arrival flags
        for (auto &R : ranks) {
            R.eng->make_current();
            tp_wait_logits_kernel<<<1, 32, 0, R.eng->stream>>>(R.lflags, R.tick, n, R.err);
        }
        return;
    }
    // in place: rank r's slice already sits at r * vocab_local
    const size_t vl = (size_t)ranks[0].model->vocab_local;
    nccl_check(nccl().GroupStart(), "ncclGroupStart");
    for (int r = 0; r < n; r++) {
        TpRank &R = ranks[(size_t)r];
        float *lg = R.seq->logits.as<float>();
        nccl_check(nccl().AllGather(lg + (size_t)r * vl, lg, vl, ncclFloat, R.comm, R.eng->stream), "ncclAllGather");
    }
    nccl_check(nccl().GroupEnd(), "ncclGroupEnd");
}

// Prompt chunks of up to 64 tokens (Seq::prefill_chunk on every shard: the multi-token GEMM streams each shard's
// weights once per chunk instead of once per token).  One host thread per rank, because a chunk is a long sequence of
// launches with an all-reduce after every K-split projection: each thread enqueues its rank's share, NCCL pairs the
// all-reduces up on the devices ([T][n_embd] doubles each: bandwidth-sized, unlike the decode step's).
bool TpModelImpl::chunks_usable() const {
    if (ranks[0].eng->flags & AIOS_ENGINE_NO_GEMM) return false;
    for (const Layer &L : ranks[0].model->layers)
        if (!gemm_q_usable(L.wo.w.type, L.wo.w.k) || !gemm_q_usable(L.w_down.w.type, L.w_down.w.k)) return false;
    return true;
}

void TpModelImpl::prefill_chunks(int n_tokens) {
    struct Reduce : ChunkReduce {
        ncclComm_t comm;
        void all_reduce(double *part, size_t count, cudaStream_t st) override {
            nccl_check(nccl().AllReduce(part, part, count, ncclDouble, ncclSum, comm, st), "ncclAllReduce");
        }
    };
    std::vector<std::exception_ptr> errs((size_t)n);
    std::vector<std::thread> threads;
    for (int r = 0; r < n; r++) {
        threads.emplace_back([&, r] {
            try {
                TpRank &R = ranks[(size_t)r];
                R.eng->make_current();
                Reduce red;
                red.comm = R.comm;
                for (int i = 0; i < n_tokens; i += GEMM_MAX_TOKENS) {
                    const int T = std::min((int)GEMM_MAX_TOKENS, n_tokens - i);
                    R.seq->prefill_chunk(i, T, n_past + i, false, R.eng->stream, &red);
                }
                // the last token's residual row is what the lm_head reads
                const int last = (n_tokens - 1) % GEMM_MAX_TOKENS;
                const size_t E = (size_t)R.model->hp.n_embd;
                CUDA_CHECK(cudaMemcpyAsync(R.seq->x.p, R.seq->cx.as<float>() + (size_t)last * E, E * 4, cudaMemcpyDeviceToDevice, R.eng->stream));
            } catch (...) {
                errs[(size_t)r] = std::current_exception();
            }
        });
    }
    for (auto &t : threads) t.join();
    for (auto &e : errs)
        if (e) std::rethrow_exception(e);
    // the logits of the last token: lm_head on every shard's slice, then the gather
    auto &lists = lists_for(true, false);
    for (int r = 0; r < n; r++) {
        TpRank &R = ranks[(size_t)r];
        R.eng->make_current();
        if (push) tp_tick_kernel<<<1, 1, 0, R.eng->stream>>>(R.tick);
        PassLaunch L = lists[(size_t)r].back();
        L.pdl = false;
        R.seq->launch_pass_item(L, R.eng->stream, 0);
    }
    gather_logits();
}

// the sampler reads pos (position of the last fed token) and step (index into out_tokens) from the device state
void TpModelImpl::set_pos_step() {
    const int32_t hdr[2] = {n_past - 1, 0};
    for (auto &r : ranks) {
        r.eng->make_current();
        CUDA_CHECK(cudaMemcpyAsync(r.seq->state.p, hdr, sizeof(hdr), cudaMemcpyHostToDevice, r.eng->stream));
    }
}

// one decode step of ONE rank, in stream order (push mode only: the ranks meet through flags, not through the host)
void TpModelImpl::rank_step(int r, const aios_sampling &sp, int *launches, bool long_attn) {
    TpRank &R = ranks[(size_t)r];
    R.eng->make_current();
    cudaStream_t st = R.eng->stream;
    const int E = hp().n_embd, per_token = 2 * hp().n_layer;
    int li = 0, n_reduce = 0;
    R.seq->launch_sample_kernel(sp, st);
    tp_tick_kernel<<<1, 1, 0, st>>>(R.tick);
    R.seq->launch_embed(false, 0, 0, st);
    li += 3;
    for (PassLaunch &L : lists_for(true, long_attn)[(size_t)r]) {
        R.seq->launch_pass_item(L, st, li++);
        if (L.reduce_after) {
            launch_wait_add(R.seq->x.as<float>(), R.slabs, R.flags, R.tick, per_token, n_reduce++, n, E, R.err, st, pdl_wait);
            li++;
        }
    }
    tp_wait_logits_kernel<<<1, 32, 0, st>>>(R.lflags, R.tick, n, R.err);
    li++;
    if (launches) *launches = li;
}

void TpModelImpl::ensure_graphs(const aios_sampling &sp, bool long_attn) {
    const int v = long_attn ? 1 : 0;
    if (graphs_built[v] && !memcmp(&sp, &graph_sp[v], sizeof(sp))) return;
    lists_for(true, long_attn);
    sync_all();
    for (int r = 0; r < n; r++) {
        TpRank &R = ranks[(size_t)r];
        R.eng->make_current();
        if (R.graph[v]) { cudaGraphExecDestroy(R.graph[v]); R.graph[v] = nullptr; }
        cudaGraph_t g = nullptr;
        CUDA_CHECK(cudaStreamBeginCapture(R.eng->stream, cudaStreamCaptureModeThreadLocal));
        try {
            rank_step(r, sp, &graph_launches, long_attn);
        } catch (...) {
            cudaStreamEndCapture(R.eng->stream, &g);
            if (g) cudaGraphDestroy(g);
            throw;
        }
        CUDA_CHECK(cudaStreamEndCapture(R.eng->stream, &g));
        const cudaError_t e = cudaGraphInstantiate(&R.graph[v], g, 0);
        cudaGraphDestroy(g);
        CUDA_CHECK(e);
    }
    graph_sp[v] = sp;
    graphs_built[v] = true;
}

// n_steps x (sample the next token from the logits, feed it); asynchronous
void TpModelImpl::decode_steps(const aios_sampling &sp, int n_steps, int *launches) {
    // step i feeds the token at position n_past + i: n_past + i + 1 rows to attend to (n_past is the caller's)
    if (use_graphs && !timing) {
        for (int i = 0; i < n_steps; i++) {
            const bool la = want_long(n_past + i + 1);
            ensure_graphs(sp, la);
            for (auto &R : ranks) {
                R.eng->make_current();
                CUDA_CHECK(cudaGraphLaunch(R.graph[la ? 1 : 0], R.eng->stream));
            }
        }
        if (launches) *launches += n_steps * graph_launches;
        return;
    }
    for (int i = 0; i < n_steps; i++) {
        sample_all(sp);
        if (launches) (*launches)++;
        token_pass(false, 0, 0, true, launches, want_long(n_past + i + 1));
    }
}

void TpModelImpl::sample_all(const aios_sampling &sp) {
    for (auto &r : ranks) {
        r.eng->make_current();
        r.seq->launch_sample_kernel(sp, r.eng->stream);
    }
}

// ---------------------------------------------------------------- TpModel
TpModel::TpModel(const std::vector<int> &devices, const std::string &name, const std::string &path, int n_ctx, int flags) {
    const int n = (int)devices.size();
    if (n < 2 || n > 8) throw std::invalid_argument("tensor parallelism takes 2 to 8 devices");
    for (int i = 0; i < n; i++)
        for (int j = 0; j < i; j++)
            if (devices[(size_t)i] == devices[(size_t)j]) throw std::invalid_argument("tensor-parallel devices must be distinct");
    impl.reset(new TpModelImpl());
    impl->n = n;
    impl->ranks.resize((size_t)n);
    for (int r = 0; r < n; r++) {
        aios_engine_opts o{};
        o.device = devices[(size_t)r];
        o.max_batch = 1;
        o.flags = flags;
        TpRank &R = impl->ranks[(size_t)r];
        R.eng.reset(new Engine(o));
        R.model.reset(new Model(R.eng.get(), name, path, n_ctx, r, n));
        R.seq.reset(new Seq(R.model.get(), /*internal=*/false));
    }
    std::vector<ncclComm_t> comms((size_t)n);
    nccl_check(nccl().CommInitAll(comms.data(), n, devices.data()), "ncclCommInitAll");
    for (int r = 0; r < n; r++) impl->ranks[(size_t)r].comm = comms[(size_t)r];
    impl->setup_push();
}
TpModel::~TpModel() = default;

int TpModel::n_ranks() const { return impl->n; }
int TpModel::n_vocab() const { return impl->hp().n_vocab; }
int TpModel::n_ctx() const { return impl->hp().n_ctx; }
int TpModel::n_past() const { return impl->n_past; }
void TpModel::info(aios_model_info *o) const {
    memset(o, 0, sizeof(*o));
    Model &M = *impl->ranks[0].model;
    strncpy(o->name, M.name.c_str(), sizeof(o->name) - 1);
    strncpy(o->arch, M.arch.c_str(), sizeof(o->arch) - 1);
    const int n = impl->n;
    o->n_layer = M.hp.n_layer; o->n_embd = M.hp.n_embd; o->n_head = M.hp.n_head * n; o->n_head_kv = M.hp.n_head_kv * n;
    o->head_dim = M.hp.head_dim; o->n_ff = M.hp.n_ff * n; o->n_vocab = M.hp.n_vocab; o->n_ctx = M.hp.n_ctx;
    o->n_rot = M.hp.n_rot; o->file_type = M.hp.file_type; o->n_tensors = (int32_t)M.file->tensors().size();
    o->max_batch = 1; o->rms_eps = M.hp.rms_eps; o->rope_freq_base = M.hp.rope_freq_base;
    o->file_bytes = (int64_t)M.file->file_bytes(); o->weight_bytes = M.weight_bytes;
    o->kv_bytes = (int64_t)(M.kv_pool.pages_in_use() * M.kv_pool.page_bytes);
    o->weight_bytes_per_token = M.weight_bytes_per_token;
    o->kv_bytes_per_position = (int64_t)M.kv_pos_bytes();
}

void TpModel::reset() {
    for (auto &r : impl->ranks) {
        r.eng->make_current();
        r.seq->reset();
    }
    impl->n_past = 0;
    impl->have_logits = false;
}

void TpModel::prefill(const int32_t *tokens, int n) {
    if (n <= 0) return;
    const HParams &hp = impl->hp();
    if (impl->n_past + n > hp.n_ctx) throw std::runtime_error("prefill would exceed the context length");
    for (int i = 0; i < n; i++)
        if (tokens[i] < 0 || tokens[i] >= hp.n_vocab) throw std::invalid_argument("token id out of range");
    for (auto &r : impl->ranks) {
        r.eng->make_current();
        CUDA_CHECK(cudaMemcpyAsync(r.seq->prompt.p, tokens, (size_t)n * 4, cudaMemcpyHostToDevice, r.eng->stream));
    }
    if (n >= 4 && impl->chunks_usable()) impl->prefill_chunks(n);
    else
        for (int i = 0; i < n; i++) impl->token_pass(true, i, impl->n_past + i, i == n - 1, nullptr, impl->want_long(impl->n_past + i + 1));
    impl->sync_all();
    impl->check_push_errors();
    impl->n_past += n;
    impl->have_logits = true;
}

int TpModel::decode(const aios_sampling &sp, int max_tokens, int32_t *out) {
    if (!impl->have_logits) throw std::runtime_error("decode needs a prefill first");
    const HParams &hp = impl->hp();
    const int room = hp.n_ctx - impl->n_past;
    const int n = std::min(max_tokens, room);
    if (n <= 0) return 0;
    impl->set_pos_step();
    impl->decode_steps(sp, n, nullptr);  // per step: token -> state.cur_token, pos++, out_tokens[step++], then the pass
    impl->sync_all();
    impl->check_push_errors();
    TpRank &R0 = impl->ranks[0];
    R0.eng->make_current();
    CUDA_CHECK(cudaMemcpy(out, R0.seq->out_tokens.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    impl->n_past += n;
    return n;
}

void TpModel::read_logits(float *out) {
    if (!impl->have_logits) throw std::runtime_error("no logits yet");
    TpRank &R0 = impl->ranks[0];
    R0.eng->make_current();
    CUDA_CHECK(cudaMemcpy(out, R0.seq->logits.p, (size_t)impl->hp().n_vocab * 4, cudaMemcpyDeviceToHost));
}

void TpModel::bench_decode(int n_tokens, aios_tp_bench *out) {
    if (!impl->have_logits) throw std::runtime_error("bench needs a prefill first");
    memset(out, 0, sizeof(*out));
    const HParams &hp = impl->hp();
    n_tokens = std::min(n_tokens, hp.n_ctx - impl->n_past - 2);
    if (n_tokens <= 0) throw std::runtime_error("no room left in the context");
    aios_sampling greedy{};
    greedy.temperature = 0.f;
    impl->set_pos_step();
    impl->decode_steps(greedy, 2, nullptr);  // warm-up (captures the graphs)
    impl->sync_all();
    impl->n_past += 2;
    out->n_past_begin = impl->n_past;
    TpRank &R0 = impl->ranks[0];
    // pass 1: the whole region, nothing else on the streams
    R0.eng->make_current();
    cudaEvent_t a, b;
    CUDA_CHECK(cudaEventCreate(&a));
    CUDA_CHECK(cudaEventCreate(&b));
    int launches = 0;
    impl->sync_all();
    R0.eng->make_current();
    CUDA_CHECK(cudaEventRecord(a, R0.eng->stream));
    impl->decode_steps(greedy, n_tokens, &launches);
    R0.eng->make_current();
    CUDA_CHECK(cudaEventRecord(b, R0.eng->stream));
    impl->sync_all();
    R0.eng->make_current();
    CUDA_CHECK(cudaEventElapsedTime(&out->ms_total, a, b));
    impl->n_past += n_tokens;
    out->n_tokens = n_tokens;
    out->kernel_launches = launches;
    out->n_ranks = impl->n;
    out->allreduce_kind = impl->push ? 1 : 0;
    out->graphs = impl->use_graphs ? 1 : 0;
    // pass 2 (a few tokens): an event pair around every all-reduce on rank 0's stream
    const int n_timed = std::min(8, hp.n_ctx - impl->n_past);
    if (n_timed > 0) {
        impl->timing = true;  // launched kernel by kernel
        impl->decode_steps(greedy, n_timed, nullptr);
        impl->sync_all();
        impl->timing = false;
        impl->n_past += n_timed;
        R0.eng->make_current();
        double ms = 0;
        for (auto &e : impl->ar_events) {
            float t = 0;
            CUDA_CHECK(cudaEventElapsedTime(&t, e.first, e.second));
            ms += t;
            cudaEventDestroy(e.first);
            cudaEventDestroy(e.second);
        }
        out->n_allreduce_per_token = (int32_t)(impl->ar_events.size() / (size_t)n_timed);
        out->us_per_allreduce = impl->ar_events.empty() ? 0.f : (float)(ms * 1e3 / (double)impl->ar_events.size());
        impl->ar_events.clear();
    }
    out->n_past_end = impl->n_past;
    cudaEventDestroy(a);
    cudaEventDestroy(b);
}

}  // namespace aios
