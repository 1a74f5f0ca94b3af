// batch.cu — the batched decode step of the continuous batcher (engine.h Batch; SURVEY.md §7 step 9, BASELINE.json
// configs[3]: "64 concurrent agent requests per GPU").  Reference requirement: Infer calls run in parallel against one
// model (runtime/src/main.rs:50-51, grpc_service.rs:86-141), and the orchestrator alone keeps 10 tasks in flight
// (agent-core/src/autonomy.rs:29-35).
#include <math.h>
#include <string.h>

#include <algorithm>

#include "common.cuh"
#include "engine.h"

namespace aios {

Batch::Batch(Model *m_) : m(m_) {
    const HParams &hp = m->hp;
    const int C = capacity();
    const int E = hp.n_embd, FF = hp.n_ff, QD = hp.n_head * hp.head_dim, KVD = hp.n_head_kv * hp.head_dim, LQ = QD + 2 * KVD;
    m->eng->make_current();
    X = DevBuf((size_t)C * E * 4); QKV = DevBuf((size_t)C * LQ * 4); ATT = DevBuf((size_t)C * QD * 4);
    H = DevBuf((size_t)C * FF * 4); GU = DevBuf((size_t)C * 2 * FF * 4); LOGITS = DevBuf((size_t)C * hp.n_vocab * 4);
    const int kmax = std::max(std::max(E, QD), FF);
    IMG_A = DevBuf(gemm_images_bytes(kmax, C)); IMG_B = DevBuf(gemm_images_bytes(kmax, C));
    POS = DevBuf((size_t)C * 4); TOK = DevBuf((size_t)C * 4); KVB = DevBuf((size_t)C * sizeof(void *));
    CFG = DevBuf((size_t)C * sizeof(SampleLane));
    mask_bytes = ((size_t)hp.n_vocab + 7) / 8;
    mask_bytes = (mask_bytes + 15) & ~(size_t)15;
    MASKS = DevBuf((size_t)C * mask_bytes);
    // rows of padding lanes (beyond n, up to the GEMM's token padding) must hold finite values
    for (DevBuf *b : {&X, &QKV, &ATT, &H, &GU}) CUDA_CHECK(cudaMemset(b->p, 0, b->bytes));
    CUDA_CHECK(cudaMemset(TOK.p, 0, TOK.bytes));
    CUDA_CHECK(cudaMallocHost(&h_tok, (size_t)C * 4));
    CUDA_CHECK(cudaMallocHost(&h_pos, (size_t)C * 4));
    CUDA_CHECK(cudaMallocHost(&h_kvb, (size_t)C * sizeof(void *)));
    CUDA_CHECK(cudaMallocHost(&h_cfg, (size_t)C * sizeof(SampleLane)));
    seqs.resize((size_t)C);
    cfg.resize((size_t)C);
    in_seq.assign((size_t)C, 0);
    graphs.assign((size_t)C + 1, nullptr);
}

Batch::~Batch() {
    for (auto g : graphs)
        if (g) cudaGraphExecDestroy(g);
    cudaFreeHost(h_tok); cudaFreeHost(h_pos); cudaFreeHost(h_kvb); cudaFreeHost(h_cfg);
}

int Batch::admit(const int32_t *prompt, int n_prompt, const aios_sampling &sp) {
    if (n >= capacity()) throw std::runtime_error("batch is full");
    const int lane = n;
    if (!seqs[(size_t)lane]) seqs[(size_t)lane].reset(new Seq(m, true));
    Seq &s = *seqs[(size_t)lane];
    s.reset();
    s.prefill(prompt, n_prompt);  // (synchronises) the last prompt token's logits are in s.logits
    s.ext_mask = nullptr;
    in_seq[(size_t)lane] = 1;     // ... and stay there until the lane takes part in a batched step
    SampleLane &c = cfg[(size_t)lane];
    c.temperature = sp.temperature; c.top_k = sp.top_k; c.top_p = sp.top_p; c.min_p = sp.min_p; c.seed = sp.seed;
    c.mask = nullptr;
    n++;
    dirty = true;
    return lane;
}

void Batch::retire(int lane) {
    if (lane < 0 || lane >= n) return;
    cudaStream_t st = m->eng->stream;
    const int last = n - 1;
    seqs[(size_t)lane]->reset();  // (synchronises) gives the KV pages back
    if (lane != last) {
        std::swap(seqs[(size_t)lane], seqs[(size_t)last]);
        const bool had_mask = cfg[(size_t)last].mask != nullptr;
        cfg[(size_t)lane] = cfg[(size_t)last];
        in_seq[(size_t)lane] = in_seq[(size_t)last];
        if (!in_seq[(size_t)last])
            CUDA_CHECK(cudaMemcpyAsync(LOGITS.as<float>() + (size_t)lane * m->hp.n_vocab, LOGITS.as<float>() + (size_t)last * m->hp.n_vocab,
                                       (size_t)m->hp.n_vocab * 4, cudaMemcpyDeviceToDevice, st));
        if (had_mask) {
            CUDA_CHECK(cudaMemcpyAsync(MASKS.as<uint8_t>() + (size_t)lane * mask_bytes, MASKS.as<uint8_t>() + (size_t)last * mask_bytes,
                                       mask_bytes, cudaMemcpyDeviceToDevice, st));
            cfg[(size_t)lane].mask = MASKS.as<uint8_t>() + (size_t)lane * mask_bytes;
        }
    }
    n--;
    dirty = true;
}

void Batch::set_mask(int lane, const uint8_t *bits) {
    if (lane < 0 || lane >= n) return;
    SampleLane &c = cfg[(size_t)lane];
    uint8_t *dst = MASKS.as<uint8_t>() + (size_t)lane * mask_bytes;
    if (!bits) {
        if (c.mask) { c.mask = nullptr; dirty = true; }
        return;
    }
    // pageable source: the copy is staged by the runtime before the call returns, so the caller's buffer is free again
    CUDA_CHECK(cudaMemcpyAsync(dst, bits, ((size_t)m->hp.n_vocab + 7) / 8, cudaMemcpyHostToDevice, m->eng->stream));
    if (c.mask != dst) { c.mask = dst; dirty = true; }
}

void Batch::upload_lane_state(cudaStream_t st) {
    for (int i = 0; i < n; i++) {
        h_pos[i] = seqs[(size_t)i]->n_past - 1;  // position of the last token fed
        h_kvb[i] = seqs[(size_t)i]->kv.as<void>();
        h_cfg[i] = cfg[(size_t)i];
    }
    CUDA_CHECK(cudaMemcpyAsync(POS.p, h_pos, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(KVB.p, h_kvb, (size_t)n * sizeof(void *), cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(CFG.p, h_cfg, (size_t)n * sizeof(SampleLane), cudaMemcpyHostToDevice, st));
}

// sample (every lane) -> embed -> layers -> logits, for the n live lanes
void Batch::launch_step(cudaStream_t st, int *n_launch) {
    const HParams &hp = m->hp;
    const int T = n, Tp = gemm_pad_tokens(T);
    const int E = hp.n_embd, FF = hp.n_ff, QD = hp.n_head * hp.head_dim, KVD = hp.n_head_kv * hp.head_dim, LQ = QD + 2 * KVD;
    float *x = X.as<float>(), *qkv = QKV.as<float>(), *att = ATT.as<float>(), *h = H.as<float>(), *gu = GU.as<float>();
    int launches = 0;
    // Launched kernel by kernel (AIOS_ENGINE_NO_GRAPH) every kernel of the step is a programmatic dependent of the one
    // before it (they all wait before touching data): TinyLlama 64 lanes 3.04 -> 2.74 ms/step.  Inside the captured
    // graph the same edges measured SLOWER (Mistral-7B 9.87 -> 10.44 ms/step, TinyLlama 2.64 -> 2.71): plain edges there.
    PdlScope pdl_scope((m->eng->flags & AIOS_ENGINE_NO_GRAPH) && !(m->eng->flags & AIOS_ENGINE_NO_PDL));
    {
        SampleBatchParams sp{};
        sp.logits = LOGITS.as<float>(); sp.n_vocab = hp.n_vocab; sp.lanes = CFG.as<SampleLane>();
        sp.pos = POS.as<int32_t>(); sp.tok = TOK.as<int32_t>();
        launch_sample_batch(sp, T, st);
        launches++;
    }
    launch_embed_rows(m->tok_embd_type, m->tok_embd_raw.p, m->tok_embd_row_bytes, E, TOK.as<int32_t>(), T, x, E, st);
    launches++;
    // activation images of `in` for weight type(s) of the segments: one image set per permutation class in use
    auto perm_class = [](int type) { return type == T_Q6_K ? 1 : type == T_Q8_0 ? 2 : 0; };
    auto matmul = [&](std::initializer_list<const WPlanes *> ws, std::initializer_list<float *> outs, const float *in, int64_t ldin,
                      const float *norm, int64_t ldo, int epi) {
        GemmMulti g{};
        g.T = T; g.epi = epi; g.ldo = ldo;
        uint8_t *img_of_class[3] = {nullptr, nullptr, nullptr};
        uint8_t *bufs[2] = {IMG_A.as<uint8_t>(), IMG_B.as<uint8_t>()};
        int n_bufs = 0;
        auto o = outs.begin();
        for (const WPlanes *w : ws) {
            const int pc = perm_class(w->type);
            if (!img_of_class[pc]) {
                if (n_bufs >= 2) throw CudaError("batched decode: more than two weight-type classes in one launch");
                img_of_class[pc] = bufs[n_bufs++];
                launch_quant_rows(w->type, in, ldin, norm, hp.rms_eps, w->k, Tp, img_of_class[pc], st);
                launches++;
            }
            g.seg[g.n_seg] = *w; g.out[g.n_seg] = *o++; g.images[g.n_seg] = img_of_class[pc];
            g.n_seg++;
        }
        launch_gemm_multi(g, st);
        launches++;
    };
    const bool batch_ok = attn_batch_usable(hp.n_head, hp.n_head_kv, hp.n_ctx, hp.head_dim);
    for (int il = 0; il < hp.n_layer; il++) {
        const Layer &L = m->layers[il];
        matmul({&L.wq.w, &L.wk.w, &L.wv.w}, {qkv, qkv + QD, qkv + QD + KVD}, x, E, L.attn_norm.as<float>(), LQ, EPI_STORE);
        if (batch_ok) {   // RoPE + KV append + attention: one CTA per (lane, kv head), every lane over its own KV range
            AttnBatchParams a{};
            a.qkv = qkv; a.out = att;
            a.rope_cos = m->rope_cos.as<float>(); a.rope_sin = m->rope_sin.as<float>();
            a.lane_pos = POS.as<int32_t>(); a.lane_kv = KVB.as<uint16_t *>();
            a.k_off = (int64_t)(((size_t)il * 2 + 0) * m->kvd()); a.v_off = (int64_t)(((size_t)il * 2 + 1) * m->kvd());
            a.lane_qkv_stride = LQ; a.lane_out_stride = QD;
            a.n_head = hp.n_head; a.n_kv = hp.n_head_kv; a.hd = hp.head_dim; a.n_rot = hp.n_rot; a.n_ctx = hp.n_ctx;
            a.kv_stride = (int32_t)m->kv_pos_elems(); a.kv_ready = 0;
            a.scale = 1.0f / sqrtf((float)hp.head_dim);
            launch_attn_batch(a, T, st);
            launches++;
        } else {  // very long contexts: the cluster kernel, lane by lane
            for (int l = 0; l < T; l++) {
                AttnParams b{};
                Seq &s = *seqs[(size_t)l];
                b.qkv = qkv + (size_t)l * LQ; b.out = att + (size_t)l * QD;
                b.k_cache = s.k_cache(il); b.v_cache = s.v_cache(il); b.kv_stride = (int32_t)m->kv_pos_elems();
                b.rope_cos = m->rope_cos.as<float>(); b.rope_sin = m->rope_sin.as<float>();
                b.n_past_ptr = POS.as<int32_t>() + l;
                b.scratch = s.attn_scratch.as<float>(); b.counters = s.attn_counters.as<uint32_t>();
                b.n_head = hp.n_head; b.n_kv = hp.n_head_kv; b.hd = hp.head_dim; b.n_rot = hp.n_rot;
                b.n_split = m->attn_splits; b.n_ctx = hp.n_ctx;
                b.scale = 1.0f / sqrtf((float)hp.head_dim);
                launch_attn_decode(b, st, false, 2);
                launches++;
            }
        }
        matmul({&L.wo.w}, {x}, att, QD, nullptr, E, EPI_ADD);
        matmul({&L.w_gate.w, &L.w_up.w}, {gu, gu + FF}, x, E, L.ffn_norm.as<float>(), 2 * FF, EPI_STORE);
        launch_swiglu_rows(gu, 2 * FF, FF, T, h, FF, st);
        launches++;
        matmul({&L.w_down.w}, {x}, h, FF, nullptr, E, EPI_ADD);
    }
    matmul({&m->output.w}, {LOGITS.as<float>()}, x, E, m->output_norm.as<float>(), hp.n_vocab, EPI_STORE);
    if (n_launch) *n_launch = launches;
}

void Batch::step() {
    if (n <= 0) return;
    m->eng->make_current();
    cudaStream_t st = m->eng->stream;
    const size_t row_bytes = (size_t)m->hp.n_vocab * 4;
    if (n <= SMALL_BATCH) {
        // few lanes: every lane through the single-sequence token pass (its own CUDA graph of GEMV kernels), which
        // streams the weights once per lane but without the multi-token kernel's per-(row, token) overhead
        for (int i = 0; i < n; i++) {
            Seq &s = *seqs[(size_t)i];
            if (!in_seq[(size_t)i]) {
                CUDA_CHECK(cudaMemcpyAsync(s.logits.p, LOGITS.as<float>() + (size_t)i * m->hp.n_vocab, row_bytes, cudaMemcpyDeviceToDevice, st));
                in_seq[(size_t)i] = 1;
            }
            const SampleLane &c = cfg[(size_t)i];
            aios_sampling sp{c.temperature, c.top_k, c.top_p, c.min_p, c.seed};
            s.ext_mask = c.mask;
            s.have_logits = true;
            if (s.decode(sp, 1, &h_tok[i]) != 1) throw std::runtime_error("context is full");
        }
        return;
    }
    for (int i = 0; i < n; i++) {
        if (in_seq[(size_t)i]) {  // joins the batched steps: its logits move into its LOGITS row
            CUDA_CHECK(cudaMemcpyAsync(LOGITS.as<float>() + (size_t)i * m->hp.n_vocab, seqs[(size_t)i]->logits.p, row_bytes, cudaMemcpyDeviceToDevice, st));
            in_seq[(size_t)i] = 0;
            dirty = true;
        }
        seqs[(size_t)i]->ensure_kv(seqs[(size_t)i]->n_past + 1);  // the token fed this step
    }
    if (dirty) { upload_lane_state(st); dirty = false; }
    const bool use_graph = !(m->eng->flags & AIOS_ENGINE_NO_GRAPH);
    if (use_graph) {
        cudaGraphExec_t &ge = graphs[(size_t)n];
        if (!ge) {
            cudaGraph_t g = nullptr;
            CUDA_CHECK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            try {
                launch_step(st, &kernels_per_step);
            } catch (...) {
                cudaStreamEndCapture(st, &g);
                if (g) cudaGraphDestroy(g);
                throw;
            }
            CUDA_CHECK(cudaStreamEndCapture(st, &g));
            cudaError_t e = cudaGraphInstantiate(&ge, g, 0);
            cudaGraphDestroy(g);
            CUDA_CHECK(e);
        }
        CUDA_CHECK(cudaGraphLaunch(ge, st));
    } else {
        launch_step(st, &kernels_per_step);
    }
    CUDA_CHECK(cudaMemcpyAsync(h_tok, TOK.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    CUDA_CHECK(cudaGetLastError());
    for (int i = 0; i < n; i++) seqs[(size_t)i]->n_past += 1;
}

}  // namespace aios
