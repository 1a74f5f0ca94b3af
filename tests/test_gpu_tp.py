"""Tensor parallelism (aios_tp_*): a model split over 2 / 4 GPUs must produce the unsplit model's logits and tokens.

The K-split projections all-reduce f64 partial sums and round once, so the comparison is bit for bit (against the
oracle, which is also what the single-GPU engine is pinned to).  Skipped on boxes with fewer GPUs.
(BASELINE.json configs[4]; reference: docs/HARDWARE.md:135-138 lists the 70B-class models this exists for.)
"""
import numpy as np
import pytest

from aios_b200 import capi
from aios_b200 import gguf_synth as gs

pytestmark = pytest.mark.gpu


def _need_gpus(n):
    if capi.device_count() < n:
        pytest.skip(f"needs {n} GPUs")


@pytest.mark.parametrize("arch,ftype,n_dev", [("tiny-gqa4", "Q4_K_M", 2), ("tiny-gqa4", "Q8_0", 2), ("tiny-tp4", "Q4_K_M", 4),
                                             ("tiny-tp4", "Q8_0", 2)])
@pytest.mark.parametrize("allreduce", ["push", "push-eager", "nccl"])
def test_tp_matches_oracle(oracle, tmp_models, monkeypatch, arch, ftype, n_dev, allreduce):
    _need_gpus(n_dev)
    monkeypatch.setenv("AIOS_TP_ALLREDUCE", allreduce.split("-")[0])
    flags = capi.ENGINE_NO_GRAPH if allreduce.endswith("eager") else 0
    path = tmp_models(arch, ftype)
    prompt = gs.synthetic_prompt(arch, 24, 5)
    om = oracle.OracleModel(path, n_ctx=128)
    ref_logits = om.eval(prompt)
    ref, _ = om.greedy(prompt, 24)
    tp = capi.TpModel(list(range(n_dev)), "tp", path, n_ctx=128, flags=flags)
    try:
        assert tp.info.n_head == gs.ARCHS[arch].n_head and tp.info.n_ff == gs.ARCHS[arch].n_ff
        tp.reset()
        tp.prefill(prompt)
        assert np.array_equal(tp.logits().view(np.uint32), ref_logits.view(np.uint32))
        assert np.array_equal(tp.decode(24), ref)
        # a second request on the same shards, prompt fed in two pieces
        tp.reset()
        tp.prefill(prompt[:10])
        tp.prefill(prompt[10:])
        assert np.array_equal(tp.decode(24), ref)
    finally:
        tp.unload()


def test_tp_stochastic_matches_single_gpu(tmp_models):
    """Every rank draws from the same logits with the same counter-based generator: the split model's sampled tokens
    are the single-GPU engine's."""
    _need_gpus(2)
    path = tmp_models("tiny-gqa4", "Q4_K_M")
    prompt = gs.synthetic_prompt("tiny-gqa4", 12, 9)
    sp = capi.make_sampling(temperature=0.9, top_k=50, top_p=0.9, min_p=0.02, seed=1234)
    eng = capi.Engine(device=0)
    try:
        m = eng.load_model("one", path, n_ctx=128)
        want = m.generate(prompt, 32, sp)
        m.unload()
    finally:
        eng.close()
    tp = capi.TpModel([0, 1], "tp", path, n_ctx=128)
    try:
        assert np.array_equal(tp.generate(prompt, 32, sp), want)
        b = tp.bench_decode(16)
        assert b.allreduce_kind == 1  # two GPUs of one box see each other: the fused all-reduce is the default
        assert b.n_ranks == 2 and b.n_tokens == 16 and b.n_allreduce_per_token == 2 * gs.ARCHS["tiny-gqa4"].n_layer
        assert b.us_per_allreduce > 0 and b.ms_total > 0
    finally:
        tp.unload()


def test_tp_rejects_uneven_split(tmp_models):
    _need_gpus(2)
    path = tmp_models("tiny", "Q4_K_M")  # n_head * head_dim = 256: half of it is no multiple of 256
    with pytest.raises(capi.AiosError) as ei:
        capi.TpModel([0, 1], "bad", path, n_ctx=64)
    assert ei.value.code == capi.E_FORMAT
    with pytest.raises(capi.AiosError):
        capi.TpModel([0, 0], "dup", path, n_ctx=64)


def test_runtime_serves_a_tensor_parallel_model(tmp_path):
    """The gRPC-facing layer on a split model: a model whose name says 70b is loaded tensor-parallel by the backend
    (`--tp 2`), Infer / StreamInfer give the text of the same file served from one GPU, HealthCheck shows both GPUs."""
    import dataclasses
    import json
    import os

    from aios_b200.runtime.inference import EngineBackend, InferenceEngine, InferParams, TpLoaded

    _need_gpus(2)
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "spm_vocab.json"), encoding="utf-8") as f:
        g = json.load(f)
    vocab = [tuple(v) for v in g["vocab"]] + [("<|im_start|>", 0.0, 3), ("<|im_end|>", 0.0, 3)]
    n_vocab = (len(vocab) + 7) // 2 * 2
    arch = dataclasses.replace(gs.ARCHS["tiny-gqa4"], n_vocab=n_vocab)
    path = str(tmp_path / "llama-3-70b-synthetic.gguf")
    gs.write_synthetic_gguf(path, arch, vocab=vocab, chat_template=gs.CHATML_TEMPLATE)
    p = InferParams(prompt="Hello, world!", system_prompt="Be brief.", max_tokens=24, temperature=0.8, seed=7)
    eng = InferenceEngine(stream_chunk_tokens=4)
    one = EngineBackend(device=0)
    try:
        lm = one.load("llama-3-70b", path, 256)
        want = eng.infer(lm, "llama-3-70b", p)
        one.unload(lm)
    finally:
        one.close()
    be = EngineBackend(devices=[0, 1], tp=2)
    try:
        lm = be.load("llama-3-70b", path, 256)
        assert isinstance(lm, TpLoaded) and be.alive(lm) and be.gpu_load() == {0: [1, 0], 1: [1, 0]}
        got = eng.infer(lm, "llama-3-70b", p)
        assert got.text == want.text and got.tokens_used == want.tokens_used
        chunks = list(eng.stream_infer(lm, "llama-3-70b", p))
        assert chunks[-1] == ("", True) and "".join(t for t, _ in chunks) == want.text
        small = be.load("tinyllama", path, 256)  # other models are still replicated, one replica per GPU
        assert not isinstance(small, TpLoaded) and len(small.replicas) == 2
        # ... and requests alternate between the replicas (least loaded, ties in turn), same text from either GPU
        texts = [eng.infer(small, "tinyllama", p).text for _ in range(4)]
        assert texts == [want.text] * 4 and small.served == [2, 2]
        be.unload(small)
        be.unload(lm)
        assert not be.alive(lm)
    finally:
        be.close()


@pytest.mark.parametrize("n_ctx,long_min", [(512, 1), (512, 50), (5000, None)])
def test_tp_long_context_attention(oracle, tmp_models, monkeypatch, n_ctx, long_min):
    """The shards switch to the position-split attention kernel like a single GPU does (forced from the first token,
    switched in mid-call, and a context beyond the latency kernel's 4096 rows as loaded): same tokens as the oracle."""
    _need_gpus(2)
    if long_min is not None:
        monkeypatch.setenv("AIOS_ATTN_LONG_MIN", str(long_min))
    path = tmp_models("tiny-gqa4", "Q4_K_M")
    prompt = gs.synthetic_prompt("tiny-gqa4", 40, 21)
    om = oracle.OracleModel(path, n_ctx=128)
    ref, _ = om.greedy(prompt, 32)
    for flags in (0, capi.ENGINE_NO_GRAPH):
        tp = capi.TpModel([0, 1], "tp", path, n_ctx=n_ctx, flags=flags)
        try:
            assert np.array_equal(tp.generate(prompt, 32), ref)
        finally:
            tp.unload()
