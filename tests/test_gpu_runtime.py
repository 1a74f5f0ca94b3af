"""GPU tests of the drop-in boundary end to end: the reference-shaped gRPC service on the real B200 engine.

A caller's view: LoadModel a GGUF, Infer / StreamInfer text, HealthCheck, UnloadModel — through a real socket,
the C ABI underneath, no CPU fallback anywhere.  Token-level parity of what the engine generates is covered by
tests/test_gpu_parity.py; here the greedy continuation the service decodes is checked against the oracle once
more through the text path (chat prompt -> tokenizer -> prefill -> decode -> detokenizer).
"""
import dataclasses
import json
import os

import grpc
import numpy as np
import pytest

from aios_b200 import capi, gguf_synth
from aios_b200.runtime import proto as pb
from aios_b200.runtime.inference import EngineBackend, InferenceEngine, InferParams
from aios_b200.runtime.service import RuntimeServer

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def chat_gguf(tmp_path_factory):
    with open(os.path.join(HERE, "golden", "spm_vocab.json"), encoding="utf-8") as f:
        g = json.load(f)
    vocab = [tuple(v) for v in g["vocab"]] + [("<|im_start|>", 0.0, 3), ("<|im_end|>", 0.0, 3)]
    arch = dataclasses.replace(gguf_synth.ARCHS["tiny"], n_vocab=len(vocab) + 6)
    d = tmp_path_factory.mktemp("chat")
    path = str(d / "tinyllama-1.1b-synthetic.gguf")
    gguf_synth.write_synthetic_gguf(path, arch, vocab=vocab, chat_template=gguf_synth.CHATML_TEMPLATE)
    return path


@pytest.fixture(scope="module")
def backend():
    be = EngineBackend(device=0)
    yield be
    be.close()


def test_model_vocab_matches_file_vocab(backend, chat_gguf):
    lm = backend.load("v", chat_gguf, 256)
    try:
        fv = capi.Vocab.open(chat_gguf)
        text = lm.vocab.chat_prompt("Be brief.", "What is the capital of France?")
        assert text == fv.chat_prompt("Be brief.", "What is the capital of France?")
        a = lm.vocab.tokenize(text, parse_special=True)
        assert np.array_equal(a, fv.tokenize(text, parse_special=True)) and a.size > 10
        assert lm.model.meta_str("tokenizer.ggml.model") == "llama"
        assert lm.model.meta_str("no.such.key") is None
    finally:
        backend.unload(lm)


def test_text_path_greedy_matches_oracle(backend, chat_gguf, oracle):
    """chat prompt -> ids -> engine greedy tokens == oracle greedy tokens on the same ids."""
    lm = backend.load("g", chat_gguf, 256)
    try:
        ids = lm.vocab.tokenize(lm.vocab.chat_prompt("", "Hello, world!"), parse_special=True)
        lm.seq.reset()
        lm.seq.prefill(ids)
        got = lm.seq.decode(24)
        om = oracle.OracleModel(chat_gguf, n_ctx=256)
        want, _margins = om.greedy(ids, 24)
        assert np.array_equal(got, want)
    finally:
        backend.unload(lm)


def test_infer_engine_on_device(backend, chat_gguf):
    lm = backend.load("m", chat_gguf, 256)
    try:
        eng = InferenceEngine(stream_chunk_tokens=4)
        p = InferParams(prompt="Hello, world!", system_prompt="Be brief.", max_tokens=24, temperature=0.8, seed=7)
        r1 = eng.infer(lm, "m", p)
        r2 = eng.infer(lm, "m", p)
        n_prompt = lm.vocab.tokenize(lm.vocab.chat_prompt(p.system_prompt, p.prompt), parse_special=True).size
        assert r1.text == r2.text and r1.tokens_used == r2.tokens_used  # same seed, same sample path
        assert n_prompt < r1.tokens_used <= n_prompt + 24 and r1.model_used == "m"
        chunks = list(eng.stream_infer(lm, "m", p))
        assert chunks[-1] == ("", True) and "".join(t for t, _ in chunks) == r1.text
        other = eng.infer(lm, "m", dataclasses.replace(p, seed=8))
        assert other.tokens_used > n_prompt
        with pytest.raises(RuntimeError, match="exceeds the available context"):
            eng.infer(lm, "m", InferParams(prompt="word " * 400))
    finally:
        backend.unload(lm)


def test_concurrent_infer_rpcs(backend, chat_gguf):
    """Eight Infer / StreamInfer RPCs at once (tonic handlers run concurrently in the reference): all are served, and a
    pinned seed gives every caller the same text as a lone call."""
    import threading
    lm = backend.load("c", chat_gguf, 256)
    try:
        eng = InferenceEngine(stream_chunk_tokens=4)
        p = InferParams(prompt="Hello, world!", max_tokens=20, temperature=0.9, seed=11)
        alone = eng.infer(lm, "c", p)
        results = [None] * 8

        def run(i):
            if i % 2:
                results[i] = "".join(t for t, _ in eng.stream_infer(lm, "c", p))
            else:
                results[i] = eng.infer(lm, "c", p).text

        ts = [threading.Thread(target=run, args=(i,)) for i in range(8)]
        for t in ts:
            t.start()
        for t in ts:
            t.join(60)
        assert results == [alone.text] * 8
    finally:
        backend.unload(lm)


def test_json_mode_on_device(backend, chat_gguf):
    """Unary Infer is grammar-constrained like the reference's json_object request (inference.rs:119-121): whatever a
    random-weight model wants to say, the text is a prefix of one JSON object, and a closed object ends the call."""
    import json
    from aios_b200.runtime.json_grammar import JsonPrefix, is_json_object_prefix
    lm = backend.load("j", chat_gguf, 256)
    try:
        eng = InferenceEngine()
        for seed in range(4):
            p = InferParams(prompt="Reply with JSON.", max_tokens=48, temperature=0.9, seed=seed, json_mode=True)
            r = eng.infer(lm, "j", p)
            raw = r.text.encode("utf-8")
            assert raw and is_json_object_prefix(raw), r.text
            st = JsonPrefix()
            st.feed(raw)
            if st.complete:
                assert isinstance(json.loads(r.text), dict)
            assert eng.infer(lm, "j", p).text == r.text  # same seed, same draw
        # the unconstrained path on the same model does not produce JSON (so the constraint did the work)
        free = eng.infer(lm, "j", InferParams(prompt="Reply with JSON.", max_tokens=48, temperature=0.9, seed=0))
        assert not is_json_object_prefix(free.text.encode("utf-8"))
    finally:
        backend.unload(lm)


def test_grpc_service_on_device(backend, chat_gguf):
    srv = RuntimeServer(backend, addr="127.0.0.1:0", model_dir=os.path.dirname(chat_gguf), health_interval_s=3600.0)
    srv.start(do_autoload=True)  # main.rs:65-132: picks the .gguf up by itself
    try:
        with grpc.insecure_channel(f"127.0.0.1:{srv.port}") as ch:
            stub = pb.AIRuntimeStub(ch)
            models = stub.ListModels(pb.Empty()).models
            assert [m.model_name for m in models] == ["tinyllama-1.1b-synthetic"] and models[0].status == "ready"
            again = stub.LoadModel(pb.LoadModelRequest(model_name="tinyllama-1.1b-synthetic", model_path=chat_gguf))
            assert again.port == models[0].port and again.loaded_at == models[0].loaded_at  # idempotent
            r = stub.Infer(pb.InferRequest(prompt="Hello, world!", intelligence_level="operational", max_tokens=16))
            assert r.model_used == "tinyllama-1.1b-synthetic" and r.tokens_used > 16 and r.latency_ms >= 0
            chunks = list(stub.StreamInfer(pb.InferRequest(model="tinyllama-1.1b-synthetic", prompt="Hi", max_tokens=12)))
            assert chunks[-1].done and not any(c.done for c in chunks[:-1])
            h = stub.HealthCheck(pb.Empty())
            assert h.healthy and h.details["loaded_models"] == "1"
            with pytest.raises(grpc.RpcError) as e:
                stub.LoadModel(pb.LoadModelRequest(model_name="broken", model_path=__file__))  # not a GGUF
            assert e.value.code() == grpc.StatusCode.INTERNAL
            assert any(m.status.startswith("error: ") for m in stub.ListModels(pb.Empty()).models)
            u = stub.UnloadModel(pb.UnloadModelRequest(model_name="tinyllama-1.1b-synthetic"))
            assert u.success and u.message == "Model 'tinyllama-1.1b-synthetic' unloaded"
            with pytest.raises(grpc.RpcError) as e:
                stub.Infer(pb.InferRequest(prompt="x"))
            assert e.value.code() == grpc.StatusCode.UNAVAILABLE
    finally:
        srv.server.stop(0).wait()
        srv._stop.set()


def test_unload_races_running_requests(backend, chat_gguf):
    """UnloadModel while Infer calls are in flight: the requests are cancelled, the scheduler retires them and the
    model really leaves HBM (the reference kills the child, which always frees it: model_manager.rs:285-345)."""
    from aios_b200.capi import make_sampling
    lm = backend.load("race", chat_gguf, 256)
    r0 = lm.replicas[0]
    toks = r0.vocab.tokenize(r0.vocab.chat_prompt("", "Hello, world!"), parse_special=True)
    reqs = [lm.submit(toks, 200, make_sampling(temperature=0.9, seed=i)) for i in range(6)]
    backend.unload(lm)  # cancels, drains, unloads — without the drain this raised AIOS_E_STATE and leaked the model
    assert not r0.model.h and not backend.alive(lm)
    for r in reqs:
        assert r.wait(1000)  # the request objects outlive the model and report that they are done
        r.free()
    with pytest.raises(RuntimeError, match="unloading"):
        r0.submit(toks, 4, make_sampling())


def test_openai_http_on_device(backend, chat_gguf):
    """The llama-server-compatible endpoint (api-gateway/src/openai.rs:148-165, runtime/src/inference.rs:100) on the
    real engine: non-streaming and SSE answers carry the same text for a pinned seed, JSON mode answers JSON."""
    import urllib.request
    from aios_b200.runtime.json_grammar import is_json_object_prefix
    from aios_b200.runtime.openai_http import OpenAIServer
    lm = backend.load("http", chat_gguf, 256)
    srv = OpenAIServer(lm, "http", 0, InferenceEngine(stream_chunk_tokens=4)).start()
    url = f"http://127.0.0.1:{srv.port}"

    def post(body):
        req = urllib.request.Request(url + "/v1/chat/completions", data=json.dumps(body).encode(), method="POST",
                                     headers={"Content-Type": "application/json"})
        return urllib.request.urlopen(req, timeout=60)

    try:
        with urllib.request.urlopen(url + "/health", timeout=5) as r:
            assert json.load(r) == {"status": "ok"}
        body = {"messages": [{"role": "system", "content": "Be brief."}, {"role": "user", "content": "Hello, world!"}],
                "max_tokens": 24, "temperature": 0.8, "seed": 5, "stream": False}
        with post(body) as r:
            obj = json.load(r)
        text = obj["choices"][0]["message"]["content"]
        u = obj["usage"]
        assert u["total_tokens"] == u["prompt_tokens"] + u["completion_tokens"] and 0 < u["completion_tokens"] <= 24
        # the same request over the engine API gives the same text (one code path under both front ends)
        direct = InferenceEngine().infer(lm, "http", InferParams(prompt="Hello, world!", system_prompt="Be brief.",
                                                                 max_tokens=24, temperature=0.8, seed=5))
        assert direct.text == text
        with post(dict(body, stream=True)) as r:
            assert r.headers["Content-Type"] == "text/event-stream"
            events = [ln[6:] for ln in r.read().decode("utf-8").split("\n") if ln.startswith("data: ")]
        assert events[-1] == "[DONE]"
        deltas = [json.loads(e)["choices"][0] for e in events[:-1]]
        assert "".join(d["delta"].get("content", "") for d in deltas) == text and deltas[-1]["finish_reason"] == "stop"
        # what InferenceEngine::infer really posts (inference.rs:114-122): JSON mode
        with post(dict(body, response_format={"type": "json_object"}, max_tokens=40)) as r:
            jtext = json.load(r)["choices"][0]["message"]["content"]
        assert jtext and is_json_object_prefix(jtext.encode("utf-8"))
    finally:
        srv.close()
        backend.unload(lm)
