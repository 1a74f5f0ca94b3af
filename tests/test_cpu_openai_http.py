"""CPU tests of the OpenAI-compatible HTTP endpoint (aios_b200/runtime/openai_http.py) with the scripted model of
test_cpu_runtime.py: the response shapes are the ones the reference's own client code parses
(runtime/src/inference.rs:37-67 structs, :161-171 fields read, :274-343 SSE loop; model_manager.rs:223 /health)."""
import json
import urllib.error
import urllib.request

import pytest

from aios_b200.runtime.inference import InferenceEngine
from aios_b200.runtime.openai_http import OpenAIServer
from aios_b200.runtime.service import RuntimeServer
from test_cpu_runtime import FakeBackend, FakeLoaded, vocab_gguf, spm_golden  # noqa: F401  (fixtures)


@pytest.fixture()
def http(vocab_gguf):  # noqa: F811
    lm = FakeLoaded("tinyllama", vocab_gguf, 512)
    srv = OpenAIServer(lm, "tinyllama", 0, InferenceEngine(stream_chunk_tokens=3)).start()
    yield f"http://127.0.0.1:{srv.port}", lm
    srv.close()


def _post(url, body):
    req = urllib.request.Request(url + "/v1/chat/completions", data=json.dumps(body).encode(), method="POST",
                                 headers={"Content-Type": "application/json"})
    return urllib.request.urlopen(req, timeout=10)


def test_health_and_models(http):
    url, _ = http
    with urllib.request.urlopen(url + "/health", timeout=5) as r:
        assert r.status == 200 and json.load(r) == {"status": "ok"}
    with urllib.request.urlopen(url + "/v1/models", timeout=5) as r:
        assert json.load(r)["data"][0]["id"] == "tinyllama"
    with pytest.raises(urllib.error.HTTPError) as e:
        urllib.request.urlopen(url + "/nope", timeout=5)
    assert e.value.code == 404


def test_chat_completion_shape_the_reference_parses(http):
    url, lm = http
    # exactly what InferenceEngine::infer posts (inference.rs:114-122)
    body = {"messages": [{"role": "system", "content": "Be brief."}, {"role": "user", "content": "Hello, world!"}],
            "max_tokens": 512, "temperature": 0.7, "stream": False, "response_format": {"type": "json_object"}}
    with _post(url, body) as r:
        assert r.status == 200
        obj = json.load(r)
    # ...and what it reads back (:161-171)
    text = obj["choices"][0]["message"]["content"]
    assert obj["choices"][0]["message"]["role"] == "assistant" and text.endswith("assistant\n")
    u = obj["usage"]
    assert u["total_tokens"] == u["prompt_tokens"] + u["completion_tokens"] and u["completion_tokens"] == 12
    assert obj["choices"][0]["finish_reason"] == "stop" and obj["object"] == "chat.completion"
    body["max_tokens"] = 5
    with _post(url, body) as r:
        obj = json.load(r)
    assert obj["usage"]["completion_tokens"] == 5 and obj["choices"][0]["finish_reason"] == "length"


def test_chat_completion_stream_parsed_like_the_reference(http):
    url, _ = http
    body = {"messages": [{"role": "user", "content": "Hello, world!"}], "max_tokens": 64, "temperature": 0.2, "stream": True}
    with _post(url, body) as r:
        assert r.headers["Content-Type"] == "text/event-stream"
        raw = r.read().decode("utf-8")
    # the reference's loop (inference.rs:274-343): lines, "data: " prefix, [DONE], delta.content, finish_reason
    text, done_seen, finished = "", False, False
    for line in raw.splitlines():
        line = line.strip()
        if not line:
            continue
        data = line[len("data: "):].strip() if line.startswith("data: ") else line
        if data == "[DONE]":
            done_seen = True
            break
        chunk = json.loads(data)
        text += chunk["choices"][0].get("delta", {}).get("content") or ""
        if chunk["choices"][0].get("finish_reason") is not None:
            finished = True
    assert done_seen and finished
    with _post(url, dict(body, stream=False)) as r:
        assert json.load(r)["choices"][0]["message"]["content"] == text


def test_bad_requests(http):
    url, _ = http
    for body in ({"messages": []}, {"messages": [{"role": "system", "content": "x"}]}, {}):
        with pytest.raises(urllib.error.HTTPError) as e:
            _post(url, body)
        assert e.value.code == 400 and "error" in json.load(e.value)
    with pytest.raises(urllib.error.HTTPError) as e:  # prompt longer than the 512-token context
        _post(url, {"messages": [{"role": "user", "content": "word " * 2000}]})
    assert e.value.code == 400


def test_runtime_server_serves_models_on_their_ports(vocab_gguf, tmp_path):  # noqa: F811
    """With openai_http=True, ModelStatus.port is a real port again ('return the OpenAI-compat port')."""
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        free_port = s.getsockname()[1]
    srv = RuntimeServer(FakeBackend(), addr="127.0.0.1:0", model_dir=str(tmp_path), health_interval_s=3600.0, openai_http=True)
    srv.start()
    try:
        from aios_b200.runtime.model_manager import LoadRequest
        st = srv.manager.load_model(LoadRequest("tinyllama", vocab_gguf, port=free_port))
        assert st.status == "ready" and st.port == free_port
        with urllib.request.urlopen(f"http://127.0.0.1:{free_port}/health", timeout=5) as r:
            assert r.status == 200
        # a second model asking for the same port fails in-band, like a llama-server that cannot bind
        with pytest.raises(RuntimeError):
            srv.manager.load_model(LoadRequest("clash", vocab_gguf, port=free_port))
        assert [m for m in srv.manager.list_models() if m.model_name == "clash"][0].status.startswith("error: ")
        srv.manager.unload_model("tinyllama")
        with pytest.raises(Exception):
            urllib.request.urlopen(f"http://127.0.0.1:{free_port}/health", timeout=2)
    finally:
        srv.stop(0)


def test_whole_conversation_goes_through_the_chat_template(http):
    """A multi-turn message list is rendered by the model's own tokenizer.chat_template (aios_chat_apply_template), not
    cut down to (system, last user turn): the prompt the scripted model is fed has the length of the full rendering."""
    url, lm = http
    body = {"messages": [{"role": "user", "content": "a"}, {"role": "assistant", "content": "beta"},
                         {"role": "user", "content": "c"}], "max_tokens": 64, "temperature": 0.2}
    with _post(url, body) as r:
        obj = json.load(r)
    want = lm.vocab.apply_chat_template([("user", "a"), ("assistant", "beta"), ("user", "c")])
    assert obj["usage"]["prompt_tokens"] == lm.vocab.tokenize(want, add_special=True, parse_special=True).size
    assert obj["usage"]["prompt_tokens"] > lm.vocab.tokenize(lm.vocab.chat_prompt("", "c"), parse_special=True).size
    assert want == ("<|im_start|>user\na<|im_end|>\n<|im_start|>assistant\nbeta<|im_end|>\n<|im_start|>user\nc<|im_end|>\n"
                    "<|im_start|>assistant\n")
