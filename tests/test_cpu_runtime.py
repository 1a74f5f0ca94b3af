"""CPU tests of the host side of the drop-in boundary: wire messages, tokenizer, ModelManager / gRPC service logic.

They restate the reference's own unit tests for this path (runtime/src/model_manager.rs:556-712,
grpc_service.rs:239-335, inference.rs:382-447) against the mirror, and drive the real gRPC server over a socket
with a fake backend (no GPU: the engine is replaced by a scripted token source; tokenizer and vocabulary are the
real host-only C code).
"""
import dataclasses
import json
import os
import sys
import threading

import grpc
import numpy as np
import pytest

from aios_b200 import capi, gguf_synth
from aios_b200.runtime import proto as pb
from aios_b200.runtime.inference import (InferenceEngine, InferParams, build_messages, effective_max_tokens,
                                         effective_temperature)
from aios_b200.runtime.model_manager import (BASE_PORT, LoadRequest, ManagedModel, ModelManager, ModelState,
                                             model_to_status, now_epoch_ms)
from aios_b200.runtime.service import RuntimeServer, autoload_plan

HERE = os.path.dirname(os.path.abspath(__file__))
REF_PY = "/root/reference/agent-core/python"


# ------------------------------------------------------------------ fixtures
@pytest.fixture(scope="module")
def spm_golden():
    with open(os.path.join(HERE, "golden", "spm_vocab.json"), encoding="utf-8") as f:
        return json.load(f)


@pytest.fixture(scope="module")
def vocab_gguf(tmp_path_factory, spm_golden):
    """A tiny GGUF whose vocabulary is the committed SentencePiece one (+ chatml control tokens)."""
    d = tmp_path_factory.mktemp("vocab")
    vocab = [tuple(v) for v in spm_golden["vocab"]] + [("<|im_start|>", 0.0, 3), ("<|im_end|>", 0.0, 3)]
    arch = dataclasses.replace(gguf_synth.ARCHS["tiny"], n_vocab=len(vocab) + 6)
    path = str(d / "tinyllama-1.1b-chat.gguf")
    gguf_synth.write_synthetic_gguf(path, arch, vocab=vocab,
                                    chat_template=gguf_synth.CHATML_TEMPLATE)
    return path


class FakeSeq:
    """Scripted engine: 'generates' the prompt's own non-special tokens back, then EOS."""

    def __init__(self, lm):
        self.lm, self.queue, self.n_past = lm, [], 0

    def reset(self):
        self.queue, self.n_past = [], 0

    def prefill(self, tokens):
        special = {1, 2, self.lm.im_start, self.lm.im_end}
        self.queue = [int(t) for t in tokens if int(t) not in special][-12:] + [2, 7, 7, 7]
        self.n_past = len(tokens)

    def decode(self, max_tokens, sampling=None):
        self.lm.samplings.append((sampling.temperature, sampling.top_k))
        out, self.queue = self.queue[:max_tokens], self.queue[max_tokens:]
        self.n_past += len(out)
        return np.asarray(out, dtype=np.int32)


class FakeLoaded:
    def __init__(self, name, path, ctx):
        self.name, self.n_ctx = name, ctx
        self.vocab = capi.Vocab.open(path)
        self.im_start = int(self.vocab.tokenize("<|im_start|>", add_special=False, parse_special=True)[0])
        self.im_end = int(self.vocab.tokenize("<|im_end|>", add_special=False, parse_special=True)[0])
        self.samplings = []
        self.seq = FakeSeq(self)
        self.lock = threading.Lock()
        self.closed = False


class FakeBackend:
    def __init__(self):
        self.loaded, self.unloaded, self.dead = [], [], set()

    def load(self, name, path, ctx):
        if not os.path.isfile(path):
            raise FileNotFoundError(f"model file {path} not found")
        lm = FakeLoaded(name, path, ctx)
        self.loaded.append(lm)
        return lm

    def unload(self, handle):
        handle.closed = True
        self.unloaded.append(handle)

    def alive(self, handle):
        return handle.name not in self.dead


def _ready(name, port=8080):
    return ManagedModel(name=name, path="/tmp/x.gguf", port=port, status=ModelState(ModelState.READY), loaded_at=1000,
                        last_used=2000, context_length=4096, handle=object())


# ------------------------------------------------------------------ wire format
def test_wire_messages_match_reference_stubs():
    """Descriptors built at import time serialise exactly like the reference's protoc output."""
    if not os.path.isdir(REF_PY):
        pytest.skip("reference stubs not present on this machine")
    sys.path.insert(0, REF_PY)
    try:
        from aios_agent.proto import common_pb2 as C, runtime_pb2 as R
    finally:
        sys.path.remove(REF_PY)
    from google.protobuf import descriptor_pb2 as dp

    def norm(desc):
        d = dp.DescriptorProto()
        desc.CopyToProto(d)
        return d.SerializeToString()

    for n in ("LoadModelRequest", "UnloadModelRequest", "ModelStatus", "ModelList", "InferRequest", "InferResponse",
              "InferChunk"):
        assert norm(getattr(R, n).DESCRIPTOR) == norm(getattr(pb, n).DESCRIPTOR), n
    for n in ("Empty", "Status", "HealthStatus"):
        assert norm(getattr(C, n).DESCRIPTOR) == norm(getattr(pb, n).DESCRIPTOR), n
    a = R.InferRequest(model="m", prompt="p", system_prompt="s", max_tokens=7, temperature=0.25,
                       intelligence_level="tactical", requesting_agent="a", task_id="t").SerializeToString()
    assert pb.InferRequest.FromString(a).SerializeToString() == a
    svc = R.DESCRIPTOR.services_by_name["AIRuntime"]
    assert svc.full_name == pb.SERVICE
    assert {m.name: (m.input_type.full_name, m.output_type.full_name, m.server_streaming) for m in svc.methods} == \
        {k: (v[0].DESCRIPTOR.full_name, v[1].DESCRIPTOR.full_name, v[2]) for k, v in pb.METHODS.items()}


def test_wire_known_bytes():
    """Field numbers / types pinned without the reference present (runtime.proto:15-62)."""
    assert pb.LoadModelRequest(model_name="a", port=3).SerializeToString() == b"\x0a\x01a\x30\x03"
    assert pb.ModelStatus(model_name="m", status="ready", port=8080, request_count=2).SerializeToString() == \
        b"\x0a\x01m\x12\x05ready\x18\x90\x3f\x30\x02"
    assert pb.InferChunk(done=True).SerializeToString() == b"\x10\x01"
    assert pb.InferResponse(text="x", tokens_used=5, latency_ms=9, model_used="m").SerializeToString() == \
        b"\x0a\x01x\x10\x05\x18\x09\x22\x01m"
    h = pb.HealthStatus(healthy=True, service="aios-runtime", details={"k": "v"})
    assert pb.HealthStatus.FromString(h.SerializeToString()).details["k"] == "v"


# ------------------------------------------------------------------ tokenizer (host-only C code)
def test_tokenizer_matches_sentencepiece(vocab_gguf, spm_golden):
    v = capi.Vocab.open(vocab_gguf)
    assert v.info.model == b"llama" and v.info.bos == 1 and v.info.eos == 2
    for c in spm_golden["cases"]:
        ids = v.tokenize(c["text"], add_special=False).tolist()
        assert ids == c["ids"], c["text"]
        assert v.detokenize(ids, strip_space=True).decode("utf-8") == c["decoded"], c["text"]
    assert v.tokenize("Hello", add_special=True)[0] == 1
    for i in (0, 1, 2, 3, 258, 300):
        piece, ty, score = v.piece(i)
        assert piece.decode("utf-8") == spm_golden["vocab"][i][0] and ty == spm_golden["vocab"][i][2]
        assert score == pytest.approx(spm_golden["vocab"][i][1])
    v.close()


def test_tokenizer_special_tokens_and_chat_prompt(vocab_gguf):
    v = capi.Vocab.open(vocab_gguf)
    assert v.info.chat_format == 0  # chatml recognised from the template's markers
    text = v.chat_prompt("You are helpful.", "Hello!")
    assert text == "<|im_start|>system\nYou are helpful.<|im_end|>\n<|im_start|>user\nHello!<|im_end|>\n<|im_start|>assistant\n"
    assert v.chat_prompt("", "Hello!") == "<|im_start|>user\nHello!<|im_end|>\n<|im_start|>assistant\n"
    ims = int(v.tokenize("<|im_start|>", add_special=False, parse_special=True)[0])
    ids = v.tokenize(text, add_special=True, parse_special=True)
    assert ids[0] == 1 and int((ids == ims).sum()) == 3
    # without parse_special the marker is ordinary text: several tokens, none of them the control id
    plain = v.tokenize("<|im_start|>", add_special=False, parse_special=False)
    assert plain.size > 1 and ims not in plain.tolist()
    assert v.is_eog(2) and v.is_eog(ims + 1) and not v.is_eog(ims) and not v.is_eog(300)
    # control tokens are skipped unless asked for
    assert v.detokenize([ims, 300]) == v.detokenize([300])
    assert v.detokenize([ims], render_special=True) == b"<|im_start|>"
    with pytest.raises(capi.AiosError) as e:
        v.detokenize([10 ** 6])
    assert e.value.code == capi.E_INVAL
    v.close()


def test_vocab_open_errors(tmp_path):
    with pytest.raises(capi.AiosError) as e:
        capi.Vocab.open(str(tmp_path / "missing.gguf"))
    assert e.value.code == capi.E_IO
    p = tmp_path / "wordpiece.gguf"
    gguf_synth.write_synthetic_gguf(str(p), "tiny", tokenizer_model="bert")
    v = capi.Vocab.open(str(p))
    with pytest.raises(capi.AiosError) as e:
        v.tokenize("hello")
    assert e.value.code == capi.E_UNSUPPORTED


# ------------------------------------------------------------------ byte-level BPE ("gpt2" vocabularies: Llama-3)
@pytest.fixture(scope="module")
def bpe_golden():
    with open(os.path.join(HERE, "golden", "bpe_vocab.json"), encoding="utf-8") as f:
        return json.load(f)


def _bpe_gguf(tmp_path, g, pre, template=None):
    n_sp = g["n_special"]
    vocab = [(p, 0.0, 3 if i < n_sp else 1) for i, p in enumerate(g["vocab"])]
    arch = dataclasses.replace(gguf_synth.ARCHS["tiny"], n_vocab=len(vocab) + 6)
    path = str(tmp_path / f"bpe-{pre}.gguf")
    gguf_synth.write_synthetic_gguf(path, arch, vocab=vocab, tokenizer_model="gpt2", merges=g["merges"],
                                    tokenizer_pre=pre, bos_id=0, eos_id=1, chat_template=template)
    return path


@pytest.mark.parametrize("pre,cases", [("llama-bpe", "cases_llama3"), ("gpt-2", "cases_gpt2")])
def test_bpe_ids_match_hf_tokenizers(bpe_golden, tmp_path, pre, cases):
    """Every golden case (contractions in both cases, digit runs, blank / newline runs, CRLF, Unicode letters, numbers
    and spaces, emoji, text that no merge covers) encodes to the ids HF `tokenizers` produced, and decodes back."""
    v = capi.Vocab.open(_bpe_gguf(tmp_path, bpe_golden, pre))
    assert len(bpe_golden[cases]) >= 30
    for text, want in bpe_golden[cases]:
        got = v.tokenize(text, add_special=False)
        assert got.tolist() == want, (text, got.tolist(), want)
        assert v.detokenize(got).decode("utf-8") == text
    # BOS: Llama-3 vocabularies get one, plain GPT-2 ones do not; special strings become their ids when asked to
    ids = v.tokenize("Hello", add_special=True)
    assert (ids[0] == 0) == (pre == "llama-bpe")
    sp = v.tokenize("<|eot_id|>Hello<|end_of_text|>", add_special=False, parse_special=True)
    assert sp[0] == 2 and sp[-1] == 1 and v.is_eog(2) and v.is_eog(1)
    assert v.detokenize(sp, render_special=True).decode() == "<|eot_id|>Hello<|end_of_text|>"
    assert v.detokenize(sp).decode() == "Hello"
    # bytes that are not UTF-8 still round-trip (byte-level vocabulary)
    raw = b"ok \xff\xfe \xc3 end"
    assert v.detokenize(v.tokenize(raw, add_special=False)) == raw


def test_llama3_chat_prompt(bpe_golden, tmp_path):
    """Llama-3's chat format, picked by the template's marker string, rendered for {system, user}
    (runtime/src/inference.rs:363-376) and tokenised with the header markers as single ids."""
    v = capi.Vocab.open(_bpe_gguf(tmp_path, bpe_golden, "llama-bpe",
                                  template="{% for m in messages %}<|start_header_id|>{{ m['role'] }}<|end_header_id|>…"))
    text = v.chat_prompt("Be brief.", "Hi")
    assert text == ("<|start_header_id|>system<|end_header_id|>\n\nBe brief.<|eot_id|>"
                    "<|start_header_id|>user<|end_header_id|>\n\nHi<|eot_id|>"
                    "<|start_header_id|>assistant<|end_header_id|>\n\n")
    ids = v.tokenize(text, add_special=True, parse_special=True).tolist()
    assert ids[0] == 0 and ids[1] == 3 and ids.count(3) == 3 and ids.count(4) == 3 and ids.count(2) == 2


# ------------------------------------------------------------------ tokenizer.chat_template evaluated (jinja_lite)
def test_chat_templates_match_jinja2(spm_golden, tmp_path):  # noqa: F811
    """Every golden rendering (five real-world template shapes + two synthetic ones x five conversations x with/without
    the generation prompt) matches what the jinja2 package produced; where jinja2 raised (raise_exception in the
    template), aios_chat_apply_template answers AIOS_E_UNSUPPORTED."""
    with open(os.path.join(HERE, "golden", "chat_templates.json"), encoding="utf-8") as f:
        g = json.load(f)
    vocab = [tuple(v) for v in spm_golden["vocab"]]
    arch = dataclasses.replace(gguf_synth.ARCHS["tiny"], n_vocab=len(vocab) + 6)
    n_ok = n_raise = 0
    for tname, src in g["templates"].items():
        path = str(tmp_path / f"tmpl-{tname}.gguf")
        # add_bos_token is on (SPM default): give the template an empty-looking BOS so that nothing is stripped here
        gguf_synth.write_synthetic_gguf(path, arch, vocab=vocab, chat_template=src)
        v = capi.Vocab.open(path)
        bos, eos = v.detokenize([1], render_special=True).decode(), v.detokenize([2], render_special=True).decode()
        assert (bos, eos) == ("<s>", "</s>")
        for case in (c for c in g["cases"] if c["template"] == tname):
            msgs = g["conversations"][case["conversation"]]
            want = case["expected"]
            if want is None:
                with pytest.raises(capi.AiosError) as e:
                    v.apply_chat_template(msgs, case["add_generation_prompt"])
                assert e.value.code == capi.E_UNSUPPORTED
                n_raise += 1
                continue
            got = v.apply_chat_template(msgs, case["add_generation_prompt"])
            if want.startswith("<s>"):
                want = want[3:]  # the tokenizer adds BOS itself: a leading BOS string from the template is dropped
            assert got == want, (tname, case["conversation"], case["add_generation_prompt"], got, want)
            n_ok += 1
        v.close()
    assert n_ok >= 60 and n_raise >= 8


def test_chat_prompt_prefers_the_files_template(spm_golden, tmp_path):  # noqa: F811
    vocab = [tuple(v) for v in spm_golden["vocab"]]
    arch = dataclasses.replace(gguf_synth.ARCHS["tiny"], n_vocab=len(vocab) + 6)

    def vocab_with(template):
        path = str(tmp_path / f"t{abs(hash(template))}.gguf")
        gguf_synth.write_synthetic_gguf(path, arch, vocab=vocab, chat_template=template)
        return capi.Vocab.open(path)

    # a template nobody's marker list knows: rendered as written
    v = vocab_with("{% for m in messages %}<{{ m.role }}>{{ m.content }}</{{ m.role }}>{% endfor %}<bot>")
    assert v.chat_prompt("S", "U") == "<system>S</system><user>U</user><bot>"
    assert v.chat_prompt("", "U") == "<user>U</user><bot>"
    # Mistral's template refuses a system role: the recognised [INST] format takes over (system text inside the turn)
    with open(os.path.join(HERE, "golden", "chat_templates.json"), encoding="utf-8") as f:
        mistral = json.load(f)["templates"]["mistral-v0.2"]
    v = vocab_with(mistral)
    assert v.chat_prompt("", "U") == "[INST] U [/INST]"
    assert v.chat_prompt("S", "U") == "[INST] S\n\nU [/INST]"
    # outside the subset (a macro): the marker-string format again, and the direct call says why
    v = vocab_with("{% macro f(x) %}{{ x }}{% endmacro %}<|im_start|>{{ f(messages[0].content) }}")
    assert v.chat_prompt("", "U") == "<|im_start|>user\nU<|im_end|>\n<|im_start|>assistant\n"
    with pytest.raises(capi.AiosError, match="macro"):
        v.apply_chat_template([("user", "U")])
    # no template at all
    path = str(tmp_path / "none.gguf")
    gguf_synth.write_synthetic_gguf(path, arch, vocab=vocab)
    v = capi.Vocab.open(path)
    with pytest.raises(capi.AiosError) as e:
        v.apply_chat_template([("user", "U")])
    assert e.value.code == capi.E_STATE


# ------------------------------------------------------------------ model_manager.rs:556-712 restated
def test_now_epoch_ms_is_positive():
    assert now_epoch_ms() > 0


def test_model_state_display():
    assert str(ModelState(ModelState.LOADING)) == "loading"
    assert str(ModelState(ModelState.READY)) == "ready"
    assert str(ModelState(ModelState.ERROR, "oops")) == "error: oops"
    assert str(ModelState(ModelState.UNLOADING)) == "unloading"


def test_select_model_empty():
    mgr = ModelManager(FakeBackend(), "/nonexistent")
    for level in ("reactive", "operational", "tactical", "strategic"):
        assert mgr.select_model_for_level(level) is None


def test_first_ready_from_partial_match():
    mgr = ModelManager(FakeBackend(), "/nonexistent")
    mgr.models["DeepSeek-R1-Distill-Qwen-8B-Q4_K_M"] = _ready("DeepSeek-R1-Distill-Qwen-8B-Q4_K_M", 8082)
    assert "DeepSeek-R1" in mgr.first_ready_from(["DeepSeek-R1-Distill-Qwen-8B"])


def test_tactical_prefers_deepseek_over_mistral():
    mgr = ModelManager(FakeBackend(), "/nonexistent")
    mgr.models["mistral-7b"] = _ready("mistral-7b")
    mgr.models["DeepSeek-R1-Distill-Qwen-8B-Q4_K_M"] = _ready("DeepSeek-R1-Distill-Qwen-8B-Q4_K_M", 8082)
    assert "DeepSeek" in mgr.select_model_for_level("tactical")
    assert mgr.select_model_for_level("operational") == "DeepSeek-R1-Distill-Qwen-8B-Q4_K_M"
    mgr.models["DeepSeek-R1-Distill-Qwen-8B-Q4_K_M"].status = ModelState(ModelState.ERROR, "x")
    assert mgr.select_model_for_level("tactical") == "mistral-7b"
    assert mgr.select_model_for_level("made-up-level") == "mistral-7b"


def test_allocate_port():
    mgr = ModelManager(FakeBackend(), "/nonexistent")
    assert mgr.allocate_port(0) == BASE_PORT
    assert mgr.allocate_port(0) == BASE_PORT + 1
    mgr = ModelManager(FakeBackend(), "/nonexistent")
    assert mgr.allocate_port(9999) == 9999
    assert mgr.allocate_port(0) == BASE_PORT  # the auto-counter did not advance


def test_list_models_empty_and_get_missing():
    mgr = ModelManager(FakeBackend(), "/nonexistent")
    assert mgr.list_models() == []
    assert mgr.get_model("nonexistent") is None
    assert mgr.model_port("nonexistent") is None


def test_model_to_status_conversion():
    m = _ready("test-model")
    m.request_count = 42
    s = model_to_status(m)
    assert (s.model_name, s.status, s.port, s.loaded_at, s.last_used, s.request_count) == \
        ("test-model", "ready", 8080, 1000, 2000, 42)


def test_load_unload_lifecycle(vocab_gguf, tmp_path):
    be = FakeBackend()
    mgr = ModelManager(be, os.path.dirname(vocab_gguf), auto_restart=False)  # the reference's own behaviour
    t0 = now_epoch_ms()
    st = mgr.load_model(LoadRequest(model_name=os.path.basename(vocab_gguf)))  # empty path -> model_dir/name
    assert st.status == "ready" and st.port == BASE_PORT and st.request_count == 0
    assert t0 <= st.loaded_at <= now_epoch_ms() and st.last_used == st.loaded_at
    assert be.loaded[0].n_ctx == 2048  # context default (model_manager.rs:167-171)
    # idempotent: no second load, same record (model_manager.rs:152-157)
    again = mgr.load_model(LoadRequest(model_name=os.path.basename(vocab_gguf), context_length=512))
    assert again.port == st.port and len(be.loaded) == 1
    # model_port: ready check + last_used stamp + request_count (model_manager.rs:367-376)
    assert mgr.model_port(st.model_name) == BASE_PORT
    assert mgr.list_models()[0].request_count == 1
    # explicit path, explicit port, explicit context
    st2 = mgr.load_model(LoadRequest("second", vocab_gguf, context_length=256, port=9000))
    assert st2.port == 9000 and be.loaded[1].n_ctx == 256
    assert [m.model_name for m in mgr.list_models()] == [st.model_name, "second"]
    # failed load: recorded in-band as "error: ..." AND raised (model_manager.rs:237-252)
    with pytest.raises(RuntimeError):
        mgr.load_model(LoadRequest("ghost", str(tmp_path / "ghost.gguf")))
    ghost = [m for m in mgr.list_models() if m.model_name == "ghost"][0]
    assert ghost.status.startswith("error: ") and mgr.model_port("ghost") is None
    # an errored model can be loaded again (only ready|loading short-circuit)
    st3 = mgr.load_model(LoadRequest("ghost", vocab_gguf))
    assert st3.status == "ready"
    # health task flips ready <-> error
    be.dead.add("second")
    mgr.health_check_all()
    assert not mgr.is_model_ready("second") and mgr.is_model_ready("ghost")
    be.dead.clear()
    mgr.health_check_all()
    assert mgr.is_model_ready("second")
    # unload
    mgr.unload_model("second")
    assert be.unloaded[0].name == "second" and be.unloaded[0].closed
    with pytest.raises(KeyError) as e:
        mgr.unload_model("second")
    assert e.value.args[0] == "Model 'second' not found"


def test_autoload_plan_sizes(tmp_path):
    """main.rs:65-132: *.gguf only, name = stem, (ctx, threads) by file size."""
    for name, size in (("small.gguf", 10), ("mid.gguf", 2_000_000_001), ("big.gguf", 4_000_000_001),
                       ("huge.gguf", 8_000_000_001), ("notes.txt", 5)):
        with open(tmp_path / name, "wb") as f:
            f.truncate(size)  # sparse
    plan = {r.model_name: r for r in autoload_plan(str(tmp_path))}
    assert set(plan) == {"small", "mid", "big", "huge"}
    assert (plan["small"].context_length, plan["small"].threads) == (2048, 2)
    assert (plan["mid"].context_length, plan["mid"].threads) == (4096, 3)
    assert (plan["big"].context_length, plan["big"].threads) == (4096, 4)
    assert (plan["huge"].context_length, plan["huge"].threads) == (8192, 4)
    assert plan["small"].model_path == str(tmp_path / "small.gguf") and plan["small"].port == 0
    assert autoload_plan(str(tmp_path / "nope")) == []


# ------------------------------------------------------------------ inference.rs:382-447 restated
def test_build_messages():
    m = build_messages("You are helpful.", "Hello!")
    assert m == [{"role": "system", "content": "You are helpful."}, {"role": "user", "content": "Hello!"}]
    m = build_messages("", "Hello!")
    assert m == [{"role": "user", "content": "Hello!"}]


def test_request_defaults():
    assert effective_max_tokens(0) == 512 and effective_max_tokens(-3) == 512 and effective_max_tokens(10) == 10
    assert effective_temperature(0.0) == pytest.approx(0.7) and effective_temperature(0.5) == 0.5


def test_infer_engine_with_scripted_model(vocab_gguf):
    lm = FakeLoaded("m", vocab_gguf, 2048)
    eng = InferenceEngine(stream_chunk_tokens=3)
    r = eng.infer(lm, "m", InferParams(prompt="Hello, world!", system_prompt="Be brief."))
    n_prompt = lm.vocab.tokenize(lm.vocab.chat_prompt("Be brief.", "Hello, world!"), parse_special=True).size
    assert r.model_used == "m" and r.latency_ms >= 0
    assert r.text.endswith("assistant\n") and "\x00" not in r.text  # the scripted echo of the prompt's tail
    assert r.tokens_used == n_prompt + 12  # prompt + completion, EOS and what follows not counted
    assert lm.samplings[0] == (pytest.approx(0.7), 40)  # temperature default + llama-server sampler defaults
    # streaming yields the same text in pieces and ends with done=True exactly once
    chunks = list(eng.stream_infer(lm, "m", InferParams(prompt="Hello, world!", system_prompt="Be brief.", temperature=0.2)))
    assert chunks[-1] == ("", True) and all(not d for _, d in chunks[:-1]) and len(chunks) >= 4
    assert "".join(t for t, _ in chunks) == r.text
    assert lm.samplings[-1][0] == pytest.approx(0.2)
    # max_tokens caps the completion
    r = eng.infer(lm, "m", InferParams(prompt="Hello, world!", max_tokens=5))
    assert r.tokens_used == lm.vocab.tokenize(lm.vocab.chat_prompt("", "Hello, world!"), parse_special=True).size + 5
    # prompt longer than the context: an error, like llama-server's 400
    small = FakeLoaded("s", vocab_gguf, 16)
    with pytest.raises(RuntimeError, match="exceeds the available context"):
        eng.infer(small, "s", InferParams(prompt="Hello, world! " * 20))


# ------------------------------------------------------------------ grpc_service.rs:239-335 restated, over a real socket
@pytest.fixture()
def server(vocab_gguf, tmp_path):
    srv = RuntimeServer(FakeBackend(), addr="127.0.0.1:0", model_dir=str(tmp_path), health_interval_s=3600.0)
    srv.start(do_autoload=True)  # empty directory: nothing to load
    ch = grpc.insecure_channel(f"127.0.0.1:{srv.port}")
    yield srv, pb.AIRuntimeStub(ch)
    ch.close()
    srv.stop(0)


def _req(**kw):
    base = dict(model="", prompt="hello", system_prompt="", max_tokens=10, temperature=0.5, intelligence_level="",
                requesting_agent="test", task_id="t1")
    base.update(kw)
    return pb.InferRequest(**base)


def test_grpc_health_and_list_empty(server):
    _, stub = server
    h = stub.HealthCheck(pb.Empty())
    assert h.healthy and h.service == "aios-runtime" and h.details["loaded_models"] == "0"
    assert h.details["total_models"] == "0" and h.message.startswith("0/0 models loaded, uptime ")
    assert list(stub.ListModels(pb.Empty()).models) == []


def test_grpc_error_codes_without_models(server):
    _, stub = server
    with pytest.raises(grpc.RpcError) as e:
        stub.Infer(_req())
    assert e.value.code() == grpc.StatusCode.UNAVAILABLE
    with pytest.raises(grpc.RpcError) as e:
        stub.Infer(_req(intelligence_level="reactive"))
    assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT
    with pytest.raises(grpc.RpcError) as e:
        stub.Infer(_req(intelligence_level="strategic"))
    assert e.value.code() == grpc.StatusCode.FAILED_PRECONDITION
    with pytest.raises(grpc.RpcError) as e:
        list(stub.StreamInfer(_req(intelligence_level="reactive")))
    assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT
    with pytest.raises(grpc.RpcError) as e:
        stub.UnloadModel(pb.UnloadModelRequest(model_name="ghost"))
    assert e.value.code() == grpc.StatusCode.INTERNAL
    assert e.value.details() == "Failed to unload model: Model 'ghost' not found"
    with pytest.raises(grpc.RpcError) as e:
        stub.LoadModel(pb.LoadModelRequest(model_name="ghost.gguf"))
    assert e.value.code() == grpc.StatusCode.INTERNAL and e.value.details().startswith("Failed to load model: ")
    # ...and the failed load is visible in-band
    assert stub.ListModels(pb.Empty()).models[0].status.startswith("error: ")


def test_grpc_full_cycle(server, vocab_gguf):
    srv, stub = server
    st = stub.LoadModel(pb.LoadModelRequest(model_name="tinyllama-1.1b-chat", model_path=vocab_gguf, context_length=1024))
    assert st.status == "ready" and st.port == BASE_PORT and st.loaded_at > 0
    h = stub.HealthCheck(pb.Empty())
    assert h.details["loaded_models"] == "1" and h.details["model:tinyllama-1.1b-chat"] == "status=ready port=8080"
    # explicit model name
    r = stub.Infer(_req(model="tinyllama-1.1b-chat", prompt="Hello, world!"))
    assert r.model_used == "tinyllama-1.1b-chat" and r.tokens_used > 0 and r.text
    # level routing: "operational" prefers a name containing tinyllama-1.1b
    assert stub.Infer(_req(intelligence_level="operational")).model_used == "tinyllama-1.1b-chat"
    # strategic falls back to None -> FailedPrecondition (no Qwen/DeepSeek/mistral loaded)
    with pytest.raises(grpc.RpcError) as e:
        stub.Infer(_req(intelligence_level="strategic"))
    assert e.value.code() == grpc.StatusCode.FAILED_PRECONDITION
    # unknown model name -> last resort: any ready model
    assert stub.Infer(_req(model="nope")).model_used == "tinyllama-1.1b-chat"
    chunks = list(stub.StreamInfer(_req(model="tinyllama-1.1b-chat", prompt="Hello, world!")))
    assert chunks[-1].done and not any(c.done for c in chunks[:-1])
    assert "".join(c.text for c in chunks) == r.text
    m = stub.ListModels(pb.Empty()).models[0]
    assert m.request_count == 4 and m.last_used >= m.loaded_at
    u = stub.UnloadModel(pb.UnloadModelRequest(model_name="tinyllama-1.1b-chat"))
    assert u.success and u.message == "Model 'tinyllama-1.1b-chat' unloaded"
    assert list(stub.ListModels(pb.Empty()).models) == []
    assert srv.manager.backend.unloaded[0].closed


def test_reference_client_stub_talks_to_the_server(server, vocab_gguf):
    """The reference's own generated client (agent-core/python/aios_agent/proto) against this server."""
    if not os.path.isdir(REF_PY):
        pytest.skip("reference stubs not present on this machine")
    sys.path.insert(0, REF_PY)
    try:
        from aios_agent.proto import common_pb2, runtime_pb2, runtime_pb2_grpc
    finally:
        sys.path.remove(REF_PY)
    srv, _ = server
    with grpc.insecure_channel(f"127.0.0.1:{srv.port}") as ch:
        stub = runtime_pb2_grpc.AIRuntimeStub(ch)
        st = stub.LoadModel(runtime_pb2.LoadModelRequest(model_name="mistral-7b-test", model_path=vocab_gguf))
        assert st.status == "ready"
        r = stub.Infer(runtime_pb2.InferRequest(prompt="Hello, world!", intelligence_level="tactical", max_tokens=8))
        assert r.model_used == "mistral-7b-test" and r.tokens_used > 8
        assert stub.HealthCheck(common_pb2.Empty()).details["loaded_models"] == "1"
        assert [c.done for c in stub.StreamInfer(runtime_pb2.InferRequest(prompt="hi", max_tokens=4))][-1] is True


# ------------------------------------------------------------------ lifecycle extras the reference's docs promise
def test_failed_unload_keeps_the_model_for_a_retry(vocab_gguf):
    """An unload that the backend refuses must not orphan the model (its weights would stay in HBM, unreachable):
    the entry stays, in its error state, and a second UnloadModel goes through."""
    class Stubborn(FakeBackend):
        refuse = True

        def unload(self, handle):
            if self.refuse:
                raise RuntimeError("model still has queued or running requests")
            super().unload(handle)

    be = Stubborn()
    mgr = ModelManager(be, model_dir=os.path.dirname(vocab_gguf))
    mgr.load_model(LoadRequest("tinyllama-1.1b-chat", model_path=vocab_gguf))
    with pytest.raises(RuntimeError, match="queued or running"):
        mgr.unload_model("tinyllama-1.1b-chat")
    st = mgr.list_models()[0]
    assert st.status.startswith("error: unload failed") and mgr.models["tinyllama-1.1b-chat"].handle is not None
    assert mgr.model_port("tinyllama-1.1b-chat") is None  # not routable meanwhile
    be.refuse = False
    mgr.unload_model("tinyllama-1.1b-chat")
    assert mgr.list_models() == [] and be.unloaded[0].closed


def test_idle_unload_spares_the_operational_model(vocab_gguf):
    """config/default-config.toml:26 `idle_unload_minutes`, docs/phases/04-AI-RUNTIME.md:208: tactical models leave
    after the idle timeout; the operational model (name contains "tinyllama") stays; a busy model stays."""
    be = FakeBackend()
    mgr = ModelManager(be, model_dir=os.path.dirname(vocab_gguf), idle_unload_minutes=5)
    for name in ("tinyllama-1.1b-chat", "mistral-7b-instruct", "qwen3-14b"):
        mgr.load_model(LoadRequest(name, model_path=vocab_gguf))
    old = now_epoch_ms() - 6 * 60_000
    for m in mgr.models.values():
        m.last_used = old
    mgr.models["qwen3-14b"].handle.load = lambda: {0: 1}  # one request in flight
    mgr.health_check_all()
    assert sorted(mgr.models) == ["qwen3-14b", "tinyllama-1.1b-chat"]
    assert [h.name for h in be.unloaded] == ["mistral-7b-instruct"]
    mgr.models["qwen3-14b"].handle.load = lambda: {0: 0}
    mgr.get_model("qwen3-14b")  # a use stamps last_used (model_manager.rs:357-365)
    mgr.health_check_all()
    assert "qwen3-14b" in mgr.models
    # off by default, as in the reference (which never unloads on its own)
    assert ModelManager(FakeBackend(), model_dir="/tmp").idle_unload_ms == 0


def test_dead_model_is_restarted(vocab_gguf):
    """docs/phases/04-AI-RUNTIME.md:210 / model_manager.rs:391-392: a model whose backend handle died is loaded again
    from the same file, keeps its port and request count, and goes back to ready; when the file is gone it stays in
    its error state after MAX_RESTARTS attempts."""
    from aios_b200.runtime.model_manager import MAX_RESTARTS
    be = FakeBackend()
    mgr = ModelManager(be, model_dir=os.path.dirname(vocab_gguf))
    st = mgr.load_model(LoadRequest("mistral-7b", model_path=vocab_gguf))
    mgr.model_port("mistral-7b")
    first = mgr.models["mistral-7b"].handle
    be.dead.add("mistral-7b")
    real_alive = be.alive
    be.alive = lambda h: h is not first and real_alive(h) or (h is not first)
    mgr.health_check_all()
    m = mgr.models["mistral-7b"]
    assert m.status.ready and m.handle is not first and first.closed and m.port == st.port and m.request_count == 1
    assert m.restarts == 0 and len(be.loaded) == 2
    # a model that cannot come back
    be.alive = lambda h: False
    m.path = "/nonexistent/model.gguf"
    for _ in range(MAX_RESTARTS + 2):
        mgr.health_check_all()
        mgr.models["mistral-7b"].handle = mgr.models["mistral-7b"].handle or object()  # keep it checkable
    m = mgr.models["mistral-7b"]
    assert str(m.status).startswith("error: ") and m.restarts == MAX_RESTARTS
    # without auto_restart the reference's behaviour: flagged, left alone
    be2 = FakeBackend()
    mgr2 = ModelManager(be2, model_dir=os.path.dirname(vocab_gguf), auto_restart=False)
    mgr2.load_model(LoadRequest("mistral-7b", model_path=vocab_gguf))
    be2.dead.add("mistral-7b")
    mgr2.health_check_all()
    assert str(mgr2.models["mistral-7b"].status).startswith("error: ") and len(be2.loaded) == 1
