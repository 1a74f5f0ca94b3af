"""CPU tests of JSON-mode constrained sampling (aios_b200/runtime/json_grammar.py + InferenceEngine._generate_json):
what llama-server's ``response_format: json_object`` grammar does for the reference's unary Infer
(runtime/src/inference.rs:119-121)."""
import json
import threading

import numpy as np
import pytest

from aios_b200 import capi
from aios_b200.runtime.inference import InferenceEngine, InferParams
from aios_b200.runtime.json_grammar import JsonPrefix, is_json_object_prefix
from test_cpu_runtime import vocab_gguf, spm_golden  # noqa: F401  (fixtures)

VALID = [
    '{}', ' {"a": 1}', '{"a":{"b":[1,2.5,-3e+7,true,false,null,"x\\n\\u00e9"]},"c":""}', '{"k":[]}', '{"k":[[],{}]}',
    '{"é": "naïve ☃"}', '{"n":0}', '{"n":-0.5E-2}', '{\n  "tool": "fs.read",\n  "args": {"path": "/etc/hosts"}\n}',
]
INVALID = ['[1]', '"x"', '{a:1}', "{'a':1}", '{"a":01}', '{"a":1,}', '{"a":.5}', '{"a":1e}', '{"a":tru}', '{"a":"\x01"}',
           '{"a":"\\x"}', '{"a" 1}', '{"a":1}}', '{"a":1} x', '{"a":-}', '{"a":[1,]}', '{"a":"\\u12G4"}']


@pytest.mark.parametrize("text", VALID)
def test_every_prefix_of_valid_json_is_accepted(text):
    raw = text.encode("utf-8")
    json.loads(text)
    st = JsonPrefix()
    for i in range(len(raw)):
        assert st.accepts(raw[i:i + 1]), (text, i)
        assert st.feed(raw[i:i + 1])
        assert is_json_object_prefix(raw[: i + 1])
    assert st.complete
    assert st.accepts(b" \n") and not st.accepts(b"x") and not st.accepts(b"")


@pytest.mark.parametrize("text", INVALID)
def test_invalid_json_is_rejected(text):
    raw = text.encode("utf-8")
    st = JsonPrefix()
    ok = st.feed(raw)
    assert not (ok and st.complete), text  # either rejected on the way, or never a closed object
    with pytest.raises(Exception):
        v = json.loads(text)
        if not isinstance(v, dict):
            raise ValueError("not an object")


def test_incomplete_and_split_multibyte():
    st = JsonPrefix()
    assert st.feed('{"a": "caf'.encode()) and not st.complete
    e = "é".encode("utf-8")
    assert st.accepts(e[:1]) and not st.accepts(e[1:])  # a lone continuation byte is not text
    assert st.feed(e[:1]) and not st.accepts(b'"') and st.feed(e[1:]) and st.feed(b'"}') and st.complete
    deep = JsonPrefix(depth_limit=3)
    assert deep.feed(b'{"a":{"b":{"c":') and not deep.accepts(b"{") and deep.accepts(b"1")


class ScriptedSeq:
    """Logits that prefer junk: the grammar has to steer.  Token preference rotates with the position."""

    def __init__(self, n_vocab, seed=0):
        self.n_vocab, self.rng, self.fed = n_vocab, np.random.default_rng(seed), []

    def reset(self):
        self.fed = []

    def prefill(self, tokens):
        self.fed.extend(int(t) for t in np.atleast_1d(tokens))

    def logits(self):
        return np.random.default_rng(len(self.fed)).standard_normal(self.n_vocab).astype(np.float32) * 4


class ScriptedLoaded:
    def __init__(self, path, n_ctx=512):
        self.vocab = capi.Vocab.open(path)
        self.n_ctx = n_ctx
        self.seq = ScriptedSeq(self.vocab.info.n_tokens)
        self.lock = threading.Lock()


def test_constrained_generation_only_emits_json_prefixes(vocab_gguf):  # noqa: F811
    lm = ScriptedLoaded(vocab_gguf)
    eng = InferenceEngine()
    seen_complete = 0
    for seed in range(6):
        p = InferParams(prompt="Give me JSON.", max_tokens=96, temperature=0.8, seed=seed, json_mode=True)
        r = eng.infer(lm, "m", p)
        raw = r.text.encode("utf-8")
        assert raw and is_json_object_prefix(raw), r.text
        st = JsonPrefix()
        st.feed(raw)
        if st.complete:  # a closed object ends the generation at once and parses
            assert isinstance(json.loads(r.text), dict)
            seen_complete += 1
        n_prompt = lm.vocab.tokenize(lm.vocab.chat_prompt("", "Give me JSON."), parse_special=True).size
        assert n_prompt < r.tokens_used <= n_prompt + 96
    # unconstrained sampling from the same logits is junk: the constraint is what made the difference
    junk = lm.vocab.detokenize(np.argsort(-lm.seq.logits())[:8])
    assert not is_json_object_prefix(junk)


def test_cpp_grammar_mask_equals_python_acceptor(vocab_gguf):  # noqa: F811
    """The C++ JSON grammar (csrc/json_mask.cpp — what the scheduler thread calls per decode step) must allow exactly the
    tokens the Python acceptor (runtime/json_grammar.py) accepts, state after state along random JSON-valid walks, and
    report completion at the same step."""
    from aios_b200 import capi
    from aios_b200.runtime.json_grammar import JsonPrefix
    vocab = capi.Vocab.open(vocab_gguf)
    n = vocab.info.n_tokens
    pieces = [vocab.detokenize([i]) for i in range(n)]
    rng = np.random.default_rng(11)
    for walk in range(6):
        g = capi.JsonGrammar(vocab)
        st = JsonPrefix()
        last = -1
        for step in range(60):
            rc, mask = g.step(last, n)
            if st.complete:
                assert rc == 2 and g.complete
                break
            want = np.array([bool(pieces[i]) and st.accepts(pieces[i]) for i in range(n)])
            got = np.unpackbits(mask, bitorder="little")[:n].astype(bool)
            assert np.array_equal(got, want), (walk, step, np.nonzero(got != want)[0][:5])
            assert rc == (0 if want.any() else 2)
            if not want.any():
                break
            # prefer closing tokens now and then so that walks end
            allowed = np.nonzero(want)[0]
            closers = [i for i in allowed if pieces[i].strip() in (b"}", b'"', b"]", b'"}')]
            last = int(rng.choice(closers)) if closers and rng.random() < 0.4 else int(rng.choice(allowed))
            assert st.feed(pieces[last])
        g.free()
    # a token the grammar did not allow is an error, not a silent state change
    g = capi.JsonGrammar(vocab)
    g.step(-1, n)
    bad = next(i for i in range(n) if pieces[i] and not JsonPrefix().accepts(pieces[i]))
    assert g.step(bad, n)[0] == -1
    g.free()
