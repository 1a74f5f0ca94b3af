"""The on-device sampler chain (sampler.cu) against its oracle restatement (oracle_sample, oracle/aios_oracle.c).

This is the chain every gRPC request reaches: the reference forwards only `temperature` (default 0.7) and `max_tokens`
(runtime/src/inference.rs:103-122) and leaves llama-server's defaults in force — top-k 40, top-p 0.95, min-p 0.05
(SURVEY.md §8a D5, Appendix C.8).  Pinned here, bit for bit: the candidate set and its order, the top-p and min-p
cut-off indices, and the drawn token for fixed (seed, counter), on > 1000 logit vectors; plus the logit mask JSON mode
uses, and the equivalence of a sequence's stochastic decode with the oracle chain run on its own logits.
"""
import numpy as np
import pytest

from aios_b200 import capi
from aios_b200 import gguf_synth as gs

pytestmark = pytest.mark.gpu


def _vectors(rng, n, V):
    """Logit vectors of different shapes: flat (random-init models), peaked, heavy ties, with -inf / large values."""
    for i in range(n):
        kind = i % 6
        if kind == 0:
            lg = rng.standard_normal(V).astype(np.float32) * 0.05          # nearly flat: top-p keeps everything
        elif kind == 1:
            lg = rng.standard_normal(V).astype(np.float32) * 4.0           # peaked
        elif kind == 2:
            lg = np.round(rng.standard_normal(V) * 2.0).astype(np.float32)  # many exact ties: index order decides
        elif kind == 3:
            lg = rng.standard_normal(V).astype(np.float32)
            lg[rng.integers(0, V, 5)] += 12.0                                # a few dominant tokens
        elif kind == 4:
            lg = rng.standard_normal(V).astype(np.float32) * 30.0
            lg[rng.integers(0, V, V // 3)] = -np.inf
        else:
            lg = (rng.random(V).astype(np.float32) - 0.5) * 1e-3
            lg[: V // 2] = lg[V // 2: 2 * (V // 2)]                          # duplicated halves
        yield lg


def test_sampler_chain_matches_oracle_on_many_logit_vectors(engine, oracle):
    rng = np.random.default_rng(1234)
    n_checked = 0
    for V, n in ((32000, 700), (1000, 300), (128256, 60)):
        for i, lg in enumerate(_vectors(rng, n, V)):
            temp = [0.7, 0.3, 1.0, 1.7][i % 4]
            top_k = [40, 40, 1, 128, 0, 7][i % 6]
            top_p = [0.95, 0.5, 1.0, 0.999][(i // 2) % 4]
            min_p = [0.05, 0.0, 0.3][(i // 3) % 3]
            seed, ctr = 1000 + i, (i * 37) % 4096
            want_tok, want_cand, want_counts = oracle.sample(lg, temp, top_k, top_p, min_p, seed, ctr)
            got_tok, got_cand, got_counts = engine.sample_logits(lg, capi.make_sampling(temp, top_k, top_p, min_p, seed), ctr)
            assert got_counts == want_counts, (V, i, got_counts, want_counts)
            assert np.array_equal(got_cand, want_cand), (V, i)
            assert got_tok == want_tok, (V, i)
            n_checked += 1
    assert n_checked >= 1000


def test_sampler_mask_and_greedy(engine, oracle):
    rng = np.random.default_rng(7)
    V = 32000
    for i in range(120):
        lg = rng.standard_normal(V).astype(np.float32) * (0.1 if i % 2 else 3.0)
        n_allowed = [1, 3, 50, 5000, V][i % 5]
        allowed = rng.choice(V, n_allowed, replace=False)
        bits = np.zeros((V + 7) // 8, np.uint8)
        np.bitwise_or.at(bits, allowed // 8, (1 << (allowed % 8)).astype(np.uint8))
        for temp in (0.0, 0.7):
            want = oracle.sample(lg, temp, 40, 0.95, 0.05, 5, i, mask=bits)
            got = engine.sample_logits(lg, capi.make_sampling(temp, 40, 0.95, 0.05, 5), i, mask=bits)
            assert got[0] == want[0] and got[2] == want[2] and np.array_equal(got[1], want[1]), (i, temp)
            assert got[0] in set(allowed.tolist())
    # greedy without mask: lowest index wins ties
    lg = np.zeros(V, np.float32)
    lg[[17, 900, 31999]] = 2.5
    assert engine.sample_logits(lg, capi.make_sampling(0.0))[0] == 17


def test_stochastic_decode_equals_oracle_chain_on_the_sequence_logits(engine, oracle, tmp_models):
    """aios_seq_step with temperature > 0: every token must be what the oracle chain draws from the logits the sequence
    held before the step, with counter = the position of the last token fed (n_past - 1)."""
    arch = "tiny-gqa4"
    m = engine.load_model("samp", tmp_models(arch, "Q4_K_M"), n_ctx=128)
    try:
        s = m.new_seq()
        try:
            prompt = gs.synthetic_prompt(arch, 9, 11)
            s.prefill(prompt)
            sp = capi.make_sampling(0.9, 40, 0.95, 0.05, seed=42)
            for _ in range(24):
                lg = s.logits()
                want, _, _ = oracle.sample(lg, 0.9, 40, 0.95, 0.05, 42, s.n_past - 1)
                assert s.step(sp) == want
            # a masked step: only even token ids allowed
            V = m.info.n_vocab
            bits = np.zeros((V + 7) // 8, np.uint8)
            bits[:] = 0x55
            s.set_mask(bits)
            for _ in range(8):
                lg = s.logits()
                want, _, _ = oracle.sample(lg, 0.9, 40, 0.95, 0.05, 42, s.n_past - 1, mask=bits)
                tok = s.step(sp)
                assert tok == want and tok % 2 == 0
            s.set_mask(None)
            lg = s.logits()
            assert s.step(capi.make_sampling(0.0)) == int(np.argmax(lg))
        finally:
            s.free()
    finally:
        m.unload()
