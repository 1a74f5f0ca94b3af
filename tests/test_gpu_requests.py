"""GPU tests of the queued-request part of the C ABI (aios_request_*): what concurrent gRPC handlers use.

The reference lets Infer calls run in parallel against one llama-server (runtime/src/main.rs:50-51 shares a stateless
InferenceEngine; llama-server queues them on its slot).  Here one scheduler thread per engine does the queueing; the
results must be exactly what the synchronous calls give.
"""
import threading
import time

import numpy as np
import pytest

from aios_b200 import capi
from aios_b200 import gguf_synth as gs

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def model(engine, tmp_models):
    m = engine.load_model("req", tmp_models("tiny-gqa4", "Q4_K_M"), n_ctx=256)
    yield m
    m.unload()


def test_requests_match_synchronous_generate(model):
    prompts = [gs.synthetic_prompt("tiny-gqa4", n, seed) for n, seed in ((5, 1), (17, 2), (40, 3), (9, 4))]
    want = [model.generate(p, 48) for p in prompts]
    reqs = [model.submit(p, 48) for p in prompts]  # all queued at once; submit does not block
    got = [r.result() for r in reqs]
    for r in reqs:
        r.free()
    for g, w in zip(got, want):
        assert np.array_equal(g, w)


def test_requests_from_many_threads_and_sync_calls_interleave(model):
    prompts = [gs.synthetic_prompt("tiny-gqa4", 8 + i, 100 + i) for i in range(8)]
    want = [model.generate(p, 40) for p in prompts]
    got = [None] * len(prompts)

    def worker(i):
        r = model.submit(prompts[i], 40)
        toks, parts = None, []
        while True:
            assert r.wait(30000)
            toks, done = r.poll(cap=7)  # small polls: tokens arrive in pieces, none lost, none repeated
            parts.append(toks)
            if done:
                break
        r.free()
        got[i] = np.concatenate(parts)

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(len(prompts))]
    for t in threads:
        t.start()
    # a synchronous call on the same engine while the scheduler is busy
    again = model.generate(prompts[0], 40)
    for t in threads:
        t.join(60)
    assert np.array_equal(again, want[0])
    for g, w in zip(got, want):
        assert g is not None and np.array_equal(g, w)


def test_request_stop_at_eog_cancel_and_free(model):
    p = gs.synthetic_prompt("tiny-gqa4", 12, 7)
    full = model.generate(p, 200)
    v = model.vocab()
    eog = [i for i, t in enumerate(full) if v.is_eog(int(t))]
    r = model.submit(p, 200, stop_at_eog=True)
    got = r.result()
    r.free()
    assert np.array_equal(got, full[: eog[0]] if eog else full)
    # cancel: the request ends early, what it produced is a prefix of the full continuation
    r = model.submit(p, 200)
    r.cancel()
    got = r.result()
    r.free()
    assert got.size < 200 and np.array_equal(got, full[: got.size])
    # free while queued / running: nothing leaks into later requests, and the model can be unloaded afterwards
    rs = [model.submit(p, 200) for _ in range(3)]
    for x in rs:
        x.free()
    r = model.submit(p, 16)
    assert np.array_equal(r.result(), full[:16])
    r.free()


def test_request_argument_errors(model):
    with pytest.raises(capi.AiosError) as e:
        model.submit(np.zeros(300, np.int32), 4)  # prompt longer than the context
    assert e.value.code == capi.E_STATE
    with pytest.raises(capi.AiosError) as e:
        model.submit(np.array([10 ** 6], np.int32), 4)
    assert e.value.code == capi.E_INVAL
    # max_tokens is clipped to the context
    p = gs.synthetic_prompt("tiny-gqa4", 250, 1)
    r = model.submit(p, 100)
    assert r.result().size == 6
    r.free()


@pytest.mark.parametrize("arch,ftype", [("tiny-gqa4", "Q4_K_M"), ("tiny-gqa8", "Q4_K_M_70B")])
def test_64_concurrent_requests_equal_one_by_one(engine, tmp_models, arch, ftype):
    """Continuous batching (BASELINE.json configs[3]: 64 concurrent agent requests per GPU): 64 requests submitted at
    once decode together — one token per request per step through the multi-token GEMM — and every one of them must
    return exactly the tokens the same prompt gives alone, greedy, bit for bit.  Prompts and lengths differ, so requests
    join and leave the batch at different steps.  The second model has eight query heads per kv head (two passes of the
    batched attention kernel) and the 70B type mix (Q5_K attn_v next to Q4_K q / k in one launch)."""
    m = engine.load_model("b64", tmp_models(arch, ftype), n_ctx=192)
    try:
        rng = np.random.default_rng(5)
        prompts = [gs.synthetic_prompt(arch, int(rng.integers(3, 60)), 1000 + i) for i in range(64)]
        n_gen = [int(rng.integers(1, 90)) for _ in range(64)]
        want = [m.generate(p, n) for p, n in zip(prompts, n_gen)]
        reqs = [m.submit(p, n) for p, n in zip(prompts, n_gen)]
        got = [r.result() for r in reqs]
        for r in reqs:
            r.free()
        for i, (g, w) in enumerate(zip(got, want)):
            assert np.array_equal(g, w), (i, len(prompts[i]), n_gen[i])
        # more requests than lanes: the rest wait for a lane and still match
        more = [m.submit(prompts[i % 64], n_gen[i % 64]) for i in range(100)]
        for i, r in enumerate(more):
            assert np.array_equal(r.result(), want[i % 64]), i
            r.free()
    finally:
        m.unload()


@pytest.mark.parametrize("ftype", ["Q4_K_M", "Q8_0"])
def test_batched_stochastic_requests_equal_sequential(engine, tmp_models, ftype):
    """Sampling in the batch is per lane (own parameters, seed, position counter): a request drawn with temperature 0.8
    gives the same tokens inside a batch of 24 as alone."""
    arch = "tiny-gqa4" if ftype == "Q4_K_M" else "tiny"
    m = engine.load_model("bs", tmp_models(arch, ftype), n_ctx=128)
    try:
        prompts = [gs.synthetic_prompt(arch, 4 + i, 50 + i) for i in range(24)]
        sps = [capi.make_sampling(0.8 if i % 3 else 0.0, 40, 0.95, 0.05, seed=7 + i) for i in range(24)]
        want = [m.generate(p, 40, sp) for p, sp in zip(prompts, sps)]
        reqs = [m.submit(p, 40, sp) for p, sp in zip(prompts, sps)]
        for i, r in enumerate(reqs):
            assert np.array_equal(r.result(), want[i]), i
            r.free()
    finally:
        m.unload()


def test_masked_request_applies_the_callback_mask_on_the_device(engine, tmp_models):
    """aios_request_submit_masked: the scheduler asks the callback for the allowed set before every step and the
    on-device sampler honours it; the callback sees every sampled token and can end the request."""
    m = engine.load_model("mk", tmp_models("tiny-gqa4", "Q4_K_M"), n_ctx=128)
    try:
        V = m.info.n_vocab
        prompt = gs.synthetic_prompt("tiny-gqa4", 9, 3)
        seen = []

        def cb(last_token, mask):
            seen.append(last_token)
            if len(seen) > 20:
                return 2  # "the object is closed"
            if len(seen) % 4 == 0:
                return 1  # an unrestricted step
            mask[:] = 0x55 if len(seen) % 2 else 0xAA  # even ids / odd ids
            return 0

        r = m.submit(prompt, 64, capi.make_sampling(0.7, 40, 0.95, 0.05, seed=3), mask_cb=cb)
        toks = r.result()
        r.free()
        assert len(toks) == 20 and seen[0] == -1 and seen[1:] == [int(t) for t in toks]
        for i, t in enumerate(toks):
            step = i + 1
            if step % 4 == 0:
                continue
            assert t % 2 == (0 if step % 2 else 1), (i, t)
        # the same masks applied step by step on a plain sequence give the same tokens
        s = m.new_seq()
        try:
            s.prefill(prompt)
            sp = capi.make_sampling(0.7, 40, 0.95, 0.05, seed=3)
            again = []
            for step in range(1, 21):
                if step % 4 == 0:
                    s.set_mask(None)
                else:
                    s.set_mask(np.full((V + 7) // 8, 0x55 if step % 2 else 0xAA, np.uint8))
                again.append(s.step(sp))
            assert again == [int(t) for t in toks]
        finally:
            s.free()
    finally:
        m.unload()
