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
