"""KV-slot accounting and per-device set-up of the C ABI.

`aios_engine_opts.max_batch` is the number of sequences the CALLER may hold per model.  The library's own sequences —
the scratch sequence behind aios_forward_logits and the request scheduler's — live in hidden slots, so an engine
created with max_batch = 1 (the default) must still serve forward_logits, generate and submit in any order.
(Reference behaviour being replaced: one llama-server child serves every kind of call, runtime/src/inference.rs:94-186.)
"""
import numpy as np
import pytest

from aios_b200 import capi
from aios_b200 import gguf_synth as gs

pytestmark = pytest.mark.gpu


def test_internal_sequences_do_not_consume_caller_slots(oracle, tmp_models):
    path = tmp_models("tiny", "Q4_K_M")
    eng = capi.Engine(device=0, max_batch=1)
    try:
        m = eng.load_model("slots", path, n_ctx=128)
        try:
            prompt = gs.synthetic_prompt("tiny", 16, 3)
            om = oracle.OracleModel(path, n_ctx=128)
            ref_logits = om.eval(prompt)
            ref, _ = om.greedy(prompt, 16)
            # the order that broke round 1's smoke(): scratch sequence first, then a temporary one, then the queue
            got_logits = m.forward_logits(prompt)
            assert np.array_equal(got_logits.view(np.uint32), ref_logits.view(np.uint32))
            assert np.array_equal(m.generate(prompt, 16), ref)
            r = m.submit(prompt, 16)
            assert np.array_equal(r.result(), ref)
            r.free()
            # the caller's own budget is intact: exactly max_batch sequences, then AIOS_E_STATE
            s = m.new_seq()
            try:
                s.prefill(prompt)
                assert np.array_equal(s.decode(16), ref)
                with pytest.raises(capi.AiosError) as ei:
                    m.new_seq()
                assert ei.value.code == capi.E_STATE
                # ... and the internal paths still work while the caller holds its slot
                assert np.array_equal(m.forward_logits(prompt).view(np.uint32), ref_logits.view(np.uint32))
                r = m.submit(prompt, 16)
                assert np.array_equal(r.result(), ref)
                r.free()
            finally:
                s.free()
            assert np.array_equal(m.generate(prompt, 16), ref)
        finally:
            m.unload()
    finally:
        eng.close()


def test_fused_decode_with_32_kv_heads(oracle, tmp_models):
    """The persistent fused kernel keeps one barrier counter per kv head next to its grid barrier: a multi-head
    model with 32 kv heads must not alias them (it deadlocked or corrupted the barriers when the counters were a
    fixed 16 + 16 words)."""
    path = tmp_models("tiny-mha32", "Q4_K_M")
    eng = capi.Engine(device=0, max_batch=1, flags=capi.ENGINE_USE_PFD)
    try:
        m = eng.load_model("mha", path, n_ctx=96)
        try:
            prompt = gs.synthetic_prompt("tiny-mha32", 12, 5)
            ref, margins = oracle.OracleModel(path, n_ctx=96).greedy(prompt, 24)
            s = m.new_seq()
            try:
                s.prefill(prompt)
                assert np.array_equal(s.decode(24), ref), margins.min()
            finally:
                s.free()
        finally:
            m.unload()
    finally:
        eng.close()


def test_two_engines_on_two_devices(oracle, tmp_models):
    """Kernel attributes (dynamic shared memory, carve-out) are per-device: a second engine on another GPU in the same
    process must get them too (round 1 guarded them with process-wide flags)."""
    if capi.device_count() < 2:
        pytest.skip("needs two GPUs")
    path = tmp_models("tiny-gqa4", "Q4_K_M")
    prompt = gs.synthetic_prompt("tiny-gqa4", 20, 9)
    ref, _ = oracle.OracleModel(path, n_ctx=128).greedy(prompt, 24)
    engines = [capi.Engine(device=d, max_batch=1) for d in (0, 1)]
    try:
        models = [e.load_model(f"m{i}", path, n_ctx=128) for i, e in enumerate(engines)]
        try:
            reqs = [m.submit(prompt, 24) for m in models]   # both GPUs work at the same time
            for r in reqs:
                assert np.array_equal(r.result(), ref)
                r.free()
            for m in models:
                assert np.array_equal(m.generate(prompt, 24), ref)
        finally:
            for m in models:
                m.unload()
    finally:
        for e in engines:
            e.close()


def test_service_backend_replicates_models_and_spreads_requests(tmp_models):
    """The multi-GPU service: EngineBackend with several devices loads one replica of a model per GPU and concurrent
    requests land on different devices (least-loaded dispatch); HealthCheck-style load report per GPU."""
    n_dev = capi.device_count()
    if n_dev < 2:
        pytest.skip("needs two GPUs")
    from aios_b200.runtime.inference import EngineBackend
    path = tmp_models("tiny-gqa4", "Q4_K_M")
    be = EngineBackend(devices=[0, 1])
    try:
        rs = be.load("m", path, 128)
        prompt = gs.synthetic_prompt("tiny-gqa4", 12, 4)
        want = rs.replicas[0].model.generate(prompt, 24)
        a = rs.submit(prompt, 24, capi.make_sampling(0.0))
        b = rs.submit(prompt, 24, capi.make_sampling(0.0))
        assert {a.device, b.device} == {0, 1}
        assert be.gpu_load() == {0: [1, 1], 1: [1, 1]}
        ra, rb = a.result(), b.result()
        a.free(); b.free()
        # stop_at_eog may cut at an end-of-generation token: compare the common prefix
        assert np.array_equal(ra, want[: len(ra)]) and np.array_equal(rb, want[: len(rb)])
        assert be.gpu_load() == {0: [1, 0], 1: [1, 0]}
        be.unload(rs)
    finally:
        for e in be.engines:
            e.close()
