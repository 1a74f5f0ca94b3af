"""aios_seq_decode_static: greedy decode for serving threads that may not allocate device memory.

Some hosts hand out all device memory at start-up and then freeze the allocator for the serving loop
(cudaMalloc / cudaFree synchronise the device and take a process-wide lock, which a latency-bound loop cannot
afford, and an embedding application may own the rest of HBM).  For them the library offers one decode entry
point whose kernel keeps its cross-CTA exchange scratch in a static device array compiled into the library,
so the call itself issues only memset / launch / memcpy.

The test drives the new entry point through the C ABI and checks it against the existing entry point
(aios_seq_decode on an engine that uses the same fused kernel with cudaMalloc'd scratch), bit for bit: token
ids, kept logits and the f16 KV cache.  It makes two calls in a row, because every call shares the one static
array and finds the previous call's contents in it.  This file sorts last on purpose: a run that stops at the
first failure has then already run every other test; it passes alone in a fresh process as well.
"""
import numpy as np
import pytest

from aios_b200 import capi
from aios_b200 import gguf_synth as gs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("arch,ftype", [("tiny", "Q4_K_M"), ("tiny-gqa4", "Q4_K_M")])
def test_static_scratch_decode_matches_allocating_decode_twice(tmp_models, arch, ftype):
    path = tmp_models(arch, ftype)
    eng = capi.Engine(device=0, max_batch=2, flags=capi.ENGINE_USE_PFD)
    try:
        model = eng.load_model("noalloc", path, n_ctx=160)
        try:
            ref, new = model.new_seq(), model.new_seq()
            try:
                prompt = gs.synthetic_prompt(arch, 24, 21)
                ref.prefill(prompt)
                new.prefill(prompt)
                n_layer = gs.ARCHS[arch].n_layer
                for call, n in enumerate((40, 40)):  # two calls in a row on the shared static array
                    want = ref.decode(n)             # existing entry point: fused kernel, cudaMalloc'd scratch
                    got = new.decode_static(n)       # new entry point: same kernel, built-in static scratch
                    assert got.shape == want.shape == (n,), call
                    assert np.array_equal(got, want), (call, np.nonzero(got != want)[0][:4])
                    assert np.array_equal(new.logits().view(np.uint32), ref.logits().view(np.uint32)), call
                    assert new.n_past == ref.n_past
                    for il in range(n_layer):
                        k1, v1 = ref.read_kv(il, 0, ref.n_past)
                        k2, v2 = new.read_kv(il, 0, new.n_past)
                        assert np.array_equal(k1, k2) and np.array_equal(v1, v2), (call, il)
                # the context limit clamps both the same way
                assert np.array_equal(new.decode_static(1000), ref.decode(1000))
                assert new.n_past == 160 and new.decode_static(5).size == 0
            finally:
                ref.free()
                new.free()
        finally:
            model.unload()
    finally:
        eng.close()


def test_static_scratch_decode_rejects_bad_calls(tmp_models):
    path = tmp_models("tiny", "Q4_K_M")
    eng = capi.Engine(device=0, max_batch=1)
    try:
        model = eng.load_model("noalloc", path, n_ctx=64)
        try:
            s = model.new_seq()
            try:
                with pytest.raises(capi.AiosError) as ei:  # nothing fed yet: no logits to sample from
                    s.decode_static(4)
                assert ei.value.code == capi.E_STATE
                import ctypes as C
                n = C.c_int32(0)
                rc = capi.lib().aios_seq_decode_static(s.h, 4, None, C.byref(n))
                assert rc == capi.E_INVAL
            finally:
                s.free()
        finally:
            model.unload()
    finally:
        eng.close()
