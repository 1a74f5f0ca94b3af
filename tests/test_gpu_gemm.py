"""GPU parity of the multi-token tensor-core GEMM (gemm_q.cu, aios_matmul): Y[T][N] must equal T oracle mat-vecs
bit for bit — the int8 MMA only changes HOW the 32-weight integer dots are formed, not their values, and the scale
arithmetic and the double-precision sums around them are the GEMV's."""
import numpy as np
import pytest

from aios_b200 import capi
from aios_b200 import gguf_synth as gs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("qt", [gs.Q4_K, gs.Q5_K, gs.Q6_K, gs.Q8_0])
@pytest.mark.parametrize("shape", [(16, 256, 8), (37, 512, 3), (64, 2048, 32), (130, 4096, 17), (70, 5632, 9), (48, 14336, 24),
                                   (33, 1024, 64), (95, 2048, 41), (1, 256, 1)])
def test_matmul_equals_oracle_matvecs(engine, oracle, qt, shape):
    n, k, t = shape
    rng = np.random.default_rng([qt, n, k, t])
    w = gs.random_blocks(qt, n * k // gs.BLOCK[qt][0], rng, 0.02)
    x = rng.standard_normal((t, k), dtype=np.float32)
    x[0, :9] = 0
    if t > 2:
        x[2] = 0  # an all-zero token
    got = engine.matmul(qt, w, n, k, x)
    assert got.shape == (t, n)
    for i in range(t):
        ref = oracle.matvec(qt, w, n, k, x[i])
        assert np.array_equal(got[i].view(np.uint32), ref.view(np.uint32)), (i, np.abs(got[i] - ref).max())


@pytest.mark.parametrize("qt", [gs.Q4_K, gs.Q5_K, gs.Q6_K, gs.Q8_0])
def test_matmul_with_rmsnorm_equals_matvec(engine, qt):
    n, k, t = 96, 2048, 12
    rng = np.random.default_rng(5)
    w = gs.random_blocks(qt, n * k // gs.BLOCK[qt][0], rng, 0.02)
    x = rng.standard_normal((t, k), dtype=np.float32) * 2
    nw = (rng.random(k, dtype=np.float32) * 0.4 + 0.8).astype(np.float32)
    got = engine.matmul(qt, w, n, k, x, norm_w=nw, eps=1e-5)
    for i in range(t):
        ref = engine.matvec(qt, w, n, k, x[i], norm_w=nw, eps=1e-5)
        assert np.array_equal(got[i].view(np.uint32), ref.view(np.uint32)), i


@pytest.mark.parametrize("arch,ftype", [("tiny", "Q4_K_M"), ("tiny-gqa4", "Q4_K_M"), ("tiny", "Q8_0"), ("tiny-gqa4", "Q4_K_M_70B"), ("tiny-gqa8", "Q4_K_M_70B")])
@pytest.mark.parametrize("n_prompt", [4, 33, 70, 121])
def test_chunked_prefill_is_bit_identical_to_token_by_token(engine, oracle, tmp_models, arch, ftype, n_prompt):
    """Seq::prefill feeds prompts in 64-token chunks through the GEMM; logits, KV rows and the continuation must be
    those of the token-by-token path (AIOS_ENGINE_NO_GEMM) and of the oracle."""
    path = tmp_models(arch, ftype)
    prompt = gs.synthetic_prompt(arch, n_prompt, 31)
    slow = capi.Engine(device=0, max_batch=1, flags=capi.ENGINE_NO_GEMM)
    m1, m2 = engine.load_model("a", path, n_ctx=128), slow.load_model("b", path, n_ctx=128)
    try:
        s1, s2 = m1.new_seq(), m2.new_seq()
        try:
            s1.prefill(prompt[:3])      # a short first call (token path), then chunks starting at position 3
            s1.prefill(prompt[3:])
            s2.prefill(prompt)
            assert s1.n_past == s2.n_past == n_prompt
            assert np.array_equal(s1.logits().view(np.uint32), s2.logits().view(np.uint32))
            om = oracle.OracleModel(path, n_ctx=128)
            ref = om.eval(prompt)
            assert np.array_equal(s1.logits().view(np.uint32), ref.view(np.uint32))
            for il in range(gs.ARCHS[arch].n_layer):
                k1, v1 = s1.read_kv(il, 0, n_prompt)
                k2, v2 = s2.read_kv(il, 0, n_prompt)
                assert np.array_equal(k1, k2) and np.array_equal(v1, v2)
            assert np.array_equal(s1.decode(16), s2.decode(16))
        finally:
            s1.free()
            s2.free()
    finally:
        m1.unload()
        m2.unload()
        slow.close()


def test_matmul_argument_errors(engine):
    w = np.zeros(4 * 256 * 4, np.uint8)
    with pytest.raises(capi.AiosError) as e:
        engine.matmul(gs.F32, w, 4, 256, np.zeros((2, 256), np.float32))  # not a quantised matrix type
    assert e.value.code == capi.E_UNSUPPORTED
    with pytest.raises(capi.AiosError) as e:
        engine.matmul(gs.Q4_K, np.zeros(144 * 4, np.uint8), 4, 256, np.zeros((65, 256), np.float32))
    assert e.value.code == capi.E_RANGE
