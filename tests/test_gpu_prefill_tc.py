"""The prefill GEMM on tcgen05 tensor cores (prefill_tc.cu, aios_matmul_tc) against exact arithmetic.

Stated numerics: weights are rounded to f16 after dequantisation, activations to f16, products exact, f32 accumulation
in tensor memory.  Tolerance against the f64 product of the oracle-dequantised weights with the f32 activations:
rms error <= 8e-4 and max error <= 4e-3 of the output's rms (two f16 roundings per product, 2^-12 rms each: ~4e-4 rms,
and the maximum over ~1e5 outputs sits at 4-5 sigma; measured 1.3e-3 .. 2.0e-3).  For Q8_0, where f16(d * q) is a single rounding that numpy reproduces, the result
must also agree to accumulation error (1e-4 of rms) with the f64 product of those f16 operands.
"""
import numpy as np
import pytest

from aios_b200 import gguf_synth as gs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("qt", [gs.Q8_0, gs.Q4_K, gs.Q5_K, gs.Q6_K])
@pytest.mark.parametrize("shape", [(128, 256, 256), (128, 512, 64), (200, 1024, 300), (384, 4096, 513), (130, 14336, 40), (64, 256, 1)])
def test_matmul_tc_within_tolerance(engine, oracle, qt, shape):
    n, k, t = shape
    rng = np.random.default_rng([qt, n, k, t])
    w = gs.random_blocks(qt, n * k // gs.BLOCK[qt][0], rng, 0.02)
    x = rng.standard_normal((t, k), dtype=np.float32)
    x[0, :9] = 0
    got = engine.matmul_tc(qt, w, n, k, x)
    assert got.shape == (t, n)
    wd = oracle.dequantize(qt, w).reshape(n, k).astype(np.float64)
    exact = x.astype(np.float64) @ wd.T
    rms = np.sqrt(np.mean(exact ** 2))
    err = np.abs(got - exact).max() / rms
    err_rms = np.sqrt(np.mean((got - exact) ** 2)) / rms
    print(f"type {qt} shape {shape}: max err / rms = {err:.2e}, rms err / rms = {err_rms:.2e}")
    assert err < 4e-3 and err_rms < 8e-4
    if qt == gs.Q8_0:
        blocks = w.reshape(-1, 34)
        d = blocks[:, :2].copy().view(np.float16).astype(np.float32)
        q = blocks[:, 2:].view(np.int8).astype(np.float32)
        w16 = (d * q).astype(np.float16).reshape(n, k).astype(np.float64)   # one rounding, like HMUL2
        x16 = x.astype(np.float16).astype(np.float64)
        ref = x16 @ w16.T
        assert np.abs(got - ref).max() / rms < 1e-4   # f32 accumulation in tensor memory over K up to 14336


def test_matmul_tc_with_rmsnorm(engine, oracle):
    n, k, t = 256, 2048, 70
    rng = np.random.default_rng(9)
    qt = gs.Q4_K
    w = gs.random_blocks(qt, n * k // 256, rng, 0.02)
    x = rng.standard_normal((t, k), dtype=np.float32) * 3
    nw = (rng.random(k, dtype=np.float32) * 0.4 + 0.8).astype(np.float32)
    got = engine.matmul_tc(qt, w, n, k, x, norm_w=nw, eps=1e-5)
    xn = x.astype(np.float64)
    xn = xn / np.sqrt((xn ** 2).mean(axis=1, keepdims=True) + 1e-5) * nw
    exact = xn @ oracle.dequantize(qt, w).reshape(n, k).astype(np.float64).T
    assert np.abs(got - exact).max() / np.sqrt(np.mean(exact ** 2)) < 4e-3


@pytest.mark.parametrize("arch,ftype", [("tiny-gqa4", "Q4_K_M"), ("tiny", "Q8_0"), ("tiny-mha32", "Q4_K_M"), ("tiny-gqa4", "Q4_K_M_70B")])
def test_tc_prefill_logits_against_exact_math(oracle, tmp_models, arch, ftype):
    """Whole path: a 300-token prompt through the tensor-core prefill (tcgen05 GEMMs, RoPE / KV append for all rows, causal
    flash attention) against an exact f64 forward pass of the dequantised model (oracle/ref_llama.py).
    Stated tolerance: the last token's logits within 5e-2 of the logit std (measured 2.4e-2; the f16 layers contribute
    a few 1e-3, most of the rest is the lm_head, which stays on the exact integer GEMV with its Q8 activations).  The
    bit-exact integer path differs from the same reference by its activation-quantisation noise (6e-2 .. 1e-1 of the
    std), so the float path is the CLOSER of the two to exact arithmetic; the test prints both."""
    from aios_b200 import capi
    from oracle import ref_llama
    path = tmp_models(arch, ftype)
    prompt = gs.synthetic_prompt(arch, 300, 77)
    hp, w = ref_llama.load_weights(path, oracle)
    exact = ref_llama.forward_logits(hp, w, prompt)[-1]
    sd = exact.std()
    eng_tc = capi.Engine(device=0, max_batch=1)
    eng_ex = capi.Engine(device=0, max_batch=1, flags=capi.ENGINE_EXACT_PREFILL)
    try:
        m_tc, m_ex = eng_tc.load_model("tc", path, n_ctx=512), eng_ex.load_model("ex", path, n_ctx=512)
        try:
            s_tc, s_ex = m_tc.new_seq(), m_ex.new_seq()
            try:
                s_tc.prefill(prompt)
                s_ex.prefill(prompt)
                l_tc, l_ex = s_tc.logits().astype(np.float64), s_ex.logits().astype(np.float64)
                e_tc, e_ex = np.abs(l_tc - exact).max() / sd, np.abs(l_ex - exact).max() / sd
                print(f"{arch} {ftype}: max |logit - exact| / std: tensor-core path {e_tc:.2e}, integer path {e_ex:.2e}")
                assert e_tc < 5e-2
                # the KV rows the two paths wrote agree to f16 rounding of slightly different f32 values
                k1, v1 = s_tc.read_kv(0, 0, 300)
                k2, v2 = s_ex.read_kv(0, 0, 300)
                kd = np.abs(k1.view(np.float16).astype(np.float32) - k2.view(np.float16).astype(np.float32)).max()
                assert kd < 0.1 * np.abs(k2.view(np.float16).astype(np.float32)).std() + 1e-3
                # decode continues from either cache; report how long the greedy continuations agree
                a, b = s_tc.decode(32), s_ex.decode(32)
                agree = int(np.argmax(np.concatenate([a != b, [True]])))
                print(f"greedy continuation: first {agree} of 32 tokens identical")
                # a second chunk on top of a cached prefix (positions 300..): attention sees the cached rows
                more = gs.synthetic_prompt(arch, 280, 78)[: 512 - 300 - 32 - 1]
                s_tc.reset(); s_tc.prefill(np.concatenate([prompt[:40]]))      # exact (short), then a long chunk at position 40
                s_tc.prefill(np.concatenate([prompt[40:], more[:100]]))
                full = ref_llama.forward_logits(hp, w, np.concatenate([prompt, more[:100]]))[-1]
                assert np.abs(s_tc.logits().astype(np.float64) - full).max() / full.std() < 1.5e-1  # exact prefix: Q8 noise
            finally:
                s_tc.free()
                s_ex.free()
        finally:
            m_tc.unload()
            m_ex.unload()
    finally:
        eng_tc.close()
        eng_ex.close()
