"""The prefill GEMM on tcgen05 tensor cores (prefill_tc.cu, aios_matmul_tc) against exact arithmetic.

Stated numerics: weights are rounded to f16 after dequantisation, activations to f16, products exact, f32 accumulation
in tensor memory.  Tolerance: 2e-3 of the output's rms against the f64 product of the oracle-dequantised weights with
the f32 activations (measured ~2e-4).  For Q8_0, where f16(d * q) is a single rounding that numpy reproduces, the result
must also agree to accumulation error (1e-5 of rms) with the f64 product of those f16 operands.
"""
import numpy as np
import pytest

from aios_b200 import gguf_synth as gs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("qt", [gs.Q8_0, gs.Q4_K, gs.Q6_K])
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
    print(f"type {qt} shape {shape}: max err / rms = {err:.2e}")
    assert err < 2e-3
    if qt == gs.Q8_0:
        blocks = w.reshape(-1, 34)
        d = blocks[:, :2].copy().view(np.float16).astype(np.float32)
        q = blocks[:, 2:].view(np.int8).astype(np.float32)
        w16 = (d * q).astype(np.float16).reshape(n, k).astype(np.float64)   # one rounding, like HMUL2
        x16 = x.astype(np.float16).astype(np.float64)
        ref = x16 @ w16.T
        assert np.abs(got - ref).max() / rms < 1e-5


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
    assert np.abs(got - exact).max() / np.sqrt(np.mean(exact ** 2)) < 2e-3
