"""GPU parity tests at the sizes BASELINE.json names.

configs[0]: TinyLlama-1.1B Q4_K_M (random-init GGUF of that architecture), 128-token prompt, 128-token greedy decode —
the reference's own CPU-runnable case.  The oracle (CPU restatement of the llama.cpp CPU path) produces the tokens; the
engine must produce the same 128 tokens and bit-identical logits at the end.  This is also the only model-level test
whose matrices take the default double-buffered GEMV kernel for K = 2048 and the single-image fallback for
K = 5632 (ffn_down) in the same pass.

configs[1] (Mistral-7B Q4_K_M, the bench workload) is compared with the oracle directly — a 40-token prompt, 16
greedy tokens and the final logits, bit for bit — and covered at full size by size-independent properties: the decode
is deterministic, a chunked decode equals a one-shot decode, the sequence snapshot/restore replays identically, and
the queued-request path returns the same tokens as the synchronous one.  (bench.py repeats the oracle comparison on its
own 384-token prompt before it times anything.)
"""
import numpy as np
import pytest

from aios_b200 import gguf_synth as gs

pytestmark = pytest.mark.gpu


def test_config0_tinyllama_128_prompt_128_greedy(engine, oracle, tmp_models):
    path = tmp_models("tinyllama-1.1b", "Q4_K_M", 0xA105)
    prompt = gs.synthetic_prompt("tinyllama-1.1b", 128, 0xA105)
    om = oracle.OracleModel(path, n_ctx=260, simd=False)
    ref, margins = om.greedy(prompt, 128)
    ref_logits = om.eval(np.concatenate([prompt, ref]).astype(np.int32))  # logits after all 256 tokens were fed
    m = engine.load_model("tl", path, n_ctx=260)
    try:
        seq = m.new_seq()
        try:
            seq.prefill(prompt)
            got = seq.decode(128)
            print(f"min top-1/top-2 margin over 128 tokens: {margins.min():.3e}")
            assert np.array_equal(got, ref)
            # decode feeds every token it samples: the kept logits are those after all 256 tokens — bit for bit
            assert seq.n_past == 256
            assert np.array_equal(seq.logits().view(np.uint32), ref_logits.view(np.uint32))
        finally:
            seq.free()
    finally:
        m.unload()


def test_config1_mistral7b_matches_oracle(engine, oracle):
    """The benchmarked model against the oracle at Mistral-7B size: every matrix shape of the headline number (K = 4096
    and 14336, Q4_K and Q6_K, the mixed q/k/v launch, 32 layers of GQA-4 attention with head_dim 128)."""
    import bench
    path = bench.ensure_model("mistral-7b", "Q4_K_M", True)
    prompt = gs.synthetic_prompt("mistral-7b", 40, bench.SEED)
    om = oracle.OracleModel(path, n_ctx=64, simd=True, threads=bench.host_threads())
    ref, margins = om.greedy(prompt, 16)
    ref_logits = om.eval(np.concatenate([prompt, ref]).astype(np.int32))  # after all 56 tokens were fed
    om.close()
    m = engine.load_model("m7o", path, n_ctx=64)
    try:
        seq = m.new_seq()
        try:
            seq.prefill(prompt)
            got = seq.decode(16)
            print(f"min top-1/top-2 margin over 16 tokens: {margins.min():.3e}")
            assert np.array_equal(got, ref)
            assert seq.n_past == 56
            assert np.array_equal(seq.logits().view(np.uint32), ref_logits.view(np.uint32))  # bit for bit
        finally:
            seq.free()
    finally:
        m.unload()


def test_config1_mistral7b_full_size_properties(engine):
    import bench
    path = bench.ensure_model("mistral-7b", "Q4_K_M", True)
    prompt = gs.synthetic_prompt("mistral-7b", 48, bench.SEED)
    m = engine.load_model("m7", path, n_ctx=160)
    try:
        a = m.generate(prompt, 96)
        b = m.generate(prompt, 96)
        assert np.array_equal(a, b)  # deterministic
        seq = m.new_seq()
        try:
            seq.prefill(prompt)
            seq.snapshot()
            parts = [seq.decode(n) for n in (1, 7, 24, 64)]  # chunked == one shot
            assert np.array_equal(np.concatenate(parts), a)
            l1 = seq.logits()
            seq.restore()
            assert np.array_equal(seq.decode(96), a)  # snapshot / restore replays identically
            assert np.array_equal(seq.logits().view(np.uint32), l1.view(np.uint32))
        finally:
            seq.free()
        r = m.submit(prompt, 96)  # the scheduler thread gives the same tokens
        assert np.array_equal(r.result(), a)
        r.free()
        info = m.info
        assert abs(info.weight_bytes_per_token - 4.294e9) < 2e6  # 4.2940 GB of weights per token
        assert info.kv_bytes_per_position == gs.kv_bytes_per_position("mistral-7b") == 131072
    finally:
        m.unload()


def test_tied_embeddings_match_oracle(oracle, tmp_path):
    """Files without `output.weight` use `token_embd.weight` as the lm_head (real-file compatibility,
    §8f-4): logits and greedy tokens against the oracle, which takes the same branch (oracle/aios_oracle.c)."""
    from aios_b200 import capi

    path = str(tmp_path / "tied.gguf")
    info = gs.write_synthetic_gguf(path, "tiny-gqa4", "Q4_K_M", 11, tied=True)
    assert "output.weight" not in info.tensors
    prompt = gs.synthetic_prompt("tiny-gqa4", 20, 3)
    om = oracle.OracleModel(path, n_ctx=128)
    ref_logits = om.eval(prompt)
    ref, _ = om.greedy(prompt, 24)
    eng = capi.Engine(device=0)
    try:
        m = eng.load_model("tied", path, n_ctx=128)
        try:
            assert np.array_equal(m.forward_logits(prompt).view(np.uint32), ref_logits.view(np.uint32))
            assert np.array_equal(m.generate(prompt, 24), ref)
        finally:
            m.unload()
    finally:
        eng.close()


@pytest.mark.parametrize("name,rms_tol,corr_min,max_tol", [("exact_logits_tinyllama.npz", 0.15, 0.99, 0.8),
                                                           ("exact_logits_mistral7b.npz", 0.25, 0.975, 1.2)])
def test_engine_against_exact_math_at_baseline_shapes(engine, name, rms_tol, corr_min, max_tol):
    """The CUDA path against the independent full-size pin (gguf-py dequantisers + float64 numpy forward pass,
    scripts/make_golden_exact.py): TinyLlama-1.1B and Mistral-7B shapes, within the activation-quantisation noise that
    separates the llama.cpp-style integer path from exact math (same bounds as the oracle's own test, tests/test_cpu.py)."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name))
    arch, ftype, seed = str(g["arch"]), str(g["ftype"]), int(g["seed"])
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else "/tmp"
    path = os.path.join(base, f"aios_pin_{arch}_{os.getpid()}.gguf")
    try:
        gs.write_synthetic_gguf(path, arch, ftype, seed)
        m = engine.load_model("pin", path, n_ctx=64)
        try:
            got = m.forward_logits(g["prompt"])
        finally:
            m.unload()
    finally:
        if os.path.exists(path):
            os.remove(path)
    exact = g["logits"]
    sd = exact.std()
    assert np.sqrt(np.mean((got - exact) ** 2)) / sd < rms_tol
    assert np.corrcoef(got, exact)[0, 1] > corr_min
    assert np.abs(got - exact).max() / sd < max_tol
