"""GPU parity tests: the CUDA path through the C ABI vs the CPU oracle on the same seeded inputs.

Bars (BASELINE.json north_star): dequantised blocks bit-exact; logits within a stated relative
tolerance; greedy tokens identical, with the minimum top-1/top-2 margin reported.

The stated logit tolerance is LOGIT_TOL = 1e-6 (max-abs-err / logit-std), and in practice the
logits are BIT-IDENTICAL: the engine mirrors the oracle's order-free numerics (one f32 term per
64-weight unit summed in double; deterministic expf; exact f16 products summed in double), so the
only possible difference is a double-rounding event of probability ~1e-7 per value.  A loose
tolerance would be meaningless here anyway: one flipped Q8 rounding re-randomises the activation
quantisation noise of all later layers (~4e-2 of the logit std on the tiny models).
"""
import os

import numpy as np
import pytest

from aios_b200 import capi
from aios_b200 import gguf_synth as gs

pytestmark = pytest.mark.gpu

QTYPES = [gs.Q8_0, gs.Q4_K, gs.Q5_K, gs.Q6_K]
LOGIT_TOL = 1e-6  # max |gpu - oracle| / std(oracle logits)
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("qt", QTYPES)
def test_dequantize_bit_exact(engine, oracle, qt):
    rng = np.random.default_rng(qt)
    blocks = gs.random_blocks(qt, 4096, rng, 0.02)
    # edge blocks: all-zero, all-ones bytes with finite scales, max 6-bit scales, sub-normal d
    blocks[0] = 0
    blocks[1] = 0xFF
    blocks[2, :] = 0x3F
    be, bb = gs.BLOCK[qt]
    if qt in (gs.Q4_K, gs.Q5_K):
        blocks[1, 0:4] = np.array([0x00, 0x3C, 0x00, 0x38], np.uint8)  # d=1.0 dmin=0.5
        blocks[3, 0:2] = np.array([0x01, 0x00], np.uint8)  # smallest sub-normal d
    elif qt == gs.Q6_K:
        blocks[1, 208:210] = np.array([0x00, 0x3C], np.uint8)
        blocks[3, 208:210] = np.array([0x01, 0x00], np.uint8)
    else:
        blocks[1, 0:2] = np.array([0x00, 0x3C], np.uint8)
        blocks[3, 0:2] = np.array([0x01, 0x00], np.uint8)
    n = blocks.shape[0] * be
    ref = oracle.dequantize(qt, blocks)
    got = engine.dequantize(qt, blocks, n)
    assert np.array_equal(ref.view(np.uint32), got.view(np.uint32))


@pytest.mark.parametrize("qt", QTYPES)
def test_dequantize_golden_gguf_py(engine, qt):
    """Committed fixture made by gguf-py (scripts/make_golden.py): edge blocks incl. max-finite and sub-normal
    fp16 scales, where the result overflows to inf / nan; NaN payloads are not compared."""
    name = {gs.Q8_0: "Q8_0", gs.Q4_K: "Q4_K", gs.Q5_K: "Q5_K", gs.Q6_K: "Q6_K"}[qt]
    g = np.load(os.path.join(GOLD, "dequant_gguf_py.npz"))
    blocks, want = g[f"{name}_blocks"], g[f"{name}_dequant_bits"].reshape(-1)
    got = engine.dequantize(qt, blocks, want.size).view(np.uint32)
    gf, wf = got.view(np.float32), want.view(np.float32)
    assert np.all((got == want) | (np.isnan(gf) & np.isnan(wf)))


def test_dequantize_rejects_ragged(engine):
    with pytest.raises(capi.AiosError) as ei:
        engine.dequantize(gs.Q4_K, np.zeros(144, np.uint8), 100)
    assert ei.value.code == capi.E_INVAL
    assert engine.dequantize(gs.Q4_K, np.zeros(0, np.uint8), 0).size == 0  # empty input


@pytest.mark.parametrize("qt", QTYPES)
@pytest.mark.parametrize("shape", [(1, 256), (37, 512), (64, 2048), (300, 4096), (50, 5632), (33, 14336)])
def test_matvec_vs_oracle(engine, oracle, qt, shape):
    n, k = shape
    rng = np.random.default_rng([qt, n, k])
    w = gs.random_blocks(qt, n * k // gs.BLOCK[qt][0], rng, 0.02)
    x = rng.standard_normal(k, dtype=np.float32)
    x[:7] = 0  # some exact zeros
    ref = oracle.matvec(qt, w, n, k, x)
    got = engine.matvec(qt, w, n, k, x)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), np.abs(got - ref).max()


@pytest.mark.parametrize("qt", [gs.Q4_K, gs.Q6_K, gs.Q8_0])
def test_matvec_all_zero_and_signed_ties(engine, oracle, qt):
    n, k = 16, 512
    rng = np.random.default_rng(3)
    w = gs.random_blocks(qt, n * k // gs.BLOCK[qt][0], rng, 0.02)
    x = np.zeros(k, np.float32)
    assert np.array_equal(engine.matvec(qt, w, n, k, x), np.zeros(n, np.float32))
    # +a and -a tie for amax: the first one decides the sign of the Q8_K scale
    x = rng.standard_normal(k, dtype=np.float32) * 0.1
    x[5], x[200] = -3.0, 3.0
    x[300], x[400] = 2.5, -2.5
    ref = oracle.matvec(qt, w, n, k, x)
    got = engine.matvec(qt, w, n, k, x)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))


@pytest.mark.parametrize("qt", [gs.Q4_K, gs.Q6_K, gs.Q8_0])
def test_matvec_with_rmsnorm(engine, oracle, qt):
    n, k = 96, 2048
    rng = np.random.default_rng(11)
    w = gs.random_blocks(qt, n * k // gs.BLOCK[qt][0], rng, 0.02)
    x = rng.standard_normal(k, dtype=np.float32) * 3
    nw = (rng.random(k, dtype=np.float32) * 0.4 + 0.8).astype(np.float32)
    eps = 1e-5
    ss = np.sum((x * x).astype(np.float64))  # f32 squares summed in double, like ggml_rms_norm
    scale = np.float32(1.0) / np.sqrt(np.float32(ss / k) + np.float32(eps), dtype=np.float32)
    xn = ((x * scale).astype(np.float32) * nw).astype(np.float32)
    ref = oracle.matvec(qt, w, n, k, xn)
    got = engine.matvec(qt, w, n, k, x, norm_w=nw, eps=eps)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), np.abs(got - ref).max()


@pytest.mark.parametrize("arch,ftype", [("tiny", "Q4_K_M"), ("tiny", "Q8_0"), ("tiny-gqa4", "Q4_K_M")])
def test_model_logits_and_kv(engine, oracle, tmp_models, arch, ftype):
    path = tmp_models(arch, ftype)
    m = engine.load_model("t", path, n_ctx=128)
    try:
        om = oracle.OracleModel(path, n_ctx=128)
        prompt = gs.synthetic_prompt(arch, 24, 5)
        ref = om.eval_all(prompt)
        seq = m.new_seq()
        try:
            for i in range(len(prompt)):
                seq.prefill(prompt[i:i + 1])
                got = seq.logits()
                err = np.abs(got - ref[i]).max() / ref[i].std()
                assert err < LOGIT_TOL, (i, err)
            # the f16 KV cache must hold exactly the oracle's bits
            import ctypes as C
            L = oracle.lib()
            L.oracle_k_cache.restype = C.POINTER(C.c_uint16)
            L.oracle_k_cache.argtypes = [C.c_void_p]
            L.oracle_v_cache.restype = C.POINTER(C.c_uint16)
            L.oracle_v_cache.argtypes = [C.c_void_p]
            a = gs.ARCHS[arch]
            kvd = a.n_head_kv * a.head_dim
            ok = np.ctypeslib.as_array(L.oracle_k_cache(om.h), shape=(a.n_layer, 128, kvd))
            ov = np.ctypeslib.as_array(L.oracle_v_cache(om.h), shape=(a.n_layer, 128, kvd))
            for il in range(a.n_layer):
                k, v = seq.read_kv(il, 0, len(prompt))
                assert np.array_equal(k, ok[il, :len(prompt)]) and np.array_equal(v, ov[il, :len(prompt)])
        finally:
            seq.free()
        # whole prompt in one call == token by token
        got = m.forward_logits(prompt)
        assert np.abs(got - ref[-1]).max() / ref[-1].std() < LOGIT_TOL
    finally:
        m.unload()


@pytest.mark.parametrize("arch,ftype", [("tiny", "Q4_K_M"), ("tiny-gqa4", "Q4_K_M"), ("tiny", "Q8_0")])
def test_greedy_tokens_match_oracle(engine, oracle, tmp_models, arch, ftype):
    path = tmp_models(arch, ftype)
    m = engine.load_model("t", path, n_ctx=256)
    try:
        om = oracle.OracleModel(path, n_ctx=256)
        prompt = gs.synthetic_prompt(arch, 32, 9)
        n_gen = 128
        ref, margins = om.greedy(prompt, n_gen)
        got = m.generate(prompt, n_gen)
        print(f"min top1-top2 margin {margins.min():.3e}")
        assert got.shape == ref.shape
        assert np.array_equal(got, ref), (np.nonzero(got != ref)[0][:4], margins.min())
    finally:
        m.unload()


def test_decode_stops_at_context_and_steps_agree(engine, tmp_models):
    path = tmp_models("tiny", "Q4_K_M")
    m = engine.load_model("t", path, n_ctx=48)
    try:
        prompt = gs.synthetic_prompt("tiny", 40, 1)
        a = m.generate(prompt, 100)
        assert a.size == 8  # clamped: n_prompt + generated <= n_ctx
        s = m.new_seq()
        try:
            s.prefill(prompt)
            b = np.array([s.step() for _ in range(8)], np.int32)
            assert np.array_equal(a, b)
            with pytest.raises(capi.AiosError):
                s.step()
        finally:
            s.free()
    finally:
        m.unload()


def test_load_errors(engine, tmp_path):
    with pytest.raises(capi.AiosError) as ei:
        engine.load_model("x", str(tmp_path / "missing.gguf"))
    assert ei.value.code == capi.E_IO
    p = tmp_path / "junk.gguf"
    p.write_bytes(b"not a gguf file at all, just some bytes....")
    with pytest.raises(capi.AiosError) as ei:
        engine.load_model("x", str(p))
    assert ei.value.code == capi.E_FORMAT


def test_loader_matches_gguf_py(engine, tmp_models):
    from gguf import GGUFReader

    path = tmp_models("tiny", "Q4_K_M")
    m = engine.load_model("tiny", path, n_ctx=64)
    try:
        r = GGUFReader(path)
        assert m.info.n_tensors == len(r.tensors)
        for i, t in enumerate(r.tensors):
            ti = m.tensor_info(i)
            assert ti.name.decode() == t.name
            assert ti.type == int(t.tensor_type)
            assert list(ti.ne)[: ti.n_dims] == [int(x) for x in t.shape]
            assert ti.offset == int(t.data_offset) - int(r.data_offset)
            assert ti.nbytes == int(t.n_bytes)
        a = gs.ARCHS["tiny"]
        assert (m.info.n_layer, m.info.n_embd, m.info.n_head, m.info.n_head_kv, m.info.n_ff, m.info.n_vocab) == \
               (a.n_layer, a.n_embd, a.n_head, a.n_head_kv, a.n_ff, a.n_vocab)
        assert m.info.weight_bytes_per_token == gs.weight_bytes_per_token("tiny", "Q4_K_M")
        assert m.info.kv_bytes_per_position == gs.kv_bytes_per_position("tiny")
    finally:
        m.unload()


@pytest.mark.parametrize("arch,ftype", [("tiny", "Q4_K_M"), ("tiny-gqa4", "Q4_K_M"), ("tiny", "Q8_0")])
def test_fused_decode_kernel_matches_oracle_and_per_op_path(engine, engine_pfd, oracle, tmp_models, arch, ftype):
    """The persistent fused decode kernel (one cooperative launch for all tokens) must give the oracle's tokens,
    the same logits bit for bit as the per-op kernels, and leave the sequence usable for further steps."""
    path = tmp_models(arch, ftype)
    om = oracle.OracleModel(path, n_ctx=256)
    prompt = gs.synthetic_prompt(arch, 20, 11)
    ref, margins = om.greedy(prompt, 96)
    m1 = engine.load_model("a", path, n_ctx=256)
    m2 = engine_pfd.load_model("b", path, n_ctx=256)
    try:
        s1, s2 = m1.new_seq(), m2.new_seq()
        try:
            s1.prefill(prompt)
            s2.prefill(prompt)
            a = s1.decode(64)
            b = s2.decode(64)            # one fused launch, 64 tokens
            r = s2.bench_decode(8)       # another fused launch continuing from there
            assert r.kernel_launches == 1, "the fused kernel was not used"
            s1.decode(8)
            assert np.array_equal(a, ref[:64]) and np.array_equal(b, ref[:64]), margins.min()
            assert np.array_equal(s1.logits().view(np.uint32), s2.logits().view(np.uint32))
            c1 = np.array([s1.step() for _ in range(4)])
            c2 = np.array([s2.step() for _ in range(4)])  # single-token fused launches
            assert np.array_equal(c1, ref[72:76]) and np.array_equal(c2, ref[72:76])
            for il in range(gs.ARCHS[arch].n_layer):
                k1, v1 = s1.read_kv(il, 0, 96)
                k2, v2 = s2.read_kv(il, 0, 96)
                assert np.array_equal(k1, k2) and np.array_equal(v1, v2)
        finally:
            s1.free()
            s2.free()
    finally:
        m1.unload()
        m2.unload()


@pytest.mark.parametrize("attn", ["split", "cluster"])
def test_long_context_attention_path_matches_oracle(oracle, tmp_models, attn):
    """Both decode-attention kernels — one wave of (head, dim-slice) CTAs (attention_split.cu, the default) and the
    cluster/DSMEM kernel (attention.cu, long contexts) — must give the oracle's logits."""
    from aios_b200 import capi
    path = tmp_models("tiny-gqa4", "Q4_K_M")
    flags = {"split": 0, "cluster": capi.ENGINE_ATTN_CLUSTER}[attn]
    engine = capi.Engine(device=0, max_batch=1, flags=flags)
    m = engine.load_model("t", path, n_ctx=1536)
    try:
        om = oracle.OracleModel(path, n_ctx=1536)
        prompt = gs.synthetic_prompt("tiny-gqa4", 40, 2)
        ref = om.eval_all(prompt)
        seq = m.new_seq()
        try:
            for i in range(len(prompt)):
                seq.prefill(prompt[i:i + 1])
                if i % 5 == 4 or i == len(prompt) - 1:
                    got = seq.logits()
                    assert np.array_equal(got.view(np.uint32), ref[i].view(np.uint32)), i  # bit for bit
        finally:
            seq.free()
    finally:
        m.unload()
        engine.close()


@pytest.mark.parametrize("flag", ["ENGINE_NO_ROPE_EPILOGUE", "ENGINE_NO_PDL", "ENGINE_NO_GRAPH"])
def test_engine_variants_are_bit_identical(engine, oracle, tmp_models, flag):
    """Every switchable variant of the token pass — RoPE / KV append in the attention kernel instead of the q/k/v
    GEMV's epilogue, no programmatic dependent launch, no CUDA graph — computes the same arithmetic: oracle tokens,
    logits equal bit for bit to the default engine's."""
    from aios_b200 import capi
    arch = "tiny-gqa4"
    path = tmp_models(arch, "Q4_K_M")
    om = oracle.OracleModel(path, n_ctx=128)
    prompt = gs.synthetic_prompt(arch, 10, 21)
    ref, margins = om.greedy(prompt, 32)
    other = capi.Engine(device=0, max_batch=1, flags=getattr(capi, flag))
    m1, m2 = engine.load_model("a", path, n_ctx=128), other.load_model("b", path, n_ctx=128)
    try:
        s1, s2 = m1.new_seq(), m2.new_seq()
        try:
            s1.prefill(prompt)
            s2.prefill(prompt)
            a, b = s1.decode(32), s2.decode(32)
            assert np.array_equal(a, ref) and np.array_equal(b, ref), margins.min()
            assert np.array_equal(s1.logits().view(np.uint32), s2.logits().view(np.uint32))
            for il in range(gs.ARCHS[arch].n_layer):
                k1, v1 = s1.read_kv(il, 0, 42)
                k2, v2 = s2.read_kv(il, 0, 42)
                assert np.array_equal(k1, k2) and np.array_equal(v1, v2)
        finally:
            s1.free()
            s2.free()
    finally:
        m1.unload()
        m2.unload()
        other.close()


@pytest.mark.parametrize("arch,n_prompt,n_ctx,long_min", [("tiny-gqa4", 700, 1024, 1), ("tiny", 37, 1024, 1), ("tiny-mha32", 150, 1024, 1),
                                                          ("tiny-gqa4", 300, 5000, None), ("tiny", 90, 6000, 1),
                                                          ("tiny-gqa4", 300, 1024, 325), ("tiny-gqa4", 300, 5000, 325),
                                                          ("tiny-gqa8", 200, 1024, 1)])
def test_long_context_attention_is_bit_identical(oracle, tmp_models, monkeypatch, arch, n_prompt, n_ctx, long_min):
    """attention_long.cu (CTA = kv head x slice of the positions, GQA heads sharing K / V rows, two launches) must give
    the latency kernel's and the oracle's results bit for bit: sums of exact f16 x f16 products in double do not care
    how the positions are cut.  AIOS_ATTN_LONG_MIN = 1 forces it from the first decoded token; the second decode call
    crosses a slice-size boundary (slices are multiples of 32 rows).  A model loaded with n_ctx beyond the latency
    kernel's 4096 rows starts on the latency kernel (sized for 4096 rows) and switches at the default threshold: the
    (300, 5000, default) case runs that configuration as loaded, below the threshold; a threshold of 325 makes the
    decode calls switch kernels (and captured graphs) in mid-call."""
    path = tmp_models(arch, "Q4_K_M")
    prompt = gs.synthetic_prompt(arch, n_prompt, 13)
    om = oracle.OracleModel(path, n_ctx=1024)
    ref, _ = om.greedy(prompt, 48)
    if long_min is not None:
        monkeypatch.setenv("AIOS_ATTN_LONG_MIN", str(long_min))
    for flags in (0, capi.ENGINE_NO_GRAPH):
        eng = capi.Engine(device=0, max_batch=1, flags=flags | capi.ENGINE_EXACT_PREFILL)
        try:
            m = eng.load_model("long", path, n_ctx=n_ctx)
            try:
                s = m.new_seq()
                s.prefill(prompt)
                got = np.concatenate([s.decode(20), s.decode(28)])
                assert np.array_equal(got, ref), (flags, got, ref)
                want_logits = om.eval(np.concatenate([prompt, ref[:-1]]))
                # the logits after feeding the 47th generated token (what the 48th was sampled from)
                s.free()
                s = m.new_seq()
                s.prefill(prompt)
                s.decode(47)
                assert np.array_equal(s.logits().view(np.uint32), want_logits.view(np.uint32))
                s.free()
            finally:
                m.unload()
        finally:
            eng.close()
