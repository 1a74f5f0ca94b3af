"""Prefill throughput: chunked (tensor-core GEMM) vs token by token.
usage: python scripts/prefill_bench.py [arch] [ftype] [n_tokens]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs  # noqa: E402
import bench  # noqa: E402

arch = sys.argv[1] if len(sys.argv) > 1 else "mistral-7b"
ftype = sys.argv[2] if len(sys.argv) > 2 else "Q4_K_M"
n = int(sys.argv[3]) if len(sys.argv) > 3 else 512
path = bench.ensure_model(arch, ftype, True)
prompt = gs.synthetic_prompt(arch, n, bench.SEED)
for name, flags in (("chunked_gemm", 0), ("token_by_token", capi.ENGINE_NO_GEMM)):
    eng = capi.Engine(0, 1, flags)
    m = eng.load_model("m", path, n_ctx=n + 8)
    seq = m.new_seq()
    seq.prefill(prompt[:64])  # warm-up (allocations, first launches)
    best = 1e9
    for _ in range(2):
        seq.reset()
        t0 = time.perf_counter()
        seq.prefill(prompt)  # synchronises before returning
        best = min(best, time.perf_counter() - t0)
    print(f"{arch} {ftype} {name:16s} {n} tokens in {best * 1e3:8.1f} ms  = {n / best:8.1f} tok/s", flush=True)
    seq.free()
    m.unload()
    eng.close()
