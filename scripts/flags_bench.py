"""A/B of engine flags on the headline workload (Mistral-7B Q4_K_M, bs 1, positions 384..512).
usage: python scripts/flags_bench.py [name=flags ...]   e.g.  default=0 attn_head=32"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs  # noqa: E402
import bench  # noqa: E402

cases = [a.split("=") for a in sys.argv[1:]] or [("default", "0"), ("attn_head", str(capi.ENGINE_ATTN_HEAD))]
path = bench.ensure_model("mistral-7b", "Q4_K_M", True)
for name, flags in cases:
    eng = capi.Engine(0, 1, int(flags))
    m = eng.load_model("m", path, n_ctx=512)
    seq = m.new_seq()
    seq.prefill(gs.synthetic_prompt("mistral-7b", 384, bench.SEED))
    seq.snapshot()
    best = 1e9
    for _ in range(3):
        seq.restore()
        r = seq.bench_decode(128)
        best = min(best, r.ms_total / r.n_tokens)
    seq.restore()
    pr = seq.profile_decode(16)
    print(f"{name:16s} ms/token {best:.4f}  tok/s {1000 / best:.1f}  launches/token {r.kernel_launches / r.n_tokens:.1f}  "
          f"| per-kernel timing: gemv {pr.ms_gemv / pr.n_tokens:.3f} ms  attention {pr.reserved[0] / pr.n_tokens:.3f} ms  "
          f"other {pr.reserved[1] / pr.n_tokens:.3f} ms", flush=True)
    seq.free()
    m.unload()
    eng.close()
