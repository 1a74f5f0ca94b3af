"""A/B of an engine environment switch on the headline workload, same process: python scripts/ab_env.py VAR v1 v2 ..."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs  # noqa: E402
import bench  # noqa: E402

var, vals = sys.argv[1], sys.argv[2:]
path = bench.ensure_model("mistral-7b", "Q4_K_M", True)
prompt = gs.synthetic_prompt("mistral-7b", 384, bench.SEED)
for rnd in range(2):
    for v in vals:
        os.environ[var] = v
        eng = capi.Engine(0, 1, capi.ENGINE_EXACT_PREFILL)
        m = eng.load_model("m", path, n_ctx=512)
        seq = m.new_seq()
        seq.prefill(prompt)
        seq.snapshot()
        for _ in range(2):
            seq.restore(); seq.bench_decode(128)
        ms = []
        for _ in range(4):
            seq.restore(); ms.append(seq.bench_decode(128).ms_total / 128)
        print(f"round {rnd} {var}={v}: {min(ms):.4f} ms/token ({1e3 / min(ms):.1f} tok/s)", flush=True)
        seq.free(); m.unload(); eng.close()
