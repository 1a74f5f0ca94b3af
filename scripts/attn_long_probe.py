"""Decode ms/token with the latency attention kernel vs the position-split one, by context length:
python scripts/attn_long_probe.py [arch]   (sets AIOS_ATTN_LONG_MIN per engine)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs  # noqa: E402
import bench  # noqa: E402

arch = sys.argv[1] if len(sys.argv) > 1 else "mistral-7b"
path = bench.ensure_model(arch, "Q4_K_M", True)
n_ctx = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
for ctx in [c for c in (256, 512, 1024, 2048, 4096, 8192) if c <= n_ctx]:
    row = []
    for mode, thr in (("split", "1000000"), ("long", "1")):
        os.environ["AIOS_ATTN_LONG_MIN"] = thr
        eng = capi.Engine(0, 1, 0)
        m = eng.load_model("m", path, n_ctx=n_ctx)
        s = m.new_seq()
        s.prefill(gs.synthetic_prompt(arch, ctx - 64, 5))
        s.bench_decode(8)
        s.free()
        s = m.new_seq()
        s.prefill(gs.synthetic_prompt(arch, ctx - 64, 5))
        r = s.bench_decode(64)
        row.append(r.ms_total / r.n_tokens)
        s.free(); m.unload(); eng.close()
    print(f"{arch} ctx {ctx}: split {row[0]:.3f} ms/token, long {row[1]:.3f} ms/token", flush=True)
