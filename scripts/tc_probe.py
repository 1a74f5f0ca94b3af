"""ms of one 4096-token prefill on the tensor-core path: python scripts/tc_probe.py [Q8_0|Q4_K_M] [tokens] [arch]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs  # noqa: E402
import bench  # noqa: E402

ftype = sys.argv[1] if len(sys.argv) > 1 else "Q8_0"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
arch = sys.argv[3] if len(sys.argv) > 3 else "mistral-7b"
path = bench.ensure_model(arch, ftype, True)
eng = capi.Engine(0, 1, 0)
m = eng.load_model("m", path, n_ctx=n + 8)
s = m.new_seq()
ms = s.bench_prefill(gs.synthetic_prompt(arch, n, 5), 3)
print(f"{arch} {ftype} prefill {n} tokens: {ms:.2f} ms = {n / ms * 1e3:.0f} tok/s")
if os.path.getsize(path) > 8e9:
    os.remove(path)
