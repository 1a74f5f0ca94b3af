"""Per-phase timeline of the persistent fused decode kernel (CTA 0, last token of a launch)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs
import bench
path = bench.ensure_model("mistral-7b", "Q4_K_M", True)
eng = capi.Engine(0, 1, capi.ENGINE_USE_PFD)
m = eng.load_model("m", path, n_ctx=512)
seq = m.new_seq()
seq.prefill(gs.synthetic_prompt("mistral-7b", 384, bench.SEED))
seq.snapshot()
for _ in range(2):
    seq.restore(); r = seq.bench_decode(32)
print("ms/token", r.ms_total / r.n_tokens)
t = seq.phase_timing().astype(np.int64)
names = ["q", "k", "v", "attn", "o", "gateup", "down"]
tot = {}
for ph in range(len(t)):
    nm = names[ph % 7] if ph < len(t) - 1 else "lm_head"
    pro, comp, bar = t[ph, 1] - t[ph, 0], t[ph, 2] - t[ph, 1], t[ph, 3] - t[ph, 2]
    if nm == "attn": pro, comp = 0, t[ph, 2] - t[ph, 0]
    a = tot.setdefault(nm, [0, 0, 0, 0]); a[0] += pro; a[1] += comp; a[2] += bar; a[3] += 1
print("phase      n   prologue_us  compute_us  barrier_us   (averages, CTA 0)")
for nm, a in tot.items():
    print(f"{nm:8s} {a[3]:3d}   {a[0]/a[3]/1e3:9.2f}  {a[1]/a[3]/1e3:9.2f}  {a[2]/a[3]/1e3:9.2f}")
print("token total us", (t[-1, 3] - t[0, 0]) / 1e3)
