import os, sys
sys.path.insert(0, "/root/repo")
from aios_b200 import capi, gguf_synth as gs
gs.write_synthetic_gguf("/tmp/t.gguf", "tiny", "Q4_K_M", 7)
eng = capi.Engine(0, 1, capi.ENGINE_USE_PFD)
m = eng.load_model("m", "/tmp/t.gguf", n_ctx=128)
seq = m.new_seq()
seq.prefill(gs.synthetic_prompt("tiny", 8, 1))
r = seq.bench_decode(4)
print("launches", r.kernel_launches, "ms", r.ms_total)
