import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs
import bench
path = bench.ensure_model("mistral-7b", "Q4_K_M", True)
eng = capi.Engine(0, 1, capi.ENGINE_NO_GRAPH)
m = eng.load_model("m", path, n_ctx=512)
seq = m.new_seq()
seq.prefill(gs.synthetic_prompt("mistral-7b", int(sys.argv[1]) if len(sys.argv) > 1 else 8, bench.SEED))
print(seq.decode(2))
