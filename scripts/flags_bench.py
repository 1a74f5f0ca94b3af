import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs
import bench
path = bench.ensure_model("mistral-7b", "Q4_K_M", True)
for name, flags in [("default", 0), ("no_l2_prefetch", capi.ENGINE_NO_L2_PREFETCH), ("no_pdl", capi.ENGINE_NO_PDL)]:
    eng = capi.Engine(0, 1, flags)
    m = eng.load_model("m", path, n_ctx=512)
    seq = m.new_seq()
    seq.prefill(gs.synthetic_prompt("mistral-7b", 384, bench.SEED))
    seq.snapshot()
    for _ in range(2):
        seq.restore(); r = seq.bench_decode(128)
    print(name, "ms/token", r.ms_total / r.n_tokens)
    seq.free(); m.unload(); eng.close()
