import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
from aios_b200 import capi, gguf_synth as gs
import bench
for arch in ("mistral-7b", "tinyllama-1.1b"):
    path = bench.ensure_model(arch, "Q4_K_M", True)
    prompts = np.stack([gs.synthetic_prompt(arch, 128, 7 + i) for i in range(64)])
    for rep in range(2):
        for flags in (0, 2):
            eng = capi.Engine(0, 1, flags)
            m = eng.load_model("m", path, n_ctx=512)
            m.bench_batch_decode(prompts[:, :8], 4, want_tokens=False)
            r, _ = m.bench_batch_decode(prompts, 96, want_tokens=False)
            print(f"{arch} flags {flags}: {r.ms_decode / r.n_steps_timed:.3f} ms/step", flush=True)
            m.unload(); eng.close()
