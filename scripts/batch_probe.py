"""Launch list of ONE batched decode step (ncu --profile-from-start off): python scripts/batch_probe.py [arch] [lanes]"""
import ctypes
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs  # noqa: E402
import bench  # noqa: E402

arch = sys.argv[1] if len(sys.argv) > 1 else "mistral-7b"
lanes = int(sys.argv[2]) if len(sys.argv) > 2 else 64
path = bench.ensure_model(arch, "Q4_K_M", True)
eng = capi.Engine(0, 1, capi.ENGINE_NO_GRAPH | int(os.environ.get("AIOS_PROBE_FLAGS", "0")))
m = eng.load_model("m", path, n_ctx=512)
prompts = np.stack([gs.synthetic_prompt(arch, 128, 7 + i) for i in range(lanes)])
rt = ctypes.CDLL("libcudart.so.12")
m.bench_batch_decode(prompts[:, :8], 3, want_tokens=False)
os.environ["AIOS_PROFILE_STEPS"] = "1"
r, _ = m.bench_batch_decode(prompts, 4, want_tokens=False)
print(f"{arch} lanes {lanes}: {r.ms_decode / r.n_steps_timed:.3f} ms/step, prefill {r.ms_prefill:.0f} ms")
m.unload()
eng.close()
