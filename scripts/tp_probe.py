"""Tensor-parallel decode probe: python scripts/tp_probe.py <arch> <n_gpus> [n_tokens] [ctx]

Writes the synthetic GGUF (random-init weights of the named shape), loads it split over n_gpus, checks that a second
load-free request repeats the first one's tokens, and times greedy decode.  One JSON line on stdout.
"""
import json
import os
import shutil
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs  # noqa: E402
import bench  # noqa: E402

arch = sys.argv[1] if len(sys.argv) > 1 else "mistral-7b"
n_gpus = int(sys.argv[2]) if len(sys.argv) > 2 else 2
n_tokens = int(sys.argv[3]) if len(sys.argv) > 3 else 128
ctx = int(sys.argv[4]) if len(sys.argv) > 4 else 512
ftype = "Q4_K_M"

path = bench.model_path(arch, ftype)
if not os.path.exists(path):
    a = gs.ARCHS[arch]
    approx = (a.n_layer * (2 * a.n_embd * a.n_head * a.head_dim + 2 * a.n_embd * a.n_head_kv * a.head_dim + 3 * a.n_embd * a.n_ff)
              + 2 * a.n_vocab * a.n_embd) * 0.6
    free = shutil.disk_usage(os.path.dirname(path)).free
    if free < approx * 1.15:
        print(json.dumps({"arch": arch, "unavailable": f"needs ~{approx / 1e9:.0f} GB for the GGUF, {free / 1e9:.0f} GB free"}))
        sys.exit(0)
t0 = time.time()
path = bench.ensure_model(arch, ftype, True)
t_write = time.time() - t0
t0 = time.time()
tp = capi.TpModel(list(range(n_gpus)), arch, path, n_ctx=ctx)
t_load = time.time() - t0
prompt = gs.synthetic_prompt(arch, 32, 7)
first = tp.generate(prompt, 16)
again = tp.generate(prompt, 16)
tp.reset()
t0 = time.time()
tp.prefill(gs.synthetic_prompt(arch, ctx // 2, 11))
t_prefill = time.time() - t0
r = tp.bench_decode(n_tokens)
rec = {"arch": arch, "ftype": ftype, "tp": n_gpus, "ctx": ctx, "n_tokens": r.n_tokens, "ms_per_token": r.ms_total / r.n_tokens,
       "tok_per_s": 1e3 * r.n_tokens / r.ms_total, "launches_per_token": r.kernel_launches / r.n_tokens,
       "allreduce": "push" if r.allreduce_kind else "nccl", "graphs": bool(r.graphs), "allreduce_per_token": r.n_allreduce_per_token, "us_per_allreduce": r.us_per_allreduce,
       "allreduce_share": r.n_allreduce_per_token * r.us_per_allreduce * 1e-3 / (r.ms_total / r.n_tokens),
       "n_past": [r.n_past_begin, r.n_past_end], "weight_gb_per_rank": tp.info.weight_bytes / 1e9,
       "prefill_tokens": ctx // 2, "prefill_s": round(t_prefill, 3), "prefill_tok_s": round((ctx // 2) / t_prefill, 1),
       "repeatable": bool(np.array_equal(first, again)), "write_s": round(t_write, 1), "load_s": round(t_load, 1)}
print(json.dumps(rec))
tp.unload()
if "--keep" not in sys.argv and os.path.getsize(path) > 8e9:
    os.remove(path)
