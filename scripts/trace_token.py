"""Per-launch timeline of one decode token on the headline workload (aios_seq_trace_decode).
usage: python scripts/trace_token.py [flags] [first_launch] [n_launches]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs  # noqa: E402
import bench  # noqa: E402

flags = int(sys.argv[1]) if len(sys.argv) > 1 else 0
first = int(sys.argv[2]) if len(sys.argv) > 2 else 6
count = int(sys.argv[3]) if len(sys.argv) > 3 else 16
path = bench.ensure_model("mistral-7b", "Q4_K_M", True)
eng = capi.Engine(0, 1, flags)
m = eng.load_model("m", path, n_ctx=512)
seq = m.new_seq()
seq.prefill(gs.synthetic_prompt("mistral-7b", 384, bench.SEED))
seq.decode(4)
tr = seq.trace_decode().astype(np.int64)
nz = tr[:, 0] > 0
t0 = tr[nz, 0].min()
print(f"{int(nz.sum())} traced launches of {len(tr)}; token span {(tr[nz, 4].max() - t0) / 1000:.1f} us")
print("launch  entry  wait_over  act_ready  first_tile  done | prod_first prod_last   (us since the first traced entry)")
names = {0: "embed"}
for i in range(first, min(first + count, len(tr))):
    r = tr[i]
    if r[0] == 0:
        print(f"{i:4d}   (not traced)")
        continue
    f = lambda v: f"{(v - t0) / 1000:9.2f}" if v > 0 else "        -"
    print(f"{i:4d} {f(r[0])} {f(r[1])} {f(r[2])} {f(r[3])} {f(r[4])} | {f(r[5])} {f(r[6])} | last CTA done {f(r[7])}")
# aggregate: per kernel (entry -> done), waits, and gaps between one kernel's done and the next one's wait_over
ent, wo, ar, ft, dn = (tr[nz, j] for j in range(5))
print(f"mean us: entry->wait_over {np.mean(wo - ent) / 1000:.2f}  wait_over->act_ready {np.mean((ar - wo)[ar > 0]) / 1000:.2f}  "
      f"act_ready->done {np.mean((dn - ar)[ar > 0]) / 1000:.2f}  done->next wait_over {np.mean(wo[1:] - dn[:-1]) / 1000:.2f}")
seq.free()
m.unload()
eng.close()
