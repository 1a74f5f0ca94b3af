"""Isolated GEMV bandwidth: one launch shape, several weight copies cycled (working set > L2), back-to-back launches."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs
eng = capi.Engine(0, 1)
shapes = [(gs.Q4_K, 28672, 4096), (gs.Q4_K, 4096, 4096), (gs.Q4_K, 6144, 4096), (gs.Q4_K, 4096, 14336),
          (gs.Q6_K, 4096, 14336), (gs.Q6_K, 32000, 4096), (gs.Q8_0, 14336, 4096)]
if len(sys.argv) > 1:
    shapes = [shapes[int(a)] for a in sys.argv[1:]]
for qt, n, k in shapes:
    nbytes = gs.tensor_nbytes((n, k), qt)
    copies = max(2, int(400e6 // nbytes) + 1)
    ms = eng.bench_gemv(qt, n, k, copies, 200)
    print(f"{gs.TYPE_NAME[qt]} {n}x{k}: {nbytes/1e6:.1f} MB, {ms*1e3:.2f} us/launch, {nbytes/ms/1e6:.0f} GB/s ({nbytes/ms/1e6/6540.8*100:.1f}% of measured peak)")
