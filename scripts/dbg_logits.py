import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs
from oracle import oracle as orc
arch = sys.argv[1] if len(sys.argv) > 1 else "tiny"
ftype = sys.argv[2] if len(sys.argv) > 2 else "Q4_K_M"
path = f"/tmp/{arch}-{ftype}.gguf"
gs.write_synthetic_gguf(path, arch, ftype, 7)
eng = capi.Engine(0, 2)
m = eng.load_model("t", path, n_ctx=128)
om = orc.OracleModel(path, n_ctx=128)
oms = orc.OracleModel(path, n_ctx=128, simd=True)
prompt = gs.synthetic_prompt(arch, 40, 5)
ref = om.eval_all(prompt)
refs = oms.eval_all(prompt)
seq = m.new_seq()
for i in range(len(prompt)):
    seq.prefill(prompt[i:i+1])
    got = seq.logits()
    print(i, "gpu-vs-oracle %.3e" % (np.abs(got-ref[i]).max()/ref[i].std()), "simd-vs-generic %.3e" % (np.abs(refs[i]-ref[i]).max()/ref[i].std()))
import ctypes as C
L = orc.lib()
L.oracle_k_cache.restype = C.POINTER(C.c_uint16)
L.oracle_k_cache.argtypes = [C.c_void_p]
L.oracle_v_cache.restype = C.POINTER(C.c_uint16)
L.oracle_v_cache.argtypes = [C.c_void_p]
a = gs.ARCHS[arch]
kvd = a.n_head_kv * a.head_dim
n_ctx = 128
ok = np.ctypeslib.as_array(L.oracle_k_cache(om.h), shape=(a.n_layer, n_ctx, kvd))
ov = np.ctypeslib.as_array(L.oracle_v_cache(om.h), shape=(a.n_layer, n_ctx, kvd))
for il in range(a.n_layer):
    k, v = seq.read_kv(il, 0, 4)
    for pos in range(4):
        kd = (k[pos].view(np.float16).astype(np.float32) - ok[il, pos].view(np.float16).astype(np.float32))
        vd = (v[pos].view(np.float16).astype(np.float32) - ov[il, pos].view(np.float16).astype(np.float32))
        print("layer", il, "pos", pos, "k maxdiff %.3e (k max %.3e)" % (np.abs(kd).max(), np.abs(ok[il,pos].view(np.float16).astype(np.float32)).max()), "v maxdiff %.3e" % np.abs(vd).max())
