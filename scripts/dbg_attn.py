import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aios_b200 import capi, gguf_synth as gs
from oracle import oracle as orc
arch="tiny1"; a = gs.ARCHS[arch]
path = f"/tmp/{arch}.gguf"
gs.write_synthetic_gguf(path, arch, "Q4_K_M", 7)
eng = capi.Engine(0, 2)
m = eng.load_model("t", path, n_ctx=128)
prompt = gs.synthetic_prompt(arch, 6, 5)
seq = m.new_seq()
H, G, D = a.n_head, a.n_head_kv, a.head_dim
for i in range(len(prompt)):
    seq.prefill(prompt[i:i+1])
    qkv = seq.read_activation(1, (H+2*G)*D)
    att = seq.read_activation(2, H*D)
    k, v = seq.read_kv(0, 0, i+1)
    K = k.view(np.float16).astype(np.float32).reshape(i+1, G, D)
    V = v.view(np.float16).astype(np.float32).reshape(i+1, G, D)
    c, s = orc.rope_table(i, D, a.rope_freq_base)
    q = qkv[:H*D].reshape(H, D).copy()
    q0, q1 = q[:, 0::2].copy(), q[:, 1::2].copy()
    q[:, 0::2] = q0*c - q1*s; q[:, 1::2] = q0*s + q1*c
    q = q.astype(np.float16).astype(np.float32)
    exp = np.zeros((H, D), np.float32)
    for h in range(H):
        g = h // (H//G)
        sc = (K[:, g, :] @ q[h]) / np.sqrt(D)
        p = np.exp(sc - sc.max()); p /= p.sum()
        exp[h] = p @ V[:, g, :]
    print(i, "att maxdiff %.3e of max %.3e" % (np.abs(att.reshape(H,D)-exp).max(), np.abs(exp).max()), "per head", np.abs(att.reshape(H,D)-exp).max(axis=1))
