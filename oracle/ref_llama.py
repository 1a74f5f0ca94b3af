"""Exact-math reference of the llama architecture in numpy float64 (test infrastructure): dequantised GGUF weights, no
activation quantisation, no rounding of K/V or probabilities.  What the tolerance tests of the tensor-core prefill path
compare with (SURVEY.md §8c tier (ii): "exact-math logits without activation quantisation")."""
import numpy as np


def load_weights(path: str, oracle):
    from gguf import GGUFReader
    r = GGUFReader(path)
    f = r.fields

    def u(key, default=None):
        if key not in f:
            return default
        fld = f[key]
        return fld.parts[fld.data[0]][0]

    hp = dict(n_layer=int(u("llama.block_count")), n_embd=int(u("llama.embedding_length")),
              n_head=int(u("llama.attention.head_count")), n_ff=int(u("llama.feed_forward_length")),
              eps=float(u("llama.attention.layer_norm_rms_epsilon", 1e-5)), base=float(u("llama.rope.freq_base", 10000.0)))
    hp["n_kv"] = int(u("llama.attention.head_count_kv", hp["n_head"]))
    hp["hd"] = hp["n_embd"] // hp["n_head"]
    hp["n_rot"] = int(u("llama.rope.dimension_count", hp["hd"]))
    w = {}
    for t in r.tensors:
        shape = [int(x) for x in t.shape]  # ne0 first
        if int(t.tensor_type) == 0:
            a = np.array(t.data, dtype=np.float32).reshape(-1)
        else:
            a = oracle.dequantize(int(t.tensor_type), np.array(t.data).reshape(-1))
        w[t.name] = a.astype(np.float64).reshape(shape[1], shape[0]) if len(shape) > 1 else a.astype(np.float64)
    return hp, w


def forward_logits(hp, w, tokens):
    """-> logits [T][V] of every position (causal), float64."""
    T = len(tokens)
    E, H, KV, hd, n_rot = hp["n_embd"], hp["n_head"], hp["n_kv"], hp["hd"], hp["n_rot"]
    x = w["token_embd.weight"][np.asarray(tokens)]

    def rms(v, g):
        return v / np.sqrt((v ** 2).mean(axis=-1, keepdims=True) + hp["eps"]) * g

    pos = np.arange(T)[:, None]
    inv = hp["base"] ** (-2.0 * np.arange(n_rot // 2) / n_rot)
    ang = pos * inv[None, :]
    cos, sin = np.cos(ang), np.sin(ang)

    def rope(v, nh):  # v [T][nh*hd]: rotate adjacent pairs of the first n_rot dims of every head
        v = v.reshape(T, nh, hd).copy()
        a, b = v[:, :, 0:n_rot:2].copy(), v[:, :, 1:n_rot:2].copy()
        v[:, :, 0:n_rot:2] = a * cos[:, None, :] - b * sin[:, None, :]
        v[:, :, 1:n_rot:2] = a * sin[:, None, :] + b * cos[:, None, :]
        return v

    mask = np.tril(np.ones((T, T), bool))
    for il in range(hp["n_layer"]):
        p = f"blk.{il}."
        xn = rms(x, w[p + "attn_norm.weight"])
        q = rope(xn @ w[p + "attn_q.weight"].T, H)
        k = rope(xn @ w[p + "attn_k.weight"].T, KV)
        v = (xn @ w[p + "attn_v.weight"].T).reshape(T, KV, hd)
        att = np.empty((T, H, hd))
        rep = H // KV
        for h in range(H):
            s = q[:, h, :] @ k[:, h // rep, :].T / np.sqrt(hd)
            s = np.where(mask, s, -np.inf)
            s = np.exp(s - s.max(axis=1, keepdims=True))
            att[:, h, :] = (s / s.sum(axis=1, keepdims=True)) @ v[:, h // rep, :]
        x = x + att.reshape(T, H * hd) @ w[p + "attn_output.weight"].T
        xn = rms(x, w[p + "ffn_norm.weight"])
        g, up = xn @ w[p + "ffn_gate.weight"].T, xn @ w[p + "ffn_up.weight"].T
        x = x + (g / (1.0 + np.exp(-g)) * up) @ w[p + "ffn_down.weight"].T
    out = w.get("output.weight", w["token_embd.weight"])
    return rms(x, w["output_norm.weight"]) @ out.T
