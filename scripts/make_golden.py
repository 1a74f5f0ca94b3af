#!/usr/bin/env python
"""Generates tests/golden/*.npz — run in the build container (needs gguf-py and transformers).

Golden sources: (i) gguf-py's numpy dequantisers (llama.cpp's own Python package;
bit-exact tier) on hand-picked edge blocks + random blocks; (ii) HF transformers LlamaForCausalLM in
fp32 on CPU fed by load_gguf_checkpoint (exact-math logits; tolerance tier) for the seeded tiny models.
The reference itself pins no logits/tokens, so these are the only external anchors available.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from aios_b200 import gguf_synth as gs  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def edge_blocks(qt, rng):
    b = gs.random_blocks(qt, 64, rng, 0.02)
    b[0] = 0
    b[1] = 0xFF
    b[2] = 0x3F
    b[3] = 0xC0  # only the high scale bits of Q4_K/Q5_K sub-blocks 4..7
    if qt in (gs.Q4_K, gs.Q5_K):
        b[1, 0:4] = np.array([0x00, 0x3C, 0x00, 0x38], np.uint8)   # d = 1.0, dmin = 0.5
        b[2, 0:4] = np.array([0xFF, 0x7B, 0xFF, 0x7B], np.uint8)   # d = dmin = 65504 (max finite fp16)
        b[3, 0:4] = np.array([0x01, 0x00, 0x01, 0x80], np.uint8)   # smallest sub-normal d, negative sub-normal dmin
        b[4, 0:4] = np.array([0x00, 0x80, 0x00, 0x00], np.uint8)   # d = -0, dmin = 0
    elif qt == gs.Q6_K:
        b[1, 208:210] = np.array([0x00, 0x3C], np.uint8)
        b[2, 208:210] = np.array([0xFF, 0x7B], np.uint8)
        b[3, 208:210] = np.array([0x01, 0x00], np.uint8)
        b[4, 208:210] = np.array([0x00, 0x80], np.uint8)
    else:
        b[1, 0:2] = np.array([0x00, 0x3C], np.uint8)
        b[2, 0:2] = np.array([0xFF, 0x7B], np.uint8)
        b[3, 0:2] = np.array([0x01, 0x00], np.uint8)
        b[4, 0:2] = np.array([0x00, 0x80], np.uint8)
    return b


def main():
    import gguf
    from gguf import quants

    os.makedirs(OUT, exist_ok=True)
    rng = np.random.default_rng(20261106)
    names = {gs.Q8_0: "Q8_0", gs.Q4_K: "Q4_K", gs.Q5_K: "Q5_K", gs.Q6_K: "Q6_K"}
    out = {}
    for qt, nm in names.items():
        blocks = edge_blocks(qt, rng)
        with np.errstate(all="ignore"):
            deq = quants.dequantize(blocks, getattr(gguf.GGMLQuantizationType, nm)).astype(np.float32)
        out[f"{nm}_blocks"] = blocks
        out[f"{nm}_dequant_bits"] = deq.view(np.uint32)  # stored as bit patterns (inf/nan safe)
    np.savez_compressed(os.path.join(OUT, "dequant_gguf_py.npz"), **out)

    # HF fp32 logits for the seeded tiny models
    import tempfile

    import torch
    from transformers import LlamaConfig, LlamaForCausalLM
    from transformers.modeling_gguf_pytorch_utils import load_gguf_checkpoint

    res = {}
    for arch, ftype, seed in [("tiny", "Q4_K_M", 7), ("tiny", "Q8_0", 7), ("tiny-gqa4", "Q4_K_M", 7)]:
        a = gs.ARCHS[arch]
        with tempfile.TemporaryDirectory() as d:
            path = os.path.join(d, "m.gguf")
            gs.write_synthetic_gguf(path, arch, ftype, seed)
            cfg = LlamaConfig(vocab_size=a.n_vocab, hidden_size=a.n_embd, intermediate_size=a.n_ff,
                              num_hidden_layers=a.n_layer, num_attention_heads=a.n_head,
                              num_key_value_heads=a.n_head_kv, rms_norm_eps=a.rms_eps, rope_theta=a.rope_freq_base,
                              max_position_embeddings=512, tie_word_embeddings=False, attention_bias=False,
                              head_dim=a.head_dim)
            model = LlamaForCausalLM(cfg).float().eval()
            ck = load_gguf_checkpoint(path, return_tensors=True, model_to_load=model)
            missing, unexpected = model.load_state_dict(ck["tensors"], strict=False)
            assert not missing and not unexpected, (missing, unexpected)
            prompt = gs.synthetic_prompt(arch, 24, 5)
            with torch.no_grad():
                logits = model(torch.tensor(prompt[None].astype(np.int64))).logits[0].numpy().astype(np.float32)
        key = f"{arch}_{ftype}_{seed}"
        res[key + "_prompt"] = prompt
        res[key + "_hf_fp32_logits"] = logits
    np.savez_compressed(os.path.join(OUT, "hf_fp32_logits.npz"), **res)
    print("wrote", os.listdir(OUT))


if __name__ == "__main__":
    main()
