"""Generate tests/golden/exact_logits_{tinyllama,mistral7b}.npz: exact-math logits at the BASELINE.json model shapes.

An independent pin of the model math at full size (SURVEY.md §8c tier (ii)): the seeded synthetic GGUF of the shape is
dequantised tensor by tensor with gguf-py's own numpy dequantisers (llama.cpp's Python package — not this repo's oracle)
and pushed through the float64 numpy forward pass of oracle/ref_llama.py (no activation quantisation, no f16 rounding of
K / V / probabilities).  The oracle and the CUDA engine must land within their quantisation noise of these logits
(tests/test_cpu.py, tests/test_gpu_configs.py).  Weights are streamed layer by layer, so Mistral-7B needs ~4 GB of RAM.

Run:  python scripts/make_golden_exact.py          (about five minutes, most of it Mistral-7B's dequantisation)
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gguf import GGUFReader  # noqa: E402
from gguf.quants import dequantize  # noqa: E402

from aios_b200 import gguf_synth as gs  # noqa: E402
from oracle import ref_llama  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
CASES = [("tinyllama-1.1b", "Q4_K_M", 7, 32, "exact_logits_tinyllama.npz"),
         ("mistral-7b", "Q4_K_M", 7, 16, "exact_logits_mistral7b.npz")]


class LazyWeights(dict):
    """name -> float64 matrix, dequantised by gguf-py when asked for and not kept."""

    def __init__(self, reader):
        super().__init__()
        self.t = {t.name: t for t in reader.tensors}

    def __contains__(self, name):
        return name in self.t

    def get(self, name, default=None):
        return self[name] if name in self.t else default

    def __getitem__(self, name):
        t = self.t[name]
        shape = [int(x) for x in t.shape]  # ne0 first
        a = dequantize(np.asarray(t.data), t.tensor_type).astype(np.float64).reshape(-1)
        return a.reshape(shape[1], shape[0]) if len(shape) > 1 else a


def main():
    for arch, ftype, seed, n_prompt, name in CASES:
        with tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None) as d:
            path = os.path.join(d, "m.gguf")
            gs.write_synthetic_gguf(path, arch, ftype, seed)
            r = GGUFReader(path)
            a = gs.ARCHS[arch]
            hp = dict(n_layer=a.n_layer, n_embd=a.n_embd, n_head=a.n_head, n_ff=a.n_ff, eps=a.rms_eps, base=a.rope_freq_base,
                      n_kv=a.n_head_kv, hd=a.head_dim, n_rot=a.head_dim)
            prompt = gs.synthetic_prompt(arch, n_prompt, seed)
            logits = ref_llama.forward_logits(hp, LazyWeights(r), prompt)[-1]
            np.savez_compressed(os.path.join(OUT, name), arch=arch, ftype=ftype, seed=seed, prompt=prompt.astype(np.int32),
                                logits=logits.astype(np.float32))
            top = np.argsort(-logits)[:2]
            print(f"{name}: {arch} {ftype} seed {seed}, {n_prompt} tokens; logits std {logits.std():.4f}, "
                  f"top-1 {top[0]} margin {(logits[top[0]] - logits[top[1]]) / logits.std():.3f} sigma")


if __name__ == "__main__":
    main()
