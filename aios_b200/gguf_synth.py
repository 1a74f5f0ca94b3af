"""Seeded random-init GGUF (v3) author for arch ``llama``.

The reference downloads real TinyLlama / Mistral GGUF files
(reference: scripts/download-models.sh:70-77); there is no network here, so tests and
bench.py author files of the same architecture, tensor-type mix and size with random
weights (SURVEY.md §8d "Synthetic inputs", Appendix A/B).  Quantised tensors are made by
drawing random *block bytes* and then overwriting the fp16 ``d``/``dmin`` fields with finite
values chosen so the dequantised weights have std ~= ``w_std`` — no quantiser is needed
and a 4.4 GB Mistral-7B file is written in well under a minute.

This module has its own minimal GGUF writer (streaming, one tensor at a time) so that the
C++ loader is tested against a file layout produced independently of gguf-py's writer,
and gguf-py's reader is used in tests to validate *this* writer.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field

import numpy as np

GGUF_MAGIC = 0x46554747
GGUF_VERSION = 3
ALIGNMENT = 32

# ggml type ids (gguf-py constants.py GGMLQuantizationType)
F32, F16, Q8_0, Q4_K, Q5_K, Q6_K = 0, 1, 8, 12, 13, 14
BLOCK = {F32: (1, 4), F16: (1, 2), Q8_0: (32, 34), Q4_K: (256, 144), Q5_K: (256, 176), Q6_K: (256, 210)}
TYPE_NAME = {F32: "F32", F16: "F16", Q8_0: "Q8_0", Q4_K: "Q4_K", Q5_K: "Q5_K", Q6_K: "Q6_K"}

# GGUFValueType
_U32, _F32, _STR, _ARR, _I32 = 4, 6, 8, 9, 5


@dataclass(frozen=True)
class Arch:
    name: str
    n_layer: int
    n_embd: int
    n_head: int
    n_head_kv: int
    head_dim: int
    n_ff: int
    n_vocab: int
    rope_freq_base: float = 10000.0
    rms_eps: float = 1e-5
    n_ctx_train: int = 4096


# the ChatML template as model cards carry it in tokenizer.chat_template
CHATML_TEMPLATE = ("{% for message in messages %}{{'<|im_start|>' + message['role'] + '\\n' + message['content'] + '<|im_end|>' + '\\n'}}"
                   "{% endfor %}{% if add_generation_prompt %}{{ '<|im_start|>assistant\\n' }}{% endif %}")

ARCHS = {
    # tiny shapes for CPU-speed parity tests (K multiples of 256 as K-quants require)
    "tiny": Arch("tiny", 2, 256, 4, 2, 64, 512, 512, 10000.0, 1e-5, 512),
    "tiny1": Arch("tiny1", 1, 256, 4, 2, 64, 512, 512, 10000.0, 1e-5, 512),
    "tiny-gqa4": Arch("tiny-gqa4", 3, 512, 8, 2, 64, 1024, 1000, 10000.0, 1e-5, 512),
    # one kv head for eight query heads (Llama-3-70B's group size): attention kernels take two passes of four heads
    "tiny-gqa8": Arch("tiny-gqa8", 2, 512, 8, 1, 64, 1024, 512, 10000.0, 1e-5, 512),
    # multi-head attention with 32 kv heads (Llama-2-7B's head layout on a small width)
    "tiny-mha32": Arch("tiny-mha32", 2, 2048, 32, 32, 64, 512, 512, 10000.0, 1e-5, 512),
    # splits four ways (tensor-parallel tests): 4 kv heads, n_head*head_dim / 4 and FF / 4 multiples of 256
    "tiny-tp4": Arch("tiny-tp4", 2, 1024, 16, 4, 64, 1024, 512, 10000.0, 1e-5, 512),
    # SURVEY.md Appendix B shapes
    "tinyllama-1.1b": Arch("tinyllama-1.1b", 22, 2048, 32, 4, 64, 5632, 32000, 10000.0, 1e-5, 2048),
    "mistral-7b": Arch("mistral-7b", 32, 4096, 32, 8, 128, 14336, 32000, 1000000.0, 1e-5, 32768),
    "llama3-70b": Arch("llama3-70b", 80, 8192, 64, 8, 128, 28672, 128256, 500000.0, 1e-5, 8192),
}

FTYPE_ID = {"Q4_K_M": 15, "Q4_K_M_70B": 15, "Q8_0": 7, "F32": 0}


def use_more_bits(i: int, n: int) -> bool:
    """llama.cpp's Q4_K_M rule for which layers get Q6_K attn_v / ffn_down."""
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def tensor_plan(arch: Arch, ftype: str) -> list[tuple[str, tuple[int, int] | tuple[int], int]]:
    """[(name, (rows, cols) or (n,), ggml_type)] in file order.  cols = K = quant-block axis (ne0)."""
    E, FF, V = arch.n_embd, arch.n_ff, arch.n_vocab
    QD, KVD = arch.n_head * arch.head_dim, arch.n_head_kv * arch.head_dim
    if ftype in ("Q4_K_M", "Q4_K_M_70B"):
        base, more = Q4_K, Q6_K
        # 70B-class bumps remaining attn_v to Q5_K ("Q4_K_M_70B": that mix on any shape, for tests)
        v70 = Q5_K if (arch.n_layer >= 80 or ftype == "Q4_K_M_70B") else Q4_K
    elif ftype == "Q8_0":
        base = more = v70 = Q8_0
    elif ftype == "F32":
        base = more = v70 = F32
    else:
        raise ValueError(f"unknown ftype {ftype}")
    plan: list = [("token_embd.weight", (V, E), base)]
    for i in range(arch.n_layer):
        mb = use_more_bits(i, arch.n_layer)
        p = f"blk.{i}."
        plan += [
            (p + "attn_norm.weight", (E,), F32),
            (p + "attn_q.weight", (QD, E), base),
            (p + "attn_k.weight", (KVD, E), base),
            (p + "attn_v.weight", (KVD, E), more if mb else v70),
            (p + "attn_output.weight", (E, QD), base),
            (p + "ffn_norm.weight", (E,), F32),
            (p + "ffn_gate.weight", (FF, E), base),
            (p + "ffn_up.weight", (FF, E), base),
            (p + "ffn_down.weight", (E, FF), more if mb else base),
        ]
    plan += [("output_norm.weight", (E,), F32), ("output.weight", (V, E), more)]
    return plan


def tensor_nbytes(shape, t: int) -> int:
    n = int(np.prod(shape))
    be, bb = BLOCK[t]
    assert shape[-1] % be == 0, (shape, t)
    return n // be * bb


def _f16_bytes(vals: np.ndarray) -> np.ndarray:
    return vals.astype(np.float16).view(np.uint8).reshape(-1, 2)


def random_blocks(t: int, n_blocks: int, rng: np.random.Generator, w_std: float) -> np.ndarray:
    """Random block bytes (n_blocks, block_bytes) of type t with finite fp16 scales."""
    bb = BLOCK[t][1]
    raw = rng.integers(0, 256, size=(n_blocks, bb), dtype=np.uint8)
    u = rng.random(n_blocks, dtype=np.float32) + 0.5  # U(0.5, 1.5)
    if t in (Q4_K, Q5_K):
        # w = d*sc*q - dmin*m ; sc,m ~ U{0..63}; q ~ U{0..15} (Q5_K: U{0..31}); dmin = E[q]*d => zero mean
        qmean = 7.5 if t == Q4_K else 15.5
        unit = 258.0 if t == Q4_K else 520.0
        d = (w_std / unit) * u
        raw[:, 0:2] = _f16_bytes(d)
        raw[:, 2:4] = _f16_bytes(d * qmean)
    elif t == Q6_K:
        d = (w_std / 1366.0) * u  # w = d*scale*(q-32), scale ~ U{-128..127}
        raw[:, 208:210] = _f16_bytes(d)
    elif t == Q8_0:
        raw[:, 2:][raw[:, 2:] == 0x80] = 0x81  # ggml never emits -128
        d = (w_std / 73.3) * u
        raw[:, 0:2] = _f16_bytes(d)
    else:
        raise ValueError(t)
    return raw


def make_tensor_bytes(shape, t: int, rng: np.random.Generator, w_std: float) -> np.ndarray:
    if t == F32:
        if len(shape) == 1:  # norm weights near 1
            return (rng.random(shape[0], dtype=np.float32) * 0.4 + 0.8).view(np.uint8)
        return (rng.standard_normal(int(np.prod(shape)), dtype=np.float32) * w_std).view(np.uint8)
    n_blocks = int(np.prod(shape)) // BLOCK[t][0]
    return random_blocks(t, n_blocks, rng, w_std).reshape(-1)


# ---------------------------------------------------------------- minimal GGUF v3 serialisation
def _s(b: str | bytes) -> bytes:
    if isinstance(b, str):
        b = b.encode("utf-8")
    return struct.pack("<Q", len(b)) + b


def _kv(key: str, vtype: int, val) -> bytes:
    out = _s(key) + struct.pack("<I", vtype)
    if vtype == _U32:
        out += struct.pack("<I", val)
    elif vtype == _I32:
        out += struct.pack("<i", val)
    elif vtype == _F32:
        out += struct.pack("<f", val)
    elif vtype == _STR:
        out += _s(val)
    elif vtype == _ARR:
        etype, items = val
        out += struct.pack("<IQ", etype, len(items))
        if etype == _STR:
            out += b"".join(_s(x) for x in items)
        elif etype == _F32:
            out += np.asarray(items, dtype="<f4").tobytes()
        elif etype == _I32:
            out += np.asarray(items, dtype="<i4").tobytes()
        else:
            raise ValueError(etype)
    else:
        raise ValueError(vtype)
    return out


@dataclass
class SynthInfo:
    path: str
    arch: Arch
    ftype: str
    seed: int
    n_tensors: int
    data_bytes: int
    file_bytes: int
    tensors: dict = field(default_factory=dict)  # name -> (type, shape, data offset rel. to data section)


def write_synthetic_gguf(path: str, arch: Arch | str, ftype: str = "Q4_K_M", seed: int = 0xA105,
                         w_std: float = 0.02, with_vocab: bool = True, vocab: list | None = None,
                         chat_template: str | None = None, tokenizer_model: str = "llama", tied: bool = False,
                         merges: list | None = None, tokenizer_pre: str | None = None,
                         bos_id: int = 1, eos_id: int = 2) -> SynthInfo:
    """``vocab`` = [(piece, score, token_type)] replaces the generated placeholder vocabulary (padded with unused
    entries up to ``arch.n_vocab``); ``chat_template`` is stored as ``tokenizer.chat_template``; ``tied`` leaves
    ``output.weight`` out (the lm_head is then ``token_embd.weight``, as in files with tied embeddings); ``merges``
    (["left right", ...]) and ``tokenizer_pre`` are the byte-level-BPE keys ``tokenizer.ggml.merges`` / ``.pre``."""
    if isinstance(arch, str):
        arch = ARCHS[arch]
    plan = tensor_plan(arch, ftype)
    if tied:
        plan = [t for t in plan if t[0] != "output.weight"]
    kvs = [
        _kv("general.architecture", _STR, "llama"),
        _kv("general.name", _STR, f"synthetic-{arch.name}-{ftype}-seed{seed}"),
        _kv("general.file_type", _U32, FTYPE_ID[ftype]),
        _kv("general.alignment", _U32, ALIGNMENT),
        _kv("llama.context_length", _U32, arch.n_ctx_train),
        _kv("llama.embedding_length", _U32, arch.n_embd),
        _kv("llama.block_count", _U32, arch.n_layer),
        _kv("llama.feed_forward_length", _U32, arch.n_ff),
        _kv("llama.vocab_size", _U32, arch.n_vocab),
        _kv("llama.attention.head_count", _U32, arch.n_head),
        _kv("llama.attention.head_count_kv", _U32, arch.n_head_kv),
        _kv("llama.attention.layer_norm_rms_epsilon", _F32, arch.rms_eps),
        _kv("llama.rope.dimension_count", _U32, arch.head_dim),
        _kv("llama.rope.freq_base", _F32, arch.rope_freq_base),
        _kv("tokenizer.ggml.model", _STR, tokenizer_model),
        _kv("tokenizer.ggml.bos_token_id", _U32, bos_id),
        _kv("tokenizer.ggml.eos_token_id", _U32, eos_id),
    ]
    if merges is not None:
        kvs.append(_kv("tokenizer.ggml.merges", _ARR, (_STR, list(merges))))
    if tokenizer_pre is not None:
        kvs.append(_kv("tokenizer.ggml.pre", _STR, tokenizer_pre))
    if chat_template is not None:
        kvs.append(_kv("tokenizer.chat_template", _STR, chat_template))
    if vocab is not None:
        if len(vocab) > arch.n_vocab:
            raise ValueError(f"vocabulary of {len(vocab)} pieces exceeds n_vocab {arch.n_vocab}")
        pad = arch.n_vocab - len(vocab)
        kvs += [
            _kv("tokenizer.ggml.tokens", _ARR, (_STR, [p for p, _, _ in vocab] + [f"<unused{i}>" for i in range(pad)])),
            _kv("tokenizer.ggml.scores", _ARR, (_F32, [float(s) for _, s, _ in vocab] + [0.0] * pad)),
            _kv("tokenizer.ggml.token_type", _ARR, (_I32, [int(t) for _, _, t in vocab] + [5] * pad)),
        ]
    elif with_vocab:
        toks = ["<unk>", "<s>", "</s>"] + [f"<0x{i:02X}>" for i in range(256)]
        toks += [f"▁t{i}" for i in range(arch.n_vocab - len(toks))]
        toks = toks[: arch.n_vocab]
        ttype = [2, 3, 3] + [6] * 256 + [1] * max(0, arch.n_vocab - 259)
        kvs += [
            _kv("tokenizer.ggml.tokens", _ARR, (_STR, toks)),
            _kv("tokenizer.ggml.scores", _ARR, (_F32, [-float(i) for i in range(arch.n_vocab)])),
            _kv("tokenizer.ggml.token_type", _ARR, (_I32, ttype[: arch.n_vocab])),
        ]
    infos, off = [], 0
    table = {}
    for name, shape, t in plan:
        dims = tuple(reversed(shape))  # ne0 (innermost / K) first
        infos.append(_s(name) + struct.pack("<I", len(dims)) + b"".join(struct.pack("<Q", d) for d in dims)
                     + struct.pack("<IQ", t, off))
        table[name] = (t, tuple(shape), off)
        nb = tensor_nbytes(shape, t)
        off += (nb + ALIGNMENT - 1) // ALIGNMENT * ALIGNMENT
    header = struct.pack("<IIQQ", GGUF_MAGIC, GGUF_VERSION, len(plan), len(kvs)) + b"".join(kvs) + b"".join(infos)
    pad = (-len(header)) % ALIGNMENT
    with open(path, "wb") as f:
        f.write(header + b"\0" * pad)
        for idx, (name, shape, t) in enumerate(plan):
            rng = np.random.default_rng([seed, idx])
            # the logits head / embeddings keep w_std too; tests report greedy margins
            buf = make_tensor_bytes(shape, t, rng, w_std)
            f.write(memoryview(buf))
            f.write(b"\0" * ((-buf.nbytes) % ALIGNMENT))
        total = f.tell()
    return SynthInfo(path, arch, ftype, seed, len(plan), off, total, table)


def synthetic_prompt(arch: Arch | str, n: int, seed: int = 0xA105) -> np.ndarray:
    if isinstance(arch, str):
        arch = ARCHS[arch]
    return np.random.default_rng([seed, 0x7073]).integers(0, arch.n_vocab, size=n, dtype=np.int32)


def weight_bytes_per_token(arch: Arch | str, ftype: str) -> int:
    """Algorithmic weight bytes one decode step streams: every matrix once + one embedding row."""
    if isinstance(arch, str):
        arch = ARCHS[arch]
    total = 0
    for name, shape, t in tensor_plan(arch, ftype):
        if name == "token_embd.weight":
            total += tensor_nbytes((1, shape[1]), t)
        else:
            total += tensor_nbytes(shape, t)
    return total


def kv_bytes_per_position(arch: Arch | str) -> int:
    if isinstance(arch, str):
        arch = ARCHS[arch]
    return arch.n_layer * 2 * arch.n_head_kv * arch.head_dim * 2


if __name__ == "__main__":
    import argparse, time

    ap = argparse.ArgumentParser()
    ap.add_argument("arch", choices=sorted(ARCHS))
    ap.add_argument("out")
    ap.add_argument("--ftype", default="Q4_K_M")
    ap.add_argument("--seed", type=lambda s: int(s, 0), default=0xA105)
    a = ap.parse_args()
    t0 = time.time()
    info = write_synthetic_gguf(a.out, a.arch, a.ftype, a.seed)
    print(f"{info.path}: {info.n_tensors} tensors, {info.file_bytes/1e9:.3f} GB in {time.time()-t0:.1f}s; "
          f"weights/token {weight_bytes_per_token(a.arch, a.ftype)/1e9:.4f} GB")
