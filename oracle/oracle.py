"""ctypes front-end for the CPU oracle (oracle/aios_oracle.c).

TEST INFRASTRUCTURE ONLY — importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package (aios_b200) never imports this.
PARITY UNPINNED: see the header of aios_oracle.c.

GGUF files are parsed with gguf-py's GGUFReader (a third, independent parser) so the
oracle does not depend on the product's C++ loader.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libaios_oracle.so")

F32, F16, Q8_0, Q4_K, Q5_K, Q6_K = 0, 1, 8, 12, 13, 14
BLOCK = {F32: (1, 4), F16: (1, 2), Q8_0: (32, 34), Q4_K: (256, 144), Q5_K: (256, 176), Q6_K: (256, 210)}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "aios_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libaios_oracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


class HParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_layer", "n_embd", "n_head", "n_head_kv", "head_dim", "n_ff",
                                         "n_vocab", "n_ctx", "n_rot")] + [("rms_eps", C.c_float),
                                                                          ("rope_freq_base", C.c_float)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        L.oracle_dequantize.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_void_p]
        L.oracle_quantize_row_q8_K.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.oracle_quantize_row_q8_0.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.oracle_matvec.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
        L.oracle_model_new.restype = C.c_void_p
        L.oracle_model_new.argtypes = [C.POINTER(HParams)]
        L.oracle_model_free.argtypes = [C.c_void_p]
        L.oracle_model_set_tensor.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int64, C.c_int64, C.c_void_p]
        L.oracle_eval.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.oracle_eval_all.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.oracle_greedy.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.oracle_rope_table.argtypes = [C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p]
        L.oracle_set_simd.argtypes = [C.c_int]
        L.oracle_sample.restype = C.c_int
        L.oracle_sample.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_float, C.c_int, C.c_float, C.c_float,
                                    C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p]
        L.oracle_uniform.restype = C.c_float
        L.oracle_uniform.argtypes = [C.c_uint64, C.c_uint64]
        L.oracle_set_threads.argtypes = [C.c_int]
        L.oracle_num_threads.restype = C.c_int
        L.oracle_f2h.restype = C.c_uint16
        L.oracle_f2h.argtypes = [C.c_float]
        L.oracle_h2f.restype = C.c_float
        L.oracle_h2f.argtypes = [C.c_uint16]
        _lib = L
    return _lib


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def dequantize(qtype: int, blocks: np.ndarray) -> np.ndarray:
    """blocks: uint8 array holding whole blocks; returns f32 [n_elems]."""
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1)
    be, bb = BLOCK[qtype]
    n = blocks.size // bb * be
    out = np.empty(n, dtype=np.float32)
    rc = lib().oracle_dequantize(qtype, _ptr(blocks), n, _ptr(out))
    assert rc == 0
    return out


def quantize_q8_K(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty(x.size // 256 * 292, dtype=np.uint8)
    lib().oracle_quantize_row_q8_K(_ptr(x), _ptr(out), x.size)
    return out


def quantize_q8_0(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty(x.size // 32 * 34, dtype=np.uint8)
    lib().oracle_quantize_row_q8_0(_ptr(x), _ptr(out), x.size)
    return out


def matvec(qtype: int, w_blocks: np.ndarray, n_rows: int, k: int, x: np.ndarray, simd: bool = False) -> np.ndarray:
    w = np.ascontiguousarray(w_blocks, dtype=np.uint8).reshape(-1)
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.empty(n_rows, dtype=np.float32)
    lib().oracle_set_simd(1 if simd else 0)
    rc = lib().oracle_matvec(qtype, _ptr(w), n_rows, k, _ptr(x), _ptr(y))
    lib().oracle_set_simd(0)
    assert rc == 0
    return y


def sample(logits: np.ndarray, temperature: float, top_k: int = 40, top_p: float = 0.95, min_p: float = 0.05,
           seed: int = 0, counter: int = 0, mask: np.ndarray | None = None):
    """The sampler chain (oracle_sample): returns (token, candidate ids sorted, (K, n after top-p, n after min-p)).
    `mask` = packed bits, bit i of byte i//8 set when token i is allowed."""
    lg = np.ascontiguousarray(logits, np.float32)
    cand = np.full(128, -1, np.int32)
    counts = np.zeros(3, np.int32)
    mk = None if mask is None else np.ascontiguousarray(mask, np.uint8)
    tok = lib().oracle_sample(_ptr(lg), lg.size, None if mk is None else _ptr(mk), float(temperature), int(top_k),
                              float(top_p), float(min_p), int(seed), int(counter), _ptr(cand), _ptr(counts))
    return int(tok), cand[:counts[0]].copy(), tuple(int(c) for c in counts)


def rope_table(pos: int, n_rot: int, freq_base: float):
    c = np.empty(n_rot // 2, np.float32)
    s = np.empty(n_rot // 2, np.float32)
    lib().oracle_rope_table(pos, n_rot, freq_base, _ptr(c), _ptr(s))
    return c, s


class OracleModel:
    """Arch-llama model evaluated by the CPU oracle straight off a GGUF file (mmap)."""

    def __init__(self, gguf_path: str, n_ctx: int = 512, simd: bool = False, threads: int = 0):
        from gguf import GGUFReader  # independent parser

        self.reader = GGUFReader(gguf_path)
        f = self.reader.fields

        def u(key, default=None):
            if key not in f:
                if default is None:
                    raise KeyError(key)
                return default
            fld = f[key]
            return fld.parts[fld.data[0]][0]

        hp = HParams()
        hp.n_layer = int(u("llama.block_count"))
        hp.n_embd = int(u("llama.embedding_length"))
        hp.n_head = int(u("llama.attention.head_count"))
        hp.n_head_kv = int(u("llama.attention.head_count_kv", hp.n_head))
        hp.n_rot = int(u("llama.rope.dimension_count", hp.n_embd // hp.n_head))
        hp.head_dim = hp.n_embd // hp.n_head
        hp.n_ff = int(u("llama.feed_forward_length"))
        hp.rms_eps = float(u("llama.attention.layer_norm_rms_epsilon", 1e-5))
        hp.rope_freq_base = float(u("llama.rope.freq_base", 10000.0))
        hp.n_ctx = n_ctx
        tensors = {t.name: t for t in self.reader.tensors}
        hp.n_vocab = int(tensors["token_embd.weight"].shape[1])
        self.hp = hp
        self.simd = simd
        L = lib()
        if threads:
            L.oracle_set_threads(threads)
        self.h = L.oracle_model_new(C.byref(hp))
        self._keep = []
        for name, t in tensors.items():
            data = t.data  # numpy memmap view over the file
            ne0 = int(t.shape[0])
            ne1 = int(t.shape[1]) if len(t.shape) > 1 else 1
            rc = L.oracle_model_set_tensor(self.h, name.encode(), int(t.tensor_type), ne0, ne1,
                                           C.c_void_p(data.ctypes.data))
            assert rc == 0, f"oracle: unknown tensor {name}"
            self._keep.append(data)

    def close(self):
        if self.h:
            lib().oracle_model_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def eval(self, tokens, n_past: int = 0) -> np.ndarray:
        """Evaluate tokens at positions n_past..; returns logits of the last token."""
        toks = np.ascontiguousarray(tokens, dtype=np.int32)
        out = np.empty(self.hp.n_vocab, dtype=np.float32)
        lib().oracle_set_simd(1 if self.simd else 0)
        rc = lib().oracle_eval(self.h, _ptr(toks), toks.size, n_past, _ptr(out))
        lib().oracle_set_simd(0)
        assert rc == 0
        return out

    def eval_all(self, tokens, n_past: int = 0) -> np.ndarray:
        toks = np.ascontiguousarray(tokens, dtype=np.int32)
        out = np.empty((toks.size, self.hp.n_vocab), dtype=np.float32)
        lib().oracle_set_simd(1 if self.simd else 0)
        rc = lib().oracle_eval_all(self.h, _ptr(toks), toks.size, n_past, _ptr(out))
        lib().oracle_set_simd(0)
        assert rc == 0
        return out

    def greedy(self, prompt, n_gen: int):
        toks = np.ascontiguousarray(prompt, dtype=np.int32)
        out = np.empty(n_gen, dtype=np.int32)
        margins = np.empty(n_gen, dtype=np.float32)
        lib().oracle_set_simd(1 if self.simd else 0)
        rc = lib().oracle_greedy(self.h, _ptr(toks), toks.size, n_gen, _ptr(out), _ptr(margins))
        lib().oracle_set_simd(0)
        assert rc == 0
        return out, margins
