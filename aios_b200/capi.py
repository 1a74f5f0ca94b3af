"""ctypes binding of include/aios_b200.h (libaios_b200.so).

This is the Python stand-in for the Rust ``-sys`` crate a maintainer of the reference's
``runtime`` crate would write; the gRPC shim (aios_b200.runtime) and the
tests call the engine only through these functions.  There is no CPU fallback: if the
shared library is missing, importing a function from it raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libaios_b200.so")

AIOS_OK, E_INVAL, E_IO, E_FORMAT, E_CUDA, E_NOMEM, E_STATE, E_RANGE, E_UNSUPPORTED = 0, -1, -2, -3, -4, -5, -6, -7, -8
TYPE_F32, TYPE_F16, TYPE_Q8_0, TYPE_Q4_K, TYPE_Q5_K, TYPE_Q6_K = 0, 1, 8, 12, 13, 14
ENGINE_NO_GRAPH = 1
ENGINE_NO_PDL = 2
ENGINE_NO_PFD = 4
ENGINE_USE_PFD = 8
ENGINE_ATTN_CLUSTER = 64
ENGINE_NO_ROPE_EPILOGUE = 256
ENGINE_NO_GEMM = 1024
ENGINE_EXACT_PREFILL = 2048


class AiosError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"aios_b200 error {code}: {msg}")
        self.code = code
        self.msg = msg


class EngineOpts(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_batch", C.c_int32), ("flags", C.c_int32), ("reserved", C.c_int32)]


class ModelInfo(C.Structure):
    _fields_ = [("name", C.c_char * 128), ("arch", C.c_char * 32)] + \
               [(n, C.c_int32) for n in ("n_layer", "n_embd", "n_head", "n_head_kv", "head_dim", "n_ff", "n_vocab",
                                         "n_ctx", "n_rot", "file_type", "n_tensors", "max_batch")] + \
               [("rms_eps", C.c_float), ("rope_freq_base", C.c_float)] + \
               [(n, C.c_int64) for n in ("file_bytes", "weight_bytes", "kv_bytes", "weight_bytes_per_token",
                                         "kv_bytes_per_position")]


class TensorInfo(C.Structure):
    _fields_ = [("name", C.c_char * 96), ("type", C.c_int32), ("n_dims", C.c_int32), ("ne", C.c_int64 * 4),
                ("offset", C.c_int64), ("nbytes", C.c_int64)]


class VocabInfo(C.Structure):
    _fields_ = [("n_tokens", C.c_int32), ("bos", C.c_int32), ("eos", C.c_int32), ("add_bos", C.c_int32),
                ("chat_format", C.c_int32), ("model", C.c_char * 32)]


class Sampling(C.Structure):
    _fields_ = [("temperature", C.c_float), ("top_k", C.c_int32), ("top_p", C.c_float), ("min_p", C.c_float),
                ("seed", C.c_uint64)]


class BatchBench(C.Structure):
    _fields_ = [("ms_decode", C.c_float), ("ms_prefill", C.c_float), ("n_lanes", C.c_int32), ("n_steps_timed", C.c_int32),
                ("kernel_launches", C.c_int32), ("n_past_begin", C.c_int32), ("n_past_end", C.c_int32),
                ("reserved", C.c_int32)]


class BenchResult(C.Structure):
    _fields_ = [("ms_total", C.c_float), ("n_tokens", C.c_int32), ("kernel_launches", C.c_int32),
                ("n_past_begin", C.c_int32), ("n_past_end", C.c_int32), ("ms_gemv", C.c_float),
                ("reserved", C.c_float * 3)]


class TpBench(C.Structure):
    _fields_ = [("ms_total", C.c_float), ("n_tokens", C.c_int32), ("n_ranks", C.c_int32), ("kernel_launches", C.c_int32),
                ("n_past_begin", C.c_int32), ("n_past_end", C.c_int32), ("n_allreduce_per_token", C.c_int32),
                ("us_per_allreduce", C.c_float), ("allreduce_kind", C.c_int32), ("graphs", C.c_int32)]


TOK_ADD_SPECIAL, TOK_PARSE_SPECIAL, TOK_RENDER_SPECIAL, TOK_STRIP_SPACE = 1, 2, 4, 8

# every symbol include/aios_b200.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    "aios_last_error": (C.c_char_p, []),
    "aios_version": (C.c_char_p, []),
    "aios_device_count": (C.c_int, [C.POINTER(C.c_int32)]),
    "aios_engine_create": (C.c_int, [C.POINTER(EngineOpts), C.POINTER(_P)]),
    "aios_engine_destroy": (None, [_P]),
    "aios_model_load": (C.c_int, [_P, C.c_char_p, C.c_char_p, C.c_int32, C.POINTER(_P)]),
    "aios_model_unload": (C.c_int, [_P]),
    "aios_model_info_get": (C.c_int, [_P, C.POINTER(ModelInfo)]),
    "aios_model_tensor_info": (C.c_int, [_P, C.c_int32, C.POINTER(TensorInfo)]),
    "aios_model_meta_str": (C.c_int, [_P, C.c_char_p, _P, C.c_int32, C.POINTER(C.c_int32)]),
    "aios_vocab_open": (C.c_int, [C.c_char_p, C.POINTER(_P)]),
    "aios_vocab_close": (None, [_P]),
    "aios_model_vocab": (C.c_int, [_P, C.POINTER(_P)]),
    "aios_vocab_info_get": (C.c_int, [_P, C.POINTER(VocabInfo)]),
    "aios_vocab_token_piece": (C.c_int, [_P, C.c_int32, _P, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                        C.POINTER(C.c_float)]),
    "aios_tokenize": (C.c_int, [_P, C.c_char_p, C.c_int32, C.c_int32, _P, C.c_int32, C.POINTER(C.c_int32)]),
    "aios_detokenize": (C.c_int, [_P, _P, C.c_int32, C.c_int32, _P, C.c_int32, C.POINTER(C.c_int32)]),
    "aios_chat_prompt": (C.c_int, [_P, C.c_char_p, C.c_char_p, _P, C.c_int32, C.POINTER(C.c_int32)]),
    "aios_chat_apply_template": (C.c_int, [_P, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p), C.c_int32, C.c_int32, _P, C.c_int32,
                                          C.POINTER(C.c_int32)]),
    "aios_token_is_eog": (C.c_int, [_P, C.c_int32, C.POINTER(C.c_int32)]),
    "aios_seq_create": (C.c_int, [_P, C.POINTER(_P)]),
    "aios_seq_free": (None, [_P]),
    "aios_seq_reset": (C.c_int, [_P]),
    "aios_seq_position": (C.c_int, [_P, C.POINTER(C.c_int32)]),
    "aios_seq_prefill": (C.c_int, [_P, _P, C.c_int32]),
    "aios_seq_decode": (C.c_int, [_P, C.POINTER(Sampling), C.c_int32, _P, C.POINTER(C.c_int32)]),
    "aios_seq_step": (C.c_int, [_P, C.POINTER(Sampling), C.POINTER(C.c_int32)]),
    "aios_seq_decode_static": (C.c_int, [_P, C.c_int32, _P, C.POINTER(C.c_int32)]),
    "aios_seq_logits": (C.c_int, [_P, _P]),
    "aios_seq_set_mask": (C.c_int, [_P, _P, C.c_int32]),
    "aios_sample_logits": (C.c_int, [_P, _P, C.c_int32, _P, C.POINTER(Sampling), C.c_uint64, C.POINTER(C.c_int32), _P, _P]),
    "aios_generate": (C.c_int, [_P, _P, C.c_int32, C.POINTER(Sampling), C.c_int32, _P, C.POINTER(C.c_int32)]),
    "aios_request_submit": (C.c_int, [_P, _P, C.c_int32, C.POINTER(Sampling), C.c_int32, C.c_int32, C.POINTER(_P)]),
    "aios_request_submit_masked": (C.c_int, [_P, _P, C.c_int32, C.POINTER(Sampling), C.c_int32, C.c_int32, _P, _P, C.POINTER(_P)]),
    "aios_json_grammar_create": (C.c_int, [_P, C.POINTER(_P)]),
    "aios_json_grammar_free": (None, [_P]),
    "aios_json_grammar_complete": (C.c_int, [_P, C.POINTER(C.c_int32)]),
    "aios_json_grammar_mask_cb": (C.c_int, [_P, C.c_int32, _P, C.c_int32]),
    "aios_request_poll": (C.c_int, [_P, _P, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "aios_request_wait": (C.c_int, [_P, C.c_int32, C.POINTER(C.c_int32)]),
    "aios_request_cancel": (C.c_int, [_P]),
    "aios_request_free": (None, [_P]),
    "aios_dequantize": (C.c_int, [_P, C.c_int32, _P, C.c_int64, _P]),
    "aios_matvec": (C.c_int, [_P, C.c_int32, _P, C.c_int64, C.c_int64, _P, _P, C.c_float, _P]),
    "aios_matmul": (C.c_int, [_P, C.c_int32, _P, C.c_int64, C.c_int64, _P, C.c_int32, _P, C.c_float, _P]),
    "aios_matmul_tc": (C.c_int, [_P, C.c_int32, _P, C.c_int64, C.c_int64, _P, C.c_int32, _P, C.c_float, _P]),
    "aios_forward_logits": (C.c_int, [_P, _P, C.c_int32, C.c_int32, _P]),
    "aios_seq_read_kv": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, _P, _P]),
    "aios_seq_read_activation": (C.c_int, [_P, C.c_int32, _P, C.c_int64]),
    "aios_seq_bench_decode": (C.c_int, [_P, C.c_int32, C.POINTER(BenchResult)]),
    "aios_seq_profile_decode": (C.c_int, [_P, C.c_int32, C.POINTER(BenchResult)]),
    "aios_seq_trace_decode": (C.c_int, [_P, _P, C.c_int32, C.POINTER(C.c_int32)]),
    "aios_seq_phase_timing": (C.c_int, [_P, _P, C.c_int32, C.POINTER(C.c_int32)]),
    "aios_seq_snapshot": (C.c_int, [_P]),
    "aios_seq_restore": (C.c_int, [_P]),
    "aios_seq_bench_prefill": (C.c_int, [_P, _P, C.c_int32, C.c_int32, C.POINTER(C.c_float)]),
    "aios_bench_batch_decode": (C.c_int, [_P, C.c_int32, _P, C.c_int32, C.c_int32, _P, C.POINTER(BatchBench)]),
    "aios_bench_gemv": (C.c_int, [_P, C.c_int32, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.POINTER(C.c_float)]),
    "aios_tp_model_load": (C.c_int, [_P, C.c_int32, C.c_char_p, C.c_char_p, C.c_int32, C.c_int32, C.POINTER(_P)]),
    "aios_tp_model_unload": (C.c_int, [_P]),
    "aios_tp_model_info": (C.c_int, [_P, C.POINTER(ModelInfo)]),
    "aios_tp_reset": (C.c_int, [_P]),
    "aios_tp_prefill": (C.c_int, [_P, _P, C.c_int32]),
    "aios_tp_decode": (C.c_int, [_P, C.POINTER(Sampling), C.c_int32, _P, C.POINTER(C.c_int32)]),
    "aios_tp_get_logits": (C.c_int, [_P, _P, C.c_int32]),
    "aios_tp_bench_decode": (C.c_int, [_P, C.c_int32, C.POINTER(TpBench)]),
}

_lib = None


def lib():
    """Load libaios_b200.so (fails loudly if it has not been built: run __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _check(rc: int):
    if rc != 0:
        raise AiosError(rc, (lib().aios_last_error() or b"").decode("utf-8", "replace"))


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def device_count() -> int:
    n = C.c_int32(0)
    _check(lib().aios_device_count(C.byref(n)))
    return n.value


def make_sampling(temperature=0.0, top_k=40, top_p=0.95, min_p=0.05, seed=0) -> Sampling:
    return Sampling(float(temperature), int(top_k), float(top_p), float(min_p), int(seed))


def _bytes_out(call) -> bytes:
    """Run a (buf, cap, len*) style entry point: size query first, then the copy."""
    ln = C.c_int32(0)
    _check(call(None, 0, C.byref(ln)))
    if ln.value == 0:
        return b""
    buf = C.create_string_buffer(ln.value)
    _check(call(buf, ln.value, C.byref(ln)))
    return buf.raw[: ln.value]


class Vocab:
    """A GGUF file's tokenizer (host only).  ``Vocab.open(path)`` needs no GPU; ``Model.vocab()`` borrows the
    loaded model's."""

    def __init__(self, handle, owned: bool, keep=None):
        self.h, self.owned, self._keep = handle, owned, keep
        self.info = VocabInfo()
        _check(lib().aios_vocab_info_get(self.h, C.byref(self.info)))

    @classmethod
    def open(cls, path: str) -> "Vocab":
        h = _P()
        _check(lib().aios_vocab_open(path.encode(), C.byref(h)))
        return cls(h, owned=True)

    def close(self):
        if self.h and self.owned:
            lib().aios_vocab_close(self.h)
        self.h = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def piece(self, i: int):
        """(piece bytes, token_type, score) of token id i."""
        ln, ty, sc = C.c_int32(0), C.c_int32(0), C.c_float(0)
        buf = C.create_string_buffer(512)
        _check(lib().aios_vocab_token_piece(self.h, i, buf, 512, C.byref(ln), C.byref(ty), C.byref(sc)))
        return buf.raw[: ln.value], ty.value, sc.value

    def tokenize(self, text: str | bytes, add_special: bool = True, parse_special: bool = False) -> np.ndarray:
        raw = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        flags = (TOK_ADD_SPECIAL if add_special else 0) | (TOK_PARSE_SPECIAL if parse_special else 0)
        n = C.c_int32(0)
        _check(lib().aios_tokenize(self.h, raw, len(raw), flags, None, 0, C.byref(n)))
        out = np.empty(max(n.value, 1), dtype=np.int32)
        _check(lib().aios_tokenize(self.h, raw, len(raw), flags, _ptr(out), out.size, C.byref(n)))
        return out[: n.value].copy()

    def detokenize(self, tokens, render_special: bool = False, strip_space: bool = False) -> bytes:
        toks = np.ascontiguousarray(tokens, dtype=np.int32)
        flags = (TOK_RENDER_SPECIAL if render_special else 0) | (TOK_STRIP_SPACE if strip_space else 0)
        return _bytes_out(lambda b, c, l: lib().aios_detokenize(self.h, _ptr(toks), toks.size, flags, b, c, l))

    def chat_prompt(self, system_prompt: str, user_prompt: str) -> str:
        return _bytes_out(lambda b, c, l: lib().aios_chat_prompt(
            self.h, (system_prompt or "").encode("utf-8"), user_prompt.encode("utf-8"), b, c, l)).decode("utf-8", "replace")

    def apply_chat_template(self, messages, add_generation_prompt: bool = True) -> str:
        """messages: [(role, content)] or [{"role":…, "content":…}] through the file's tokenizer.chat_template
        (AiosError E_UNSUPPORTED if the template is outside the supported Jinja subset or raises, E_STATE without one)."""
        pairs = [(m["role"], m["content"]) if isinstance(m, dict) else tuple(m) for m in messages]
        n = len(pairs)
        roles = (C.c_char_p * max(n, 1))(*[r.encode("utf-8") for r, _ in pairs])
        contents = (C.c_char_p * max(n, 1))(*[c.encode("utf-8") for _, c in pairs])
        gen = 1 if add_generation_prompt else 0
        return _bytes_out(lambda buf, cap, ln: lib().aios_chat_apply_template(self.h, roles, contents, n, gen, buf, cap, ln)).decode("utf-8")

    def is_eog(self, token: int) -> bool:
        r = C.c_int32(0)
        _check(lib().aios_token_is_eog(self.h, int(token), C.byref(r)))
        return bool(r.value)


class JsonGrammar:
    """JSON-object grammar over a vocabulary (aios_json_grammar): the allowed-token mask of every step, computed in C++.
    One per request; pass it to ``Model.submit(json_grammar=...)`` and free it afterwards."""

    def __init__(self, vocab: Vocab):
        self.h = _P()
        self._vocab = vocab
        _check(lib().aios_json_grammar_create(vocab.h, C.byref(self.h)))

    def step(self, last_token: int, n_vocab: int):
        """-> (rc, mask): advance by last_token (-1 first) and return the allowed set (the call the scheduler makes)."""
        mask = np.zeros((n_vocab + 7) // 8, np.uint8)
        rc = lib().aios_json_grammar_mask_cb(self.h, int(last_token), _ptr(mask), mask.size)
        return rc, mask

    @property
    def complete(self) -> bool:
        c = C.c_int32(0)
        _check(lib().aios_json_grammar_complete(self.h, C.byref(c)))
        return bool(c.value)

    def free(self):
        if self.h:
            lib().aios_json_grammar_free(self.h)
            self.h = _P()


class Engine:
    def __init__(self, device: int = 0, max_batch: int = 1, flags: int = 0):
        self.h = _P()
        opts = EngineOpts(device, max_batch, flags, 0)
        _check(lib().aios_engine_create(C.byref(opts), C.byref(self.h)))

    def close(self):
        if self.h:
            lib().aios_engine_destroy(self.h)
            self.h = _P()

    def dequantize(self, qtype: int, blocks: np.ndarray, n: int) -> np.ndarray:
        blocks = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1)
        out = np.empty(n, dtype=np.float32)
        _check(lib().aios_dequantize(self.h, qtype, _ptr(blocks), n, _ptr(out)))
        return out

    def matvec(self, qtype: int, w_blocks: np.ndarray, n_rows: int, k: int, x: np.ndarray, norm_w=None,
               eps: float = 1e-5) -> np.ndarray:
        w = np.ascontiguousarray(w_blocks, dtype=np.uint8).reshape(-1)
        x = np.ascontiguousarray(x, dtype=np.float32)
        y = np.empty(n_rows, dtype=np.float32)
        nw = None
        if norm_w is not None:
            nw = np.ascontiguousarray(norm_w, dtype=np.float32)
        _check(lib().aios_matvec(self.h, qtype, _ptr(w), n_rows, k, _ptr(x), _ptr(nw) if nw is not None else None,
                                 eps, _ptr(y)))
        return y

    def matmul(self, qtype: int, w_blocks: np.ndarray, n_rows: int, k: int, x: np.ndarray, norm_w=None,
               eps: float = 1e-5) -> np.ndarray:
        """Y[T][n_rows] for a chunk x[T][k] of 1..32 tokens through the tensor-core GEMM (aios_matmul)."""
        w = np.ascontiguousarray(w_blocks, dtype=np.uint8).reshape(-1)
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, k)
        y = np.empty((x.shape[0], n_rows), dtype=np.float32)
        nw = np.ascontiguousarray(norm_w, dtype=np.float32) if norm_w is not None else None
        _check(lib().aios_matmul(self.h, qtype, _ptr(w), n_rows, k, _ptr(x), x.shape[0],
                                 _ptr(nw) if nw is not None else None, eps, _ptr(y)))
        return y

    def matmul_tc(self, qtype: int, w_blocks: np.ndarray, n_rows: int, k: int, x: np.ndarray, norm_w=None,
                  eps: float = 1e-5) -> np.ndarray:
        """Y[T][n_rows] through the tcgen05 prefill GEMM (aios_matmul_tc): float arithmetic, stated tolerance."""
        w = np.ascontiguousarray(w_blocks, dtype=np.uint8).reshape(-1)
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, k)
        y = np.empty((x.shape[0], n_rows), dtype=np.float32)
        nw = np.ascontiguousarray(norm_w, dtype=np.float32) if norm_w is not None else None
        _check(lib().aios_matmul_tc(self.h, qtype, _ptr(w), n_rows, k, _ptr(x), x.shape[0],
                                    _ptr(nw) if nw is not None else None, eps, _ptr(y)))
        return y

    def sample_logits(self, logits: np.ndarray, sampling: Sampling, counter: int = 0, mask: np.ndarray | None = None):
        """One run of the on-device sampler chain on host logits (aios_sample_logits): returns (token, candidate ids
        best first, (K, n after top-p, n after min-p))."""
        lg = np.ascontiguousarray(logits, dtype=np.float32)
        mk = np.ascontiguousarray(mask, dtype=np.uint8) if mask is not None else None
        tok = C.c_int32(0)
        cand = np.empty(128, dtype=np.int32)
        counts = np.zeros(3, dtype=np.int32)
        _check(lib().aios_sample_logits(self.h, _ptr(lg), lg.size, _ptr(mk) if mk is not None else None,
                                        C.byref(sampling), int(counter), C.byref(tok), _ptr(cand), _ptr(counts)))
        return tok.value, cand[: counts[0]].copy(), tuple(int(c) for c in counts)

    def bench_gemv(self, qtype: int, n_rows: int, k: int, n_copies: int, iters: int) -> float:
        ms = C.c_float(0)
        _check(lib().aios_bench_gemv(self.h, qtype, n_rows, k, n_copies, iters, C.byref(ms)))
        return ms.value

    def load_model(self, name: str, path: str, n_ctx: int = 0) -> "Model":
        return Model(self, name, path, n_ctx)


class Model:
    def __init__(self, engine: Engine, name: str, path: str, n_ctx: int = 0):
        self.engine = engine
        self.h = _P()
        _check(lib().aios_model_load(engine.h, name.encode(), path.encode(), n_ctx, C.byref(self.h)))
        self.info = ModelInfo()
        _check(lib().aios_model_info_get(self.h, C.byref(self.info)))

    def unload(self):
        if self.h:
            _check(lib().aios_model_unload(self.h))
            self.h = _P()

    def tensor_info(self, i: int) -> TensorInfo:
        t = TensorInfo()
        _check(lib().aios_model_tensor_info(self.h, i, C.byref(t)))
        return t

    def meta_str(self, key: str, default: str | None = None) -> str | None:
        try:
            return _bytes_out(lambda b, c, l: lib().aios_model_meta_str(self.h, key.encode(), b, c, l)).decode(
                "utf-8", "replace")
        except AiosError:
            return default

    def vocab(self) -> "Vocab":
        """The model's tokenizer (borrowed handle: lives as long as the model)."""
        h = _P()
        _check(lib().aios_model_vocab(self.h, C.byref(h)))
        return Vocab(h, owned=False, keep=self)

    def new_seq(self) -> "Seq":
        return Seq(self)

    def submit(self, prompt, max_tokens: int, sampling: Sampling | None = None, stop_at_eog: bool = False,
               mask_cb=None, json_grammar: "JsonGrammar | None" = None) -> "Request":
        """Queue a generation on the engine's scheduler thread (aios_request_submit); returns at once.
        ``mask_cb(last_token, mask) -> int`` (optional): called by the scheduler before every step of this request with
        the token sampled in the previous step (-1 first) and a zeroed uint8 array of ceil(n_vocab / 8) bytes to fill
        with the allowed set of the next token; returns 0 (mask written), 1 (no restriction), 2 (finish now)."""
        toks = np.ascontiguousarray(prompt, dtype=np.int32)
        sp = sampling if sampling is not None else Sampling()
        h = _P()
        flags = REQ_STOP_AT_EOG if stop_at_eog else 0
        if json_grammar is not None:  # the C++ grammar drives the mask: no Python on the scheduler thread
            cb = C.cast(lib().aios_json_grammar_mask_cb, _P)
            _check(lib().aios_request_submit_masked(self.h, _ptr(toks), toks.size, C.byref(sp), max_tokens, flags, cb,
                                                    json_grammar.h, C.byref(h)))
            r = Request(h, self)
            r._keep = json_grammar
            return r
        if mask_cb is None:
            _check(lib().aios_request_submit(self.h, _ptr(toks), toks.size, C.byref(sp), max_tokens, flags, C.byref(h)))
            return Request(h, self)

        def tramp(_user, last_token, mask_ptr, n_bytes):
            try:
                arr = np.ctypeslib.as_array(mask_ptr, shape=(n_bytes,))
                return int(mask_cb(int(last_token), arr))
            except Exception:  # never unwind into the scheduler thread
                return -1

        cfun = MASK_CB(tramp)
        _check(lib().aios_request_submit_masked(self.h, _ptr(toks), toks.size, C.byref(sp), max_tokens, flags,
                                                C.cast(cfun, _P), None, C.byref(h)))
        r = Request(h, self)
        r._keep = cfun  # the C callback must outlive the request
        return r

    def bench_batch_decode(self, prompts: np.ndarray, n_steps: int, want_tokens: bool = True):
        """aios_bench_batch_decode: prompts[n_lanes][n_prompt] -> (BatchBench, tokens[n_lanes][n_steps] or None)."""
        pr = np.ascontiguousarray(prompts, dtype=np.int32)
        out = np.empty((pr.shape[0], n_steps), dtype=np.int32) if want_tokens else None
        r = BatchBench()
        _check(lib().aios_bench_batch_decode(self.h, pr.shape[0], _ptr(pr), pr.shape[1], n_steps,
                                             _ptr(out) if out is not None else None, C.byref(r)))
        return r, out

    def forward_logits(self, tokens, n_past: int = 0) -> np.ndarray:
        toks = np.ascontiguousarray(tokens, dtype=np.int32)
        out = np.empty(self.info.n_vocab, dtype=np.float32)
        _check(lib().aios_forward_logits(self.h, _ptr(toks), toks.size, n_past, _ptr(out)))
        return out

    def generate(self, prompt, max_tokens: int, sampling: Sampling | None = None) -> np.ndarray:
        toks = np.ascontiguousarray(prompt, dtype=np.int32)
        out = np.empty(max(max_tokens, 1), dtype=np.int32)
        n = C.c_int32(0)
        sp = sampling if sampling is not None else Sampling()
        _check(lib().aios_generate(self.h, _ptr(toks), toks.size, C.byref(sp), max_tokens, _ptr(out), C.byref(n)))
        return out[: n.value].copy()


REQ_STOP_AT_EOG = 1
MASK_CB = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int32, C.POINTER(C.c_uint8), C.c_int32)


class Request:
    """A queued generation: poll() never blocks, wait() blocks until there is something to poll."""

    def __init__(self, handle, model: Model):
        self.h, self.model = handle, model

    def poll(self, cap: int = 256):
        """-> (new tokens, done).  Raises AiosError if the request failed (after its tokens were taken)."""
        out = np.empty(max(cap, 1), dtype=np.int32)
        n, done = C.c_int32(0), C.c_int32(0)
        _check(lib().aios_request_poll(self.h, _ptr(out), cap, C.byref(n), C.byref(done)))
        return out[: n.value].copy(), bool(done.value)

    def wait(self, timeout_ms: int = -1) -> bool:
        r = C.c_int32(0)
        _check(lib().aios_request_wait(self.h, timeout_ms, C.byref(r)))
        return bool(r.value)

    def cancel(self):
        _check(lib().aios_request_cancel(self.h))

    def result(self, timeout_ms: int = 60000) -> np.ndarray:
        """Block until the request is finished; all its tokens."""
        parts = []
        while True:
            if not self.wait(timeout_ms):
                raise TimeoutError("request did not make progress")
            toks, done = self.poll()
            parts.append(toks)
            if done:
                return np.concatenate(parts) if parts else np.empty(0, np.int32)

    def free(self):
        if self.h:
            lib().aios_request_free(self.h)
            self.h = _P()


class Seq:
    def __init__(self, model: Model):
        self.model = model
        self.h = _P()
        _check(lib().aios_seq_create(model.h, C.byref(self.h)))

    def free(self):
        if self.h:
            lib().aios_seq_free(self.h)
            self.h = _P()

    def reset(self):
        _check(lib().aios_seq_reset(self.h))

    @property
    def n_past(self) -> int:
        n = C.c_int32(0)
        _check(lib().aios_seq_position(self.h, C.byref(n)))
        return n.value

    def prefill(self, tokens):
        toks = np.ascontiguousarray(tokens, dtype=np.int32)
        _check(lib().aios_seq_prefill(self.h, _ptr(toks), toks.size))

    def decode(self, max_tokens: int, sampling: Sampling | None = None) -> np.ndarray:
        out = np.empty(max(max_tokens, 1), dtype=np.int32)
        n = C.c_int32(0)
        sp = sampling if sampling is not None else Sampling()
        _check(lib().aios_seq_decode(self.h, C.byref(sp), max_tokens, _ptr(out), C.byref(n)))
        return out[: n.value].copy()

    def decode_static(self, max_tokens: int) -> np.ndarray:
        """Greedy decode that allocates no device memory (aios_seq_decode_static): the fused kernel's exchange
        scratch is the library's built-in static device array."""
        out = np.empty(max(max_tokens, 1), dtype=np.int32)
        n = C.c_int32(0)
        _check(lib().aios_seq_decode_static(self.h, max_tokens, _ptr(out), C.byref(n)))
        return out[: n.value].copy()

    def step(self, sampling: Sampling | None = None) -> int:
        t = C.c_int32(0)
        sp = sampling if sampling is not None else Sampling()
        _check(lib().aios_seq_step(self.h, C.byref(sp), C.byref(t)))
        return t.value

    def set_mask(self, mask: np.ndarray | None):
        """Logit mask for the sampler: packed bits (bit i of byte i//8 = token i allowed); None clears it."""
        if mask is None:
            _check(lib().aios_seq_set_mask(self.h, None, 0))
            return
        mk = np.ascontiguousarray(mask, dtype=np.uint8)
        _check(lib().aios_seq_set_mask(self.h, _ptr(mk), mk.size))

    def logits(self) -> np.ndarray:
        out = np.empty(self.model.info.n_vocab, dtype=np.float32)
        _check(lib().aios_seq_logits(self.h, _ptr(out)))
        return out

    def read_kv(self, layer: int, pos0: int, n: int):
        kvd = self.model.info.n_head_kv * self.model.info.head_dim
        k = np.empty((n, kvd), dtype=np.uint16)
        v = np.empty((n, kvd), dtype=np.uint16)
        _check(lib().aios_seq_read_kv(self.h, layer, pos0, n, _ptr(k), _ptr(v)))
        return k, v

    def read_activation(self, which: int, n: int) -> np.ndarray:
        out = np.empty(n, dtype=np.float32)
        _check(lib().aios_seq_read_activation(self.h, which, _ptr(out), n))
        return out

    def bench_prefill(self, tokens, iters: int = 3) -> float:
        """ms per prefill of `tokens` from position 0 (aios_seq_bench_prefill, CUDA events)."""
        toks = np.ascontiguousarray(tokens, dtype=np.int32)
        ms = C.c_float(0)
        _check(lib().aios_seq_bench_prefill(self.h, _ptr(toks), toks.size, iters, C.byref(ms)))
        return ms.value

    def profile_decode(self, n_tokens: int) -> BenchResult:
        r = BenchResult()
        _check(lib().aios_seq_profile_decode(self.h, n_tokens, C.byref(r)))
        return r

    def trace_decode(self, cap: int = 4096) -> np.ndarray:
        """[launch][8] u64 %globaltimer marks of one greedy token (see aios_seq_trace_decode)."""
        out = np.zeros((cap, 8), dtype=np.uint64)
        n = C.c_int32(0)
        _check(lib().aios_seq_trace_decode(self.h, _ptr(out), cap, C.byref(n)))
        return out[: n.value].copy()

    def phase_timing(self) -> np.ndarray:
        out = np.zeros((4096, 4), dtype=np.uint64)
        n = C.c_int32(0)
        _check(lib().aios_seq_phase_timing(self.h, _ptr(out), 4096, C.byref(n)))
        return out[: n.value].copy()

    def snapshot(self):
        _check(lib().aios_seq_snapshot(self.h))

    def restore(self):
        _check(lib().aios_seq_restore(self.h))

    def bench_decode(self, n_tokens: int) -> BenchResult:
        r = BenchResult()
        _check(lib().aios_seq_bench_decode(self.h, n_tokens, C.byref(r)))
        return r


class TpModel:
    """One model split over several GPUs of this node (aios_tp_*): row/K-split projections, NCCL all-reduce of the
    f64 partial sums, every rank sampling the same token from the gathered logits."""

    def __init__(self, devices, name: str, path: str, n_ctx: int = 0, flags: int = 0):
        devs = np.ascontiguousarray(devices, dtype=np.int32)
        self.h = _P()
        _check(lib().aios_tp_model_load(_ptr(devs), devs.size, name.encode(), path.encode(), n_ctx, flags, C.byref(self.h)))
        self.info = ModelInfo()
        _check(lib().aios_tp_model_info(self.h, C.byref(self.info)))
        self.n_ranks = int(devs.size)

    def unload(self):
        if self.h:
            _check(lib().aios_tp_model_unload(self.h))
            self.h = _P()

    def reset(self):
        _check(lib().aios_tp_reset(self.h))

    def prefill(self, tokens):
        toks = np.ascontiguousarray(tokens, dtype=np.int32)
        _check(lib().aios_tp_prefill(self.h, _ptr(toks), toks.size))

    def decode(self, max_tokens: int, sampling: Sampling | None = None) -> np.ndarray:
        out = np.empty(max(max_tokens, 1), dtype=np.int32)
        n = C.c_int32(0)
        sp = sampling if sampling is not None else Sampling()
        _check(lib().aios_tp_decode(self.h, C.byref(sp), max_tokens, _ptr(out), C.byref(n)))
        return out[: n.value].copy()

    def logits(self) -> np.ndarray:
        out = np.empty(self.info.n_vocab, dtype=np.float32)
        _check(lib().aios_tp_get_logits(self.h, _ptr(out), out.size))
        return out

    def generate(self, prompt, max_tokens: int, sampling: Sampling | None = None) -> np.ndarray:
        self.reset()
        self.prefill(prompt)
        return self.decode(max_tokens, sampling)

    def bench_decode(self, n_tokens: int) -> TpBench:
        r = TpBench()
        _check(lib().aios_tp_bench_decode(self.h, n_tokens, C.byref(r)))
        return r
