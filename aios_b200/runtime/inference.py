"""InferenceEngine — host-side mirror of ``runtime/src/inference.rs`` whose HTTP call to llama-server is replaced
by the B200 engine behind the C ABI (``aios_b200.capi``).

Reference behaviour kept (file:line): defaults ``max_tokens <= 0 -> 512`` and ``temperature <= 0 -> 0.7``
(inference.rs:103-112, :203-212); messages = optional system + user (``build_messages`` :363-376);
``tokens_used = prompt + completion`` (llama-server's ``usage.total_tokens`` read at :167-171); ``latency_ms`` =
wall milliseconds around the whole call (:138, :158); ``model_used`` = the resolved name (:180-185); the stream
ends with ``done: true`` (:290-299, :345-352).  What the llama-server child did per request — chat template,
tokenizer, prefill, the decode loop, the sampler chain (top-k 40, top-p 0.95, min-p 0.05, temperature; SURVEY.md
§8a D2-D5) — happens in the engine; the sampler runs on the device so a decode step moves 4 bytes to the host.

Differences, on purpose: ``stream_infer`` really streams (the reference buffers the whole SSE body first,
inference.rs:261).  JSON mode (``response_format: json_object``, :119-121): requests with
``json_mode`` go through the request queue with a JSON-object grammar (csrc/json_mask.cpp) whose allowed-token mask
the on-device sampler applies every step — n_vocab / 8 bytes down, 4 bytes up per token, no Python per token.
Models without an engine behind them (scripted test doubles) keep the host-side chain over ``json_grammar.JsonPrefix``.

There is no CPU fallback: ``EngineBackend`` needs libaios_b200.so and a B200.
"""
from __future__ import annotations

import codecs
import logging
import os
import threading
import time
from dataclasses import dataclass
from typing import Iterator, Optional

import numpy as np

log = logging.getLogger("aios_b200.runtime")

DEFAULT_MAX_TOKENS = 512  # inference.rs:103-107
DEFAULT_TEMPERATURE = 0.7  # inference.rs:108-112
STREAM_CHUNK_TOKENS = 8  # tokens per device round trip while streaming


@dataclass
class InferParams:
    """The fields of InferRequest (runtime.proto:41-50) the engine consumes."""
    prompt: str
    system_prompt: str = ""
    max_tokens: int = 0
    temperature: float = 0.0
    requesting_agent: str = ""
    task_id: str = ""
    seed: Optional[int] = None  # not on the wire; tests pin it
    json_mode: bool = False     # response_format json_object: what unary Infer always asks for (inference.rs:119-121)
    # the whole conversation, when the caller has one (the OpenAI-compatible endpoint): [(role, content)], rendered by
    # the model's own chat template; prompt / system_prompt are then only the fallback
    messages: Optional[list] = None


@dataclass
class InferResult:
    """InferResponse (runtime.proto:52-57)."""
    text: str
    tokens_used: int
    latency_ms: int
    model_used: str


def build_messages(system_prompt: str, user_prompt: str) -> list[dict]:
    """inference.rs:363-376"""
    msgs = []
    if system_prompt:
        msgs.append({"role": "system", "content": system_prompt})
    msgs.append({"role": "user", "content": user_prompt})
    return msgs


def effective_max_tokens(v: int) -> int:
    return v if v > 0 else DEFAULT_MAX_TOKENS


def effective_temperature(v: float) -> float:
    return v if v > 0.0 else DEFAULT_TEMPERATURE


class LoadedModel:
    """One model resident in HBM: what a llama-server child was for the reference."""

    def __init__(self, engine, name: str, path: str, context_length: int):
        self.name = name
        self.model = engine.load_model(name, path, context_length)
        self.vocab = self.model.vocab()
        self.n_ctx = self.model.info.n_ctx
        self.seq = self.model.new_seq()  # for direct (non-queued) use: tests, prompt caching
        self.lock = threading.Lock()
        self._live: set = set()  # queued / running requests that have not been freed yet
        self._mu = threading.Lock()
        self.closing = False

    def submit(self, tokens, max_tokens: int, sampling, json_mode: bool = False):
        """Queue a generation on the engine's scheduler; stops at the first end-of-generation token.  json_mode: the
        request carries a JSON-object grammar whose per-step mask the on-device sampler applies."""
        with self._mu:
            if self.closing:
                raise RuntimeError(f"model '{self.name}' is unloading")
            g = None
            if json_mode:
                from ..capi import JsonGrammar
                g = JsonGrammar(self.vocab)
            try:
                if g is None:
                    req = self.model.submit(tokens, max_tokens, sampling, stop_at_eog=True)
                else:
                    req = self.model.submit(tokens, max_tokens, sampling, json_grammar=g)
            except Exception:
                if g is not None:
                    g.free()
                raise
            live = _LiveRequest(req, g, self)
            self._live.add(live)
            return live

    def close(self, drain_timeout_s: float = 10.0):
        """Unload.  A request that is still queued or running is cancelled first and the unload waits until the
        engine's scheduler has retired it (the reference's unload SIGTERMs the child and waits 10 s,
        model_manager.rs:285-345): an UnloadModel that races an Infer must free the model, not leave it in HBM."""
        from ..capi import AiosError, E_STATE
        with self._mu:
            self.closing = True
            live = list(self._live)
        for r in live:
            try:
                r.cancel()
            except Exception:
                pass
        deadline = time.monotonic() + drain_timeout_s
        with self.lock:
            if self.seq is not None:
                self.seq.free()
                self.seq = None
            while True:
                try:
                    self.model.unload()
                    return
                except AiosError as e:
                    if e.code != E_STATE or time.monotonic() >= deadline:
                        with self._mu:
                            self.closing = False  # still loaded: the caller may retry the unload
                        raise
                    time.sleep(0.005)


class _LiveRequest:
    """A queued request as the runtime holds it: known to its model until freed (so that an unload can cancel it), and
    freeing its JSON grammar, if it has one, with it."""

    def __init__(self, req, grammar, owner):
        self._req, self._g, self._owner = req, grammar, owner

    def __getattr__(self, name):
        return getattr(self._req, name)

    def free(self):
        owner, self._owner = self._owner, None
        if owner is not None:
            with owner._mu:
                owner._live.discard(self)
        self._req.free()
        if self._g is not None:
            self._g.free()
            self._g = None


class _TpSeq:
    """The sequence interface the inference layer drives (reset / prefill / decode / logits) on a tensor-parallel model."""

    def __init__(self, tp):
        self.tp = tp

    def reset(self):
        self.tp.reset()

    def prefill(self, tokens):
        self.tp.prefill(tokens)

    def decode(self, max_tokens: int, sampling=None):
        return self.tp.decode(max_tokens, sampling)

    def logits(self):
        return self.tp.logits()


class TpLoaded:
    """A model split over several GPUs (``aios_tp_*``: 70B-class models, docs/HARDWARE.md:135-138).  One request at
    a time under ``lock`` (the tensor-parallel group has no request scheduler yet); same attributes as LoadedModel
    otherwise, so Infer / StreamInfer / the HTTP endpoint run on it unchanged."""

    def __init__(self, capi, devices: list, name: str, path: str, context_length: int, flags: int = 0):
        self.name, self.devices = name, list(devices)
        self.tp = capi.TpModel(self.devices, name, path, context_length, flags)
        self.vocab = capi.Vocab.open(path)
        self.n_ctx = self.tp.info.n_ctx
        self.seq = _TpSeq(self.tp)
        self.lock = threading.Lock()
        self.model = None
        self.replicas = []

    def load(self) -> dict:
        busy = 1 if self.lock.locked() else 0
        return {d: busy for d in self.devices}

    def alive(self) -> bool:
        return bool(self.tp.h)

    def close(self):
        with self.lock:  # waits for the request that is running, like the reference's SIGTERM + wait
            self.tp.unload()
            self.vocab.close()


class EngineBackend:
    """``Backend`` of ModelManager on the real engine: one ``aios_engine`` per GPU of this process.

    ``tp`` (default ``AIOS_TP``, 0 = off): models whose lower-cased name contains one of ``tp_models`` (default
    ``AIOS_TP_MODELS`` or "70b") are not replicated but split ``tp`` ways over the first ``tp`` devices.

    ``devices``: CUDA ordinals to serve on; default = ``AIOS_GPUS`` ("0,2,3" or "all") if set, else ``[device]``.
    With several devices LoadModel puts one full replica of the model on every GPU and each request goes to the
    least-loaded replica (``aios_b200.replicas.ReplicaSet``; SURVEY.md §8e: replicas only, no collective)."""

    def __init__(self, device: int = 0, flags: int = 0, devices: Optional[list] = None, tp: Optional[int] = None,
                 tp_models: Optional[tuple] = None):
        from .. import capi
        self.capi = capi
        self.flags = flags
        self.tp = int(os.environ.get("AIOS_TP", "0") or 0) if tp is None else int(tp)
        if tp_models is None:
            tp_models = tuple(x for x in os.environ.get("AIOS_TP_MODELS", "70b").lower().split(",") if x)
        self.tp_models = tuple(x.lower() for x in tp_models)
        if devices is None:
            env = os.environ.get("AIOS_GPUS", "").strip()
            if env.lower() == "all":
                devices = list(range(capi.device_count()))
            elif env:
                devices = [int(x) for x in env.split(",") if x.strip()]
            else:
                devices = [device]
        self.devices = list(devices)
        # caller-visible sequences per model: one for direct use (tests, prompt caching) and a spare; the scheduler's
        # batch lanes and the scratch sequence are not counted
        self.engines = [capi.Engine(device=d, max_batch=2, flags=flags) for d in self.devices]
        self.engine = self.engines[0]
        self._sets: list = []

    def load(self, name: str, path: str, context_length: int):
        if not os.path.isfile(path):
            raise FileNotFoundError(f"model file {path} not found")
        if self.tp > 1 and any(x in name.lower() for x in self.tp_models):
            if len(self.devices) < self.tp:
                raise RuntimeError(f"tensor parallelism {self.tp} needs {self.tp} GPUs, this backend serves {self.devices}")
            handle = TpLoaded(self.capi, self.devices[: self.tp], name, path, context_length, self.flags)
            self._sets.append(handle)
            return handle
        from ..replicas import ReplicaSet
        loaded = []
        try:
            for e in self.engines:
                loaded.append(LoadedModel(e, name, path, context_length))
        except Exception:
            for lm in loaded:
                lm.close()
            raise
        rs = ReplicaSet(loaded, self.devices)
        self._sets.append(rs)
        return rs

    def unload(self, handle) -> None:
        if handle in self._sets:
            self._sets.remove(handle)
        handle.close()

    def alive(self, handle) -> bool:
        if isinstance(handle, TpLoaded):
            return handle.alive()
        return all(bool(r.model.h) for r in handle.replicas)

    def gpu_load(self) -> dict:
        """device -> (models resident, requests in flight), for HealthCheck."""
        out = {d: [0, 0] for d in self.devices}
        for rs in self._sets:
            for d, n in rs.load().items():
                out[d][0] += 1
                out[d][1] += n
        return out

    def close(self):
        for e in self.engines:
            e.close()


class InferenceEngine:
    """Stateless, shared by all handlers, like the reference's ``Arc<InferenceEngine>`` (main.rs:51)."""

    def __init__(self, stream_chunk_tokens: int = STREAM_CHUNK_TOKENS):
        self.stream_chunk_tokens = stream_chunk_tokens

    # -- what llama-server did before the forward passes: template + tokenize
    def _prepare(self, lm: LoadedModel, p: InferParams):
        text = None
        if p.messages and hasattr(lm.vocab, "apply_chat_template"):
            try:
                text = lm.vocab.apply_chat_template(p.messages, True)
            except Exception:  # no template in the file, outside the supported subset, or the template raised
                text = None
        if text is None:
            text = lm.vocab.chat_prompt(p.system_prompt, p.prompt)
        tokens = lm.vocab.tokenize(text, add_special=True, parse_special=True)
        if tokens.size >= lm.n_ctx:
            raise RuntimeError(f"the request exceeds the available context size ({tokens.size} prompt tokens, "
                               f"context {lm.n_ctx})")
        max_tokens = min(effective_max_tokens(p.max_tokens), lm.n_ctx - int(tokens.size))
        from ..capi import make_sampling
        seed = p.seed if p.seed is not None else int.from_bytes(os.urandom(8), "little")
        sampling = make_sampling(temperature=effective_temperature(p.temperature), seed=seed)
        return tokens, max_tokens, sampling

    def _generate(self, lm: LoadedModel, p: InferParams, chunk: int) -> Iterator[tuple[np.ndarray, int, bool]]:
        """Yields (new completion tokens without the end-of-generation token, n_prompt, finished)."""
        tokens, max_tokens, sampling = self._prepare(lm, p)
        if hasattr(lm, "submit"):
            # the engine's own scheduler thread (aios_request_*): concurrent RPCs are batched there (continuous
            # batching, one token per request per step); with several GPUs the least-loaded replica takes the request
            req = lm.submit(tokens, max_tokens, sampling)
            try:
                while True:
                    if not req.wait(120000):  # the reference's HTTP timeout (inference.rs:81)
                        raise TimeoutError("generation made no progress for 120 s")
                    out, done = req.poll(chunk)
                    if out.size or done:
                        yield out, int(tokens.size), done
                    if done:
                        return
            finally:
                req.free()
        with lm.lock:
            lm.seq.reset()
            lm.seq.prefill(tokens)
            left = max_tokens
            while left > 0:
                out = lm.seq.decode(min(chunk, left), sampling)
                if out.size == 0:
                    break
                left -= int(out.size)
                stop = next((i for i, t in enumerate(out) if lm.vocab.is_eog(int(t))), None)
                if stop is not None:
                    yield out[:stop], int(tokens.size), True
                    return
                yield out, int(tokens.size), left <= 0
            yield np.empty(0, np.int32), int(tokens.size), True

    # -- JSON mode: llama-server turns response_format json_object into a grammar that masks the sampler's candidates
    #    (SURVEY.md §8f-3).  Here the sampler chain runs on the host for these requests: per step the logits come back
    #    (128 KB D2H), candidates are walked best first and kept only if json_grammar accepts their piece.
    @staticmethod
    def _pieces(lm) -> list:
        pieces = getattr(lm, "_piece_bytes", None)
        if pieces is None:
            n = lm.vocab.info.n_tokens
            pieces = [lm.vocab.detokenize([i]) for i in range(n)]
            lm._piece_bytes = pieces
        return pieces

    def _generate_json(self, lm, p: InferParams):
        """-> (completion tokens, n_prompt).  Stops when the JSON object closes, at max_tokens, or when nothing fits.

        On the real engine the request goes through the scheduler like any other (it shares the batch with whatever
        else is running): the JSON grammar runs in C++ on the scheduler thread (aios_json_grammar_mask_cb), the allowed
        set goes to the GPU as a bit mask (n_vocab / 8 bytes per step) and the on-device sampler applies it — the
        logits never leave the GPU and no Python runs per token.  The host chain below is what scripted test models
        (no engine) use."""
        from .json_grammar import JsonPrefix
        tokens, max_tokens, sampling = self._prepare(lm, p)
        if getattr(lm, "model", None) is not None and hasattr(lm, "submit") and hasattr(lm.vocab, "h"):
            req = lm.submit(tokens, max_tokens, sampling, json_mode=True)
            try:
                out = req.result(120000)
            finally:
                req.free()
            return np.asarray(out, np.int32), int(tokens.size)
        pieces = self._pieces(lm)
        rng = np.random.default_rng(sampling.seed & 0xFFFFFFFF)
        temp, top_k, top_p, min_p = float(sampling.temperature), int(sampling.top_k), float(sampling.top_p), float(sampling.min_p)
        state, out = JsonPrefix(), []
        with lm.lock:
            lm.seq.reset()
            lm.seq.prefill(tokens)
            for _ in range(max_tokens):
                logits = lm.seq.logits()
                n_look = min(256, logits.size)
                order = np.argpartition(-logits, n_look - 1)[:n_look]
                order = order[np.argsort(-logits[order], kind="stable")]
                cand = [int(t) for t in order if state.accepts(pieces[int(t)])][: max(top_k, 1)]
                if not cand:  # rare: nothing in the best 256 fits — look at the whole vocabulary
                    for t in np.argsort(-logits, kind="stable"):
                        if state.accepts(pieces[int(t)]):
                            cand = [int(t)]
                            break
                if not cand:
                    break
                lg = logits[cand].astype(np.float64)
                pr = np.exp(lg - lg[0])
                keep = len(cand)
                if top_p < 1.0:
                    keep = min(keep, int(np.searchsorted(np.cumsum(pr / pr.sum()), top_p) + 1))
                if min_p > 0.0:
                    keep = min(keep, max(1, int((pr >= min_p * pr[0]).sum())))
                pr = np.exp((lg[:keep] - lg[0]) / max(temp, 1e-6))
                tok = cand[int(rng.choice(keep, p=pr / pr.sum()))]
                state.feed(pieces[tok])
                out.append(tok)
                if state.complete:
                    break
                lm.seq.prefill(np.array([tok], np.int32))
        return np.asarray(out, np.int32), int(tokens.size)

    def infer(self, lm: LoadedModel, model_name: str, p: InferParams) -> InferResult:
        """inference.rs:94-186"""
        start = time.monotonic()
        if p.json_mode and hasattr(lm.seq, "logits"):
            comp, n_prompt = self._generate_json(lm, p)
            text = lm.vocab.detokenize(comp).decode("utf-8", "replace")
            return InferResult(text=text, tokens_used=n_prompt + int(comp.size),
                               latency_ms=int((time.monotonic() - start) * 1000), model_used=model_name)
        log.info("Sending inference request model=%s max_tokens=%d temperature=%.3g agent=%s task=%s", model_name,
                 effective_max_tokens(p.max_tokens), effective_temperature(p.temperature), p.requesting_agent, p.task_id)
        done_tokens, n_prompt = [], 0
        for toks, n_prompt, finished in self._generate(lm, p, chunk=64):
            done_tokens.append(toks)
            if finished:
                break
        comp = np.concatenate(done_tokens) if done_tokens else np.empty(0, np.int32)
        text = lm.vocab.detokenize(comp).decode("utf-8", "replace")
        latency_ms = int((time.monotonic() - start) * 1000)
        return InferResult(text=text, tokens_used=n_prompt + int(comp.size), latency_ms=latency_ms, model_used=model_name)

    def stream_infer(self, lm: LoadedModel, model_name: str, p: InferParams) -> Iterator[tuple[str, bool]]:
        """inference.rs:194-356: yields (text, done); the last item is always done=True."""
        utf8 = codecs.getincrementaldecoder("utf-8")("replace")
        for toks, _, finished in self._generate(lm, p, chunk=self.stream_chunk_tokens):
            text = utf8.decode(lm.vocab.detokenize(toks), final=finished) if toks.size or finished else ""
            if finished:
                if text:
                    yield text, False
                yield "", True
                return
            if text:
                yield text, False
        yield "", True
