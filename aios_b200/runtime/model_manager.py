"""ModelManager — host-side mirror of ``runtime/src/model_manager.rs`` with the llama-server child replaced by an
in-process model handle on the B200 engine.

Same public surface and observable behaviour as the reference (file:line cited per method); what changes is what
a "managed model" is: the reference spawns ``llama-server --model … --port …`` and polls ``/health``
(model_manager.rs:187-277), here ``backend.load(name, path, ctx)`` maps the GGUF into HBM through the C ABI
(``aios_model_load``) and returns a handle.  ``port`` stays in ``ModelStatus`` (model_manager.rs:543) as a stable
pseudo-port from the same allocator (BASE_PORT 8080, :69, :137-144) — nothing listens on it.

Differences, on purpose (SURVEY.md §8b "Threading"): the table lock is NOT held while a model loads, so routing
and inference on other models keep running during a load (the reference blocks them for up to 120 s,
grpc_service.rs:40-41); a second LoadModel for a name that is still loading gets ``status: "loading"`` back,
which is what the reference's idempotence rule (:152-157) says for that state.
"""
from __future__ import annotations

import logging
import os
import threading
import time
from dataclasses import dataclass, field
from typing import Any, Optional, Protocol

log = logging.getLogger("aios_b200.runtime")

BASE_PORT = 8080  # model_manager.rs:69
DEFAULT_MODEL_DIR = "/var/lib/aios/models/"  # model_manager.rs:118-120
DEFAULT_CONTEXT = 2048  # model_manager.rs:167-171
DEFAULT_THREADS = 4  # model_manager.rs:173 (kept in the record; the GPU engine has no use for it)


def now_epoch_ms() -> int:
    """model_manager.rs:100-105"""
    return int(time.time() * 1000)


class ModelState:
    """model_manager.rs:24-40: the four states and their display strings."""
    LOADING, READY, ERROR, UNLOADING = "loading", "ready", "error", "unloading"

    def __init__(self, kind: str, msg: str = ""):
        self.kind, self.msg = kind, msg

    def __str__(self) -> str:
        return f"error: {self.msg}" if self.kind == self.ERROR else self.kind

    @property
    def ready(self) -> bool:
        return self.kind == self.READY


@dataclass
class ManagedModel:
    """model_manager.rs:42-56, with ``process: Option<Child>`` replaced by ``handle``."""
    name: str
    path: str
    port: int
    status: ModelState
    loaded_at: int
    last_used: int
    request_count: int = 0
    context_length: int = DEFAULT_CONTEXT
    gpu_layers: int = 0
    threads: int = DEFAULT_THREADS
    handle: Any = None
    restarts: int = 0  # consecutive automatic restarts that have not yet produced a healthy model
    http: Any = None  # the model's OpenAI-compatible endpoint on `port`, when the manager was given an http_factory
    loaded_evt: threading.Event = field(default_factory=threading.Event, repr=False)


@dataclass
class ModelStatusRecord:
    """Plain-data ModelStatus (runtime.proto:28-35); ``model_to_status`` at model_manager.rs:539-548."""
    model_name: str
    status: str
    port: int
    loaded_at: int
    last_used: int
    request_count: int


@dataclass
class LoadRequest:
    """LoadModelRequest (runtime.proto:15-22)."""
    model_name: str
    model_path: str = ""
    context_length: int = 0
    gpu_layers: int = 0
    threads: int = 0
    port: int = 0


class Backend(Protocol):
    """What stands where the llama-server child stood."""

    def load(self, name: str, path: str, context_length: int) -> Any: ...
    def unload(self, handle: Any) -> None: ...
    def alive(self, handle: Any) -> bool: ...


def model_to_status(m: ManagedModel) -> ModelStatusRecord:
    return ModelStatusRecord(m.name, str(m.status), int(m.port), m.loaded_at, m.last_used, m.request_count)


MAX_RESTARTS = 3  # consecutive failed restarts after which a model is left in its error state


class ModelManager:
    def __init__(self, backend: Backend, model_dir: Optional[str] = None, http_factory=None,
                 idle_unload_minutes: Optional[float] = None, idle_exempt: tuple = ("tinyllama",),
                 auto_restart: bool = True):
        # idle_unload_minutes: config/default-config.toml:26 (`idle_unload_minutes = 5`) and
        # docs/phases/04-AI-RUNTIME.md:208 ("unload tactical models after idle timeout"); None reads
        # AIOS_IDLE_UNLOAD_MINUTES, 0 = never.  Models whose lower-cased name contains one of `idle_exempt` stay (the
        # operational model must always be there, 04-AI-RUNTIME.md:206).  auto_restart: 04-AI-RUNTIME.md:210 /
        # model_manager.rs:391-392 ("Automatic restart can be added later"): a model whose handle died is reloaded.
        if idle_unload_minutes is None:
            idle_unload_minutes = float(os.environ.get("AIOS_IDLE_UNLOAD_MINUTES", "0") or 0)
        self.idle_unload_ms = int(idle_unload_minutes * 60_000)
        self.idle_exempt = tuple(s.lower() for s in idle_exempt)
        self.auto_restart = auto_restart
        # http_factory(name, handle, port) -> object with .close(): serves the loaded model on its port the way the
        # llama-server child did; None = the port stays a pseudo-port
        self.http_factory = http_factory
        # model_manager.rs:115-131
        self.backend = backend
        self.model_dir = model_dir or os.environ.get("AIOS_MODEL_DIR", DEFAULT_MODEL_DIR)
        self.models: dict[str, ManagedModel] = {}
        self.next_port = BASE_PORT
        self.lock = threading.RLock()
        log.info("ModelManager initialised model_dir=%s", self.model_dir)

    # ------------------------------------------------------------------ port allocation (model_manager.rs:134-144)
    def allocate_port(self, requested: int) -> int:
        if requested > 0:
            return requested
        port = self.next_port
        self.next_port += 1
        return port

    # ------------------------------------------------------------------ load (model_manager.rs:149-277)
    def load_model(self, req: LoadRequest) -> ModelStatusRecord:
        name = req.model_name
        with self.lock:
            existing = self.models.get(name)
            if existing is not None and existing.status.kind in (ModelState.READY, ModelState.LOADING):
                log.info("Model %s already loaded, returning current status", name)
                return model_to_status(existing)
            path = req.model_path if req.model_path else os.path.join(self.model_dir, name)
            port = self.allocate_port(req.port & 0xFFFF)  # `req.port as u16`
            ctx = req.context_length if req.context_length > 0 else DEFAULT_CONTEXT
            threads = req.threads if req.threads > 0 else DEFAULT_THREADS
            now = now_epoch_ms()
            managed = ManagedModel(name=name, path=path, port=port, status=ModelState(ModelState.LOADING),
                                   loaded_at=now, last_used=now, context_length=ctx, gpu_layers=req.gpu_layers,
                                   threads=threads)
            self.models[name] = managed
        log.info("Loading model %s path=%s port=%d ctx=%d", name, path, port, ctx)
        http = None
        try:
            handle = self.backend.load(name, path, ctx)
            if self.http_factory is not None:
                try:
                    http = self.http_factory(name, handle, port)
                except Exception:
                    self.backend.unload(handle)
                    raise
        except Exception as e:  # the reference's "exited early" branch: recorded in-band AND returned as Err
            msg = str(e)
            with self.lock:
                managed.status = ModelState(ModelState.ERROR, msg)
                managed.loaded_evt.set()
            log.error("Model %s failed to load: %s", name, msg)
            raise RuntimeError(msg) from e
        with self.lock:
            if managed.status.kind == ModelState.UNLOADING or self.models.get(name) is not managed:
                # unloaded while it was loading: drop the fresh handle
                if http is not None:
                    http.close()
                self.backend.unload(handle)
                managed.loaded_evt.set()
                return model_to_status(managed)
            managed.handle = handle
            managed.http = http
            managed.status = ModelState(ModelState.READY)
            managed.loaded_evt.set()
            log.info("Model %s is ready", name)
            return model_to_status(managed)

    # ------------------------------------------------------------------ unload (model_manager.rs:285-345)
    def unload_model(self, name: str) -> None:
        """The reference kills the child, which always frees the model.  Here the backend cancels and drains what is
        in flight before it unloads; if that fails the entry stays, in its error state and with its handle, so that
        the unload can be retried (nothing is left unreachable in HBM)."""
        with self.lock:
            model = self.models.get(name)
            if model is None:
                raise KeyError(f"Model '{name}' not found")
            model.status = ModelState(ModelState.UNLOADING)
            handle = model.handle
            http, model.http = model.http, None
        log.info("Unloading model %s", name)
        try:
            if http is not None:
                http.close()
            if handle is not None:
                self.backend.unload(handle)
        except Exception as e:
            with self.lock:
                model.status = ModelState(ModelState.ERROR, f"unload failed: {e}")
            log.error("Model %s failed to unload: %s", name, e)
            raise
        with self.lock:
            model.handle = None
            if self.models.get(name) is model:
                del self.models[name]
        log.info("Model %s unloaded", name)

    # ------------------------------------------------------------------ list / get (model_manager.rs:352-376)
    def list_models(self) -> list[ModelStatusRecord]:
        with self.lock:
            return [model_to_status(m) for m in self.models.values()]

    def get_model(self, name: str) -> Optional[ManagedModel]:
        with self.lock:
            m = self.models.get(name)
            if m is not None:
                m.last_used = now_epoch_ms()
            return m

    def model_port(self, name: str) -> Optional[int]:
        """Ready check, ``last_used`` stamp, ``request_count += 1`` (model_manager.rs:367-376)."""
        with self.lock:
            m = self.get_model(name)
            if m is not None and m.status.ready:
                m.request_count += 1
                return m.port
            return None

    def model_handle(self, name: str) -> Any:
        """The engine handle that replaces the port as the coupling to the backend."""
        with self.lock:
            m = self.models.get(name)
            return m.handle if m is not None and m.status.ready else None

    # ------------------------------------------------------------------ health (model_manager.rs:393-447)
    def health_check_all(self) -> None:
        """Reference: child liveness + GET /health per model, flipping Ready <-> Error.  Here: ask the backend
        whether the handle is still usable (a sticky CUDA error is the in-process version of a dead child).  On top of
        the reference, the two behaviours its phase document promises for this loop (04-AI-RUNTIME.md:246-250):
        a dead model is restarted, and a model idle for longer than the timeout is unloaded."""
        restart, idle = [], []
        now = now_epoch_ms()
        with self.lock:
            for m in self.models.values():
                if m.status.kind in (ModelState.LOADING, ModelState.UNLOADING) or m.handle is None:
                    continue
                try:
                    ok = self.backend.alive(m.handle)
                except Exception as e:
                    ok, why = False, str(e)
                else:
                    why = "engine handle no longer usable"
                if ok and m.status.kind == ModelState.ERROR:
                    log.info("Model %s recovered", m.name)
                    m.status = ModelState(ModelState.READY)
                    m.restarts = 0
                elif not ok and m.status.kind != ModelState.ERROR:
                    log.warning("Model %s health check failed: %s", m.name, why)
                    m.status = ModelState(ModelState.ERROR, why)
                if not ok and self.auto_restart and m.restarts < MAX_RESTARTS:
                    restart.append(m)
                elif ok and m.status.ready and self.idle_unload_ms > 0 and now - m.last_used >= self.idle_unload_ms \
                        and not any(x in m.name.lower() for x in self.idle_exempt) and not self._busy(m.handle):
                    idle.append(m.name)
        for m in restart:
            self._restart(m)
        for name in idle:
            log.info("Model %s idle for %d min: unloading", name, self.idle_unload_ms // 60_000)
            try:
                self.unload_model(name)
            except Exception as e:  # stays in the table in its error state
                log.warning("idle unload of %s failed: %s", name, e)

    @staticmethod
    def _busy(handle) -> bool:
        load = getattr(handle, "load", None)
        try:
            return bool(load and sum(load().values()) > 0)
        except Exception:
            return False

    def _restart(self, m: ManagedModel) -> None:
        """Crash recovery: drop what is left of the dead handle and load the same file again on the same port."""
        m.restarts += 1
        log.warning("Restarting model %s (attempt %d of %d)", m.name, m.restarts, MAX_RESTARTS)
        with self.lock:
            if self.models.get(m.name) is not m:
                return
            old, m.handle = m.handle, None
            m.status = ModelState(ModelState.LOADING)
        try:
            self.backend.unload(old)
        except Exception as e:
            log.warning("could not release the dead handle of %s: %s", m.name, e)
        try:
            handle = self.backend.load(m.name, m.path, m.context_length)
        except Exception as e:
            with self.lock:
                m.status = ModelState(ModelState.ERROR, f"restart failed: {e}")
            log.error("Model %s failed to restart: %s", m.name, e)
            return
        with self.lock:
            if self.models.get(m.name) is not m or m.status.kind == ModelState.UNLOADING:
                self.backend.unload(handle)
                return
            m.handle = handle
            if m.http is not None and hasattr(m.http, "handle"):
                m.http.handle = handle  # the model's HTTP endpoint keeps its port and serves the new handle
            m.status = ModelState(ModelState.READY)
            m.loaded_at = now_epoch_ms()
            m.restarts = 0
        log.info("Model %s restarted", m.name)

    # ------------------------------------------------------------------ level routing (model_manager.rs:462-518)
    LEVELS = {
        "operational": ["tinyllama-1.1b", "DeepSeek-R1-Distill-Qwen-8B", "mistral-7b"],
        "tactical": ["DeepSeek-R1-Distill-Qwen-8B", "Qwen3-14B", "mistral-7b", "tinyllama-1.1b"],
        "strategic": ["Qwen3-14B", "DeepSeek-R1-Distill-Qwen-8B", "mistral-7b"],
    }

    def select_model_for_level(self, level: str) -> Optional[str]:
        if level == "reactive":
            return None  # heuristics, no LLM
        if level in self.LEVELS:
            return self.first_ready_from(self.LEVELS[level])
        log.warning("Unknown intelligence level %r, falling back to first ready model", level)
        return self.first_ready_model()

    def first_ready_from(self, candidates: list[str]) -> Optional[str]:
        with self.lock:
            for cand in candidates:
                c = cand.lower()
                for name, m in self.models.items():
                    if m.status.ready and c in name.lower():
                        return name
        return None

    def is_model_ready(self, name: str) -> bool:
        with self.lock:
            m = self.models.get(name)
            return m is not None and m.status.ready

    def first_ready_model(self) -> Optional[str]:
        with self.lock:
            for m in self.models.values():
                if m.status.ready:
                    return m.name
        return None
