"""gRPC service ``aios.runtime.AIRuntime`` on ``[::]:50055`` — mirror of ``runtime/src/grpc_service.rs`` and
``runtime/src/main.rs`` in front of the B200 engine.

A caller of the reference (Python ``AIRuntimeStub`` agent-core/python/aios_agent/base.py:192-197, Rust
``AiRuntimeClient`` agent-core/src/clients.rs:100-110) talks to this server unchanged: same service and message
names on the wire (``aios_b200.runtime.proto``), same status strings, messages and gRPC status codes:

* LoadModel   -> ModelStatus, errors ``Internal("Failed to load model: …")``      grpc_service.rs:33-48
* UnloadModel -> Status{success, "Model '<n>' unloaded"}, ``Internal("Failed to unload model: …")``  :53-72
* ListModels  -> ModelList                                                        :77-82
* Infer / StreamInfer: resolve model (explicit name -> level routing -> any ready model), ``InvalidArgument`` for
  level "reactive", ``FailedPrecondition`` for "strategic", ``Unavailable`` when nothing is loaded,
  ``Internal("Inference failed: …")``                                            :87-141, :187-233
* HealthCheck -> HealthStatus{healthy:true, service:"aios-runtime", details{loaded_models,total_models,
  model:<n>}}                                                                     :146-178

Start-up auto-loads every ``*.gguf`` in ``$AIOS_MODEL_DIR`` with the reference's size-based context choice
(main.rs:65-132) and runs the periodic health task (main.rs:55-63).
"""
from __future__ import annotations

import argparse
import logging
import os
import threading
import time
from concurrent import futures
from typing import Optional

import grpc

from . import proto as pb
from .inference import InferenceEngine, InferParams
from .model_manager import LoadRequest, ModelManager

log = logging.getLogger("aios_b200.runtime")

LISTEN_ADDR = "[::]:50055"  # main.rs:140
HEALTH_CHECK_INTERVAL_S = 30.0  # main.rs HEALTH_CHECK_INTERVAL


def _status_msg(s) -> "pb.ModelStatus":
    return pb.ModelStatus(model_name=s.model_name, status=s.status, port=s.port, loaded_at=s.loaded_at,
                          last_used=s.last_used, request_count=s.request_count)


class AIRuntimeService:
    def __init__(self, model_manager: ModelManager, inference_engine: InferenceEngine):
        self.model_manager = model_manager
        self.inference_engine = inference_engine
        self.start_time = time.monotonic()

    # ---------------------------------------------------------------- grpc_service.rs:33-48
    def LoadModel(self, req, context):
        log.info("gRPC LoadModel model=%s", req.model_name)
        try:
            st = self.model_manager.load_model(LoadRequest(req.model_name, req.model_path, req.context_length,
                                                           req.gpu_layers, req.threads, req.port))
        except Exception as e:
            log.error("LoadModel failed: %s", e)
            context.abort(grpc.StatusCode.INTERNAL, f"Failed to load model: {e}")
        return _status_msg(st)

    # ---------------------------------------------------------------- grpc_service.rs:53-72
    def UnloadModel(self, req, context):
        log.info("gRPC UnloadModel model=%s", req.model_name)
        try:
            self.model_manager.unload_model(req.model_name)
        except KeyError as e:
            context.abort(grpc.StatusCode.INTERNAL, f"Failed to unload model: {e.args[0]}")
        except Exception as e:
            log.error("UnloadModel failed: %s", e)
            context.abort(grpc.StatusCode.INTERNAL, f"Failed to unload model: {e}")
        return pb.Status(success=True, message=f"Model '{req.model_name}' unloaded")

    # ---------------------------------------------------------------- grpc_service.rs:77-82
    def ListModels(self, req, context):
        models = self.model_manager.list_models()
        log.info("gRPC ListModels count=%d", len(models))
        return pb.ModelList(models=[_status_msg(m) for m in models])

    # ---------------------------------------------------------------- grpc_service.rs:187-233
    def resolve_model(self, req, context):
        """-> (engine handle, model name); aborts the RPC with the reference's status codes."""
        mgr = self.model_manager
        with mgr.lock:
            if req.model:
                if mgr.model_port(req.model) is not None:
                    return mgr.model_handle(req.model), req.model
                log.warning("Requested model %s not ready, trying level routing", req.model)
            if req.intelligence_level:
                name = mgr.select_model_for_level(req.intelligence_level)
                if name is not None and mgr.model_port(name) is not None:
                    return mgr.model_handle(name), name
                if req.intelligence_level == "reactive":
                    context.abort(grpc.StatusCode.INVALID_ARGUMENT,
                                  "Reactive level does not require LLM inference — handle with heuristics")
                if req.intelligence_level == "strategic":
                    context.abort(grpc.StatusCode.FAILED_PRECONDITION,
                                  "Strategic level requires external API — route via api-gateway")
            for m in mgr.list_models():
                if m.status == "ready" and mgr.model_port(m.model_name) is not None:
                    return mgr.model_handle(m.model_name), m.model_name
        context.abort(grpc.StatusCode.UNAVAILABLE, "No model available for inference.  Load a model first with LoadModel.")

    @staticmethod
    def _params(req) -> InferParams:
        return InferParams(prompt=req.prompt, system_prompt=req.system_prompt, max_tokens=req.max_tokens,
                           temperature=req.temperature, requesting_agent=req.requesting_agent, task_id=req.task_id)

    # ---------------------------------------------------------------- grpc_service.rs:87-113
    def Infer(self, req, context):
        log.info("gRPC Infer model=%s level=%s agent=%s task=%s", req.model, req.intelligence_level,
                 req.requesting_agent, req.task_id)
        handle, name = self.resolve_model(req, context)
        try:
            params = self._params(req)
            params.json_mode = True  # the reference's unary call always sets response_format json_object (inference.rs:119-121)
            r = self.inference_engine.infer(handle, name, params)
        except Exception as e:
            log.error("Inference failed model=%s: %s", name, e)
            context.abort(grpc.StatusCode.INTERNAL, f"Inference failed: {e}")
        return pb.InferResponse(text=r.text, tokens_used=r.tokens_used, latency_ms=r.latency_ms, model_used=r.model_used)

    # ---------------------------------------------------------------- grpc_service.rs:118-141
    def StreamInfer(self, req, context):
        log.info("gRPC StreamInfer model=%s level=%s agent=%s task=%s", req.model, req.intelligence_level,
                 req.requesting_agent, req.task_id)
        handle, name = self.resolve_model(req, context)
        it = self.inference_engine.stream_infer(handle, name, self._params(req))
        try:
            first = next(it)  # prompt errors surface before the first chunk, like the reference's HTTP status check
        except StopIteration:
            first = ("", True)
        except Exception as e:
            log.error("Stream inference failed model=%s: %s", name, e)
            context.abort(grpc.StatusCode.INTERNAL, f"Stream inference failed: {e}")
        try:
            yield pb.InferChunk(text=first[0], done=first[1])
            if not first[1]:
                for text, done in it:
                    if not context.is_active():
                        log.warning("Stream receiver dropped model=%s", name)
                        break
                    yield pb.InferChunk(text=text, done=done)
        except Exception as e:
            log.error("Stream read error model=%s: %s", name, e)
            context.abort(grpc.StatusCode.INTERNAL, f"Stream read error: {e}")
        finally:
            it.close()

    # ---------------------------------------------------------------- grpc_service.rs:146-178
    def HealthCheck(self, req, context):
        models = self.model_manager.list_models()
        loaded = sum(1 for m in models if m.status == "ready")
        total = len(models)
        uptime = int(time.monotonic() - self.start_time)
        details = {"loaded_models": str(loaded), "total_models": str(total)}
        for m in models:
            details[f"model:{m.model_name}"] = f"status={m.status} port={m.port}"
        gpu_load = getattr(self.model_manager.backend, "gpu_load", None)
        if gpu_load:  # one engine per GPU: resident replicas and requests in flight on each
            for dev, (n_models, n_req) in sorted(gpu_load().items()):
                details[f"gpu:{dev}"] = f"models={n_models} inflight={n_req}"
        log.info("gRPC HealthCheck loaded=%d total=%d uptime=%d", loaded, total, uptime)
        return pb.HealthStatus(healthy=True, service="aios-runtime",
                               message=f"{loaded}/{total} models loaded, uptime {uptime}s", uptime_seconds=uptime,
                               details=details)


def generic_handler(service: AIRuntimeService) -> grpc.GenericRpcHandler:
    handlers = {}
    for name, (req_cls, resp_cls, stream) in pb.METHODS.items():
        mk = grpc.unary_stream_rpc_method_handler if stream else grpc.unary_unary_rpc_method_handler
        handlers[name] = mk(getattr(service, name), request_deserializer=req_cls.FromString,
                            response_serializer=resp_cls.SerializeToString)
    return grpc.method_handlers_generic_handler(pb.SERVICE, handlers)


def autoload_plan(model_dir: str) -> list[LoadRequest]:
    """main.rs:65-132: every ``*.gguf`` of the directory, name = file stem, context/threads by file size."""
    plan = []
    if not os.path.isdir(model_dir):
        log.info("Model directory %s not found, skipping auto-load", model_dir)
        return plan
    for entry in os.scandir(model_dir):
        stem, ext = os.path.splitext(entry.name)
        if ext != ".gguf" or not entry.is_file():
            continue
        size = entry.stat().st_size
        if size > 8_000_000_000:
            ctx, threads = 8192, 4
        elif size > 4_000_000_000:
            ctx, threads = 4096, 4
        elif size > 2_000_000_000:
            ctx, threads = 4096, 3
        else:
            ctx, threads = 2048, 2
        plan.append(LoadRequest(model_name=stem, model_path=entry.path, context_length=ctx, gpu_layers=0,
                                threads=threads, port=0))
    return plan


def autoload(mgr: ModelManager) -> None:
    for req in autoload_plan(mgr.model_dir):
        try:
            st = mgr.load_model(req)
            log.info("Model auto-loaded model=%s status=%s port=%d", req.model_name, st.status, st.port)
        except Exception as e:
            log.error("Failed to auto-load %s: %s", req.model_name, e)


class RuntimeServer:
    """The running process of main.rs: service + health task + listener."""

    def __init__(self, backend, addr: str = LISTEN_ADDR, model_dir: Optional[str] = None, max_workers: int = 16,
                 health_interval_s: float = HEALTH_CHECK_INTERVAL_S, inference_engine: Optional[InferenceEngine] = None,
                 openai_http: bool = False, idle_unload_minutes: Optional[float] = None, auto_restart: bool = True):
        engine = inference_engine or InferenceEngine()
        factory = None
        if openai_http:  # every loaded model also answers OpenAI-compatible HTTP on its ModelStatus.port
            from .openai_http import OpenAIServer
            factory = lambda name, handle, port: OpenAIServer(handle, name, port, engine).start()  # noqa: E731
        self.manager = ModelManager(backend, model_dir, http_factory=factory, idle_unload_minutes=idle_unload_minutes,
                                    auto_restart=auto_restart)
        self.service = AIRuntimeService(self.manager, engine)
        self.server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers))
        self.server.add_generic_rpc_handlers((generic_handler(self.service),))
        self.port = self.server.add_insecure_port(addr)
        if self.port == 0:
            raise OSError(f"cannot listen on {addr}")
        self._stop = threading.Event()
        self._health = threading.Thread(target=self._health_loop, args=(health_interval_s,), daemon=True)

    def _health_loop(self, interval: float):
        while not self._stop.wait(interval):
            self.manager.health_check_all()

    def start(self, do_autoload: bool = True):
        self._health.start()
        if do_autoload:
            autoload(self.manager)
        self.server.start()
        log.info("aiOS AI Runtime (B200 engine) listening on port %d", self.port)
        return self

    def stop(self, grace: float = 1.0):
        self._stop.set()
        self.server.stop(grace).wait()
        for m in self.manager.list_models():
            try:
                self.manager.unload_model(m.model_name)
            except Exception as e:  # pragma: no cover
                log.error("unload %s on shutdown: %s", m.model_name, e)


def main(argv=None):
    ap = argparse.ArgumentParser(description="aios-runtime drop-in on the B200 engine")
    ap.add_argument("--addr", default=LISTEN_ADDR)
    ap.add_argument("--device", type=int, default=int(os.environ.get("LOCAL_RANK", "0")))
    ap.add_argument("--gpus", default=None, help='CUDA ordinals to serve on, e.g. "0,1,2,3" or "all" (default: $AIOS_GPUS, '
                                                 "else --device): one engine and one replica of every model per GPU, "
                                                 "requests go to the least-loaded replica")
    ap.add_argument("--tp", type=int, default=None,
                    help="split 70B-class models (name contains one of $AIOS_TP_MODELS, default \"70b\") this many ways "
                         "over the first GPUs instead of replicating them (default $AIOS_TP, 0 = off)")
    ap.add_argument("--model-dir", default=None)
    ap.add_argument("--openai-http", action="store_true",
                    help="serve every loaded model on its port (8080, 8081, ...) with the OpenAI-compatible HTTP API")
    ap.add_argument("--idle-unload-minutes", type=float, default=None,
                    help="unload a model that has not been used for this long (config/default-config.toml: "
                         "idle_unload_minutes; default $AIOS_IDLE_UNLOAD_MINUTES, 0 = never); models whose name "
                         'contains "tinyllama" (the operational model) stay')
    ap.add_argument("--no-auto-restart", action="store_true", help="flag a dead model as errored and leave it (the "
                                                                   "reference's behaviour) instead of reloading it")
    args = ap.parse_args(argv)
    logging.basicConfig(level=logging.INFO, format="%(asctime)s %(levelname)s %(message)s")
    from .inference import EngineBackend
    log.info("aiOS AI Runtime starting...")
    if args.gpus:
        os.environ["AIOS_GPUS"] = args.gpus
    srv = RuntimeServer(EngineBackend(device=args.device, tp=args.tp), addr=args.addr, model_dir=args.model_dir,
                        openai_http=args.openai_http, idle_unload_minutes=args.idle_unload_minutes,
                        auto_restart=not args.no_auto_restart).start()
    try:
        srv.server.wait_for_termination()
    except KeyboardInterrupt:
        srv.stop()


if __name__ == "__main__":
    main()
