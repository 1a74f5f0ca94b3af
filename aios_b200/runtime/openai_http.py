"""OpenAI-compatible HTTP endpoint of one loaded model — what a llama-server child exposed on ``ModelStatus.port``.

SURVEY.md §8f rank 1.  Consumers in the reference: ``InferenceEngine`` itself (``POST /v1/chat/completions``, request
shape ``ChatCompletionRequest`` runtime/src/inference.rs:20-36, response fields read at :161-171 and, streaming,
:274-343), the api-gateway's ``local`` provider (api-gateway/src/openai.rs:148-165) and ``ModelManager``'s
``GET /health`` probe (runtime/src/model_manager.rs:223, :428).  Shapes served here are the ones those parsers read:

* ``GET /health`` -> 200 ``{"status": "ok"}``
* ``POST /v1/chat/completions`` with ``{"messages": [{"role", "content"}...], "max_tokens", "temperature", "stream"}``
  -> ``{"choices": [{"index": 0, "message": {"role": "assistant", "content": ...}, "finish_reason": "stop"|"length"}],
  "usage": {"prompt_tokens", "completion_tokens", "total_tokens"}, ...}``; with ``"stream": true`` server-sent events
  ``data: {"choices": [{"index": 0, "delta": {"content": ...}, "finish_reason": null}]}`` ..., a last event whose
  ``finish_reason`` is set, then ``data: [DONE]``.

The first ``system`` message and the last ``user`` message are rendered through the model's chat format (one turn,
like the reference sends, inference.rs:363-376).  ``response_format: {"type": "json_object"}`` switches the non-streaming call to
grammar-constrained sampling (``json_grammar.py``).  Binds 127.0.0.1 only, like the reference's ``--host 127.0.0.1`` (model_manager.rs:197-198).
"""
from __future__ import annotations

import json
import logging
import threading
import time
import uuid
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

from .inference import InferenceEngine, InferParams

log = logging.getLogger("aios_b200.runtime")


def _messages_to_params(body: dict) -> InferParams:
    msgs = body.get("messages")
    if not isinstance(msgs, list) or not msgs:
        raise ValueError("'messages' must be a non-empty list")
    system = next((m.get("content", "") for m in msgs if m.get("role") == "system"), "")
    users = [m.get("content", "") for m in msgs if m.get("role") == "user"]
    if not users:
        raise ValueError("no message with role 'user'")
    rf = body.get("response_format")
    # the conversation as sent goes through the model's chat template; (system, last user turn) is the fallback
    convo = [(str(m.get("role", "")), str(m.get("content", "") or "")) for m in msgs if isinstance(m, dict)]
    return InferParams(prompt=str(users[-1]), system_prompt=str(system or ""), max_tokens=int(body.get("max_tokens") or 0),
                       temperature=float(body.get("temperature") or 0.0), seed=body.get("seed"),
                       json_mode=isinstance(rf, dict) and rf.get("type") == "json_object",
                       messages=convo if len(convo) == len(msgs) else None)


class OpenAIServer:
    """``OpenAIServer(model_handle, name, port).start()``; ``port=0`` picks a free one (``.port`` tells which)."""

    def __init__(self, handle, model_name: str, port: int, engine: InferenceEngine | None = None):
        self.handle, self.model_name = handle, model_name
        self.engine = engine or InferenceEngine()
        outer = self

        class Handler(BaseHTTPRequestHandler):
            protocol_version = "HTTP/1.1"

            def log_message(self, fmt, *args):  # quiet: the gRPC layer logs requests
                log.debug("http %s", fmt % args)

            def _json(self, code: int, obj: dict):
                raw = json.dumps(obj).encode("utf-8")
                self.send_response(code)
                self.send_header("Content-Type", "application/json")
                self.send_header("Content-Length", str(len(raw)))
                self.end_headers()
                self.wfile.write(raw)

            def do_GET(self):
                if self.path.split("?")[0] == "/health":
                    self._json(200, {"status": "ok"})
                elif self.path.split("?")[0] == "/v1/models":
                    self._json(200, {"object": "list", "data": [{"id": outer.model_name, "object": "model"}]})
                else:
                    self._json(404, {"error": {"message": f"no route {self.path}", "type": "not_found"}})

            def do_POST(self):
                if self.path.split("?")[0] != "/v1/chat/completions":
                    self._json(404, {"error": {"message": f"no route {self.path}", "type": "not_found"}})
                    return
                try:
                    n = int(self.headers.get("Content-Length") or 0)
                    body = json.loads(self.rfile.read(n) or b"{}")
                    params = _messages_to_params(body)
                except Exception as e:
                    self._json(400, {"error": {"message": str(e), "type": "invalid_request_error"}})
                    return
                if body.get("stream"):
                    outer._stream(self, params)
                else:
                    outer._complete(self, params)

        self.httpd = ThreadingHTTPServer(("127.0.0.1", port), Handler)
        self.httpd.daemon_threads = True
        self.port = self.httpd.server_address[1]
        self.thread = threading.Thread(target=self.httpd.serve_forever, kwargs={"poll_interval": 0.1}, daemon=True)

    def start(self) -> "OpenAIServer":
        self.thread.start()
        return self

    def close(self):
        self.httpd.shutdown()
        self.httpd.server_close()

    # ------------------------------------------------------------------ non-streaming
    def _complete(self, h, params: InferParams):
        try:
            pieces, n_prompt, n_comp = [], 0, 0
            limit = params.max_tokens if params.max_tokens > 0 else 512
            if params.json_mode and hasattr(self.handle.seq, "logits"):  # grammar-constrained (json_grammar.py)
                toks, n_prompt = self.engine._generate_json(self.handle, params)
                pieces.append(toks)
                n_comp = int(toks.size)
            else:
                for toks, n_prompt, finished in self.engine._generate(self.handle, params, chunk=64):
                    pieces.append(toks)
                    n_comp += int(toks.size)
                    if finished:
                        break
            import numpy as np
            comp = np.concatenate(pieces) if pieces else np.empty(0, np.int32)
            text = self.handle.vocab.detokenize(comp).decode("utf-8", "replace")
        except Exception as e:
            h._json(400 if "context" in str(e) else 500, {"error": {"message": str(e), "type": "server_error"}})
            return
        h._json(200, {
            "id": "chatcmpl-" + uuid.uuid4().hex[:24], "object": "chat.completion", "created": int(time.time()),
            "model": self.model_name,
            "choices": [{"index": 0, "message": {"role": "assistant", "content": text},
                         "finish_reason": "length" if n_comp >= limit else "stop"}],
            "usage": {"prompt_tokens": n_prompt, "completion_tokens": n_comp, "total_tokens": n_prompt + n_comp}})

    # ------------------------------------------------------------------ server-sent events
    def _stream(self, h, params: InferParams):
        cid, created = "chatcmpl-" + uuid.uuid4().hex[:24], int(time.time())

        def event(delta: dict, finish):
            obj = {"id": cid, "object": "chat.completion.chunk", "created": created, "model": self.model_name,
                   "choices": [{"index": 0, "delta": delta, "finish_reason": finish}]}
            raw = b"data: " + json.dumps(obj).encode("utf-8") + b"\n\n"
            h.wfile.write(hex(len(raw))[2:].encode() + b"\r\n" + raw + b"\r\n")
            h.wfile.flush()

        it = self.engine.stream_infer(self.handle, self.model_name, params)
        try:
            first = next(it)  # prompt errors before any header is sent
        except StopIteration:
            first = ("", True)
        except Exception as e:
            h._json(400 if "context" in str(e) else 500, {"error": {"message": str(e), "type": "server_error"}})
            return
        h.send_response(200)
        h.send_header("Content-Type", "text/event-stream")
        h.send_header("Cache-Control", "no-cache")
        h.send_header("Transfer-Encoding", "chunked")
        h.end_headers()
        try:
            event({"role": "assistant"}, None)
            text, done = first
            while True:
                if text:
                    event({"content": text}, None)
                if done:
                    break
                text, done = next(it)
            event({}, "stop")
            tail = b"data: [DONE]\n\n"
            h.wfile.write(hex(len(tail))[2:].encode() + b"\r\n" + tail + b"\r\n0\r\n\r\n")
            h.wfile.flush()
        except (BrokenPipeError, ConnectionResetError):
            log.warning("stream receiver dropped model=%s", self.model_name)
        finally:
            it.close()
