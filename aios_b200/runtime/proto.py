"""Wire messages of ``aios.runtime.AIRuntime`` built at import time, without protoc.

The reference's contract is ``agent-core/proto/runtime.proto:6-62`` plus ``aios.common.{Empty,Status,
HealthStatus}`` (``agent-core/proto/common.proto:4-9``, ``:113-119``).  This image has no ``protoc`` /
``grpc_tools``, so the same descriptors are assembled with ``descriptor_pb2`` (field names, numbers and types
copied from the .proto contract — that *is* the wire format) and the message classes come from the protobuf
runtime.  ``tests/test_cpu_runtime.py`` checks them byte-for-byte against the reference's generated stubs when
``/root/reference`` is present.
"""
from __future__ import annotations

from google.protobuf import descriptor_pb2 as dp
from google.protobuf import descriptor_pool, message_factory

_F = dp.FieldDescriptorProto
_T = {"string": _F.TYPE_STRING, "int32": _F.TYPE_INT32, "int64": _F.TYPE_INT64, "float": _F.TYPE_FLOAT,
      "bool": _F.TYPE_BOOL}


def _msg(fd, name, fields):
    m = fd.message_type.add()
    m.name = name
    for fname, ftype, num, *rest in fields:
        f = m.field.add()
        f.name = fname
        f.number = num
        f.label = _F.LABEL_REPEATED if (rest and rest[0] == "repeated") else _F.LABEL_OPTIONAL
        if ftype in _T:
            f.type = _T[ftype]
        else:
            f.type = _F.TYPE_MESSAGE
            f.type_name = ftype
    return m


def _build():
    pool = descriptor_pool.DescriptorPool()
    common = dp.FileDescriptorProto()
    common.name = "aios_b200/common.proto"
    common.package = "aios.common"
    common.syntax = "proto3"
    _msg(common, "Empty", [])
    _msg(common, "Status", [("success", "bool", 1), ("message", "string", 2)])
    hs = _msg(common, "HealthStatus", [("healthy", "bool", 1), ("service", "string", 2), ("message", "string", 3),
                                       ("uptime_seconds", "int64", 4),
                                       ("details", ".aios.common.HealthStatus.DetailsEntry", 5, "repeated")])
    entry = hs.nested_type.add()
    entry.name = "DetailsEntry"
    entry.options.map_entry = True
    for fname, num in (("key", 1), ("value", 2)):
        f = entry.field.add()
        f.name, f.number, f.label, f.type = fname, num, _F.LABEL_OPTIONAL, _F.TYPE_STRING
    pool.Add(common)

    rt = dp.FileDescriptorProto()
    rt.name = "aios_b200/runtime.proto"
    rt.package = "aios.runtime"
    rt.syntax = "proto3"
    rt.dependency.append(common.name)
    _msg(rt, "LoadModelRequest", [("model_name", "string", 1), ("model_path", "string", 2), ("context_length", "int32", 3),
                                  ("gpu_layers", "int32", 4), ("threads", "int32", 5), ("port", "int32", 6)])
    _msg(rt, "UnloadModelRequest", [("model_name", "string", 1)])
    _msg(rt, "ModelStatus", [("model_name", "string", 1), ("status", "string", 2), ("port", "int32", 3),
                             ("loaded_at", "int64", 4), ("last_used", "int64", 5), ("request_count", "int64", 6)])
    _msg(rt, "ModelList", [("models", ".aios.runtime.ModelStatus", 1, "repeated")])
    _msg(rt, "InferRequest", [("model", "string", 1), ("prompt", "string", 2), ("system_prompt", "string", 3),
                              ("max_tokens", "int32", 4), ("temperature", "float", 5), ("intelligence_level", "string", 6),
                              ("requesting_agent", "string", 7), ("task_id", "string", 8)])
    _msg(rt, "InferResponse", [("text", "string", 1), ("tokens_used", "int32", 2), ("latency_ms", "int64", 3),
                               ("model_used", "string", 4)])
    _msg(rt, "InferChunk", [("text", "string", 1), ("done", "bool", 2)])
    svc = rt.service.add()
    svc.name = "AIRuntime"
    for name, i, o, stream in [("LoadModel", ".aios.runtime.LoadModelRequest", ".aios.runtime.ModelStatus", False),
                               ("UnloadModel", ".aios.runtime.UnloadModelRequest", ".aios.common.Status", False),
                               ("ListModels", ".aios.common.Empty", ".aios.runtime.ModelList", False),
                               ("Infer", ".aios.runtime.InferRequest", ".aios.runtime.InferResponse", False),
                               ("StreamInfer", ".aios.runtime.InferRequest", ".aios.runtime.InferChunk", True),
                               ("HealthCheck", ".aios.common.Empty", ".aios.common.HealthStatus", False)]:
        m = svc.method.add()
        m.name, m.input_type, m.output_type, m.server_streaming = name, i, o, stream
    pool.Add(rt)
    return pool


_POOL = _build()


def _cls(full_name):
    return message_factory.GetMessageClass(_POOL.FindMessageTypeByName(full_name))


Empty = _cls("aios.common.Empty")
Status = _cls("aios.common.Status")
HealthStatus = _cls("aios.common.HealthStatus")
LoadModelRequest = _cls("aios.runtime.LoadModelRequest")
UnloadModelRequest = _cls("aios.runtime.UnloadModelRequest")
ModelStatus = _cls("aios.runtime.ModelStatus")
ModelList = _cls("aios.runtime.ModelList")
InferRequest = _cls("aios.runtime.InferRequest")
InferResponse = _cls("aios.runtime.InferResponse")
InferChunk = _cls("aios.runtime.InferChunk")

SERVICE = "aios.runtime.AIRuntime"
# method -> (request class, response class, server streaming)
METHODS = {
    "LoadModel": (LoadModelRequest, ModelStatus, False),
    "UnloadModel": (UnloadModelRequest, Status, False),
    "ListModels": (Empty, ModelList, False),
    "Infer": (InferRequest, InferResponse, False),
    "StreamInfer": (InferRequest, InferChunk, True),
    "HealthCheck": (Empty, HealthStatus, False),
}


class AIRuntimeStub:
    """Client stub with the same attribute names as the reference's generated ``AIRuntimeStub``
    (agent-core/python/aios_agent/proto/runtime_pb2_grpc.py), for tests and examples."""

    def __init__(self, channel):
        for name, (req, resp, stream) in METHODS.items():
            path = f"/{SERVICE}/{name}"
            mk = channel.unary_stream if stream else channel.unary_unary
            setattr(self, name, mk(path, request_serializer=req.SerializeToString, response_deserializer=resp.FromString))
