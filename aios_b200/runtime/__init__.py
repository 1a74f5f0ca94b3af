"""Host-side mirror of the reference's ``runtime`` crate (runtime/src/{main,grpc_service,model_manager,inference}.rs)
serving ``aios.runtime.AIRuntime`` on top of the B200 engine's C ABI.  See INTEGRATION.md."""
from .inference import EngineBackend, InferenceEngine, InferParams, InferResult, build_messages  # noqa: F401
from .model_manager import LoadRequest, ModelManager, ModelState, now_epoch_ms  # noqa: F401
from .service import AIRuntimeService, RuntimeServer, autoload_plan  # noqa: F401
