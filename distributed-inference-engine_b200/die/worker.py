"""Worker: holds loaded models and answers inference requests; drop-in for the request-processing part
of the reference's Worker (src/worker.py:135-184).  The asyncio TCP/JSON server, signal handling and
psutil metrics of the reference (src/worker.py:41-133,186-253) are transport / ops plumbing outside the
hot path: in the B200 build requests reach the owner GPU through the NCCL all-to-all (die/exchange.py)."""
from __future__ import annotations

import logging
import time
from typing import Any, Dict

from .config import ModelConfig
from .mock_models import FakeModel

logger = logging.getLogger(__name__)


class Worker:
    def __init__(self, worker_id: str, host: str = "0.0.0.0", port: int = 0):
        self.worker_id = worker_id
        self.host = host
        self.port = port
        self.models: Dict[str, FakeModel] = {}
        self._start_time = time.time()
        self._request_count = 0
        self._error_count = 0

    def load_model(self, config: ModelConfig, model: FakeModel = None) -> bool:
        """``model`` (additive): a pre-built FakeModel subclass, e.g. LinearModel, for this config."""
        if config.model_name in self.models:
            return True
        try:
            self.models[config.model_name] = model if model is not None else FakeModel(config)
            return True
        except Exception as e:  # noqa: BLE001
            logger.error("Error loading model %r: %s", config.model_name, e)
            return False

    def unload_model(self, model_name: str) -> bool:
        return self.models.pop(model_name, None) is not None

    async def _process_request(self, request: Dict[str, Any]) -> Dict[str, Any]:
        if not isinstance(request, dict):
            return {"error": "Request must be a JSON object", "success": False}
        model_name, inputs = request.get("model"), request.get("inputs")
        if not model_name or inputs is None:
            return {"error": "Missing required fields: model and inputs are required", "success": False}
        if model_name not in self.models:
            return {"error": f"Model '{model_name}' not found", "success": False}
        try:
            outputs = await self.models[model_name].predict(inputs)
            return {"model": model_name, "outputs": outputs, "worker_id": self.worker_id, "success": True}
        except Exception as e:  # noqa: BLE001 - the reference never raises from here
            return {"error": str(e), "success": False}

    def infer(self, model_name: str, packed_batch):
        """Batched forward on a device batch (north-star alias): no envelope, no latency simulation."""
        return self.models[model_name].forward(packed_batch)

    async def start(self) -> int:
        raise NotImplementedError("the TCP/JSON transport of src/worker.py:41-133 is out of scope for the B200 hot path")

    def get_metrics(self) -> Dict[str, Any]:
        return {"worker_id": self.worker_id, "start_time": self._start_time, "uptime": time.time() - self._start_time,
                "request_count": self._request_count, "error_count": self._error_count,
                "loaded_models": list(self.models), "model_metrics": {n: m.get_metrics() for n, m in self.models.items()}}
