"""Mock-model forwards.

``FakeModel`` is the reference's model (src/mock_models/fake_model.py:12-83): ``predict`` echoes its
input inside a metadata envelope.  For CUDA tensors / ``PackedBatch`` the echo is the K8 kernel
(``die_echo_fwd``); any other Python object is returned as the same object, as in the reference.
``LinearModel`` and ``MLPModel`` are the dense mock models of BASELINE.json configs[1] and [4]; the
reference has no arithmetic for them (oracle/models.py defines it).
"""
from __future__ import annotations

import asyncio
import time
from typing import Any, Dict, Optional

from ..config import ModelConfig


def _is_cuda_tensor(x) -> bool:
    try:
        import torch
        return isinstance(x, torch.Tensor) and x.is_cuda
    except ImportError:  # pragma: no cover
        return False


class FakeModel:
    def __init__(self, config: ModelConfig, *, simulate_latency: bool = True):
        self.config = config
        self.model_name = config.model_name
        self.batch_size = config.batch_size
        self.max_batch_size = config.max_batch_size
        self.input_schema = config.input_schema or {}
        self.output_schema = config.output_schema or {}
        self.simulate_latency = simulate_latency  # the reference always sleeps 50-150 ms (fake_model.py:47)
        self.request_count = 0
        self.error_count = 0
        self.total_latency = 0.0
        self.last_inference_time = 0.0

    def forward(self, inputs: Any) -> Any:
        """The model proper: identity.  Device data goes through the echo kernel."""
        from ..batcher import PackedBatch
        from .. import ops
        if isinstance(inputs, PackedBatch):
            return PackedBatch(ops.echo_fwd(inputs.data), inputs.lengths, inputs.host_lengths)
        if _is_cuda_tensor(inputs):
            return ops.echo_fwd(inputs)
        return inputs

    async def predict(self, inputs: Any) -> Dict[str, Any]:
        start = time.time()
        self.request_count += 1
        try:
            if self.simulate_latency:
                await asyncio.sleep(0.05 + (time.time() % 0.1))
            return {"model": self.model_name, "output": self.forward(inputs),
                    "metadata": {"batch_size": len(inputs) if isinstance(inputs, (list, tuple)) else 1,
                                 "timestamp": time.time(), "request_id": f"req_{int(time.time() * 1000)}"}}
        except Exception:
            self.error_count += 1
            raise
        finally:
            self.total_latency += time.time() - start
            self.last_inference_time = time.time()

    def get_metrics(self) -> Dict[str, Any]:
        return {"model_name": self.model_name, "request_count": self.request_count, "error_count": self.error_count,
                "avg_latency": (self.total_latency / self.request_count) if self.request_count > 0 else 0,
                "last_inference_time": self.last_inference_time, "batch_size": self.batch_size,
                "max_batch_size": self.max_batch_size}


class LinearModel(FakeModel):
    """Y = X @ W.T + b on fp32 rows (configs[1]); ``precision`` picks the CUDA-core or tcgen05 kernel."""

    def __init__(self, config: ModelConfig, weight, bias, precision: int = 0, **kw):
        super().__init__(config, **kw)
        self.weight, self.bias, self.precision = weight.contiguous(), bias.contiguous(), precision

    def forward(self, inputs: Any) -> Any:
        from ..batcher import PackedBatch
        from .. import ops
        if isinstance(inputs, PackedBatch):
            y = ops.linear_fwd(inputs.data, self.weight, self.bias, self.precision)
            n = y.shape[1]
            return PackedBatch(y, inputs.lengths.new_full(inputs.lengths.shape, n), [n] * len(inputs.host_lengths))
        return ops.linear_fwd(inputs, self.weight, self.bias, self.precision)


class MLPModel(FakeModel):
    """Y = relu(X W1^T + b1) W2^T + b2, bf16 storage, fp32 accumulation (configs[4])."""

    def __init__(self, config: ModelConfig, w1, b1, w2, b2, **kw):
        super().__init__(config, **kw)
        self.w1, self.b1, self.w2, self.b2 = w1.contiguous(), b1.contiguous(), w2.contiguous(), b2.contiguous()

    def forward(self, inputs: Any) -> Any:
        from .. import ops
        return ops.mlp_fwd_bf16(inputs, self.w1, self.b1, self.w2, self.b2)
