"""Batch mock of the reference (src/mock_models/mock_inference.py:12-73): simulated latency, one result
dict per input in input order, deterministic fault injection for the first ceil(error_rate*B) items.
Pure host-side string formatting -- there is no arithmetic here to move to the GPU."""
from __future__ import annotations

import asyncio
import logging
import time
from typing import Any, Dict, List

logger = logging.getLogger(__name__)


async def mock_batch_inference(model_name: str, version: str, inputs: List[Any], **kwargs) -> List[Dict[str, Any]]:
    latency_ms = kwargs.get("latency_ms", 100)
    error_rate = kwargs.get("error_rate", 0.0)
    n = len(inputs)
    await asyncio.sleep(latency_ms / 1000.0)
    results: List[Dict[str, Any]] = []
    for i, x in enumerate(inputs):
        base = {"model": model_name, "version": version, "input_id": i}
        if error_rate > 0 and (i / n) < error_rate:
            results.append({"success": False, "error": f"Simulated error processing input {i}", **base})
        else:
            results.append({"success": True, "result": f"Processed by {model_name}:{version} - {x}", **base,
                            "metadata": {"processing_time_ms": latency_ms, "batch_size": n, "batch_position": i,
                                         "timestamp": time.time()}})
    return results
