"""Model-load configuration type, same fields as the reference's ModelConfig (src/config.py:12-20)."""
from dataclasses import dataclass
from typing import Any, Dict, Optional


@dataclass
class ModelConfig:
    model_name: str
    model_path: str
    batch_size: int = 1
    max_batch_size: int = 32
    input_schema: Optional[Dict[str, Any]] = None
    output_schema: Optional[Dict[str, Any]] = None
