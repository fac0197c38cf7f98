"""Model-load configuration type of the drop-in package.

`ModelConfig` carries what `Worker.load_model` is told about a model: the same six fields, defaults and order as the
reference's dataclass (src/config.py:12-20), so that `ModelConfig(name, path)` and keyword construction behave
identically.  The field table below is the single place they are spelled out.
"""
from dataclasses import field, make_dataclass
from typing import Any, Dict, Optional

_Schema = Optional[Dict[str, Any]]

# (name, type, default) -- positional order matters to callers
_FIELDS = (
    ("model_name", str, None),
    ("model_path", str, None),
    ("batch_size", int, 1),
    ("max_batch_size", int, 32),
    ("input_schema", _Schema, None),
    ("output_schema", _Schema, None),
)
_REQUIRED = 2  # model_name and model_path have no default

ModelConfig = make_dataclass(
    "ModelConfig",
    [(n, t) if i < _REQUIRED else (n, t, field(default=d)) for i, (n, t, d) in enumerate(_FIELDS)],
)
ModelConfig.__module__ = __name__
ModelConfig.__doc__ = "Configuration for a model (model_name, model_path, batch_size=1, max_batch_size=32, input_schema=None, output_schema=None)."
