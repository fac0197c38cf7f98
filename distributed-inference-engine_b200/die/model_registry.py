"""Model / version / shard metadata with GPU hash-sharding.

Drop-in for the reference's ModelRegistry (src/model_registry.py).  The per-request work --
``shards[int(md5(key).hexdigest(), 16) % len(shards)]`` (src/model_registry.py:159-161) -- runs in
the K1 CUDA kernel (csrc/route.cu) through ``die_route_md5``; everything else here is control-plane
bookkeeping that stays on the host.
"""
from __future__ import annotations

import hashlib
import json
from dataclasses import dataclass, field
from enum import Enum, auto
from typing import Any, Dict, List, Optional, Sequence, Set, Tuple


class ModelStatus(Enum):
    LOADING = auto()
    READY = auto()
    UPDATING = auto()
    FAILED = auto()
    UNLOADING = auto()


@dataclass
class ModelShard:
    shard_id: int
    worker_id: str
    status: ModelStatus = ModelStatus.READY
    load: float = 0.0
    metadata: Dict[str, Any] = field(default_factory=dict)

    def to_dict(self) -> Dict[str, Any]:
        return {"shard_id": self.shard_id, "worker_id": self.worker_id, "status": self.status.name,
                "load": self.load, "metadata": self.metadata}


@dataclass
class ModelVersion:
    version: str
    model_path: str
    input_schema: Dict[str, Any]
    output_schema: Dict[str, Any]
    batch_size: int = 1
    max_batch_size: int = 32
    quantized: bool = False
    shards: List[ModelShard] = field(default_factory=list)
    metadata: Dict[str, Any] = field(default_factory=dict)

    def to_dict(self) -> Dict[str, Any]:
        return {"version": self.version, "model_path": self.model_path, "batch_size": self.batch_size,
                "max_batch_size": self.max_batch_size, "quantized": self.quantized,
                "shards": [s.to_dict() for s in self.shards], "input_schema": self.input_schema,
                "output_schema": self.output_schema, "metadata": self.metadata}


class ModelRegistry:
    def __init__(self) -> None:
        self._models: Dict[str, Dict[str, ModelVersion]] = {}
        self._worker_models: Dict[str, Set[Tuple[str, str]]] = {}
        self._model_hashes: Dict[str, str] = {}

    # -- registration (src/model_registry.py:86-147) ------------------------------------------
    def register_model(self, model_name: str, version: str, model_path: str, input_schema: Dict[str, Any],
                       output_schema: Dict[str, Any], batch_size: int = 1, max_batch_size: int = 32,
                       quantized: bool = False, metadata: Optional[Dict[str, Any]] = None) -> None:
        mv = ModelVersion(version=version, model_path=model_path, input_schema=input_schema,
                          output_schema=output_schema, batch_size=batch_size, max_batch_size=max_batch_size,
                          quantized=quantized, metadata=metadata or {})
        self._models.setdefault(model_name, {})[version] = mv
        self._update_model_hash(model_name, version)

    def add_shard(self, model_name: str, version: str, shard_id: int, worker_id: str,
                  metadata: Optional[Dict[str, Any]] = None) -> ModelShard:
        mv = self.get_model_version(model_name, version)
        if mv is None:
            raise ValueError(f"Model {model_name} version {version} not found")
        for s in mv.shards:
            if s.shard_id == shard_id:
                return s
        shard = ModelShard(shard_id=shard_id, worker_id=worker_id, metadata=metadata or {})
        mv.shards.append(shard)
        self._worker_models.setdefault(worker_id, set()).add((model_name, version))
        return shard

    # -- the hot op (src/model_registry.py:149-161) -------------------------------------------
    def get_shard_for_key(self, model_name: str, version: str, key: str) -> Optional[ModelShard]:
        out = self.get_shards_for_keys(model_name, version, [key])
        return out[0] if out else None

    def get_shards_for_keys(self, model_name: str, version: str, keys: Sequence[str]) -> List[Optional[ModelShard]]:
        """Batched get_shard_for_key: one kernel launch for all keys."""
        mv = self.get_model_version(model_name, version)
        if mv is None or not mv.shards:
            return [None] * len(keys)
        if not keys:
            return []
        from . import ops
        shard_pos, _, _ = ops.route_md5(ops.KeyBatch.from_strings(keys), len(mv.shards), want_digests=False)
        return [mv.shards[p] for p in shard_pos.tolist()]

    # -- lookups -------------------------------------------------------------------------------
    def get_model_version(self, model_name: str, version: str) -> Optional[ModelVersion]:
        return self._models.get(model_name, {}).get(version)

    def list_models(self) -> List[str]:
        return list(self._models)

    def list_versions(self, model_name: str) -> List[str]:
        return list(self._models.get(model_name, {}))

    def get_worker_models(self, worker_id: str) -> List[Tuple[str, str]]:
        return list(self._worker_models.get(worker_id, set()))

    def _update_model_hash(self, model_name: str, version: str) -> None:
        d = self._models[model_name][version].to_dict()
        d["shards"] = []  # shard placement does not change the model's identity
        blob = json.dumps(d, sort_keys=True).encode()
        self._model_hashes[f"{model_name}:{version}"] = hashlib.md5(blob).hexdigest()

    def get_model_hash(self, model_name: str, version: str) -> Optional[str]:
        return self._model_hashes.get(f"{model_name}:{version}")

    # -- (de)serialisation (src/model_registry.py:192-249) --------------------------------------
    def to_dict(self) -> Dict[str, Any]:
        return {"models": {m: {v: mv.to_dict() for v, mv in vs.items()} for m, vs in self._models.items()}}

    @classmethod
    def from_dict(cls, data: Dict[str, Any]) -> "ModelRegistry":
        reg = cls()
        for model_name, versions in data.get("models", {}).items():
            for version, md in versions.items():
                shards = [ModelShard(shard_id=s["shard_id"], worker_id=s["worker_id"],
                                     status=ModelStatus[s["status"]], load=s["load"],
                                     metadata=s.get("metadata", {})) for s in md.get("shards", [])]
                reg._models.setdefault(model_name, {})[version] = ModelVersion(
                    version=md["version"], model_path=md["model_path"], input_schema=md["input_schema"],
                    output_schema=md["output_schema"], batch_size=md.get("batch_size", 1),
                    max_batch_size=md.get("max_batch_size", 32), quantized=md.get("quantized", False),
                    shards=shards, metadata=md.get("metadata", {}))
                for s in shards:
                    reg._worker_models.setdefault(s.worker_id, set()).add((model_name, version))
        return reg
