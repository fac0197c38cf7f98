"""B200-native batched request dispatch (route -> probe -> pack -> infer -> unpack).

Drop-in mirror of the reference's hot-path classes; the compute runs in libdie.so (sm_100a CUDA),
bound through ctypes (die/_cabi.py, include/die.h).  Importing this package needs neither torch nor
a GPU; calling a compute method without the built library or without a CUDA device raises.
"""
from .batcher import Batch, BatchedRequest, Batcher, PackedBatch
from .config import ModelConfig
from .model_registry import ModelRegistry, ModelShard, ModelStatus, ModelVersion
from .router import Router, WorkerHealth, WorkerInfo
from .worker import Worker
from .mock_models import FakeModel, LinearModel, MLPModel, mock_batch_inference

__all__ = ["Batch", "BatchedRequest", "Batcher", "PackedBatch", "ModelConfig", "ModelRegistry", "ModelShard",
           "ModelStatus", "ModelVersion", "Router", "WorkerHealth", "WorkerInfo", "Worker", "FakeModel",
           "LinearModel", "MLPModel", "mock_batch_inference", "KVCache", "KVStore", "create_kv_store", "CacheEntry",
           "DispatchPipeline", "Coordinator"]


def __getattr__(name):  # torch-dependent modules load lazily
    if name in ("KVCache", "KVStore", "create_kv_store", "CacheEntry", "DeviceKV"):
        from . import kvstore
        return getattr(kvstore, name)
    if name == "DispatchPipeline":
        from .pipeline import DispatchPipeline
        return DispatchPipeline
    if name == "Coordinator":
        from .coordinator import Coordinator
        return Coordinator
    raise AttributeError(name)
