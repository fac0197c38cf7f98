"""B200-native batched request dispatch (route -> probe -> pack -> infer -> unpack).

Drop-in mirror of the reference's hot-path classes; the compute runs in libdie.so (sm_100a CUDA),
bound through ctypes (die/_cabi.py, include/die.h).  Importing this package needs neither torch nor
a GPU; calling a compute method without the built library or without a CUDA device raises.
"""
import os as _os

# Spin-waiting kernels (grid barrier of the persistent step, cross-GPU barrier) must not be resident while the driver
# loads a module lazily at some kernel's first launch; the sharded step wants its streams on separate hardware queues.
# Both are read when CUDA initialises, so they are set at import unless the process already chose otherwise.
_os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

from .batcher import Batch, BatchedRequest, Batcher, PackedBatch
from .config import ModelConfig
from .model_registry import ModelRegistry, ModelShard, ModelStatus, ModelVersion
from .router import Router, WorkerHealth, WorkerInfo
from .worker import Worker
from .mock_models import FakeModel, LinearModel, MLPModel, mock_batch_inference

__all__ = ["Batch", "BatchedRequest", "Batcher", "PackedBatch", "ModelConfig", "ModelRegistry", "ModelShard",
           "ModelStatus", "ModelVersion", "Router", "WorkerHealth", "WorkerInfo", "Worker", "FakeModel",
           "LinearModel", "MLPModel", "mock_batch_inference", "KVCache", "KVStore", "create_kv_store", "CacheEntry",
           "DispatchPipeline", "Coordinator", "DeviceBatcher"]


def __getattr__(name):  # torch-dependent modules load lazily
    if name in ("KVCache", "KVStore", "create_kv_store", "CacheEntry", "DeviceKV"):
        from . import kvstore
        return getattr(kvstore, name)
    if name == "DispatchPipeline":
        from .pipeline import DispatchPipeline
        return DispatchPipeline
    if name == "DeviceBatcher":
        from .device_batcher import DeviceBatcher
        return DeviceBatcher
    if name == "Coordinator":
        from .coordinator import Coordinator
        return Coordinator
    raise AttributeError(name)
