"""Mock models: the reference's echo model and batch mock, plus the dense models BASELINE.json names."""
from .fake_model import FakeModel, LinearModel, MLPModel
from .mock_inference import mock_batch_inference

__all__ = ["FakeModel", "LinearModel", "MLPModel", "mock_batch_inference"]
