import json
import os
import sys

import pytest

# The one-GPU loopback tests of the sharded step run several ranks' streams on one device and their flag barriers
# wait on each other: give every stream its own hardware queue (must be set before CUDA is initialised).
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")  # a kernel's first launch must not wait for a spinning barrier

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_DIR = os.path.join(ROOT, "distributed-inference-engine_b200")
for p in (ROOT, PKG_DIR):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden():
    return load_golden


def pytest_collection_modifyitems(config, items):
    # GPU tests must fail loudly, not skip, on a GPU box; on a CPU-only box they are
    # deselected with -m "not gpu" by the driver.  Nothing to do here.
    return
