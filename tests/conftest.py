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


def replay_ticks(case, tick_fn):
    """Drives `tick_fn(now, arrivals, force) -> [(key, ids, trigger)]` with a golden case's ticks and checks every
    flushed batch (key, members in order), when it happened relative to the ticks, and the final stop()."""
    want = case["flushed"]
    got = []
    for t in case["ticks"]:
        out = tick_fn(t["now"], [(k, rid) for k, rid in t["arrivals"]], False)
        n_latency = sum(1 for _k, _ids, trig in out if trig == "latency")
        assert [trig for _k, _ids, trig in out] == ["latency"] * n_latency + ["size"] * (len(out) - n_latency)
        # the reference's timers had all fired when the tick's first request arrived
        assert len(got) + n_latency == t["flushed_before"], t["now"]
        got.extend(out)
    before_stop = len(got)
    got.extend(tick_fn(case["ticks"][-1]["now"], [], True))
    assert all(trig == "forced" for _k, _ids, trig in got[before_stop:])
    assert [(k, ids) for k, ids, _t in got] == [(w["key"], w["ids"]) for w in want]
    return before_stop
