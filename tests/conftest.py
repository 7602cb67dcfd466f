"""pytest configuration: the `gpu` marker and shared fixtures.

`-m "not gpu"` runs in the CPU-only container (oracle, golden vectors, host logic, C-ABI export
check, gloo world_size-2).  `-m gpu` runs on a B200 and is the parity suite proper: every test
there calls the CUDA path through the C ABI and compares with the oracle.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    import oracle
    return oracle.Oracle()


@pytest.fixture(scope="session")
def golden_rows():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "known_answers.json")) as f:
        return json.load(f)["rows"]


@pytest.fixture(scope="session")
def golden_random():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "random_vectors.npz"))
