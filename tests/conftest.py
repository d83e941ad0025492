"""pytest configuration: markers, import path, shared helpers."""
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: takes more than a few seconds")


def pkg(sub: str = ""):
    """Import ``copilot-for-consensus_b200[.sub]`` (the directory name is not an identifier)."""
    return importlib.import_module("copilot-for-consensus_b200" + ("." + sub if sub else ""))


def bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (b.astype(np.uint32) << 16).view(np.float32)


@pytest.fixture(scope="session")
def search_golden():
    return np.load(os.path.join(GOLDEN, "search_golden.npz"))


@pytest.fixture(scope="session")
def encode_golden():
    return np.load(os.path.join(GOLDEN, "encode_golden.npz"))
