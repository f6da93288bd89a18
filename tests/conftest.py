import importlib.util
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def load_package():
    """Import n-body-simulation_b200/ (hyphenated directory) as module `nbody_b200`."""
    if "nbody_b200" in sys.modules:
        return sys.modules["nbody_b200"]
    path = os.path.join(ROOT, "n-body-simulation_b200", "__init__.py")
    spec = importlib.util.spec_from_file_location(
        "nbody_b200", path, submodule_search_locations=[os.path.dirname(path)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["nbody_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def rel_err(a, b):
    """per-body relative error of 3-vectors: |a-b| / |b| (b = oracle)"""
    a, b = np.asarray(a)[:, :3], np.asarray(b)[:, :3]
    den = np.linalg.norm(b, axis=1)
    den = np.where(den > 0, den, 1.0)
    return np.linalg.norm(a - b, axis=1) / den


@pytest.fixture(scope="session")
def nb():
    return load_package()


@pytest.fixture(scope="session")
def oracle_lib():
    import oracle
    return oracle.load_oracle()
