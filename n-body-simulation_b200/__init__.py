"""n-body-simulation_b200 -- B200-native engine behind the reference's exhaustive-path interface.

The directory name is not a Python identifier; load it with
``importlib`` (see tests/conftest.py: ``load_package()``) or run the C host
programs under ``bin/``.  Contents: ``csrc/`` (CUDA kernels + C ABI),
``host/`` (C host programs and I/O), ``nbody.py`` (ctypes host mirror).
"""
from .nbody import *  # noqa: F401,F403
from . import nbody  # noqa: F401
