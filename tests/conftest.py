"""pytest configuration: markers and import paths.

``-m "not gpu"`` runs on the CPU build container (oracle vs goldens, host logic, C-ABI
symbol export).  ``-m gpu`` tests are the parity tests proper: they call the CUDA path
through the C-ABI on a real B200 and compare against ``oracle/`` (test infrastructure).
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "nano-vllm-hw3_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a)")


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_golden.npz"))
