"""pytest configuration: marker registration, import paths, shared fixtures.

`-m "not gpu"` runs on the CPU container: oracle vs golden vectors / reference, host logic, C-ABI symbols.
`-m gpu` runs on a B200: CUDA path through the C ABI vs the oracle and the golden vectors.
"""
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT, ROOT / "tests"):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))

GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def golden_generated():
    return np.load(GOLDEN / "qp_generated.npz")


@pytest.fixture(scope="session")
def golden_fixed():
    return np.load(GOLDEN / "qp_fixed.npz")


@pytest.fixture(scope="session")
def golden_box():
    return np.load(GOLDEN / "qp_box_synth.npz")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle_py
    oracle_py.oracle_lib()
    return oracle_py
