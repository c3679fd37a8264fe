"""pytest configuration: the `gpu` marker and import paths.

CPU suite (`-m "not gpu"`): oracle vs golden vectors, host logic, C-ABI symbol check, the
product's row arithmetic compiled for the host (tests/host_harness) vs the oracle.
GPU suite (`-m gpu`): parity of the CUDA path (through the C ABI) with the oracle.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "ac-solver-caltech_b200")
for p in (PKG, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session")
def oracle():
    import ac_oracle

    ac_oracle.build()
    return ac_oracle


@pytest.fixture(scope="session")
def cuda():
    import torch

    if not torch.cuda.is_available():
        pytest.fail("a test marked `gpu` was run without a CUDA device")
    import acx  # noqa: F401  (fails loudly when libacx.so is missing)

    return torch.device("cuda:0")
