"""Shared fixtures. Markers: `gpu` = needs a B200 (run by the driver with -m gpu on the GPU box).

Libraries:
  prod    terra_b200/libterra_b200.so   the product (CUDA). Loads on a CPU-only box; compute aborts.
  oracle  oracle/libterra_oracle.so     CPU restatement (TEST INFRASTRUCTURE, the checker)
  ref     oracle/_ref/libterra_ref.so   the unmodified reference, compiled from /root/reference when
                                        present; the prebuilt .so travels to the GPU box. Tests that
                                        need it skip when it is absent; the committed fixtures in
                                        tests/golden/ (made from it by tests/golden/make_golden.py)
                                        pin the oracle either way.
Nothing here reads /root/reference at run time.
"""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
GOLDEN = REPO / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


def _harness_api():
    from oracle.binding import harness_api      # terra_b200.api + the checker bindings (oracle/binding.py) as one namespace
    return harness_api()


def _ensure_built():
    api = _harness_api()
    if not api.PRODUCT_LIB.exists():
        subprocess.run(["sh", str(REPO / "build.sh")], check=True)
    if not api.ORACLE_LIB.exists() or (Path("/root/reference/src").is_dir() and not api.REFERENCE_LIB.exists()):
        subprocess.run(["make", "-C", str(REPO / "oracle"), "all"], check=True)


@pytest.fixture(scope="session")
def api():
    _ensure_built()
    return _harness_api()


@pytest.fixture(scope="session")
def scenes(api):
    from terra_b200 import scenes as _scenes
    return _scenes


@pytest.fixture(scope="session")
def prod(api):
    return api.TerraLib()


@pytest.fixture(scope="session")
def oracle(api):
    return api.OracleLib()


@pytest.fixture(scope="session")
def ref(api):
    if not api.REFERENCE_LIB.exists():
        pytest.skip("oracle/_ref/libterra_ref.so not built (no /root/reference here); golden fixtures cover the pin")
    return api.TerraLib(api.REFERENCE_LIB)


@pytest.fixture(scope="session")
def have_gpu(prod):
    return prod.terra_b200_device_count() > 0


@pytest.fixture()
def gpu(prod):
    if prod.terra_b200_device_count() <= 0:
        pytest.fail("this test is marked gpu and needs a CUDA device: the product has no CPU fallback")
    return prod


def golden(name):
    return np.load(GOLDEN / name)
