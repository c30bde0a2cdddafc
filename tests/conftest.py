import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """the plain-C restatement (oracle/liboracle.so); built on demand"""
    import subprocess
    import pyoracle
    if not pyoracle.Oracle.available("oracle"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), os.path.join(ROOT, "oracle", "liboracle.so")])
    return pyoracle.Oracle("oracle")


@pytest.fixture(scope="session")
def ref():
    """the reference's own headers compiled into oracle/_ref (skips when not built)"""
    import pyoracle
    if not pyoracle.Oracle.available("ref"):
        pytest.skip("oracle/_ref/libnvbio_ref.so not built (needs /root/reference)")
    return pyoracle.Oracle("ref")


def pytest_collection_modifyitems(config, items):
    """GPU tests need a CUDA device; without one they are reported as skipped (run them with -m gpu on a B200)"""
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
