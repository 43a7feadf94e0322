import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # build the CPU checker (oracle + refprobe when /root/reference is present); cheap when up to date
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=False, capture_output=True)


@pytest.fixture(scope="session")
def orc():
    from cpu_checker import Checker
    return Checker("orc")


@pytest.fixture(scope="session")
def ref():
    from cpu_checker import Checker, have_ref
    if not have_ref():
        pytest.skip("oracle/_ref/libavr_refprobe.so not built (needs /root/reference)")
    return Checker("ref")


@pytest.fixture(scope="session")
def ctx():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from acceleratedvolrenderer_b200.api import Context
    c = Context(0)
    yield c
    c.close()
