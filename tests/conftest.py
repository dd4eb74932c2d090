import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Native pieces are built in-tree by __graft_entry__.build(); make is a no-op when fresh."""
    import subprocess

    if not os.path.exists(os.path.join(ROOT, "oracle", "liborc_oracle.so")):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])
    if not os.path.exists(os.path.join(ROOT, "rpt_b200", "librptx.so")):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "rpt_b200", "csrc")])
