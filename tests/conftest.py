import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import rt_oracle
    rt_oracle.build()
    rt_oracle.lib()
    return rt_oracle


@pytest.fixture(scope="session")
def rtlib():
    import ray_tracer_b200 as rt
    if not os.path.exists(rt._lib.LIB_PATH):
        rt.build()
    return rt
