import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def product_lib():
    """libraytracing_b200.so built in-tree (make); tests never fall back to anything else."""
    import cuda_raytracing_b200 as rt
    if not os.path.exists(rt.LIB_PATH):
        import subprocess
        subprocess.check_call(["make", "-s", "-C", ROOT, "cuda_raytracing_b200/libraytracing_b200.so"])
    return rt


@pytest.fixture(scope="session")
def oracle():
    import oracleabi
    oracleabi.lib()
    return oracleabi


@pytest.fixture(scope="session")
def gpu_ctx(product_lib):
    ctx = product_lib.Context(0)
    yield ctx
    ctx.close()
