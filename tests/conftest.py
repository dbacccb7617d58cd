"""pytest configuration: registers the `gpu` marker and puts the package / oracle loaders on sys.path.

Layout reminder: the product lives in velociraptor-stf_b200/ (ctypes binding `vrgpu`, generator `vrsynth`),
the checkers in oracle/ (`vroracle`: our C restatement + the compiled reference).  Only tests may import both.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for sub in ("velociraptor-stf_b200", "oracle"):
    p = os.path.join(ROOT, sub)
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    import vroracle
    return vroracle.Oracle()


@pytest.fixture(scope="session")
def ref():
    import vroracle
    if not vroracle.Ref.available("serial"):
        pytest.skip("oracle/_ref/libvrref.so not built (needs /root/reference at build time)")
    return vroracle.Ref("serial")


@pytest.fixture(scope="session")
def gpu_ctx():
    import vrgpu
    ctx = vrgpu.Context(0)
    yield ctx
    ctx.close()
