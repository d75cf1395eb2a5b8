"""pytest configuration: the `gpu` marker and shared fixtures.

CPU suite (`-m "not gpu"`): oracle vs golden vectors, host logic, C-ABI symbol export.
GPU suite (`-m gpu`): parity of the CUDA path (through the C ABI) against the oracle.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU checker: the reference-built library when present, else the plain-C port."""
    from oracle import pyoracle
    if not os.path.exists(pyoracle.PORT_LIB):
        pyoracle.build("port")
    return pyoracle.load_best()


@pytest.fixture(scope="session")
def port_oracle():
    from oracle import pyoracle
    if not os.path.exists(pyoracle.PORT_LIB):
        pyoracle.build("port")
    return pyoracle.load_port()


@pytest.fixture(scope="session")
def ref_oracle():
    from oracle import pyoracle
    o = pyoracle.load_ref()
    if o is None:
        pytest.skip("oracle/_ref/libref_oracle.so not built (needs /root/reference)")
    return o


@pytest.fixture(scope="session")
def engine():
    """One rte_ctx on cuda:0; raises (does not skip) when the CUDA library is unusable on a GPU box."""
    from raytracingengine_b200.engine import Engine
    e = Engine(0)
    yield e
    e.close()
