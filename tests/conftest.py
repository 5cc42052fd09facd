import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def b200():
    from b200_import import load_package
    return load_package()


@pytest.fixture(scope="session")
def oracle_mod():
    from b200_import import load_oracle
    return load_oracle()


@pytest.fixture(scope="session")
def oracle(oracle_mod):
    return oracle_mod.Oracle("own")


@pytest.fixture(scope="session")
def ref_oracle(oracle_mod):
    if not oracle_mod.have_reference() and not os.path.exists(os.path.join(ROOT, "oracle/_ref/libref_oracle.so")):
        pytest.skip("reference tree / oracle/_ref not available on this machine")
    return oracle_mod.Oracle("ref")


@pytest.fixture(scope="session")
def ref_libs(oracle_mod):
    if not oracle_mod.have_reference() and not os.path.exists(os.path.join(ROOT, "oracle/_ref/libref_mc.so")):
        pytest.skip("reference tree / oracle/_ref not available on this machine")
    return oracle_mod.RefLibs()


def has_gpu():
    try:
        from b200_import import load_package
        return load_package().lib().gnss_b200_device_count() > 0
    except Exception:
        return False
