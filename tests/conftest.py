import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import __graft_entry__ as ge  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    return ge.load_package()


@pytest.fixture(scope="session")
def fec(pkg):
    return pkg.fec


@pytest.fixture(scope="session")
def orc():
    return ge.load_oracle()


@pytest.fixture(scope="session")
def oracle(orc):
    return orc.Oracle()


@pytest.fixture(scope="session")
def ref(orc):
    """The reference itself (oracle/_ref/libref_fec.so): built in the container that has /root/reference."""
    if not orc.Ref.available():
        if not orc.build_ref():
            pytest.skip("oracle/_ref/libref_fec.so not built (no /root/reference here)")
    return orc.Ref()


@pytest.fixture(scope="session")
def engine(fec):
    eng = fec.FecEngine(device=0, harq_slots=1024, max_batch_cbs=1024)
    yield eng
    eng.close()


@pytest.fixture(scope="session")
def engine_generic(fec):
    eng = fec.FecEngine(device=0, harq_slots=1024, max_batch_cbs=1024, kernel_variant=1)
    yield eng
    eng.close()
