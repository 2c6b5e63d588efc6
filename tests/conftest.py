import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    return oracle_lib.oracle()


@pytest.fixture(scope="session")
def acero():
    """oracle/librpt_oracle_acero.so — checker of the README-described filter variant (parity unpinned)."""
    import oracle_lib
    return oracle_lib.acero()


@pytest.fixture(scope="session")
def ref():
    """oracle/_ref/librpt_ref.so — the reference's own sources, when the prebuilt file is present."""
    import oracle_lib
    r = oracle_lib.ref()
    if r is None:
        pytest.skip("oracle/_ref/librpt_ref.so not built (needs /root/reference)")
    return r


@pytest.fixture(scope="session")
def rpt():
    import rpt_b200
    rpt_b200.lib()
    return rpt_b200
