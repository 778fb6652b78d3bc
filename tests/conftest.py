import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib

    oracle_lib.build_oracle()
    return oracle_lib.Oracle()


@pytest.fixture(scope="session")
def ref():
    """The reference's own code (oracle/_ref). Built here when /root/reference exists; prebuilt elsewhere."""
    import oracle_lib

    if os.path.isdir("/root/reference/sources/map"):
        oracle_lib.build_ref()
    if not oracle_lib.Ref.available():
        pytest.skip("oracle/_ref/libmazetd_ref.so not built (needs /root/reference)")
    return oracle_lib.Ref()
