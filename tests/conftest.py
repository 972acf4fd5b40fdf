import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def fhe():
    import phantom_fhe_boot_b200 as m
    if not os.path.exists(os.path.join(ROOT, "phantom-fhe-boot_b200", "libb200fhe.so")):
        m.build()
    m.lib()
    return m
