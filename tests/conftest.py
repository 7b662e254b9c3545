import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def oracle_lib():
    import oracle_lib
    return oracle_lib.load_oracle()


@pytest.fixture(scope="session")
def oracle_fma_lib():
    """the reference sources rebuilt with floating-point contraction on (oracle/Makefile)"""
    import oracle_lib
    return oracle_lib.load_oracle("fma")


@pytest.fixture(scope="session")
def product_lib():
    from grackle_b200 import capi
    return capi.load_product()


@pytest.fixture(scope="session")
def synth_table():
    from grackle_b200 import synth_tables
    return synth_tables.ensure()


@pytest.fixture(scope="session")
def metal_table_2008():
    """cloudy_metals_2008_3D.h5 (old-style rank-3 table the freefall_metal golden run used):
    a copy under tests/golden/ travels to the GPU box; the reference checkout is the
    source in the dev container"""
    for p in (os.path.join(ROOT, "tests", "golden", "cloudy_metals_2008_3D.h5"),
              "/root/reference/grackle_data_files/input/cloudy_metals_2008_3D.h5"):
        if os.path.exists(p):
            return p
    pytest.skip("cloudy_metals_2008_3D.h5 not available")
