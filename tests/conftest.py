import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # The tests exercise the built library, the `lda` executable and the compiled oracle.  They are
    # built artefacts (not in git): a checkout that has not run __graft_entry__.build() yet gets them
    # here.  Nothing is rebuilt when they exist.
    need = [os.path.join(ROOT, "oni_ml_b200", "liboni_lda.so"), os.path.join(ROOT, "oni_ml_b200", "bin", "lda"),
            os.path.join(ROOT, "oracle", "_build", "liboni_oracle.so")]
    if not all(os.path.exists(p) for p in need):
        import __graft_entry__
        __graft_entry__.build()


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def c0_corpus():
    from oni_ml_b200.corpus import read_model_dat
    return read_model_dat(os.path.join(GOLDEN, "c0_model.dat"))
