import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


GOLDEN = os.path.join(ROOT, "tests", "golden")
REF_ROOT = "/root/reference"


@pytest.fixture(scope="session")
def oracle():
    from oracle.oracle_py import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    """The unmodified reference; only where oracle/_ref was built (it travels to the GPU box prebuilt)."""
    from oracle.oracle_py import Reference, build
    if not Reference.available() and os.path.isdir(REF_ROOT):
        build(ref=True)
    if not Reference.available():
        pytest.skip("oracle/_ref/libmargin_ref.so not built")
    return Reference()


@pytest.fixture(scope="session")
def golden_model():
    from margin_b200 import model

    def load(name):
        return model.load_flat_model(os.path.join(GOLDEN, "model_%s.bin" % name))
    return load


@pytest.fixture(scope="session")
def gpu_aligner(golden_model):
    """PairHmmAligner on cuda:0 with the r94-g360 model; fails loudly when the CUDA library is missing."""
    from margin_b200.aligner import PairHmmAligner
    pm = golden_model("r94g360")
    return PairHmmAligner(pm.models, pm.p, device=0)
