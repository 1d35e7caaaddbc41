import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """CPU oracle library (test infrastructure; built on demand with plain g++)."""
    from tests.harness import load_oracle
    return load_oracle()


@pytest.fixture(scope="session")
def product():
    """The product: sparknet_b200/csrc/libccaffe.so.  No fallback: missing library == failure."""
    from sparknet_b200.library import load_library
    # Tests arbitrate in the fp32 FFMA validation mode unless they ask for TF32 explicitly
    # (sparknet_set_precision(state, 1)); the library's own default is TF32.
    os.environ["SPARKNET_PRECISION"] = "ffma"
    lib = load_library()
    lib.set_mode_gpu()
    return lib
