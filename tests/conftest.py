import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_sessionstart(session):
    """Built artefacts are not in git: in a fresh checkout build them once (what __graft_entry__.build() does - the CUDA
    library, the C++ front end, the oracle and, where /root/reference exists, the compiled reference)."""
    need = [os.path.join(ROOT, "gcta_b200", "libgcta_b200.so"), os.path.join(ROOT, "gcta_b200", "gcta64_b200"),
            os.path.join(ROOT, "oracle", "libgrm_oracle.so")]
    if all(os.path.exists(p) for p in need):
        return
    try:
        import __graft_entry__
        __graft_entry__.build()
    except Exception as e:   # the tests that need the missing piece will say so
        print(f"conftest: build failed: {e}", file=sys.stderr)


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
