"""Shared fixtures.  `-m "not gpu"` runs on a CPU-only box: oracle vs the
reference's own compiled sources and golden vectors, host logic, ABI symbols,
gloo multi-process plumbing.  `-m gpu` tests are the parity tests proper and go
through the C ABI of libamrvr_b200.so on a B200."""
import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle.oracle import Oracle

    return Oracle()


@pytest.fixture(scope="session")
def ref():
    from oracle.oracle import Ref, RefUnavailable

    try:
        return Ref()
    except (RefUnavailable, OSError) as exc:  # GPU box without the prebuilt _ref
        pytest.skip(f"reference build unavailable: {exc}")


@pytest.fixture(scope="session")
def lib():
    from amrvolumerenderer_b200 import ctypes_api as api

    if not api.LIB_PATH.exists():
        import __graft_entry__ as g

        g.build()
    return api.load_library()


@pytest.fixture(scope="session")
def ctx(lib):
    from amrvolumerenderer_b200.engine import Context

    c = Context(device=0)
    yield c
    c.close()


GOLDEN = ROOT / "tests" / "golden"
