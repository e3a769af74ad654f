import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_sessionstart(session):
    """Built artefacts are not in the history: compile what is missing or stale before the first test (nvcc
    cross-compiles without a GPU; the checker libraries need g++ only).  Best effort — a box without the toolchain
    runs with whatever travelled with the snapshot, and the tests that need a missing library say so."""
    try:
        import improved_body_parts_b200 as ibp
        ibp.build_library(force=False)
    except Exception as e:  # noqa: BLE001
        sys.stderr.write("conftest: libibp_b200.so not (re)built: %s\n" % e)
    try:
        from oracle import build_oracle
        build_oracle.build_oracle()
        build_oracle.build_ref()
    except Exception as e:  # noqa: BLE001
        sys.stderr.write("conftest: checker libraries not (re)built: %s\n" % e)


@pytest.fixture(scope="session")
def oracle():
    from oracle.cpu_lib import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    from oracle.cpu_lib import Reference, reference_available
    if not reference_available():
        pytest.skip("oracle/_ref/libpafprocess_ref.so not built and /root/reference absent")
    return Reference()
