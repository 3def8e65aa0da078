import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def ctx():
    """One device context for the whole GPU session (eps_ctx)."""
    from epsilon_b200.solver import Context
    c = Context(0)
    yield c
    c.close()


@pytest.fixture(scope="session")
def ref():
    """The compiled reference solver (oracle/_ref/epsilon_ref); test infrastructure only."""
    from oracle import ref_runner
    if not ref_runner.available():
        pytest.skip("oracle/_ref/epsilon_ref not built (make -C oracle)")
    return ref_runner
