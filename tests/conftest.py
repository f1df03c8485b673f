import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import __graft_entry__ as entry  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    return entry.load_package()


@pytest.fixture(scope="session")
def synth(pkg):
    return pkg.synth


@pytest.fixture(scope="session")
def oracle_mod():
    from oracle import score_oracle

    score_oracle.build()
    return score_oracle


@pytest.fixture(scope="session")
def sched(pkg):
    """A context on cuda:0; GPU tests only."""
    s = pkg.Scheduler(0)
    yield s
    s.close()
