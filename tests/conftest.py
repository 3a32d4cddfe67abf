import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as o
    o.build()
    return o


def pytest_terminal_summary(terminalreporter):
    """How often the extended-precision referee of tests/cases.close_or_exact decided a point."""
    from tests import cases
    if not cases.REFEREE_LOG:
        terminalreporter.write_line("referee (cases.close_or_exact): never used")
        return
    terminalreporter.write_line(f"referee (cases.close_or_exact) used in {len(cases.REFEREE_LOG)} check(s):")
    for test, n, k in cases.REFEREE_LOG:
        terminalreporter.write_line(f"  {k} point(s) of n = {n}: {test}")
