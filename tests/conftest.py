import os
import sys

import pytest

# The oracle's OpenMP regions are short and its threads spin at their barriers: with as many threads as cores any other runnable process on the box
# turns that into a 10 x slowdown of single tests (seen on the shared CPU box: 15 s -> 160 s).  Half the cores keeps the CPU suite's time stable;
# bench.py's CPU legs set their own count (all cores) and restore this one.  Read by libgomp when liboracle.so is loaded, i.e. after this module.
os.environ.setdefault("OMP_NUM_THREADS", str(max(2, (os.cpu_count() or 4) // 2)))

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ufemism2.0_b200"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure the host library, the oracle and (if nvcc is present) the CUDA library are built."""
    import __graft_entry__ as g
    g.build(quiet=True)
