import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


# the resident-tile planner declines trees with little work per phase (jt_flat_kernel is faster there); the tests want the tiny
# brute-force nets to go through jt_resident_kernel as well (read at plan creation by fastbn_b200/csrc/resident.cpp; the tests of
# the rule itself remove it).
os.environ.setdefault("FBN_RES_MIN_WORK", "0")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Native pieces are built in-tree by __graft_entry__.build(); build on demand when a test session starts cold."""
    lib = os.path.join(ROOT, "fastbn_b200", "libfastbn_b200.so")
    emu = os.path.join(ROOT, "oracle", "_build", "libplan_emu.so")
    orc = os.path.join(ROOT, "oracle", "_build", "libjt_oracle.so")
    if not (os.path.exists(lib) and os.path.exists(emu) and os.path.exists(orc)):
        import __graft_entry__
        __graft_entry__.build()


@pytest.fixture(scope="session")
def alarm_tree():
    import numpy as np
    z = np.load(os.path.join(GOLDEN, "alarm_tree_ref.npz"))
    return {k: z[k] for k in z.files}


def load_golden(name):
    import numpy as np
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: z[k] for k in z.files}
