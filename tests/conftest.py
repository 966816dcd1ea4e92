import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_system(name, **over):
    from maniac_mc_b200.system import System
    s = System.load(os.path.join(GOLDEN, name + ".npz"))
    for k, v in over.items():
        setattr(s, k, v)
    return s


def ideal_mixture(ratio=2.0):
    """The reference's two-species example with every interaction switched off and swap moves only: N_A + N_B is fixed
    and detailed balance of mc_acceptance_probability_swap (monte_carlo_utils.f90:228-268) with a uniformly picked type
    makes N_A binomial with p = phi_A / (phi_A + phi_B).  Returns (system, indices of the two active types, p_A)."""
    import numpy as np
    s = load_system("zif8_ch4o_h2o", init_sk=1, deletion_mode=1)
    s.charges[:] = 0.0; s.epsilon[:] = 0.0
    s.p_translation, s.p_rotation, s.p_swap = 0.0, 0.0, 1.0
    act = np.flatnonzero(s.is_active == 1)
    s.fugacity[act[0]] = ratio * 1e-3; s.fugacity[act[1]] = 1e-3
    return s, act, ratio / (ratio + 1.0)


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


def rel_err(a, b, floor=1.0):
    a = np.asarray(a, dtype=float); b = np.asarray(b, dtype=float)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor)))
