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


# ---- move-record comparison (GPU vs oracle, oracle vs fixture) ----------------------------------------------------
REL_TOL = 1e-10    # BASELINE.json north_star: energies within 1e-10 relative in FP64
FLOOR_FRAC = 1e-3  # a component smaller than 1e-3 |E_total| of its own record is measured against that floor
FLOOR_K = 1e-6     # absolute floor in Kelvin (all-zero records)


def record_errors(a, b):
    """Per-component relative errors of the old/new energy vectors of two record arrays (b = the reference side).
    err_i = |a_i - b_i| / max(|b_i|, FLOOR_FRAC * |b_total|, FLOOR_K): every component (LJ, real Coulomb, reciprocal,
    self, intra, total) is held to the tolerance on its OWN magnitude, not on the largest one of the record; the floor only
    keeps components that are (near) zero next to the record's total from dividing by nothing.
    Returns (max error over old_e, max over new_e, max relative error of prob)."""
    out = []
    for key in ("old_e", "new_e"):
        x, y = np.asarray(a[key], dtype=float), np.asarray(b[key], dtype=float)
        if x.size == 0:
            out.append(0.0); continue
        scale = np.maximum(np.maximum(np.abs(y), FLOOR_FRAC * np.abs(y[..., 5:6])), FLOOR_K)
        out.append(float(np.max(np.abs(x - y) / scale)))
    pa, pb = np.asarray(a["prob"], dtype=float), np.asarray(b["prob"], dtype=float)
    out.append(float(np.max(np.abs(pa - pb) / np.maximum(pb, 1e-3))) if pa.size else 0.0)
    return tuple(out)


def compare_records(a, b, tol=REL_TOL):
    """identical move sequence and accept/reject decisions; energies per component within `tol`; returns the errors"""
    assert np.array_equal(a["kind"], b["kind"])
    assert np.array_equal(a["res"], b["res"]) and np.array_equal(a["mol"], b["mol"])
    assert np.array_equal(a["accepted"], b["accepted"]), "accept/reject sequence differs"
    e_old, e_new, e_p = record_errors(a, b)
    assert e_old < tol, f"old_e per-component relative error {e_old:.3e} >= {tol:g}"
    assert e_new < tol, f"new_e per-component relative error {e_new:.3e} >= {tol:g}"
    assert e_p < 1e-9, f"acceptance probability relative error {e_p:.3e}"
    return e_old, e_new, e_p
