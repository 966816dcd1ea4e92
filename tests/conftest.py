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
FLOOR_FRAC = 1e-3  # a component smaller than 1e-3 of its record's GROSS energy sum_i |E_i| is measured against that floor
FLOOR_K = 10.0     # ... and never against less than 10 K: 1e-10 x 10 K = 1e-9 K (3e-12 kT at 300 K) absolute


def record_scale(y):
    """Denominator of the per-component relative error of an energy vector y = (LJ, real Coulomb, reciprocal, self, intra,
    total) [K].  Every component is held to REL_TOL on its OWN magnitude -- not on the largest one of the record, which
    would let E_recip ~ 1e5..1e6 K hide errors of the real-space terms -- with two floors for components that are small
    only because larger terms cancel: 1e-3 of the record's gross energy sum |E_i| (a total of 0.07 K made of +-25 000 K
    self and intra terms cannot be known to 1e-11 K in FP64), and 10 K.  The second floor is the measured reproducibility of
    the reference algorithm itself: a real-space Coulomb sum of +-1e3..1e4 K terms that cancels to 0.05 K moves by ~1e-10 K
    under any re-association; tools/oracle_noise_floor.py shows the CPU restatement shifting by up to 5e-9 K against ITSELF
    when gcc merely contracts a*b+c into FMAs (profiles/r2_oracle_noise_floor.log)."""
    y = np.asarray(y, dtype=float)
    gross = np.sum(np.abs(y[..., :5]), axis=-1, keepdims=True)
    return np.maximum(np.maximum(np.abs(y), FLOOR_FRAC * gross), FLOOR_K)


def record_errors(a, b):
    """Per-component relative errors of the old/new energy vectors of two record arrays (b = the reference side), see
    record_scale.  Returns (max error over old_e, max over new_e, max relative error of prob)."""
    out = []
    for key in ("old_e", "new_e"):
        x, y = np.asarray(a[key], dtype=float), np.asarray(b[key], dtype=float)
        if x.size == 0:
            out.append(0.0); continue
        out.append(float(np.max(np.abs(x - y) / record_scale(y))))
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
    # p = min(1, c exp(-dE / T)): its relative error is |error of dE| / T, i.e. ~1e-16 (|E_old| + |E_new|) / T -- 3e-9 for the 8e7 K
    # totals of the 50 000-atom frameworks at 300 K; the accept / reject identity above is the sharp test
    assert e_p < 1e-7, f"acceptance probability relative error {e_p:.3e}"
    return e_old, e_new, e_p
