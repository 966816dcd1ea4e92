"""Triclinic cells.  The reference cannot serve as the parity target for a tilted box (its
ApplyPBC / ComputeAtomPhase / ComputeDistance mix row and column conventions), so the CPU restatement implements one
consistent convention and is pinned PHYSICALLY here against the orthorhombic path that the reference's known answers
do pin: a tilted cell and an orthorhombic supercell of the same lattice describe the same crystal, so a converged
Ewald energy of the supercell is (number of cells) x the energy of the tilted cell."""
import itertools

import numpy as np

from maniac_mc_b200.system import System
from oracle import oracle as orc

L = 12.0


def dipole_cell(tilt, n_mol=6, seed=3, kmax=(28, 28, 28), alpha=1.0, cutoff=5.5, box=(L, L, L), com=None, off=None, cap=32):
    rng = np.random.default_rng(seed)
    box = np.asarray(box, dtype=np.float64)
    if com is None:
        s = rng.random((n_mol, 3))
        H = np.array([[box[0], tilt[0], tilt[1]], [0.0, box[1], tilt[2]], [0.0, 0.0, box[2]]])
        com = s @ H.T - box / 2
        v = rng.normal(size=(n_mol, 3)); v *= 0.6 / np.linalg.norm(v, axis=1, keepdims=True)
        off = np.stack([v, -v], axis=1).reshape(-1, 3)
    n_mol = len(com)
    return System(natoms=np.array([2], dtype=np.int32), is_active=np.array([1], dtype=np.int32),
                  count=np.array([n_mol], dtype=np.int32), capacity=np.array([max(cap, n_mol)], dtype=np.int32),
                  charges=np.array([[0.4, -0.4]]), types=np.array([[1, 2]], dtype=np.int32),
                  sigma=np.array([[3.0, 2.8], [2.8, 2.6]]), epsilon=np.array([[60.0, 45.0], [45.0, 35.0]]),
                  com=com, offsets=off, box_lo=-box / 2, box_len=box, alpha=alpha, cutoff=cutoff, temperature=300.0,
                  fugacity=np.array([1e-3]), kmax=list(kmax), tilt=np.asarray(tilt, dtype=np.float64),
                  translation_step=1.5, rotation_step=0.8, p_translation=0.3, p_rotation=0.3, init_sk=1, deletion_mode=1)


def lattice_shifts(tilt, box, super_len):
    """the lattice points of the tilted cell inside the orthorhombic supercell"""
    a = np.array([box[0], 0, 0]); b = np.array([tilt[0], box[1], 0]); c = np.array([tilt[1], tilt[2], box[2]])
    seen = {}
    for i, j, k in itertools.product(range(-8, 9), repeat=3):
        p = i * a + j * b + k * c
        w = p - super_len * np.floor(p / super_len + 1e-9)
        seen[tuple(np.round(w, 6))] = w
    return np.array(list(seen.values()))


def test_tilted_cell_equals_orthorhombic_supercell():
    tilt = (L / 2, L / 2, L / 2)                    # b = (l/2, l, 0), c = (l/2, l/2, l): 2b - a and 4c - 2a - (2b - a) are orthogonal
    tri = dipole_cell(tilt)
    e_tri = orc.Oracle(tri).system_energy()
    sup = np.array([L, 2 * L, 4 * L])
    shifts = lattice_shifts(tilt, (L, L, L), sup)
    assert len(shifts) == 8                         # the supercell holds 8 cells
    com = np.concatenate([tri.com + s for s in shifts]); off = np.tile(tri.offsets, (8, 1))
    ortho = dipole_cell((0.0, 0.0, 0.0), kmax=(28, 56, 112), box=sup, com=com, off=off, cap=64)
    e_sup = orc.Oracle(ortho).system_energy()
    assert np.all(np.abs(e_tri[:5]) > 1e-3)         # every component is exercised (LJ, real, reciprocal, self, intra)
    assert np.allclose(e_sup, 8 * e_tri, rtol=1e-9, atol=1e-9), (e_sup / 8, e_tri)
    # and a tilted cell is NOT its untilted namesake
    e_flat = orc.Oracle(dipole_cell((0.0, 0.0, 0.0), com=tri.com, off=tri.offsets)).system_energy()
    assert abs(e_flat[5] - e_tri[5]) > 1e-3 * abs(e_tri[5])


def test_lattice_translation_leaves_the_energy_unchanged():
    tilt = (3.0, -2.0, 4.5)
    s = dipole_cell(tilt, kmax=(10, 10, 10), alpha=0.45)
    e0 = orc.Oracle(s).system_energy()
    a = np.array([L, 0, 0]); b = np.array([tilt[0], L, 0]); c = np.array([tilt[1], tilt[2], L])
    for vec in (a, b, c, a - b + 2 * c):
        t = s.copy(); t.com = s.com.copy(); t.com[2] += vec; t.com[4] -= vec
        assert np.allclose(orc.Oracle(t).system_energy(), e0, rtol=1e-11, atol=1e-9)


def test_mc_trajectory_in_a_tilted_cell_tracks_the_energy():
    """600 MC steps (all move types) in a tilted cell: the tracked energy equals a fresh single-point evaluation, the
    molecules stay inside the cell (ApplyPBC / insertion in fractional coordinates), and the moves are accepted."""
    tilt = (3.0, -2.0, 4.5)
    s = dipole_cell(tilt, kmax=(8, 8, 8), alpha=0.42, n_mol=10)
    o = orc.Oracle(s)
    o.system_energy()
    o.run(600, seed=21, records=False)
    tracked = o.energy()
    cnt, com, off = o.state()
    c = o.counters()
    assert c["translations"] > 0 and c["rotations"] > 0 and c["creations"] > 0 and c["deletions"] > 0, c
    fresh_sys = s.copy(); fresh_sys.count = cnt.copy(); fresh_sys.com = com; fresh_sys.offsets = off
    fresh = orc.Oracle(fresh_sys).system_energy()
    assert np.allclose(tracked, fresh, rtol=1e-9, atol=1e-7)
    H = np.array([[L, tilt[0], tilt[1]], [0.0, L, tilt[2]], [0.0, 0.0, L]])
    frac = np.linalg.solve(H, (com - s.box_lo).T).T
    assert frac.min() > -1e-9 and frac.max() < 1 + 1e-9


def test_zero_tilt_is_the_orthorhombic_path_bit_for_bit():
    s = dipole_cell((0.0, 0.0, 0.0), kmax=(8, 8, 8), alpha=0.42)
    t = s.copy(); t.tilt = np.array([0.0, 0.0, 0.0])
    a, b = orc.Oracle(s), orc.Oracle(t)
    a.system_energy(); b.system_energy()
    ra, rb = a.run(200, seed=5), b.run(200, seed=5)
    assert np.array_equal(ra, rb)


def test_reference_tilted_example_loads_and_runs():
    """`mc-topology/testcase-energy/DIPOLE-orthorhombic` is, its name notwithstanding, a 40 A cell with tilt factors
    (-20, -20, 20): 71 two-site dipoles (bond length 2, `bond_coeff 1 0 2`).  The fixture was read from the reference's
    files by tools/make_golden.py; molecules repaired across the tilted boundaries have their bond length back."""
    from conftest import load_system
    s = load_system("dipole_tri", init_sk=1, deletion_mode=1)
    assert np.array_equal(s.tilt, [-20.0, -20.0, 20.0]) and s.count.tolist() == [71] and s.kmax == [8, 7, 6]
    bond = np.linalg.norm(s.offsets[0::2] - s.offsets[1::2], axis=1)
    assert np.allclose(bond, 2.0, atol=0.3)                      # a thermal snapshot of a harmonic bond with r0 = 2
    H = np.array([[40.0, -20.0, -20.0], [0.0, 40.0, 20.0], [0.0, 0.0, 40.0]])
    frac = np.linalg.solve(H, (s.com - s.box_lo).T).T
    assert frac.min() >= 0 and frac.max() < 1
    o = orc.Oracle(s); e0 = o.system_energy()
    assert np.all(np.isfinite(e0)) and abs(e0[1]) > 0 and abs(e0[2]) > 0
    o.run(500, seed=2, records=False)
    cnt, com, off = o.state()
    t = s.copy(); t.count = cnt.copy(); t.com = com; t.offsets = off
    assert np.allclose(o.energy(), orc.Oracle(t).system_energy(), rtol=1e-9, atol=1e-7)
