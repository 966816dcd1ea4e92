"""Per-block output files: the reference's Fortran edit descriptors restated, checked against rows
written out by hand from `write_utils.f90:12-188`, and the per-chain directory scheme.  Host logic only."""
import os

import numpy as np

from maniac_mc_b200 import writers as W
from maniac_mc_b200.system import System


def test_fortran_edit_descriptors():
    assert W._fint(42, 10) == "        42" and W._fint(-7, 4) == "  -7" and W._fint(123456, 4) == "****"
    assert W._ffix(1.5, 16, 6) == "        1.500000" and W._ffix(-0.25, 12, 7) == "  -0.2500000"
    assert W._ffix(123456.789, 8, 3) == "********" and W._ffix(0.5, 4, 3) == ".500" and W._ffix(-0.5, 5, 3) == "-.500"
    assert np.array_equal(W._nint([0.5, -0.5, 1.5, 2.5, -2.5, 0.49]), [1, -1, 2, 3, -3, 0])     # not banker's rounding
    assert np.allclose(W.wrap_into_box([[15.0, -15.0, 5.0]], np.array([20.0, 20.0, 20.0])), [[-5.0, 5.0, 5.0]])
    assert np.allclose(W.wrap_into_box([[10.0, -10.0, 30.0]], np.array([20.0, 20.0, 20.0])), [[-10.0, 10.0, -10.0]])


def test_rows_follow_write_energy_and_count():
    e_k = np.array([100.0, -200.0, 50.0, -1000.0, 25.0, -1025.0])          # Kelvin, C-ABI order
    row = W.energy_row(3, e_k)
    k = W.KB_KCALMOL
    want = "%10d %16.6f %16.6f %16.6f %16.6f %16.6f %16.6f" % (3, -1025 * k, 50 * k, 100 * k, -200 * k, -1000 * k, 25 * k)
    assert row == want and len(row) == 10 + 6 * 17
    assert W.number_row(12, 345) == "        12        345"
    # mmc_counters struct order: translations, rotations, creations, deletions, then the four trial counts, swaps, trial_swaps
    ctr = [1, 3, 5, 7, 2, 4, 6, 8, 0, 0]
    lit = W.moves_row(9, ctr).split()
    # write_utils.f90:182-186: trial_translations, translations, trial_creations, creations, trial_deletions, deletions,
    # trial_rotations, deletions (the last column repeats the deletions)
    assert lit == ["9", "2", "1", "6", "5", "8", "7", "4", "7"]
    assert W.moves_row(9, ctr, literal=False).split()[-1] == "3"
    by_name = dict(zip(W.COUNTER_FIELDS, ctr))
    assert W.moves_row(9, by_name) == W.moves_row(9, ctr)
    from maniac_mc_b200.system import Counters
    assert tuple(f for f, _ in Counters._fields_) == W.COUNTER_FIELDS      # the row is indexed by the struct's own order
    assert len(W.moves_row(9, ctr)) == 9 * 12 + 8
    assert W.MOVES_HEADER.split()[1:] == ["Block", "Trial_Trans", "Trans_Moves", "Trial_Create", "Create_Moves", "Trial_Delete",
                                          "Delete_Moves", "Trial_Rotate", "Rotate_Moves", "Trial_BigMove", "Big_Moves"]


def tiny_system():
    # residue 0: 2-atom framework "molecule" (inactive), residue 1: 2-atom guest (active)
    return System(natoms=np.array([2, 2], dtype=np.int32), is_active=np.array([0, 1], dtype=np.int32),
                  count=np.array([1, 2], dtype=np.int32), capacity=np.array([1, 4], dtype=np.int32),
                  charges=np.zeros((2, 2)), types=np.array([[1, 2], [3, 3]], dtype=np.int32),
                  sigma=np.ones((3, 3)), epsilon=np.ones((3, 3)),
                  com=np.array([[9.0, 0.0, 0.0], [11.0, 1.0, -11.0], [0.0, 0.0, 0.0]]),
                  offsets=np.array([[2.0, 0, 0], [-2.0, 0, 0], [0.5, 0, 0], [-0.5, 0, 0], [0, 0.5, 0], [0, -0.5, 0]]),
                  box_lo=np.array([-10.0, -10.0, -10.0]), box_len=np.array([20.0, 20.0, 20.0]), alpha=0.3, cutoff=8.0,
                  temperature=300.0, fugacity=np.array([0.0, 1e-5]), kmax=[3, 3, 3], names=["fw", "gas"])


def test_lammpstrj_frame_wraps_like_the_reference():
    s = tiny_system()
    txt = W.lammpstrj_frame(50, s.box_len, s.types, s.is_active, s.natoms, s.count, s.com, s.offsets).splitlines()
    assert txt[:5] == ["ITEM: TIMESTEP", "        50", "ITEM: NUMBER OF ATOMS", "         6", "ITEM: BOX BOUNDS pp pp pp"]
    assert txt[5] == "   -10.00000000     10.00000000" and txt[8] == "ITEM: ATOMS id type x y z"
    rows = [l.split() for l in txt[9:]]
    assert [r[:2] for r in rows] == [["1", "1"], ["2", "2"], ["3", "3"], ["4", "3"], ["5", "3"], ["6", "3"]]
    xyz = np.array([[float(v) for v in r[2:]] for r in rows])
    # framework atoms wrapped one by one: 9+2 = 11 -> -9; 9-2 = 7
    assert np.allclose(xyz[0], [-9, 0, 0]) and np.allclose(xyz[1], [7, 0, 0])
    # active molecule: centre (11, 1, -11) -> (-9, 1, 9), atoms stay attached (not wrapped again)
    assert np.allclose(xyz[2], [-8.5, 1, 9]) and np.allclose(xyz[3], [-9.5, 1, 9])
    assert txt[9] == "     1    1   -9.0000000    0.0000000    0.0000000"


def test_chain_writer_files_and_directories(tmp_path):
    s = tiny_system()
    d = W.chain_directory(str(tmp_path), 7, 3)
    assert d.endswith("p007_r03")
    w = W.ChainWriter(d, s, nb_block=10)
    e = np.arange(6.0)
    w.write_block(0, s.count, s.com, s.offsets, e, np.zeros(10, dtype=np.int64))
    w.write_block(1, np.array([1, 0]), s.com[:1], s.offsets[:2], e, np.arange(10))     # empty block: no number_ row
    w.write_block(2, np.array([1, 1]), s.com[:2], s.offsets[:4], e, np.arange(10))
    assert sorted(os.listdir(d)) == ["energy.dat", "moves.dat", "number_gas.dat", "trajectory.lammpstrj"]
    en = open(os.path.join(d, "energy.dat")).read().splitlines()
    assert en[0] == W.ENERGY_HEADER and len(en) == 4 and en[2].split()[0] == "1"
    nm = open(os.path.join(d, "number_gas.dat")).read().splitlines()
    assert nm == [W.NUMBER_HEADER, W.number_row(0, 2), W.number_row(2, 1)]
    tr = open(os.path.join(d, "trajectory.lammpstrj")).read().splitlines()
    assert tr.count("ITEM: TIMESTEP") == 3 and tr[1] == "        10"                  # literal: nb_block on every frame
    # block 0 starts the files again (STATUS='REPLACE')
    w.write_block(0, s.count, s.com, s.offsets, e, np.zeros(10, dtype=np.int64))
    assert len(open(os.path.join(d, "energy.dat")).read().splitlines()) == 2


def test_unpack_chain_merges_device_slots_with_the_host_framework():
    s = tiny_system()
    gcom = np.arange(12.0).reshape(4, 3)         # capacity 4 slots of the active residue
    goff = np.arange(24.0).reshape(8, 3)
    com, off = W.unpack_chain(s, np.array([1, 3]), gcom, goff)
    assert com.shape == (4, 3) and off.shape == (8, 3)
    assert np.array_equal(com[0], s.com[0]) and np.array_equal(com[1:], gcom[:3])
    assert np.array_equal(off[:2], s.offsets[:2]) and np.array_equal(off[2:], goff[:6])
