"""GPU parity at trajectory length: the pytest face of tools/long_parity.py.

`pytest -m gpu` runs every case at MANIAC_PARITY_SCALE (default 0.05) of the SURVEY 8d lengths so that the whole GPU suite
stays within minutes; `python tools/long_parity.py` (scale 1: C1 1e4, C2 and C3 2e5 moves per quirk setting) is run once
per round on the GPU box and its log is committed under profiles/.  Every run must show ZERO accept/reject flips and a
per-component relative error below 1e-10."""
import ctypes
import os
import sys
import threading

import numpy as np
import pytest

from conftest import ROOT, compare_records, load_system

sys.path.insert(0, os.path.join(ROOT, "tools"))
pytestmark = pytest.mark.gpu
SCALE = float(os.environ.get("MANIAC_PARITY_SCALE", "0.05"))


def check_rows(rows):
    for r in rows:
        assert r["sequence_mismatches"] == 0, r
        assert r["accept_reject_flips"] == 0, r
        assert max(r["max_rel_err_old_e"], r["max_rel_err_new_e"]) < 1e-10, r
        assert r["max_rel_err_prob"] < 1e-7, r


def test_strict_and_batched_trajectories_against_oracle():
    """C1 h2o_ads, C2 500-water box, C3 ZIF-8 + H2O, both quirk settings, strict mode and batched tape replay"""
    import long_parity as lp
    rows = lp.run_all(scale=SCALE, chunk=2000, with_bench=False, log=lambda m: print(m, flush=True))
    assert len(rows) == 18
    assert min(r["moves_compared"] for r in rows) >= 200
    check_rows(rows)


def test_bench_system_moves_against_oracle():
    """>= 200 oracle-compared moves on the EXACT system bench.py times (bench.make_system: 50 000-atom framework,
    K = 37 782, tile cull on), strict and batched; init_sk = 0 so that neither side needs the O(N K) start-up."""
    import long_parity as lp
    name, s, n, start = lp.bench_case(240)
    assert s.num_kvectors == 37782 and int(s.natoms[np.flatnonzero(s.is_active != 1)[0]]) * int(s.count[np.flatnonzero(s.is_active != 1)[0]]) == 50000
    tape = np.random.default_rng(77).random(lp.DRAWS_PER_MOVE * n + 64)
    ref, offs, ckpt = lp.oracle_reference(s, n, n, tape, start_energy=start)
    assert set(np.unique(ref["kind"])) >= {1, 2, 3, 4}
    rows = [fn(s, ref, offs, ckpt, n, tape, f"{name} / {mode}", start_energy=start).row()
            for mode, fn in (("strict", lp.run_strict), ("strict fused", lp.run_strict_fused), ("batched tape", lp.run_batched))]
    for r in rows:
        print(r)
        assert r["moves_compared"] == n
    check_rows(rows)


def test_h2o_gcmc_mean_n_three_way():
    """Insertion / deletion acceptance pinned on the only reference-held number that exercises it: LAMMPS' N(H2O) series
    of testcase-adsorption/H2O (number.dat, mean 15.23 over 101 samples at 0.5 atm, 300 K).  256 device chains follow
    LAMMPS' protocol length (100 samples x 100 exchange attempts = 200 MC steps of the 0.25/0.25/0.5 mix, physical quirk
    switches).  Three-way: (1) chains 0..3 reproduce the CPU oracle's N series for the same Philox stream EXACTLY;
    (2) LAMMPS' time average is statistically one of ours: |N_lammps - ensemble mean| < 3 sigma of the per-chain time
    averages (the vapour is supersaturated at 0.5 atm, so the per-run spread, not the ensemble's standard error, is the
    yardstick -- see tests/test_oracle_golden.py::lammps_gcmc_series); (3) the ensemble mean is above the ideal-gas
    loading phi V = 12.23 and the ensemble's standard error is reported."""
    import test_oracle_golden as T
    from maniac_mc_b200.lib import ManiacGPU
    from oracle import oracle as orc
    lam = T.lammps_gcmc_series()
    s = load_system("h2o_ads", init_sk=1, deletion_mode=1)
    C_, seed = 256, 2024
    g = ManiacGPU(s, n_chains=C_)
    g.system_energy_all() if hasattr(g, "system_energy_all") else [g.system_energy(c) for c in range(C_)]
    cnt = np.zeros((C_, s.nres), dtype=np.int32)
    series = np.zeros((T.GCMC_SAMPLES, C_), dtype=np.int32)
    ref = {}

    def work(i):
        o = orc.Oracle(s); o.system_energy(); ns = []
        for b in range(T.GCMC_SAMPLES):
            o.run(T.GCMC_MOVES_PER_SAMPLE, seed=seed, chain=i, step0=T.GCMC_MOVES_PER_SAMPLE * b, records=False)
            ns.append(int(o.counts()[0]))
        ref[i] = ns
    ths = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    [t.start() for t in ths]
    for b in range(T.GCMC_SAMPLES):
        g.run_batch_async(T.GCMC_MOVES_PER_SAMPLE, seed=seed)
        g.download_chains(cnt, None, None, None, None)
        series[b] = cnt[:, 0]
    [t.join() for t in ths]
    for i in range(4):
        assert series[:, i].tolist() == ref[i], f"chain {i}: device N series differs from the oracle's"
    avg = series.mean(axis=0)                                   # per-chain time averages, LAMMPS' estimator
    mean, sigma = float(avg.mean()), float(avg.std(ddof=1))
    z = (lam[1:].mean() - mean) / sigma
    print(f"<N> GPU ensemble {mean:.3f} +- {sigma / np.sqrt(C_):.3f} (per-run sigma {sigma:.2f}), LAMMPS {lam[1:].mean():.3f}, z = {z:+.2f}, "
          f"fraction of chains further from the mean than LAMMPS: {np.mean(np.abs(avg - mean) >= abs(lam[1:].mean() - mean)):.2f}")
    assert abs(z) < 3.0
    assert mean > float(s.fugacity[0]) * s.volume
    g.close()


def test_overflow_and_short_tape_are_reported_on_every_path():
    """a creation at count == capacity used to be dropped silently on the isotherm / snapshot paths, and a tape
    shorter than the draws consumed ran into the next chain's tape.  Both now set a sticky device flag that EVERY host
    entry point handing back chain data reports (the reference aborts: CheckMoleculeIndex -> AbortRun)."""
    from maniac_mc_b200 import isotherm as iso
    from maniac_mc_b200.lib import ManiacGPU, ManiacGpuError
    s = load_system("h2o_ads", init_sk=1, deletion_mode=1, p_translation=0.0, p_rotation=0.0)
    s.capacity = (s.count + 2).astype(np.int32)
    s.fugacity = s.fugacity * 1e6                                  # insertions always accepted: the slots fill at once
    g = ManiacGPU(s, n_chains=2)
    for c in range(2):
        g.system_energy(c)
    cnt = np.zeros((2, s.nres), dtype=np.int32)
    g.run_batch_async(200, seed=3)
    with pytest.raises(ManiacGpuError, match="capacity"):
        g.download_chains(cnt, None, None, None, None)
    assert np.all(cnt[:, 0] == s.capacity[0])                      # the arrays were still filled
    g.download_chains(cnt, None, None, None, None)                 # reported once, then cleared
    g.run_batch_async(50, seed=3); g.snapshot_begin(0)
    with pytest.raises(ManiacGpuError, match="capacity"):
        g.snapshot_wait(0)
    with pytest.raises(ManiacGpuError, match="capacity"):
        g.synchronize()                                            # still set on the device: the snapshot only carried a copy
    g.close()
    sw = iso.IsothermSweep(s, n_points=2, replicas=1, lo_atm=1e3, hi_atm=1e4)
    with pytest.raises(ManiacGpuError, match="capacity"):
        for _ in range(4):
            sw.run_block(100, seed=1)
    sw.close()
    s2 = load_system("zif8_h2o", init_sk=1, deletion_mode=1)
    g = ManiacGPU(s2, n_chains=2); g.system_energy(0); g.system_energy(1)
    tape = np.random.default_rng(0).random((2, 40))
    with pytest.raises(ManiacGpuError, match="tape"):
        g.run_batch(50, tape=tape)                                 # 50 steps need > 40 uniforms
    g.close()


def test_moves_dat_columns_by_counter_name(tmp_path):
    """moves.dat columns against the device counters BY NAME (write_utils.f90:182-186)"""
    from maniac_mc_b200 import writers as W
    from maniac_mc_b200.lib import ManiacGPU
    from maniac_mc_b200.system import Counters, N_COUNTERS
    s = load_system("zif8_h2o", init_sk=1, deletion_mode=1)
    g = ManiacGPU(s, n_chains=2); g.system_energy(0); g.system_energy(1)
    ws = [W.ChainWriter(W.chain_directory(str(tmp_path), c, 0), s, nb_block=2) for c in range(2)]
    pipe = W.BlockPipeline(g, ws); pipe.write_initial(); pipe.run(2, 400, seed=9)
    g.synchronize()
    ctr = (Counters * 2)()
    assert ctypes.sizeof(ctr) == 2 * N_COUNTERS * 8
    g.download_chains(None, None, None, None, ctypes.addressof(ctr))
    for c in range(2):
        k = ctr[c].as_dict()
        cols = (tmp_path / f"p{c:03d}_r00" / "moves.dat").read_text().splitlines()[-1].split()
        want = [2, k["trial_translations"], k["translations"], k["trial_creations"], k["creations"], k["trial_deletions"],
                k["deletions"], k["trial_rotations"], k["deletions"]]
        assert [int(x) for x in cols] == want
        assert k["trial_translations"] > 0 and k["trial_rotations"] > 0 and k["trial_translations"] != k["trial_rotations"]
    g.close()
