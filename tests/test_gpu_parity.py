"""GPU parity tests (B200): every result goes through the C ABI of libmaniacgpu.so and is compared with the CPU
oracle on the same inputs.  FP64 tolerances: 1e-10 relative on energies (north star), 1e-12 where the sum order
is the only difference; accept/reject sequences must be identical."""
import numpy as np
import pytest

from conftest import load_system, rel_err
from maniac_mc_b200 import builders
from maniac_mc_b200.driver import StrictDriver, TapeRng
from maniac_mc_b200.lib import ManiacGPU, ManiacGpuError
from maniac_mc_b200.system import (E_INTRA, E_RECIP, E_TOTAL, MMC_CREATION, MMC_DELETION, MMC_ROTATION,
                                   MMC_TRANSLATION)
from oracle import oracle as orc

pytestmark = pytest.mark.gpu
TOL = 1e-10   # BASELINE.json north_star: energies within 1e-10 relative in FP64


def compare_records(a, b, tol=TOL):
    assert np.array_equal(a["kind"], b["kind"])
    assert np.array_equal(a["res"], b["res"]) and np.array_equal(a["mol"], b["mol"])
    assert np.array_equal(a["accepted"], b["accepted"]), "accept/reject sequence differs"
    scale = np.maximum(np.abs(b["new_e"]).max(axis=1, keepdims=True), 1.0)
    assert np.max(np.abs(a["old_e"] - b["old_e"]) / scale) < tol
    assert np.max(np.abs(a["new_e"] - b["new_e"]) / scale) < tol
    assert np.max(np.abs(a["prob"] - b["prob"])) < 1e-8


@pytest.mark.parametrize("name", ["lj_gas", "h2o_gas", "h2o_ads", "zif8_h2o"])
def test_system_energy(name):
    """ComputeSystemEnergy: configs[0] of BASELINE.json (the reference's own CPU-runnable cases)."""
    s = load_system(name)
    g = ManiacGPU(s); e = g.system_energy()
    ref = orc.Oracle(s).system_energy()
    assert rel_err(e, ref) < 1e-12
    g.close()


def test_kvectors_and_weights():
    s = load_system("h2o_gas"); g = ManiacGPU(s); o = orc.Oracle(s)
    for a, b in zip(g.kvectors(), o.kvectors()):
        if a.dtype.kind == "i":
            assert np.array_equal(a, b)
        else:
            np.testing.assert_allclose(a, b, rtol=1e-14)
    g.close()


def test_full_structure_factor():
    s = load_system("zif8_h2o", init_sk=1); g = ManiacGPU(s); o = orc.Oracle(s)
    g.system_energy(); o.system_energy()
    np.testing.assert_allclose(g.sk(), o.sk(), atol=1e-10)
    s0 = load_system("zif8_h2o"); g0 = ManiacGPU(s0); g0.system_energy()
    assert np.all(g0.sk() == 0)   # Q1 default
    g.close(); g0.close()


def test_device_erfc_accuracy():
    """The sweep's Coulomb kernel g(r^2) = erfc(alpha r)/r (degree-9 piecewise table in r^2, built for this alpha;
    slow path outside the table): absolute error at the rounding floor everywhere, relative where terms matter."""
    from scipy.special import erfc
    g = ManiacGPU(load_system("lj_gas"))
    x = np.concatenate([np.linspace(1e-3, 7.5, 20001), 1e-3 + np.random.default_rng(0).random(20000) * 27.0])
    d = g.dbg_erfc(x); ref = erfc(x)
    assert np.max(np.abs(d - ref)) < 2e-15
    m = x <= 1.5
    assert np.max(np.abs(d[m] - ref[m]) / ref[m]) < 1e-13
    g.close()


def test_table_lookup():
    s = load_system("lj_gas"); g = ManiacGPU(s); o = orc.Oracle(s)
    r = np.concatenate([[-1.0, 0.0, s.cutoff, s.cutoff + 1], np.random.default_rng(1).random(500) * s.cutoff])
    for which in range(3):
        np.testing.assert_allclose(g.dbg_table_lookup(which, r), o.table_lookup(which, r), rtol=1e-13)
    g.close()


def test_system_energy_table_mode():
    s = load_system("lj_gas", use_table=1)
    g = ManiacGPU(s); assert rel_err(g.system_energy(), orc.Oracle(s).system_energy()) < 1e-12
    g.close()


@pytest.mark.parametrize("name", ["zif8_h2o", "h2o_gas", "lj_gas"])
def test_pair_self_intra_singlemol(name):
    s = load_system(name); g = ManiacGPU(s); o = orc.Oracle(s)
    r = int(np.argmax(s.is_active))
    for m in range(min(int(s.count[r]), 5)):
        a = g.dbg_pair_singlemol(r, m); b = o.pair_singlemol(r, m)
        assert rel_err(a, b) < 1e-12
        assert abs(g.dbg_intra_singlemol(r, m) - o.intra_singlemol(r, m)) <= 1e-12 * max(1.0, abs(o.intra_singlemol(r, m)))
    assert g.dbg_self_singlemol(r) == pytest.approx(o.self_singlemol(r), rel=1e-14)
    g.close()


@pytest.mark.parametrize("name,init_sk,deletion", [("zif8_h2o", 0, 0), ("zif8_h2o", 1, 1), ("h2o_ads", 0, 0), ("h2o_ads", 1, 1)])
def test_strict_mode_trajectory(name, init_sk, deletion):
    """Strict mode: host draws the uniforms and makes the proposals (as the Fortran driver does), one
    mmc_trial_energy + mmc_trial_commit per move; same tape through the oracle."""
    s = load_system(name, init_sk=init_sk, deletion_mode=deletion)
    tape = np.random.default_rng(42).random(6000)
    o = orc.Oracle(s); o.system_energy(); ref = o.run(400, tape=tape)
    g = ManiacGPU(s); drv = StrictDriver(s, g); drv.ComputeSystemEnergy()
    rng = TapeRng(tape); got = drv.MonteCarloLoop(rng, 1, 400)
    compare_records(got, ref)
    assert rng.pos == o.tape_used
    cnt, com, off, e = g.sync_to_host()
    c2, com2, off2 = o.state()
    assert np.array_equal(cnt, c2)
    np.testing.assert_allclose(com, com2, atol=1e-12); np.testing.assert_allclose(off, off2, atol=1e-12)
    assert rel_err(e, o.energy()) < TOL and rel_err(drv.energy, o.energy()) < TOL
    np.testing.assert_allclose(g.sk(), o.sk(), atol=1e-9)
    g.close()


@pytest.mark.parametrize("name,init_sk,deletion", [("zif8_h2o", 0, 0), ("h2o_ads", 1, 1), ("lj_gas", 0, 0)])
def test_batched_tape_replay(name, init_sk, deletion):
    """Batched mode with on-device proposals replaying recorded streams: 3 chains, 3 different tapes."""
    s = load_system(name, init_sk=init_sk, deletion_mode=deletion)
    n = 300
    tape = np.random.default_rng(7).random((3, 12 * n))
    g = ManiacGPU(s, n_chains=3)
    for c in range(3):
        g.system_energy(c)
    recs = g.run_batch(n, tape=tape, records=True)
    for c in range(3):
        o = orc.Oracle(s); o.system_energy(); ref = o.run(n, tape=tape[c])
        compare_records(recs[c], ref)
        assert g.tape_used(c) == o.tape_used
        assert g.counters[c] == o.counters()
        cnt, com, off, e = g.sync_to_host(c); c2, com2, off2 = o.state()
        assert np.array_equal(cnt, c2)
        np.testing.assert_allclose(com, com2, atol=1e-12); np.testing.assert_allclose(off, off2, atol=1e-12)
        assert rel_err(e, o.energy()) < TOL
        np.testing.assert_allclose(g.sk(c), o.sk(), atol=1e-9)
    g.close()


def test_batched_philox_matches_oracle_and_graph_path():
    """Device Philox4x32-10 stream == oracle's; 300 steps = 2 CUDA-graph chunks of 128 + 44 direct steps,
    then a second batch continues the step counter."""
    s = load_system("zif8_h2o", init_sk=1, deletion_mode=1)
    g = ManiacGPU(s, n_chains=4)
    for c in range(4):
        g.system_energy(c)
    r1 = g.run_batch(300, seed=1234, records=True)
    r2 = g.run_batch(130, seed=1234, records=True)
    for c in (0, 3):
        o = orc.Oracle(s); o.system_energy()
        ref1 = o.run(300, seed=1234, chain=c, step0=0); ref2 = o.run(130, seed=1234, chain=c, step0=300)
        compare_records(r1[c], ref1); compare_records(r2[c], ref2)
    assert not np.array_equal(r1[0]["kind"], r1[1]["kind"])   # chains are independent
    g.close()


def test_per_chain_parameters():
    """Isotherm sweep: each chain has its own fugacity/temperature."""
    s = load_system("zif8_h2o", init_sk=1, deletion_mode=1)
    g = ManiacGPU(s, n_chains=2)
    fug = s.fugacity.copy(); fug[0] *= 1e-3
    g.set_chain_params(1, fug, 350.0)
    g.system_energy(0); g.system_energy(1)
    recs = g.run_batch(200, seed=9, records=True)
    o = orc.Oracle(s); o.set_params(fug, 350.0); o.system_energy()
    compare_records(recs[1], o.run(200, seed=9, chain=1))
    g.close()


def test_tracked_equals_fresh_after_batch():
    """Size-independent identity: tracked energies and A(k) after a batch == a fresh single-point evaluation."""
    s = load_system("zif8_h2o", init_sk=1, deletion_mode=1)
    g = ManiacGPU(s, n_chains=2); g.system_energy(0); g.system_energy(1)
    g.run_batch(500, seed=3)
    for c in range(2):
        _, _, _, tracked = g.sync_to_host(c); a_tracked = g.sk(c)
        fresh = g.system_energy(c)
        assert rel_err(tracked, fresh) < 1e-9
        np.testing.assert_allclose(a_tracked, g.sk(c), atol=1e-9)
    g.close()


def test_set_and_get_chain_state_roundtrip():
    s = load_system("h2o_ads"); g = ManiacGPU(s, n_chains=2)
    s2 = load_system("h2o_gas")
    cnt = np.array([40], dtype=np.int32)
    g.set_chain_state(1, cnt, s2.com[:40], s2.offsets[:160])
    c, com, off = g.get_chain_state(1)
    assert c[0] == 40 and np.array_equal(com, s2.com[:40]) and np.array_equal(off, s2.offsets[:160])
    c0, com0, _ = g.get_chain_state(0)
    assert c0[0] == s.count[0] and np.array_equal(com0, s.com)
    s3 = s.copy(); s3.count = cnt; s3.com = s2.com[:40]; s3.offsets = s2.offsets[:160]
    assert rel_err(g.system_energy(1), orc.Oracle(s3).system_energy()) < 1e-12
    g.close()


def test_error_behaviour():
    s = load_system("lj_gas"); g = ManiacGPU(s)
    with pytest.raises(ManiacGpuError, match="not an active"):
        g.trial_energy(MMC_TRANSLATION, 1, 0, np.zeros(3), np.zeros((1, 3)))   # residue B is inactive
    with pytest.raises(ManiacGpuError, match="out of range"):
        g.trial_energy(MMC_TRANSLATION, 0, 999, np.zeros(3), np.zeros((1, 3)))
    with pytest.raises(ManiacGpuError, match="Unknown move_type"):
        g.trial_energy(9, 0, 0, np.zeros(3), np.zeros((1, 3)))
    with pytest.raises(ManiacGpuError, match="no trial pending"):
        g.trial_commit(1)
    g.trial_energy(MMC_DELETION, 0, 0)
    with pytest.raises(ManiacGpuError, match="already pending"):
        g.trial_energy(MMC_DELETION, 0, 1)
    g.trial_commit(0)
    # capacity: CheckMoleculeIndex analogue
    s2 = load_system("h2o_ads"); s2.capacity = s2.count.copy()
    g2 = ManiacGPU(s2)
    with pytest.raises(ManiacGpuError, match="exceeds the maximum"):
        g2.trial_energy(MMC_CREATION, 0, int(s2.count[0]), np.zeros(3), np.zeros((4, 3)))
    g.close(); g2.close()


def test_empty_box_and_refill():
    """Edge case of config 2: delete every molecule, then insert into the empty box (slot-1 template, Q9)."""
    s = load_system("h2o_ads", init_sk=1, deletion_mode=1, p_translation=0.0, p_rotation=0.0)
    s.count = np.array([2], dtype=np.int32); s.com = s.com[:2]; s.offsets = s.offsets[:8]
    s.fugacity = s.fugacity * 1e-6     # deletions always accepted at first
    n = 120
    rng = np.random.default_rng(0); tape = rng.random((1, 12 * n))
    g = ManiacGPU(s); g.system_energy()
    o = orc.Oracle(s); o.system_energy()
    recs = g.run_batch(n, tape=tape, records=True); ref = o.run(n, tape=tape[0])
    compare_records(recs[0], ref)
    assert (ref["kind"] == 0).any()      # deletions attempted on an empty box are no-ops
    g.close()


def test_large_framework_identities():
    """C4-sized system (50k-atom framework, K~7.4k): parity on a few moves against the oracle (seconds of CPU), and the
    size-independent identity tracked == fresh after a batch."""
    s = builders.synthetic_framework(n_framework=50000, n_guests=64, box=96.0, capacity=128)
    s.init_sk, s.deletion_mode = 1, 1
    g = ManiacGPU(s, n_chains=2)
    e0 = g.system_energy(0); g.system_energy(1)
    o = orc.Oracle(s)
    assert rel_err(e0, o.system_energy()) < 1e-11
    tape = np.random.default_rng(1).random((2, 200))
    recs = g.run_batch(12, tape=tape, records=True)
    compare_records(recs[0], o.run(12, tape=tape[0]))
    g.run_batch(200, seed=5)
    for c in range(2):
        _, _, _, tracked = g.sync_to_host(c)
        assert rel_err(tracked, g.system_energy(c)) < 1e-9
    g.close()
