"""GPU parity tests (B200): every result goes through the C ABI of libmaniacgpu.so and is compared with the CPU
oracle on the same inputs.  FP64 tolerances: 1e-10 relative on energies (north star), 1e-12 where the sum order
is the only difference; accept/reject sequences must be identical."""
import os

import numpy as np
import pytest

from conftest import compare_records, ideal_mixture, load_system, rel_err
from maniac_mc_b200 import builders
from maniac_mc_b200.driver import StrictDriver, TapeRng
from maniac_mc_b200.lib import ManiacGPU, ManiacGpuError
from maniac_mc_b200.system import (E_INTRA, E_RECIP, E_TOTAL, MMC_CREATION, MMC_DELETION, MMC_ROTATION, N_COUNTERS,
                                   MMC_TRANSLATION)
from oracle import oracle as orc

pytestmark = pytest.mark.gpu
TOL = 1e-10   # BASELINE.json north_star: energies within 1e-10 relative in FP64


@pytest.mark.parametrize("name", ["lj_gas", "h2o_gas", "h2o_ads", "zif8_h2o", "mfi_co2", "slit", "zif8_ch4o_h2o"])
def test_system_energy(name):
    """ComputeSystemEnergy: configs[0] of BASELINE.json (the reference's own CPU-runnable cases)."""
    s = load_system(name)
    g = ManiacGPU(s); e = g.system_energy()
    ref = orc.Oracle(s).system_energy()
    assert rel_err(e, ref) < 1e-12
    g.close()


@pytest.mark.parametrize("name,init_sk", [("zif8_h2o", 1), ("zif8_h2o", 0), ("h2o_gas", 1), ("slit", 1), ("zif8_ch4o_h2o", 1), ("dipole_tri", 1)])
def test_system_energy_all_equals_per_chain(name, init_sk):
    """mmc_system_energy_all (framework S(k), framework pair and self terms once per handle + the guests of every chain in one
    launch set) against the independent per-chain evaluation and the oracle: energies 1e-12 (sum order only), A(k) 1e-10."""
    s = load_system(name, init_sk=init_sk)
    C = 3
    g = ManiacGPU(s, n_chains=C)
    # give chain 1 a different guest population so that the chains differ
    r = int(np.flatnonzero(s.is_active == 1)[0]); mo, ao = s.mol_offsets(); na = int(s.natoms[r])
    if int(s.count[r]) > 1:
        cnt = s.count.copy(); cnt[r] -= 1
        keep_m = np.ones(len(s.com), dtype=bool); keep_m[mo[r] + int(s.count[r]) - 1] = False
        keep_a = np.ones(len(s.offsets), dtype=bool); keep_a[ao[r] + (int(s.count[r]) - 1) * na: ao[r] + int(s.count[r]) * na] = False
        g.set_chain_state(1, cnt, s.com[keep_m], s.offsets[keep_a])
    e_all = g.system_energy_all()
    a_all = [g.sk(c) for c in range(C)]
    tracked = [g.sync_to_host(c)[3] for c in range(C)]
    for c in range(C):
        assert np.array_equal(tracked[c], e_all[c])
        e_one = g.system_energy(c)
        assert rel_err(e_all[c], e_one) < 1e-12
        np.testing.assert_allclose(a_all[c], g.sk(c), atol=1e-10)
    assert rel_err(e_all[0], orc.Oracle(s).system_energy()) < 1e-12
    if int(s.count[r]) > 1:
        assert not np.array_equal(e_all[0], e_all[1])
    e2 = g.system_energy_all()            # second call: cached framework terms
    assert np.array_equal(e2, e_all)
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


@pytest.mark.parametrize("name,init_sk,deletion", [("zif8_h2o", 0, 0), ("zif8_h2o", 1, 1), ("h2o_ads", 0, 0), ("h2o_ads", 1, 1),
                                                   ("mfi_co2", 1, 1), ("slit", 0, 0)])
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


@pytest.mark.parametrize("name,init_sk,deletion", [("zif8_h2o", 0, 0), ("zif8_h2o", 1, 1), ("mfi_co2", 1, 1), ("zif8_ch4o_h2o", 1, 1), ("dipole_tri", 1, 1)])
def test_strict_mode_fused_trial_step(name, init_sk, deletion):
    """mmc_trial_step: energies + acceptance probability + accept test (with the uniform the host drew) + commit in ONE launch,
    results through host-mapped memory.  Same tape through the oracle: identical accept sequence, energies 1e-10; and the
    state left on the device equals the oracle's."""
    s = load_system(name, init_sk=init_sk, deletion_mode=deletion)
    if name == "zif8_ch4o_h2o":
        s.p_translation, s.p_rotation, s.p_swap = 0.25, 0.25, 0.3
    tape = np.random.default_rng(43).random(6000)
    o = orc.Oracle(s); o.system_energy(); ref = o.run(400, tape=tape)
    g = ManiacGPU(s); drv = StrictDriver(s, g, fused=True); drv.ComputeSystemEnergy()
    l0 = g.launch_count()
    rng = TapeRng(tape); got = drv.MonteCarloLoop(rng, 1, 400)
    assert g.launch_count() - l0 == int((ref["kind"] > 0).sum())          # exactly one launch per attempted move
    compare_records(got, ref)
    assert rng.pos == o.tape_used
    cnt, com, off, e = g.sync_to_host(); c2, com2, off2 = o.state()
    assert np.array_equal(cnt, c2) and np.array_equal(drv.count, c2)
    np.testing.assert_allclose(com, com2, atol=1e-11); np.testing.assert_allclose(off, off2, atol=1e-12)
    assert rel_err(e, o.energy()) < TOL and rel_err(drv.energy, o.energy()) < TOL
    np.testing.assert_allclose(g.sk(), o.sk(), atol=1e-9)
    g.close()


@pytest.mark.parametrize("name,init_sk,deletion", [("zif8_h2o", 0, 0), ("h2o_ads", 1, 1), ("lj_gas", 0, 0), ("mfi_co2", 0, 0),
                                                   ("slit", 1, 1)])
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


@pytest.mark.parametrize("which,P", [("zif8_h2o", 2), ("zif8_h2o", 4), ("zif8_h2o", 8), ("zif8_ch4o_h2o", 4), ("tilted_framework", 4), ("water_box", 8)])
def test_cluster_per_chain_kernel(which, P, monkeypatch):
    """k_chain_run_cl: a thread-block cluster of P CTAs shares one chain (few chains must still fill the GPU).  Same Philox
    streams through (a) the cluster kernel, (b) the one-CTA kernel (MANIAC_CLUSTER=0) and (c) the oracle: identical accept
    sequences, records within 1e-12 of the one-CTA kernel (only the association of the P partial sums differs) and 1e-10 of
    the oracle; 300 steps = 2 graph replays of 128 + 44 direct; then a tape replay on the SAME cluster handle, which must
    take the one-CTA kernel and still continue the chain correctly."""
    if which == "zif8_h2o":
        s = load_system("zif8_h2o", init_sk=1, deletion_mode=1)
    elif which == "zif8_ch4o_h2o":
        s = load_system("zif8_ch4o_h2o", init_sk=1, deletion_mode=1); s.p_translation, s.p_rotation, s.p_swap = 0.25, 0.25, 0.3
    elif which == "water_box":
        s = water_box_500(); s.init_sk, s.deletion_mode = 1, 1
    else:
        s = tilted_framework(n_fw=3000, box=48.0, tilt=(7.0, -5.0, 9.0))
    n, seed, C = 300, 4242, 3
    monkeypatch.setenv("MANIAC_CLUSTER", str(P))
    g = ManiacGPU(s, n_chains=C)
    assert g.chain_launch_config() == (512, P)
    monkeypatch.setenv("MANIAC_CLUSTER", "0")
    g1 = ManiacGPU(s, n_chains=C)
    assert g1.chain_launch_config()[1] == 1
    g.system_energy_all(); g1.system_energy_all()
    rc, r1 = g.run_batch(n, seed=seed, records=True), g1.run_batch(n, seed=seed, records=True)
    ctr_c, ctr_1 = list(g.counters), list(g1.counters)
    oracles = []
    for c in range(C):
        compare_records(rc[c], r1[c], tol=1e-12)
        assert ctr_c[c] == ctr_1[c]
        o = orc.Oracle(s); o.system_energy()
        compare_records(rc[c], o.run(n, seed=seed, chain=c, step0=0))
        cnt, com, off, e = g.sync_to_host(c); c2, com2, off2 = o.state()
        assert np.array_equal(cnt, c2)
        np.testing.assert_allclose(com, com2, atol=1e-11); np.testing.assert_allclose(off, off2, atol=1e-12)
        assert rel_err(e, o.energy()) < TOL
        np.testing.assert_allclose(g.sk(c), o.sk(), atol=1e-9)
        oracles.append(o)
    # tape replay continues every chain on the cluster handle (one-CTA kernel): 140 steps = 1 chunk of 128 + 12
    tape = np.random.default_rng(3).random((C, 13 * 140))
    rt = g.run_batch(140, tape=tape, records=True)
    for c in range(C):
        compare_records(rt[c], oracles[c].run(140, tape=tape[c]))
        assert g.tape_used(c) == oracles[c].tape_used
    # ... and a Philox batch after it goes back to the cluster kernel with the step counter carried on
    r3 = g.run_batch(60, seed=seed, records=True)
    compare_records(r3[1], oracles[1].run(60, seed=seed, chain=1, step0=n + 140))
    assert (rc["accepted"] == 1).sum() > n // 4
    g.close(); g1.close()


@pytest.mark.parametrize("which,P", [("zif8_h2o", 0), ("zif8_ch4o_h2o", 0), ("tilted", 0), ("zif8_h2o", 4), ("zif8_ch4o_h2o", 2)])
def test_framework_energy_cache(which, P, monkeypatch):
    """ChainState.fw_pair: the batched kernels (one CTA per chain, P = 0, and a cluster of P CTAs per chain) keep, per guest molecule, the framework part of its pair energy and
    re-sums only the guest-guest part of the old geometry (the framework never moves).  The same Philox streams with the
    cache (default) and without it (MANIAC_FW_CACHE=0, the reference's re-summation, ComputeOldEnergy
    monte_carlo_utils.f90:347-396): identical accept sequences and counters, records within 1e-12 of each other (the split
    changes only the association of the sums) and 1e-10 of the oracle.  Then every path that changes a geometry behind the
    cache's back -- a strict-mode move (mmc_trial_step), a re-upload of one chain, mmc_upload_chains -- followed by another
    batch: a stale entry would show up as an energy difference between the two handles and against a fresh total."""
    if which == "zif8_h2o":
        s = load_system("zif8_h2o", init_sk=1, deletion_mode=1)
    elif which == "zif8_ch4o_h2o":
        s = load_system("zif8_ch4o_h2o", init_sk=1, deletion_mode=1); s.p_translation, s.p_rotation, s.p_swap = 0.25, 0.25, 0.3
    else:
        s = tilted_framework(n_fw=3000, box=48.0, tilt=(7.0, -5.0, 9.0))
    n, seed, C = 400, 777, 3
    monkeypatch.setenv("MANIAC_CLUSTER", str(P))   # both kernels, whatever the library would pick for so few chains
    g = ManiacGPU(s, n_chains=C)
    assert g.chain_launch_config()[1] == max(P, 1)
    monkeypatch.setenv("MANIAC_FW_CACHE", "0")
    g0 = ManiacGPU(s, n_chains=C)
    monkeypatch.delenv("MANIAC_FW_CACHE")
    handles = (g, g0)
    for h in handles:
        h.system_energy_all()
    step0 = 0

    def batch(k):
        nonlocal step0
        ra, rb = g.run_batch(k, seed=seed, records=True), g0.run_batch(k, seed=seed, records=True)
        for c in range(C):
            compare_records(ra[c], rb[c], tol=1e-12)
        assert list(g.counters) == list(g0.counters)
        step0 += k
        return ra

    ra = batch(n)
    assert (ra["accepted"] == 1).sum() > n // 4 and set(np.unique(ra["kind"])) >= {1, 2, 3, 4}
    o = orc.Oracle(s); o.system_energy()
    compare_records(ra[1], o.run(n, seed=seed, chain=1, step0=0))
    # (a) strict-mode moves on chain 0 of both handles: translate molecule 0 of the first populated active residue, always accepted (u = 0)
    for k in range(3):
        for h in handles:
            cnt, com, off, e = h.sync_to_host(0)
            r = int(np.flatnonzero((s.is_active == 1) & (cnt > 0))[0])
            mo, ao, na = int(cnt[:r].sum()), int((cnt[:r] * s.natoms[:r]).sum()), int(s.natoms[r])
            old, new, prob, acc = h.trial_step(MMC_TRANSLATION, r, 0, com_new=com[mo] + np.array([0.31, -0.22, 0.17]) * (k + 1),
                                               offs_new=off[ao:ao + na], u=0.0, chain=0)
            assert acc or prob == 0.0
    batch(150)
    # (b) chain 1's state uploaded into chain 2 (mmc_set_chain_state), totals recomputed
    for h in handles:
        cnt, com, off, e = h.sync_to_host(1)
        h.set_chain_state(2, cnt, com, off); h.system_energy(2)
    batch(150)
    # (c) every chain re-uploaded at once (mmc_upload_chains).  First WHAT WAS DOWNLOADED: the geometry is bit for bit the
    # device's, so every cache entry survives (this is bench.py's end-to-end step); then the chains rotated by one: every
    # entry of a populated slot whose geometry changed is dropped
    gm, ga = g.chain_capacity()
    known = [g.fw_cache(c) for c in range(C)]
    assert sum(int(np.isfinite(k[:, 0]).sum()) for k in known) > 3 and not np.isfinite(g0.fw_cache(0)).any()
    for h in handles:
        cnt = np.zeros((C, s.nres), dtype=np.int32); com = np.zeros((C, gm, 3)); off = np.zeros((C, ga, 3))
        h.download_chains(cnt, com, off, None, None)
        h.upload_chains(cnt, com, off)
    for c in range(C):
        assert np.array_equal(g.fw_cache(c), known[c], equal_nan=True)
    batch(100)
    for h in handles:
        h.download_chains(cnt, com, off, None, None)
        h.upload_chains(np.roll(cnt, 1, axis=0), np.roll(com, 1, axis=0), np.roll(off, 1, axis=0)); h.system_energy_all()
    moved = np.any(np.roll(com, 1, axis=0) != com, axis=2)                       # [C, g_mols]: slot geometry changed by the rotation
    for c in range(C):
        assert not np.isfinite(g.fw_cache(c)[moved[c], 0]).any()
    batch(150)
    for c in range(C):
        e_tracked = g.sync_to_host(c)[3]
        assert rel_err(e_tracked, g0.sync_to_host(c)[3]) < 1e-12
        assert rel_err(e_tracked, g.system_energy(c)) < TOL
    g.close(); g0.close()


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


# ---- BASELINE configs[1]: rigid-water GCMC insertion/deletion in an otherwise empty periodic box ----------------
def water_box_500():
    s = builders.synthetic_water_box(n_mol=500, box=30.0, cutoff=14.0, tol=1e-5, seed=1, capacity=1024,
                                     fugacity_atm=0.05, p_translation=0.25, p_rotation=0.25)
    assert s.num_kvectors == 518          # alpha = 0.199263, kmax = 6
    return s


@pytest.mark.parametrize("init_sk,deletion", [(0, 0), (1, 1)])
def test_c2_water_box_strict_mode(init_sk, deletion):
    """~500 TIP4P molecules, no framework: the guest-guest part of the real-space sweep carries all the work.
    Strict mode (host proposals from the tape) against the oracle move for move, both quirk settings."""
    s = water_box_500(); s.init_sk, s.deletion_mode = init_sk, deletion
    tape = np.random.default_rng(11).random(4000)
    o = orc.Oracle(s); o.system_energy(); ref = o.run(300, tape=tape)
    g = ManiacGPU(s); drv = StrictDriver(s, g); drv.ComputeSystemEnergy()
    rng = TapeRng(tape); got = drv.MonteCarloLoop(rng, 1, 300)
    compare_records(got, ref)
    assert rng.pos == o.tape_used
    assert set(np.unique(ref["kind"])) >= {MMC_CREATION, MMC_DELETION, MMC_TRANSLATION, MMC_ROTATION}
    cnt, com, off, e = g.sync_to_host(); c2, com2, off2 = o.state()
    assert np.array_equal(cnt, c2)
    np.testing.assert_allclose(com, com2, atol=1e-12); np.testing.assert_allclose(off, off2, atol=1e-12)
    assert rel_err(e, o.energy()) < TOL
    g.close()


def test_c2_water_box_batched_and_identities():
    """Batched mode on the same box: on-device proposals from recorded streams == oracle; afterwards the tracked
    energies equal a fresh single-point evaluation and tracked A(k) equals the freshly summed S(k)."""
    s = water_box_500(); s.init_sk, s.deletion_mode = 1, 1
    n = 250
    tape = np.random.default_rng(12).random((2, 12 * n))
    g = ManiacGPU(s, n_chains=2)
    for c in range(2):
        g.system_energy(c)
    recs = g.run_batch(n, tape=tape, records=True)
    for c in range(2):
        o = orc.Oracle(s); o.system_energy()
        compare_records(recs[c], o.run(n, tape=tape[c]))
        assert g.counters[c] == o.counters()
    g.run_batch(2000, seed=3)
    for c in range(2):
        _, _, _, tracked = g.sync_to_host(c)
        a_tracked = g.sk(c)
        fresh = g.system_energy(c)          # recomputes S(k) from the coordinates
        assert rel_err(tracked, fresh) < 1e-9
        np.testing.assert_allclose(a_tracked, g.sk(c), atol=1e-8)
    g.close()


def test_culled_sweep_equals_every_pair_sweep():
    """coulomb_xcut < 0 evaluates every minimum-image pair literally as the reference loop does (Q3); the default
    tile-culled sweep (alpha r < 6) must agree to rounding on a framework large enough for the cull to bite."""
    s = builders.synthetic_framework(n_framework=20000, n_guests=48, box=72.0, cutoff=12.0, tol=1e-6, seed=9, capacity=96)
    s.init_sk, s.deletion_mode = 1, 1
    full = s.copy(); full.coulomb_xcut = -1.0
    tape = np.random.default_rng(5).random((1, 600))
    out = []
    for sys_ in (s, full):
        g = ManiacGPU(sys_); g.system_energy()
        out.append(g.run_batch(40, tape=tape, records=True)[0])
        work = g.pair_work()
        out.append(work)
        g.close()
    compare_records(out[0], out[2], tol=1e-12)
    assert out[1][0] < 0.5 * out[3][0]        # the culled sweep formed far fewer distances
    o = orc.Oracle(s); o.system_energy()
    compare_records(out[2], o.run(40, tape=tape[0]))


def test_c4_large_k_full_size_identities():
    """BASELINE configs[3] at bench size (50 000-atom framework, r_c = 12, tol 1e-6 -> K = 37 782): too big for the
    oracle to step through in seconds, so parity goes through size-independent identities: tracked energies == fresh
    single-point energies after a batch, tracked A(k) == fresh S(k), insertion followed by deletion of the same
    molecule restores A(k), and a rejected move leaves every component untouched."""
    s = builders.synthetic_framework(n_framework=50000, n_guests=200, box=96.0, cutoff=12.0, tol=1e-6, seed=4, capacity=512,
                                     p_translation=0.4, p_rotation=0.4)
    s.init_sk, s.deletion_mode = 1, 1
    assert s.num_kvectors == 37782
    g = ManiacGPU(s, n_chains=3)
    for c in range(3):
        g.system_energy(c)
    a0 = g.sk(0); e0 = g.sync_to_host(0)[3]
    # insertion then deletion of the same molecule (strict entry points), chain 0
    n0 = int(s.count[0])
    com = np.array([11.0, 47.0, 80.5]); off = s.offsets[:4].copy()
    old, new = g.trial_energy(MMC_CREATION, 0, n0, com, off, chain=0); g.trial_commit(True, chain=0)
    old2, new2 = g.trial_energy(MMC_DELETION, 0, n0, chain=0); g.trial_commit(True, chain=0)
    np.testing.assert_allclose(g.sk(0), a0, atol=1e-10)
    assert rel_err(new[E_TOTAL], old2[E_TOTAL]) < 1e-11 and rel_err(new2[E_RECIP], old[E_RECIP]) < 1e-11
    assert rel_err(g.sync_to_host(0)[3], e0) < 1e-11
    # a rejected translation changes nothing
    a1 = g.sk(0); e1 = g.sync_to_host(0)[3]
    g.trial_energy(MMC_TRANSLATION, 0, 3, s.com[3] + 0.3, s.offsets[12:16], chain=0); g.trial_commit(False, chain=0)
    np.testing.assert_array_equal(g.sk(0), a1)
    np.testing.assert_array_equal(g.sync_to_host(0)[3], e1)
    # batched moves on all chains, then tracked == fresh
    g.run_batch(300, seed=21)
    for c in range(3):
        _, _, _, tracked = g.sync_to_host(c)
        a_tracked = g.sk(c)
        fresh = g.system_energy(c)
        assert rel_err(tracked, fresh) < 1e-9
        np.testing.assert_allclose(a_tracked, g.sk(c), atol=1e-7)
    g.close()


# ---- k-sharded mode (SURVEY 8e optional mode): two emulated ranks on one GPU, the all-reduce replaced by a sum -----
@pytest.mark.parametrize("which", ["zif8_h2o", "tilted_framework"])
def test_k_sharded_steps_match_oracle_and_unsharded(which):
    from maniac_mc_b200.sharded import emulate_ranks
    s = load_system("zif8_h2o", init_sk=1, deletion_mode=1) if which == "zif8_h2o" else tilted_framework(n_fw=3000, box=48.0, tilt=(7.0, -5.0, 9.0))
    n, seed, C = 80, 4321, 2
    shards, step, _, _ = emulate_ranks(s, world=2, n_chains=C)
    recs = np.stack([step(seed, records=True) for _ in range(n)], axis=1)
    g = ManiacGPU(s, n_chains=C)
    for c in range(C):
        g.system_energy(c)
    whole = g.run_batch(n, seed=seed, records=True)
    for c in range(C):
        o = orc.Oracle(s); o.system_energy()
        compare_records(recs[c], o.run(n, seed=seed, chain=c, step0=0))
        compare_records(recs[c], whole[c], tol=1e-12)
        # the shards' owned k ranges tile the k list and their A(k) equal the unsharded A(k)
        a_ref = g.sk(c); covered = 0
        for sh in shards:
            k0, a = sh.owned_sk(c)
            assert k0 == covered
            np.testing.assert_allclose(a, a_ref[k0:k0 + len(a)], atol=1e-10)
            covered += len(a)
        assert covered == g.nk
        for sh in shards:                      # every rank holds the same coordinates and tracked energies
            cnt, com, off, e = sh.gpu.sync_to_host(c); c2, com2, off2, e2 = g.sync_to_host(c)
            assert np.array_equal(cnt, c2)
            np.testing.assert_allclose(com, com2, atol=1e-12); np.testing.assert_allclose(off, off2, atol=1e-12)
            assert rel_err(e, e2) < 1e-11
    for sh in shards:
        sh.close()
    g.close()


def test_k_sharded_strict_trial_and_errors():
    from maniac_mc_b200.sharded import emulate_ranks
    s = load_system("zif8_h2o", init_sk=1, deletion_mode=1)
    shards, _, trial_energy, trial_commit = emulate_ranks(s, world=3)
    g = ManiacGPU(s); g.system_energy()
    n0 = int(s.count[0]); mo, ao = s.mol_offsets()
    com = s.com[mo[0]] + np.array([0.4, -0.2, 0.1]); off = s.offsets[ao[0]:ao[0] + int(s.natoms[0])]
    for kind, mol, cn, of, accept in ((MMC_TRANSLATION, 0, com, off, True), (MMC_CREATION, n0, com + 3.0, off, True),
                                      (MMC_DELETION, 1, None, None, False), (MMC_DELETION, n0, None, None, True)):
        o1, n1 = trial_energy(kind, 0, mol, cn, of); o2, n2 = g.trial_energy(kind, 0, mol, cn, of)
        assert rel_err(o1, o2) < 1e-12 and rel_err(n1, n2) < 1e-12
        trial_commit(accept); g.trial_commit(accept)
    with pytest.raises(ManiacGpuError, match="shard"):
        shards[0].gpu.trial_energy(MMC_DELETION, 0, 0)
    with pytest.raises(ManiacGpuError, match="shard"):
        shards[1].gpu.run_batch(4, seed=1)
    for sh in shards:
        sh.close()
    g.close()


# ---- swap move on the reference's two-species example ZIF8-CH4O-H2O ---------------------------------
def two_species(**over):
    s = load_system("zif8_ch4o_h2o", init_sk=1, deletion_mode=1, **over)
    s.p_translation, s.p_rotation, s.p_swap = 0.25, 0.25, 0.3
    return s


def test_swap_strict_mode_trajectory():
    s = two_species()
    tape = np.random.default_rng(8).random(6000)
    o = orc.Oracle(s); o.system_energy(); ref = o.run(400, tape=tape)
    g = ManiacGPU(s); drv = StrictDriver(s, g); drv.ComputeSystemEnergy()
    rng = TapeRng(tape); got = drv.MonteCarloLoop(rng, 1, 400)
    compare_records(got, ref)
    assert rng.pos == o.tape_used
    sw = ref[ref["kind"] == 5]
    assert len(sw) > 40 and 0 < sw["accepted"].sum() < len(sw)
    cnt, com, off, e = g.sync_to_host(); c2, com2, off2 = o.state()
    assert np.array_equal(cnt, c2) and np.array_equal(drv.count, c2)
    np.testing.assert_allclose(com, com2, atol=1e-12); np.testing.assert_allclose(off, off2, atol=1e-12)
    assert rel_err(e, o.energy()) < TOL and rel_err(drv.energy, o.energy()) < TOL
    np.testing.assert_allclose(g.sk(), o.sk(), atol=1e-9)
    assert {k: drv.counter[k] for k in ("swaps", "trial_swaps")} == {k: o.counters()[k] for k in ("swaps", "trial_swaps")}
    g.close()


def test_swap_batched_modes():
    """on-device proposals: recorded streams and Philox, then tracked == fresh; two emulated k-shards agree too"""
    from maniac_mc_b200.sharded import emulate_ranks
    s = two_species()
    n = 300
    tape = np.random.default_rng(9).random((2, 12 * n))
    g = ManiacGPU(s, n_chains=2)
    for c in range(2):
        g.system_energy(c)
    recs = g.run_batch(n, tape=tape, records=True)
    for c in range(2):
        o = orc.Oracle(s); o.system_energy()
        compare_records(recs[c], o.run(n, tape=tape[c]))
        assert g.tape_used(c) == o.tape_used and g.counters[c] == o.counters()
        assert (recs[c]["kind"] == 5).sum() > 30
        cnt, com, off, e = g.sync_to_host(c); c2, com2, off2 = o.state()
        assert np.array_equal(cnt, c2)
        np.testing.assert_allclose(com, com2, atol=1e-12); np.testing.assert_allclose(off, off2, atol=1e-12)
        np.testing.assert_allclose(g.sk(c), o.sk(), atol=1e-9)
    r2 = g.run_batch(200, seed=77, records=True)
    shards, step, _, _ = emulate_ranks(s, world=2, n_chains=2)
    for c in range(2):
        _, _, _, tracked = g.sync_to_host(c)
        assert rel_err(tracked, g.system_energy(c)) < 1e-9
    # sharded steps from the initial state with the same Philox seed == oracle
    rs = np.stack([step(77, records=True) for _ in range(120)], axis=1)
    for c in range(2):
        o = orc.Oracle(s); o.system_energy()
        compare_records(rs[c], o.run(120, seed=77, chain=c, step0=0))
    assert (rs["kind"] == 5).sum() > 20
    for sh in shards:
        sh.close()
    g.close()


def test_swap_detailed_balance_ideal_mixture():
    """SURVEY 8f.1 asks for a detailed-balance check of the swap move (there is no reference trajectory): swap moves only
    in a non-interacting two-species mixture keep N_A + N_B and make N_A binomial(N, phi_A / (phi_A + phi_B)) on every
    chain of the batched loop.  Same system and expectation as tests/test_oracle_golden.py's CPU check."""
    s, act, pA = ideal_mixture(2.0)
    n_tot = int(s.count[act].sum())
    C = 64
    g = ManiacGPU(s, n_chains=C)
    for c in range(C):
        g.set_chain_id(c, 500 + c); g.system_energy(c)
    g.run_batch(400, seed=23)
    cnt = np.zeros((C, s.nres), dtype=np.int32)
    samples = []
    for b in range(150):
        g.run_batch_async(20, seed=23)
        g.download_chains(cnt, None, None, None, None)
        assert np.all(cnt[:, act].sum(axis=1) == n_tot)
        samples.append(cnt[:, act[0]].copy())
    n = np.array(samples, dtype=np.float64)
    mean, var = n.mean(), n.var()
    assert abs(mean - n_tot * pA) < 0.02 * n_tot * pA, (mean, n_tot * pA)          # 9600 samples, ~1 per 20 moves independent
    assert abs(var - n_tot * pA * (1 - pA)) < 0.1 * n_tot * pA * (1 - pA), (var, n_tot * pA * (1 - pA))
    e = np.zeros((C, 6)); g.download_chains(None, None, None, e, None)
    assert np.all(np.abs(e[:, :3]) < 1e-9)
    g.close()


def test_swap_error_behaviour():
    s = two_species()
    g = ManiacGPU(s); g.system_energy()
    com = np.zeros(3); off6 = np.zeros((6, 3)); off4 = np.zeros((4, 3))
    with pytest.raises(ManiacGpuError, match="mmc_trial_energy_swap"):
        g.trial_energy(5, 1, 0, com, off4)                      # a swap needs both residue types
    with pytest.raises(ManiacGpuError, match="different active"):
        g.trial_energy_swap(1, 0, 1, com, off6)
    with pytest.raises(ManiacGpuError, match="different active"):
        g.trial_energy_swap(1, 0, 0, com, off6)                  # the framework is not an active type
    with pytest.raises(ManiacGpuError, match="out of range"):
        g.trial_energy_swap(1, 99, 2, com, off4)
    g.close()


def test_isotherm_sweep_single_rank():
    """configs[4] host logic on one GPU: 8 fugacities x 2 replicas of ZIF-8 + H2O as 16 chains; the gathered table is
    complete, the block averages follow the device counters, and loading does not decrease along the fugacity axis at the ends."""
    from maniac_mc_b200 import isotherm as iso
    s = load_system("zif8_h2o", init_sk=1, deletion_mode=1)
    sw = iso.IsothermSweep(s, n_points=8, replicas=2, lo_atm=1e-2, hi_atm=1e6)
    for b in range(6):
        sw.run_block(400, seed=5)
    t = sw.table()
    assert t.shape == (2, 8, 4 + 10) and not np.isnan(t).any()
    assert np.all(t[:, :, 3] == 6)
    trials = t[:, :, 8:12].sum(axis=2) + t[:, :, 13]
    assert np.all(trials <= 6 * 400) and np.all(trials > 0)                 # steps that select a molecule in an empty box are no-ops
    curve = iso.isotherm(t)
    assert curve[-1, 0] > curve[0, 0]                                       # more water at 1e6 atm than at 1e-2 atm
    for c, cid in enumerate(sw.ids):                                         # tracked == fresh on every chain
        _, _, _, tracked = sw.gpu.sync_to_host(c)
        assert rel_err(tracked, sw.gpu.system_energy(c)) < 1e-9
    sw.close()


def test_sweep_is_independent_of_the_dealing():
    """Chains carry GLOBAL Philox ids (mmc_set_chain_id): the 6 chains of a sweep give bit-identical records whether one
    handle runs all of them or two handles run the interleaved halves a 2-GPU job would deal out; and they equal the
    oracle run with the same (seed, chain id)."""
    from maniac_mc_b200 import isotherm as iso
    s = load_system("zif8_h2o", init_sk=1, deletion_mode=1)
    n, seed = 150, 99

    def run(world, rank):
        ids = iso.deal_chains(3, 2, world, rank)
        g = ManiacGPU(s, n_chains=len(ids))
        for c, cid in enumerate(ids):
            g.set_chain_id(c, cid.seed); g.system_energy(c)
        recs = g.run_batch(n, seed=seed, records=True)
        g.close()
        return {cid.index: recs[c] for c, cid in enumerate(ids)}

    whole = run(1, 0)
    halves = {**run(2, 0), **run(2, 1)}
    assert sorted(halves) == sorted(whole) == list(range(6))
    for gidx in whole:
        assert whole[gidx].tobytes() == halves[gidx].tobytes()
    cid = iso.deal_chains(3, 2)[4]
    o = orc.Oracle(s); o.system_energy()
    compare_records(whole[4], o.run(n, seed=seed, chain=cid.seed, step0=0))


def test_ideal_gas_limit_is_poisson():
    """Physics check of the whole batched GCMC loop (selection, Philox stream, insertion/deletion acceptance with the
    reference's N-after-update convention, monte_carlo_utils.f90:199-214): with every interaction switched off the
    stationary distribution of N is Poisson with mean = variance = phi V."""
    s = builders.synthetic_water_box(n_mol=8, box=30.0, cutoff=14.0, tol=1e-5, seed=2, capacity=256, fugacity_atm=20.0,
                                     p_translation=0.1, p_rotation=0.1)
    s.charges[:] = 0.0; s.epsilon[:] = 0.0
    s.init_sk, s.deletion_mode = 1, 1
    expect = float(s.fugacity[0]) * s.volume
    assert 12.0 < expect < 15.0
    C = 64
    g = ManiacGPU(s, n_chains=C)
    for c in range(C):
        g.set_chain_id(c, 1000 + c); g.system_energy(c)
    g.run_batch(3000, seed=17)                                   # equilibration
    cnt = np.zeros((C, s.nres), dtype=np.int32)
    samples = []
    for b in range(120):
        g.run_batch_async(100, seed=17)
        g.download_chains(cnt, None, None, None, None)
        samples.append(cnt[:, 0].copy())
    n = np.array(samples, dtype=np.float64)                      # [blocks, chains]
    mean, var = n.mean(), n.var()
    assert abs(mean - expect) < 0.03 * expect, (mean, expect)    # ~7700 weakly correlated samples: s.e. ~ 0.05
    assert abs(var - expect) < 0.12 * expect, (var, expect)
    e = np.zeros((C, 6)); g.download_chains(None, None, None, e, None)
    assert np.all(np.abs(e[:, :3]) < 1e-9)                       # no interaction energy anywhere
    g.close()


# ---- reservoir box: insertion geometry from a pool, pool bookkeeping on accept ----------------------
def test_reservoir_strict_and_batched():
    from test_oracle_golden import with_reservoir
    s = with_reservoir()
    n = 400
    tape = np.random.default_rng(21).random(8000)
    o = orc.Oracle(s); o.system_energy(); ref = o.run(n, tape=tape)
    # strict mode: the host owns the pool (the library is told nothing about it)
    bare = s.copy(); bare.reservoir_count = None; bare.reservoir_offsets = None
    g = ManiacGPU(bare); drv = StrictDriver(s, g); drv.ComputeSystemEnergy()
    rng = TapeRng(tape); got = drv.MonteCarloLoop(rng, 1, n)
    compare_records(got, ref)
    assert rng.pos == o.tape_used
    assert len(drv.rsv[0]) == o.reservoir()[0][0]
    g.close()
    # batched mode: pools on the device, one per chain
    tapes = np.stack([tape, np.random.default_rng(22).random(8000)])
    g = ManiacGPU(s, n_chains=2)
    for c in range(2):
        g.system_energy(c)
    recs = g.run_batch(n, tape=tapes, records=True)
    for c in range(2):
        oc = orc.Oracle(s); oc.system_energy(); rc = oc.run(n, tape=tapes[c])
        compare_records(recs[c], rc)
        assert g.tape_used(c) == oc.tape_used and g.counters[c] == oc.counters()
        cnt_g, off_g = g.reservoir(c); cnt_o, off_o = oc.reservoir()
        assert np.array_equal(cnt_g, cnt_o)
        np.testing.assert_allclose(off_g, off_o, atol=1e-12)
        cnt, com, off, e = g.sync_to_host(c); c2, com2, off2 = oc.state()
        assert np.array_equal(cnt, c2)
        np.testing.assert_allclose(off, off2, atol=1e-12)
        assert rel_err(e, oc.energy()) < TOL
    r2 = g.run_batch(300, seed=5, records=True)       # Philox stream
    oc = orc.Oracle(s); oc.system_energy(); oc.run(n, tape=tapes[1])
    compare_records(r2[1], oc.run(300, seed=5, chain=1, step0=n))
    g.close()


def test_adjust_step_sizes_per_block():
    """MonteCarloLoop's per-block AdjustMoveStepSizes (monte_carlo_utils.f90:99-129) in batched mode: blocks of 700 steps with
    recalibration after each == the oracle doing the same; the step sizes follow the reference's literal rule."""
    s = load_system("zif8_h2o", init_sk=1, deletion_mode=1, p_translation=0.45, p_rotation=0.45)
    g = ManiacGPU(s, n_chains=2)
    for c in range(2):
        g.system_energy(c)
    o = orc.Oracle(s); o.system_energy(); o.set_recalibrate(1)
    for b in range(4):
        recs = g.run_batch(700, seed=31, records=True)
        g.adjust_step_sizes()
        ref = o.run(700, seed=31, chain=1, step0=700 * b); o.adjust_step_sizes()
        compare_records(recs[1], ref)
        t, r = g.step_sizes(1); t2, r2 = o.step_sizes()
        assert t == t2 and r == r2
    assert (t, r) != (s.translation_step, s.rotation_step)
    g.close()


def test_snapshot_is_the_state_at_its_place_in_the_stream(tmp_path):
    """mmc_snapshot_begin/_wait return exactly what mmc_download_chains would have returned at that point of
    the stream, although two more batches were queued behind the snapshot before the host waited for it; both slots
    can be in flight; and BlockPipeline writes the reference's per-block files from them, one directory per chain."""
    from maniac_mc_b200 import writers as W
    s = load_system("zif8_h2o", init_sk=1, deletion_mode=1)
    C = 6
    g = ManiacGPU(s, n_chains=C)
    for c in range(C):
        g.system_energy(c)
    g_mols, g_atoms = g.chain_capacity()

    def download():
        cnt = np.zeros((C, s.nres), dtype=np.int32); com = np.zeros((C, g_mols, 3)); off = np.zeros((C, g_atoms, 3))
        e = np.zeros((C, 6)); ctr = np.zeros((C, N_COUNTERS), dtype=np.int64)
        g.download_chains(cnt, com, off, e, ctr)
        return cnt, com, off, e, ctr

    # reference sequence with synchronous downloads
    g.run_batch_async(300, seed=3); a = download()
    g.run_batch_async(300, seed=3); b = download()
    g.run_batch_async(300, seed=3); g.synchronize()
    g.close()
    # same sequence, snapshots queued and read late
    g = ManiacGPU(s, n_chains=C)
    for c in range(C):
        g.system_energy(c)
    g.run_batch_async(300, seed=3); g.snapshot_begin(0)
    g.run_batch_async(300, seed=3); g.snapshot_begin(1)
    g.run_batch_async(300, seed=3)
    sa, sb = g.snapshot_wait(0), g.snapshot_wait(1)
    for want, got in ((a, sa), (b, sb)):
        for w, x in zip(want, got):
            assert np.array_equal(w, x)
    assert not np.array_equal(sa[4], sb[4])                         # the counters moved between the two
    with pytest.raises(RuntimeError):
        g.snapshot_begin(2)
    g.close()
    g = ManiacGPU(s, n_chains=C)
    with pytest.raises(RuntimeError):
        g.snapshot_wait(0)                                          # nothing pending
    for c in range(C):
        g.system_energy(c)
    ws = [W.ChainWriter(W.chain_directory(str(tmp_path), c, 0), s, nb_block=3) if c != 4 else None for c in range(C)]
    pipe = W.BlockPipeline(g, ws)
    pipe.write_initial()
    last = pipe.run(3, 300, seed=3)
    g.synchronize()
    assert np.array_equal(last[0], download()[0]) and np.array_equal(last[3], download()[3])
    assert sorted(os.listdir(tmp_path)) == ["p000_r00", "p001_r00", "p002_r00", "p003_r00", "p005_r00"]
    d = tmp_path / "p002_r00"
    en = (d / "energy.dat").read_text().splitlines()
    assert en[0] == W.ENERGY_HEADER and [l.split()[0] for l in en[1:]] == ["0", "1", "2", "3"]
    assert en[-1] == W.energy_row(3, last[3][2])
    mv = (d / "moves.dat").read_text().splitlines()
    assert mv[-1] == W.moves_row(3, last[4][2]) and int(mv[-1].split()[1]) > 0
    tr = (d / "trajectory.lammpstrj").read_text().splitlines()
    assert tr.count("ITEM: TIMESTEP") == 4
    n_last = int((last[0][2] * s.natoms).sum())
    assert int(tr[-(n_last + 9) + 3]) == n_last                     # NUMBER OF ATOMS of the last frame
    g.close()


def test_isotherm_run_blocks_equals_blockwise_sync_and_writes_per_chain_files(tmp_path):
    """The pipelined block loop (snapshots read one block late, files written under the next block) gives the same
    averages table as the synchronous loop, and one directory of reference-format files per (point, replica)."""
    from maniac_mc_b200 import isotherm as iso
    from maniac_mc_b200 import writers as W
    s = load_system("zif8_h2o", init_sk=1, deletion_mode=1)
    a = iso.IsothermSweep(s, n_points=4, replicas=2, lo_atm=1.0, hi_atm=1e5)
    for b in range(4):
        a.run_block(250, seed=11)
    ta = a.table(); a.close()
    b = iso.IsothermSweep(s, n_points=4, replicas=2, lo_atm=1.0, hi_atm=1e5)
    b.run_blocks(4, 250, seed=11, output_root=str(tmp_path), trajectory=True, literal=False)
    tb = b.table(); b.close()
    assert np.array_equal(ta, tb)
    assert sorted(os.listdir(tmp_path)) == [f"p{p:03d}_r{r:02d}" for p in range(4) for r in range(2)]
    d = tmp_path / "p003_r01"
    rows = (d / "energy.dat").read_text().splitlines()
    assert len(rows) == 6 and rows[0] == W.ENERGY_HEADER
    tr = (d / "trajectory.lammpstrj").read_text().splitlines()
    assert [tr[i + 1].strip() for i, l in enumerate(tr) if l == "ITEM: TIMESTEP"] == ["0", "1", "2", "3", "4"]
    num = (d / f"number_res{b.res + 1}.dat").read_text().splitlines()
    if len(num) == 6:                                              # a row per block: the chain was never empty
        assert abs(np.mean([int(l.split()[1]) for l in num[2:]]) - tb[1, 3, 0]) < 1e-12   # the rows ARE the table's samples


# ---- triclinic cells (the CPU restatement is pinned physically in tests/test_triclinic_oracle.py) ----
def tilted_dipoles(**over):
    from test_triclinic_oracle import dipole_cell
    s = dipole_cell((3.0, -2.0, 4.5), kmax=(8, 8, 8), alpha=0.42, n_mol=10)
    for k, v in over.items():
        setattr(s, k, v)
    return s


def tilted_framework(n_fw=6000, box=60.0, tilt=(9.0, -7.0, 12.0)):
    """a framework big enough for the tile cull and the single-shift path (reach + molecule < L/2) in a tilted cell"""
    s = builders.synthetic_framework(n_framework=n_fw, n_guests=40, box=box, cutoff=12.0, tol=1e-5, capacity=128)
    s.tilt = np.array(tilt); s.init_sk = 1; s.deletion_mode = 1
    return s


def reference_dipoles():
    """the reference's own tilted example (testcase-energy/DIPOLE-orthorhombic, tilt -20 -20 20)"""
    return load_system("dipole_tri", init_sk=1, deletion_mode=1)


def test_triclinic_single_point_energies():
    for s in (tilted_dipoles(), tilted_framework(), reference_dipoles()):
        g = ManiacGPU(s); e = g.system_energy()
        assert rel_err(e, orc.Oracle(s).system_energy()) < 1e-11
        kv_g, kv_o = g.kvectors(), orc.Oracle(s).kvectors()
        for a, b in zip(kv_g, kv_o):
            np.testing.assert_allclose(a, b, rtol=1e-13)
        g.close()


@pytest.mark.parametrize("which", ["dipoles", "framework", "reference"])
def test_triclinic_strict_mode_trajectory(which):
    s = tilted_dipoles() if which == "dipoles" else tilted_framework() if which == "framework" else reference_dipoles()
    n = 150 if which == "framework" else 400
    tape = np.random.default_rng(8).random(16 * n)
    o = orc.Oracle(s); o.system_energy(); ref = o.run(n, tape=tape)
    g = ManiacGPU(s); drv = StrictDriver(s, g); drv.ComputeSystemEnergy()
    rng = TapeRng(tape); got = drv.MonteCarloLoop(rng, 1, n)
    compare_records(got, ref)
    assert rng.pos == o.tape_used and ref["accepted"].sum() > n // 10
    cnt, com, off, e = g.sync_to_host()
    c2, com2, off2 = o.state()
    assert np.array_equal(cnt, c2)
    np.testing.assert_allclose(com, com2, atol=1e-11); np.testing.assert_allclose(off, off2, atol=1e-12)
    assert rel_err(e, o.energy()) < TOL and rel_err(e, g.system_energy()) < 1e-9
    np.testing.assert_allclose(g.sk(), o.sk(), atol=1e-9)
    g.close()


@pytest.mark.parametrize("which", ["dipoles", "framework", "reference"])
def test_triclinic_batched_matches_oracle(which):
    """device proposals (fractional insertion, fractional ApplyPBC) + fused chain kernel in a tilted cell"""
    s = tilted_dipoles() if which == "dipoles" else tilted_framework() if which == "framework" else reference_dipoles()
    n = 120 if which == "framework" else 300
    g = ManiacGPU(s, n_chains=3)
    for c in range(3):
        g.system_energy(c)
    recs = g.run_batch(n, seed=17, records=True)
    for c in range(3):
        o = orc.Oracle(s); o.system_energy()
        compare_records(recs[c], o.run(n, seed=17, chain=c))
        _, _, _, tracked = g.sync_to_host(c)
        assert rel_err(tracked, o.energy()) < TOL
    g.close()


def test_triclinic_cull_equals_every_pair():
    """the tile cull + single-shift path in a tilted cell against the every-pair sweep of the same library"""
    s = tilted_framework()
    a = ManiacGPU(s, n_chains=2); s2 = s.copy(); s2.coulomb_xcut = -1.0; b = ManiacGPU(s2, n_chains=2)
    for c in range(2):
        a.system_energy(c); b.system_energy(c)
    ra, rb = a.run_batch(150, seed=4, records=True), b.run_batch(150, seed=4, records=True)
    for c in range(2):
        compare_records(ra[c], rb[c], tol=1e-12)
    wa, wb = a.pair_work(), b.pair_work()
    assert wa[0] < 0.5 * wb[0]                     # the cull really dropped most of the framework
    a.close(); b.close()
