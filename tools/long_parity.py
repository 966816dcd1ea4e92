"""Trajectory-length parity of the CUDA path against the CPU oracle (SURVEY 8d lengths).

Every case runs ONE recorded uniform stream through the oracle (the reference's sequential chain restated in C) and through
the library twice -- strict mode (host proposals, one mmc_trial_energy + mmc_trial_commit per move, what the unmodified
Fortran move routines would call; and its one-launch form mmc_trial_step) and batched mode (device proposals replaying the tape) -- and reports per run:
moves compared, the largest per-component relative error of the old/new energy vectors (conftest.record_errors), the
largest relative error of the acceptance probability and the NUMBER OF ACCEPT/REJECT FLIPS.  A flip is not hidden: the
two probabilities at the flip are printed (the uniform lies between them) and the GPU chain is resynchronised with the
oracle's checkpoint at the next chunk boundary, so the rest of the trajectory is still compared.

    python tools/long_parity.py [--scale 1.0] [--out profiles/r2_long_parity.log]      (needs a GPU; a few minutes)

Cases: C1 h2o_ads 100 x 100 moves; C2 ~500 TIP4P in L = 30 A, 2e5 moves; C3 ZIF-8 + H2O, 2e5 moves; each
under both quirk settings (reference defaults init_sk=0 / deletion=reference, and the physical switches); C4 the exact
bench system (bench.make_system: 50 000-atom framework, K = 37 782, tile cull on), 240 moves, init_sk=0 so that neither
side needs the O(N K) start-up.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

DRAWS_PER_MOVE = 13   # upper bound of uniforms one MC step can consume (9 + 3 reservoir draws + accept)


def oracle_reference(s, n_moves, chunk, tape, start_energy=True):
    """runs the oracle over the tape in chunks; returns (records, tape offsets per chunk, checkpoints per chunk END)"""
    from oracle import oracle as orc
    o = orc.Oracle(s)
    if start_energy:
        o.system_energy()
    else:
        o.prepare_fourier()     # phase tables of every molecule, A(k) = 0, tracked energies = 0 (what the device chain starts from)
    recs, offs, ckpt = [], [0], []
    pos = 0
    for a in range(0, n_moves, chunk):
        n = min(chunk, n_moves - a)
        recs.append(o.run(n, tape=tape[pos:pos + DRAWS_PER_MOVE * n + 16]))
        pos += int(o.tape_used)
        offs.append(pos)
        ckpt.append((o.state(), o.energy().copy(), o.sk().copy()))
    return np.concatenate(recs), offs, ckpt


class Tally:
    def __init__(self, label):
        self.label, self.moves, self.flips = label, 0, 0
        self.e_old = self.e_new = self.e_p = 0.0
        self.max_abs = 0.0     # largest absolute difference [K] over components below 1e6 K (overlapping insertions excluded)
        self.worst = None      # (error, which vector, move, kind, component, reference value, absolute difference, reference total)
        self.flip_notes = []
        self.mismatch = 0      # chunks whose move sequence differed for another reason than a flip (must stay 0)

    def add(self, got, ref, base):
        """compares one chunk; returns True when the GPU chain must be resynchronised"""
        from conftest import record_errors
        n = len(ref)
        same = (got["kind"] == ref["kind"]) & (got["res"] == ref["res"]) & (got["mol"] == ref["mol"]) & (got["accepted"] == ref["accepted"])
        good = n if same.all() else int(np.argmin(same))     # moves before the first divergence
        if good:
            eo, en, ep = record_errors(got[:good], ref[:good])
            self.e_old, self.e_new, self.e_p = max(self.e_old, eo), max(self.e_new, en), max(self.e_p, ep)
            from conftest import record_scale
            for key in ("old_e", "new_e"):       # where the largest error sits (for the log)
                x, y = got[key][:good], ref[key][:good]
                err = np.abs(x - y) / record_scale(y)
                self.max_abs = max(self.max_abs, float(np.max(np.where(np.abs(y) < 1e6, np.abs(x - y), 0.0))))
                i, c = np.unravel_index(int(np.argmax(err)), err.shape)
                if self.worst is None or err[i, c] > self.worst[0]:
                    self.worst = (float(err[i, c]), key, int(base + i), int(ref["kind"][i]), int(c), float(y[i, c]), float(abs(x[i, c] - y[i, c])), float(y[i, 5]))
        self.moves += good
        if good == n:
            return False
        g, r = got[good], ref[good]
        if g["kind"] == r["kind"] and g["res"] == r["res"] and g["mol"] == r["mol"] and g["accepted"] != r["accepted"]:
            self.flips += 1
            self.flip_notes.append(f"flip at move {base + good}: p_gpu={g['prob']:.17g} p_oracle={r['prob']:.17g} "
                                   f"|dp|={abs(g['prob'] - r['prob']):.3e} (the accept uniform lies between them)")
        else:
            self.mismatch += 1
            self.flip_notes.append(f"move sequence differs at move {base + good}: gpu kind/res/mol {g['kind']}/{g['res']}/{g['mol']} "
                                   f"oracle {r['kind']}/{r['res']}/{r['mol']}")
        return True

    def row(self):
        return {"run": self.label, "moves_compared": self.moves, "max_rel_err_old_e": self.e_old, "max_rel_err_new_e": self.e_new,
                "max_rel_err_prob": self.e_p, "max_abs_diff_K": self.max_abs, "accept_reject_flips": self.flips, "sequence_mismatches": self.mismatch,
                "worst": None if self.worst is None else dict(zip(("rel_err", "vector", "move", "kind", "component", "ref_value_K", "abs_diff_K", "ref_total_K"), self.worst)),
                "notes": self.flip_notes}


def system_at(s, ckpt):
    """a copy of `s` whose guests are the oracle checkpoint's"""
    (cnt, com, off), _, _ = ckpt
    t = s.copy()
    mo, ao = s.mol_offsets()
    coms, offs = [], []
    im = ia = 0
    for r in range(s.nres):
        n, na = int(cnt[r]), int(s.natoms[r])
        coms.append(com[im:im + n]); offs.append(off[ia:ia + n * na])
        im += n; ia += n * na
    t.count = np.asarray(cnt, dtype=np.int32).copy()
    t.com = np.concatenate(coms) if coms else np.zeros((0, 3)); t.offsets = np.concatenate(offs) if offs else np.zeros((0, 3))
    return t


def run_strict(s, ref, offs, ckpt, chunk, tape, label, start_energy=True, fused=False):
    from maniac_mc_b200.driver import StrictDriver, TapeRng
    from maniac_mc_b200.lib import ManiacGPU
    t = Tally(label)
    g = ManiacGPU(s); drv = StrictDriver(s, g, fused=fused)
    if start_energy:
        drv.ComputeSystemEnergy()
    n_moves = len(ref)
    for ci, a in enumerate(range(0, n_moves, chunk)):
        n = min(chunk, n_moves - a)
        rng = TapeRng(tape); rng.pos = offs[ci]
        got = drv.MonteCarloLoop(rng, 1, n)
        if t.add(got, ref[a:a + n], a):
            g.close()
            s2 = system_at(s, ckpt[ci])
            g = ManiacGPU(s2); drv = StrictDriver(s2, g, fused=fused)
            g.system_energy()
            g.dbg_set_tracked(ckpt[ci][1], ckpt[ci][2])
            drv.energy = ckpt[ci][1].copy()
    g.close()
    return t


def run_strict_fused(s, ref, offs, ckpt, chunk, tape, label, start_energy=True):
    """strict mode through mmc_trial_step: one launch per move, the accept test on the device with the host's uniform"""
    return run_strict(s, ref, offs, ckpt, chunk, tape, label, start_energy, fused=True)


def run_batched(s, ref, offs, ckpt, chunk, tape, label, start_energy=True):
    from maniac_mc_b200.lib import ManiacGPU
    t = Tally(label)
    g = ManiacGPU(s)
    if start_energy:
        g.system_energy()
    n_moves = len(ref)
    for ci, a in enumerate(range(0, n_moves, chunk)):
        n = min(chunk, n_moves - a)
        tp = tape[offs[ci]:offs[ci] + DRAWS_PER_MOVE * n + 16]
        got = g.run_batch(n, tape=tp[None, :], records=True)[0]
        if t.add(got, ref[a:a + n], a):
            cnt, com, off = ckpt[ci][0]
            g.set_chain_state(0, cnt, com, off)
            g.system_energy()
            g.dbg_set_tracked(ckpt[ci][1], ckpt[ci][2])
    g.close()
    return t


def cases(scale):
    from conftest import load_system
    from maniac_mc_b200 import builders
    out = []
    n1, n2, n3 = max(200, int(1e4 * scale)), max(400, int(2e5 * scale)), max(400, int(2e5 * scale))
    for q, (isk, dm) in (("reference quirks", (0, 0)), ("physical", (1, 1))):
        out.append((f"C1 h2o_ads [{q}]", load_system("h2o_ads", init_sk=isk, deletion_mode=dm), n1, True))
        w = builders.synthetic_water_box(n_mol=500, box=30.0, cutoff=14.0, tol=1e-5, seed=1, capacity=1024, fugacity_atm=0.05,
                                         p_translation=0.25, p_rotation=0.25)
        w.init_sk, w.deletion_mode = isk, dm
        out.append((f"C2 water box 500 [{q}]", w, n2, True))
        out.append((f"C3 zif8_h2o [{q}]", load_system("zif8_h2o", init_sk=isk, deletion_mode=dm), n3, True))
    return out


def bench_case(n_moves=240):
    """the exact system bench.py times (K = 37 782), init_sk = 0: tracked A(k) starts at zero on both sides, so neither
    needs the single-point start-up (oracle: ~30 s of CPU); the per-move work is the bench's"""
    import bench
    args = argparse.Namespace(framework=50000, guests=200)
    s = bench.make_system(args)
    s.init_sk, s.deletion_mode = 0, 1
    return ("C4 bench system c4_large_k (K=37782) [init_sk=0]", s, n_moves, False)


def run_all(scale=1.0, chunk=2000, with_bench=True, log=print, seed=2025):
    todo = cases(scale)
    if with_bench:
        todo.append(bench_case())
    rng = np.random.default_rng(seed)
    tapes = [rng.random(DRAWS_PER_MOVE * n + 64) for (_, _, n, _) in todo]
    refs = [None] * len(todo)

    def work(i):
        name, s, n, start = todo[i]
        t0 = time.time()
        refs[i] = oracle_reference(s, n, min(chunk, n), tapes[i], start_energy=start)
        log(f"# oracle {name}: {n} moves in {time.time() - t0:.1f} s")
    ths = [threading.Thread(target=work, args=(i,)) for i in range(len(todo))]
    [t.start() for t in ths]
    rows = []
    for i, (name, s, n, start) in enumerate(todo):
        ths[i].join()
        ref, offs, ckpt = refs[i]
        acc = float(ref["accepted"].mean())
        for mode, fn in (("strict", run_strict), ("strict fused", run_strict_fused), ("batched tape", run_batched)):
            t0 = time.time()
            t = fn(s, ref, offs, ckpt, min(chunk, n), tapes[i], f"{name} / {mode}", start_energy=start)
            r = t.row(); r["seconds"] = round(time.time() - t0, 1); r["accept_ratio"] = round(acc, 4)
            rows.append(r)
            log(json.dumps(r))
        refs[i] = None
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=1.0, help="fraction of the SURVEY 8d trajectory lengths")
    ap.add_argument("--chunk", type=int, default=2000)
    ap.add_argument("--no-bench-system", action="store_true")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    fh = open(a.out, "w") if a.out else None

    def log(msg):
        print(msg, flush=True)
        if fh:
            fh.write(msg + "\n"); fh.flush()
    log(f"# long parity run, scale {a.scale}, chunk {a.chunk}: GPU (C ABI) vs CPU oracle on identical uniform streams")
    rows = run_all(a.scale, a.chunk, not a.no_bench_system, log)
    bad = [r for r in rows if r["accept_reject_flips"] or r["sequence_mismatches"] or max(r["max_rel_err_old_e"], r["max_rel_err_new_e"]) >= 1e-10]
    log(f"# {len(rows)} runs, {sum(r['moves_compared'] for r in rows)} moves compared, "
        f"{sum(r['accept_reject_flips'] for r in rows)} accept/reject flips, "
        f"max per-component relative error {max(max(r['max_rel_err_old_e'], r['max_rel_err_new_e']) for r in rows):.3e}")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
