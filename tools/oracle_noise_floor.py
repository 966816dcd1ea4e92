"""How reproducible is the reference algorithm itself in FP64?  (the yardstick of tests/conftest.py::record_scale)

Builds the CPU restatement twice -- the normal build (gcc -O2, no FMA contraction on x86-64) and one where gcc may contract
a*b+c into fused multiply-adds (-mfma -ffp-contract=fast): same algorithm, same libm, same loop order, only the rounding of
some products changes -- runs both over one recorded uniform stream and reports how far the per-move energy vectors move:
absolute [K] and under the parity metric.  No GPU involved.

    python tools/oracle_noise_floor.py [moves] > profiles/r2_oracle_noise_floor.log
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)


def run(which, so, moves, out):
    code = f"""
import sys, numpy as np
sys.path.insert(0, {ROOT!r}); sys.path.insert(0, {os.path.join(ROOT, 'tests')!r})
from oracle import oracle as orc
if {so!r}:
    orc._SO = {so!r}; orc.build = lambda force=False: orc._SO
from conftest import load_system
res = {{}}
for name in ('zif8_h2o', 'h2o_ads', 'mfi_co2'):
    for isk, dm in ((0, 0), (1, 1)):
        s = load_system(name, init_sk=isk, deletion_mode=dm)
        tape = np.random.default_rng(5).random(13 * {moves} + 64)
        o = orc.Oracle(s); o.system_energy(); res[(name, isk)] = o.run({moves}, tape=tape)
np.save({out!r}, res, allow_pickle=True)
"""
    subprocess.check_call([sys.executable, "-c", code])


def main():
    moves = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
    from conftest import record_scale
    with tempfile.TemporaryDirectory() as d:
        so = os.path.join(d, "liboracle_fma.so")
        subprocess.check_call(["gcc", "-O2", "-std=c11", "-fPIC", "-mfma", "-ffp-contract=fast", "-shared", "-o", so,
                               os.path.join(ROOT, "oracle", "maniac_oracle.c"), "-lm"])
        run("std", "", moves, os.path.join(d, "a.npy")); run("fma", so, moves, os.path.join(d, "b.npy"))
        a = np.load(os.path.join(d, "a.npy"), allow_pickle=True).item(); b = np.load(os.path.join(d, "b.npy"), allow_pickle=True).item()
    print(f"# CPU restatement vs ITSELF built with -mfma -ffp-contract=fast, {moves} moves per case, same uniform stream")
    for k in a:
        x, y = a[k], b[k]
        same = (x["kind"] == y["kind"]) & (x["accepted"] == y["accepted"])
        good = len(x) if same.all() else int(np.argmin(same))
        worst_rel, worst_abs = 0.0, 0.0
        for key in ("old_e", "new_e"):
            X, Y = x[key][:good], y[key][:good]
            worst_rel = max(worst_rel, float(np.max(np.abs(X - Y) / record_scale(Y))))
            worst_abs = max(worst_abs, float(np.max(np.where(np.abs(Y) < 1e6, np.abs(X - Y), 0.0))))
        print(f"{k[0]:10s} init_sk={k[1]}: {good} moves with identical accept sequence, max per-component error under the parity "
              f"metric {worst_rel:.2e}, max absolute shift {worst_abs:.2e} K (components below 1e6 K)")


if __name__ == "__main__":
    main()
