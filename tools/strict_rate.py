#!/usr/bin/env python
"""Strict-mode latency (mmc_trial_step, one launch per move) on the reference's ZIF-8 example and on the bench system; the same
loop bench.py's strict_mode object times (C loop inside the library, wall clock)."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from maniac_mc_b200.lib import ManiacGPU  # noqa: E402
from maniac_mc_b200.system import System  # noqa: E402

z = System.load(os.path.join(ROOT, "tests", "golden", "zif8_h2o.npz")); z.init_sk, z.deletion_mode = 1, 1
big = bench.make_system(argparse.Namespace(framework=50000, guests=200))
for name, s, n in (("zif8_h2o", z, 4000), ("c4_large_k", big, 2000)):
    g = ManiacGPU(s, n_chains=1); g.system_energy_all()
    g.strict_loop(200, seed=1, fused=True)
    best = min(g.strict_loop(n, seed=2 + i, fused=True)[0] for i in range(3))
    print(f"{name}: {1e6 * best / n:.2f} us per move (mmc_trial_step, best of 3 x {n})")
    g.close()
