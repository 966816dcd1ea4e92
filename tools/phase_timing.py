"""Per-phase cycle breakdown of one trial of the fused chain kernel on the bench workload (needs a -DMMC_PHASE_TIMING build:
MANIAC_GPU_LIB=<that .so> python tools/phase_timing.py [chains]).  Thread 0 of every CTA stamps clock64 at the phase
boundaries; the sums over all CTAs and trials are divided by the trial count."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from maniac_mc_b200.lib import ManiacGPU  # noqa: E402

chains = int(sys.argv[1]) if len(sys.argv) > 1 else 592
workload = sys.argv[2] if len(sys.argv) > 2 else "c4_large_k"
if workload == "c4_large_k":
    s = bench.make_system(argparse.Namespace(framework=50000, guests=200))
else:       # a fixture of tests/golden, e.g. zif8_h2o (the C5 isotherm's system)
    from maniac_mc_b200.system import System
    s = System.load(os.path.join(ROOT, "tests", "golden", workload + ".npz")); s.init_sk, s.deletion_mode = 1, 1
g = ManiacGPU(s, n_chains=chains)
print(f"# workload {workload}: K = {g.nk}")
for c in range(chains):
    g.set_chain_params(c, bench.chain_fugacity(s, c), s.temperature); g.set_chain_id(c, c)
g.system_energy_all()
g.run_batch_async(256, seed=7); g.synchronize()          # warm-up: the chains relax towards their loading
p0 = g.phase_cycles(); w0 = g.pair_work()
g.run_batch_async(256, seed=7); g.synchronize()
p1 = g.phase_cycles(); w1 = g.pair_work()
n = p1["trials"] - p0["trials"]
print(f"# {chains} chains x 256 steps, {n} trials timed; cycles per trial of one CTA (thread 0's clock64 between barriers)")
tot = 0
for k in ("propose", "stage", "real_space", "sum4", "reciprocal", "sum2", "finalize_commit"):
    v = (p1[k] - p0[k]) / max(n, 1); tot += v
    print(f"{k:18s} {v:12.0f} cycles")
print(f"{'total':18s} {tot:12.0f} cycles = {tot / 1.965e3:.1f} us at 1965 MHz")
print(f"# per trial: {(w1[0] - w0[0]) / max(n, 1):.0f} distances formed, {(w1[1] - w0[1]) / max(n, 1):.0f} pairs inside reach"
      "   (-DMMC_RS_SPLIT builds: sum4 = tile cull, sum2 = framework tiles, real_space = the rest of that phase)")
g.close()
