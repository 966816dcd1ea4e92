#!/usr/bin/env python
"""Few chains per GPU: moves/s of the batched kernel for each cluster size (MANIAC_CLUSTER = 0 one CTA per chain, 2 / 4 / 8 =
k_chain_run_cl), device-timed on the library's stream.

    python tools/few_chains.py            # zif8_h2o at 512 / 256 / 128 / 64 / 32 chains (= the per-GPU share of the fixed-size C5
                                          # sweep on 1 / 2 / 4 / 8 / 16 GPUs) and c4_large_k at 148 and 64 chains
The strong-scaling efficiency of the C5 sweep on N GPUs follows as N x rate(512 / N chains) / rate(512 chains): chains never
exchange data, so a GPU's rate depends only on its own chain count."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
from maniac_mc_b200.lib import ManiacGPU  # noqa: E402
from maniac_mc_b200.system import System  # noqa: E402


def rate(s, chains, cluster, steps, reps=3):
    os.environ["MANIAC_CLUSTER"] = str(cluster)
    g = ManiacGPU(s, n_chains=chains)
    for c in range(chains):
        g.set_chain_params(c, bench.chain_fugacity(s, c * (512 // chains) if chains <= 512 else c), s.temperature); g.set_chain_id(c, c)
    g.system_energy_all()
    cfg = g.chain_launch_config()
    stream = torch.cuda.ExternalStream(g.stream())
    g.run_batch_async(steps, seed=5); g.run_batch_async(steps, seed=5); g.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        g.run_batch_async(steps, seed=5)
    e1.record(stream); g.synchronize()
    ms = e0.elapsed_time(e1)
    g.close()
    return chains * steps * reps / (ms * 1e-3), cfg


def main():
    z = System.load(os.path.join(ROOT, "tests", "golden", "zif8_h2o.npz")); z.init_sk, z.deletion_mode = 1, 1
    big = bench.make_system(argparse.Namespace(framework=50000, guests=200))
    out = []
    for name, s, counts, steps in (("zif8_h2o", z, (512, 256, 128, 64, 32), 1024), ("c4_large_k", big, (148, 64), 256)):
        for chains in counts:
            row = {"system": name, "chains": chains}
            for cl in (0, 2, 4, 8):
                if cl and cl * chains > 8 * 148:
                    continue
                r, cfg = rate(s, chains, cl, steps)
                row[f"cluster_{cl}"] = round(r)
                row[f"cfg_{cl}"] = cfg
            best = max((v, k) for k, v in row.items() if k.startswith("cluster_"))
            row["best"] = best[1]
            out.append(row)
            print(json.dumps(row), flush=True)
    base = next(r for r in out if r["system"] == "zif8_h2o" and r["chains"] == 512)
    b1 = max(v for k, v in base.items() if k.startswith("cluster_"))
    print("# fixed-size C5 sweep (512 chains of zif8_h2o in total): strong-scaling efficiency N x rate(512/N) / rate(512)")
    for r in out:
        if r["system"] == "zif8_h2o" and r["chains"] < 512:
            n = 512 // r["chains"]
            one = r["cluster_0"]; bst = max(v for k, v in r.items() if k.startswith("cluster_"))
            print(f"#   {n:2d} GPUs ({r['chains']:3d} chains each): one CTA per chain {n * one / b1:.2f}, best cluster ({r['best']}) {n * bst / b1:.2f}")


if __name__ == "__main__":
    main()
