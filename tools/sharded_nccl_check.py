#!/usr/bin/env python
"""k-sharded mode over real NCCL: launch with torchrun on G GPUs of one box.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tools/sharded_nccl_check.py [--chains 8] [--steps 200]

Every rank holds a shard of the SAME c4_large_k system (50 000-atom framework, K = 37 782).  Checks: all ranks end up
with identical move records (same accept/reject sequence, same energies), rank 0 compares them with an unsharded
handle run from the same Philox seed, and the per-step time (device events, max over ranks) is printed.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from maniac_mc_b200 import builders
    from maniac_mc_b200.lib import ManiacGPU
    from maniac_mc_b200.sharded import ShardedSystem
    p = argparse.ArgumentParser()
    p.add_argument("--chains", type=int, default=8)
    p.add_argument("--steps", type=int, default=200)
    a = p.parse_args()
    rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{lr}"))
    s = builders.synthetic_framework(n_framework=50000, n_guests=200, box=96.0, cutoff=12.0, tol=1e-6, seed=4, capacity=512,
                                     p_translation=0.4, p_rotation=0.4)
    s.init_sk, s.deletion_mode = 1, 1
    sh = ShardedSystem(s, n_chains=a.chains, rank=rank, world=world, device=lr)
    seed = 777
    recs = sh.run(a.steps, seed, records=True)                 # [chains, steps]
    # timing: steps without records
    for _ in range(10):
        sh.step(seed)
    sh.gpu.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_t = 200
    e0.record(sh.stream)
    for _ in range(n_t):
        sh.step(seed)
    e1.record(sh.stream)
    sh.gpu.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=f"cuda:{lr}")
    # every rank must hold identical records
    key = np.concatenate([recs["kind"].ravel(), recs["accepted"].ravel()]).astype(np.float64)
    en = recs["new_e"].reshape(-1)
    t = torch.from_numpy(np.concatenate([key, en])).to(f"cuda:{lr}")
    if world > 1:
        tmax, tmin = t.clone(), t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX); dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        identical = bool(torch.equal(tmax, tmin))
    else:
        identical = True
    out = None
    if rank == 0:
        g = ManiacGPU(s, n_chains=a.chains, device=lr)
        for c in range(a.chains):
            g.system_energy(c)
        whole = g.run_batch(a.steps, seed=seed, records=True)
        same_seq = bool(np.array_equal(whole["kind"], recs["kind"]) and np.array_equal(whole["accepted"], recs["accepted"]))
        scale = np.maximum(np.abs(whole["new_e"]).max(axis=2, keepdims=True), 1.0)
        err = float(np.max(np.abs(whole["new_e"] - recs["new_e"]) / scale))
        g.close()
        out = {"world": world, "chains": a.chains, "steps": a.steps, "k_vectors": int(sh.gpu.nk), "ranks_identical": identical,
               "same_accept_sequence_as_unsharded": same_seq, "max_rel_energy_diff_vs_unsharded": err,
               "ms_per_step_all_chains": float(ms) / n_t, "moves_per_s": a.chains * n_t / (float(ms) * 1e-3),
               "collective": f"all_reduce sum f64 x {6 * a.chains} per MC step (NCCL)" if world > 1 else "none (world 1)"}
        print(json.dumps(out))
        assert identical and same_seq and err < 1e-10
    sh.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
