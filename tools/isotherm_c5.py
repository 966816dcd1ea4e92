#!/usr/bin/env python
"""BASELINE configs[4]: adsorption isotherm sweep, 64 pressure points x replicas of ZIF-8 + H2O (configs[2],
mc-topology testcase-adsorption/ZIF8-H2O) partitioned over the GPUs of one box.

    python tools/isotherm_c5.py [--replicas 8] [--blocks 20] [--steps 500]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29540 tools/isotherm_c5.py

Chains are dealt to ranks (maniac_mc_b200/isotherm.py); no data-path collective; one all-gather of the per-chain
averages at the end.  Prints the isotherm (<N> per fugacity, replica mean +- s.e.) and moves/s (device time, max over ranks).
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
    from maniac_mc_b200 import isotherm as iso
    from maniac_mc_b200.system import System
    p = argparse.ArgumentParser()
    p.add_argument("--points", type=int, default=64)
    p.add_argument("--replicas", type=int, default=8)
    p.add_argument("--blocks", type=int, default=20)
    p.add_argument("--steps", type=int, default=500)
    p.add_argument("--system", default=os.path.join(ROOT, "tests", "golden", "zif8_h2o.npz"))
    p.add_argument("--files", default=None, help="directory: also time the block loop with the reference's per-block files "
                   "written per chain, synchronously and pipelined behind the next block (wall clock, rank 0 prints)")
    p.add_argument("--trajectory", action="store_true", help="with --files: also write trajectory.lammpstrj frames")
    a = p.parse_args()
    rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{lr}"))
    s = System.load(a.system)
    s.init_sk, s.deletion_mode = 1, 1
    sw = iso.IsothermSweep(s, n_points=a.points, replicas=a.replicas, lo_atm=1e-2, hi_atm=1e4, world=world, rank=rank, device=lr)
    stream = torch.cuda.ExternalStream(sw.gpu.stream(), device=lr)
    sw.run_block(a.steps, seed=1)                     # equilibration / warm-up block, not averaged
    sw.avg = iso.BlockAverages(len(sw.ids))
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for b in range(a.blocks):
        sw.run_block(a.steps, seed=1)
    e1.record(stream)
    sw.gpu.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=f"cuda:{lr}")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    table = sw.table()
    if rank == 0:
        curve = iso.isotherm(table)
        moves = a.points * a.replicas * a.blocks * a.steps
        acc = table[:, :, 4:8].sum() / max(table[:, :, 8:12].sum(), 1.0)
        print(json.dumps({"workload": "c5_isotherm_zif8_h2o", "n_gpus": world, "chains": a.points * a.replicas, "blocks": a.blocks,
                          "steps_per_block": a.steps, "moves_per_s": moves / (float(ms) * 1e-3), "ms": float(ms),
                          "acceptance": float(acc), "complete": bool(not np.isnan(table).any())}))
        f = iso.fugacity_points(a.points, 1e-2, 1e4)
        for i in range(0, a.points, max(1, a.points // 16)):
            print(f"  f = {f[i]:10.3e} atm   <N> = {curve[i, 0]:8.3f} +- {curve[i, 1]:.3f}")
    if a.files:
        import time
        from maniac_mc_b200 import writers as W
        out = {}
        for mode in ("no_files", "files_sync", "files_pipelined"):
            ws = [None] * sw.gpu.n_chains
            if mode != "no_files":
                for c, cid in enumerate(sw.ids):
                    ws[c] = W.ChainWriter(W.chain_directory(os.path.join(a.files, mode, f"rank{rank}"), cid.point, cid.replica), s,
                                          nb_block=a.blocks, trajectory=a.trajectory)
            pipe = W.BlockPipeline(sw.gpu, ws)
            sw.gpu.synchronize()
            t0 = time.perf_counter()
            pipe.run(a.blocks, a.steps, seed=1, pipelined=(mode != "files_sync"))
            sw.gpu.synchronize()
            out[mode] = time.perf_counter() - t0
        if rank == 0:
            print(json.dumps({"per_block_files": {k: round(v, 4) for k, v in out.items()}, "unit": "s wall per %d blocks" % a.blocks,
                              "chains_on_rank0": len(sw.ids), "trajectory": a.trajectory}))
    sw.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
