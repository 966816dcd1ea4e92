"""k-sharded mode: ONE large system (or a few chains of it) spread over the GPUs of a box ("optional
single-huge-system mode"; BASELINE north_star: "an NCCL allreduce over NVLink only when a single very large system's
k-vector set is sharded across GPUs").

The reference has no counterpart (single-threaded CPU code).  Every rank builds a handle from the SAME system with
`shard_rank = rank, shard_count = world`: it keeps all coordinates, makes the same proposal from the same Philox
counters, sweeps 1/world of the framework tiles and owns 1/world of the (kx, ky) rows of A(k).  Per MC step of all
chains the data path has exactly one exchange: an all-reduce (sum, FP64) of 6 doubles per chain -- the partial
(LJ_old, Coul_old, LJ_new, Coul_new, sum fW|A|^2, sum fW|A'|^2) -- enqueued on the library's stream between
`mmc_shard_step_begin` and `mmc_shard_step_end`.  All ranks receive bit-identical sums, draw the same accept uniform and
commit the same way, so no broadcast follows.

`reduce` is injectable so that the protocol can be exercised on one GPU (the tests emulate two ranks with two handles
and a plain tensor sum) and on CPU-only hosts for the host logic.
"""
from __future__ import annotations

import numpy as np


class ShardedSystem:
    """One rank's shard of a k-sharded set of chains."""

    def __init__(self, system, n_chains=1, rank=0, world=1, device=0, reduce=None):
        import torch
        from .lib import ManiacGPU   # raises if the CUDA library is missing: no CPU fallback
        self.torch = torch
        self.rank, self.world = rank, world
        self.gpu = ManiacGPU(system, n_chains=n_chains, device=device, shard_rank=rank, shard_count=world)
        self.stream = torch.cuda.ExternalStream(self.gpu.stream(), device=device)
        self.sums = torch.zeros((n_chains, 6), dtype=torch.float64, device=f"cuda:{device}")
        self.reduce = reduce if reduce is not None else self._nccl_all_reduce
        for c in range(n_chains):
            self.gpu.system_energy(c)       # replicated: every rank computes the full single-point energy and S(k)

    def _nccl_all_reduce(self, t):
        import torch.distributed as dist
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)   # NCCL on the library's stream (current stream below)

    # -- batched sharded step: proposal on the device, one all-reduce of 6 x chains doubles -----------------------
    def step_begin(self, seed):
        self.gpu.shard_step_begin(seed, self.sums.data_ptr())

    def step_end(self, records=False):
        return self.gpu.shard_step_end(self.sums.data_ptr(), records=records)

    def step(self, seed, records=False):
        with self.torch.cuda.stream(self.stream):
            self.step_begin(seed)
            self.reduce(self.sums)
            return self.step_end(records=records)

    def run(self, n_steps, seed, records=False):
        out = [self.step(seed, records=records) for _ in range(n_steps)]
        self.gpu.synchronize()
        return np.stack(out, axis=1) if records else None

    # -- strict sharded trial: host proposal, host decision --------------------------------------------------------
    def trial_energy(self, kind, res, mol, com_new=None, offs_new=None, chain=0):
        with self.torch.cuda.stream(self.stream):
            s = self.sums[chain]
            self.gpu.shard_trial_begin(kind, res, mol, com_new, offs_new, s.data_ptr(), chain=chain)
            self.reduce(s)
            return self.gpu.shard_trial_end(s.data_ptr(), chain=chain)

    def trial_commit(self, accept, chain=0):
        self.gpu.trial_commit(accept, chain=chain)

    def owned_sk(self, chain=0):
        """(k_first, A(k) of the owned k-vectors): only these entries are kept current on this rank"""
        k0, n = self.gpu.shard_range()
        return k0, self.gpu.sk(chain)[k0:k0 + n]

    def close(self):
        self.gpu.close()


def emulate_ranks(system, world, n_chains=1, device=0):
    """`world` shards on ONE GPU whose all-reduce is a plain sum over the shards' buffers -- the protocol of a real
    multi-GPU run with the collective replaced (tests, single-GPU debugging)."""
    import torch
    shards = [ShardedSystem(system, n_chains=n_chains, rank=r, world=world, device=device, reduce=lambda t: None)
              for r in range(world)]

    def step(seed, records=False):
        for s in shards:
            s.step_begin(seed)
        for s in shards:
            s.gpu.synchronize()
        total = torch.stack([s.sums for s in shards]).sum(dim=0)
        torch.cuda.synchronize()
        out = None
        for s in shards:
            s.sums.copy_(total)
            torch.cuda.synchronize()
            out = s.step_end(records=records)
        return out

    def trial_energy(kind, res, mol, com_new=None, offs_new=None, chain=0):
        for s in shards:
            s.gpu.shard_trial_begin(kind, res, mol, com_new, offs_new, s.sums[chain].data_ptr(), chain=chain)
            s.gpu.synchronize()
        total = torch.stack([s.sums[chain] for s in shards]).sum(dim=0)
        torch.cuda.synchronize()
        res_ = None
        for s in shards:
            s.sums[chain].copy_(total)
            torch.cuda.synchronize()
            res_ = s.gpu.shard_trial_end(s.sums[chain].data_ptr(), chain=chain)
        return res_

    def trial_commit(accept, chain=0):
        for s in shards:
            s.trial_commit(accept, chain=chain)

    return shards, step, trial_energy, trial_commit
