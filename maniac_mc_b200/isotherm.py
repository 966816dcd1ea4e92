"""Isotherm sweep: pressure points x replicas as independent chains dealt to the GPUs of one box (SURVEY 8e,
BASELINE configs[4]).

The reference runs ONE chain per process with one fugacity per active residue (`input_parser.f90:446-452`,
`prepare_utils.f90:68`) and has one global output path; an isotherm is a shell loop over input files.  Here the
points x replicas are the chains of one `mmc_init` handle per GPU.  Chains share only read-only data, so there is
NO data-path collective: chain g goes to rank g mod world (interleaving puts low- and high-loading points on every
GPU), every rank runs `mmc_run_batch` on its own chains, and the only exchange is one all-gather of the per-chain
averages (a few kB) at the end -- `gather_table`.

Host logic only; the device work is `ManiacGPU.run_batch_async`.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .system import N_COUNTERS


def fugacity_points(n_points=64, lo_atm=1.0e-2, hi_atm=1.0e4):
    """log-spaced fugacities in atm"""
    if n_points == 1:
        return np.array([lo_atm])
    return lo_atm * (hi_atm / lo_atm) ** (np.arange(n_points) / (n_points - 1))


@dataclass(frozen=True)
class ChainId:
    index: int     # global chain index g = replica * n_points + point
    point: int
    replica: int

    @property
    def seed(self):
        """chain seed = 1000 * point + replica"""
        return 1000 * self.point + self.replica


def deal_chains(n_points, replicas, world=1, rank=0):
    """Global chains handled by `rank`: g = replica * n_points + point, g mod world == rank.
    Consecutive points land on different ranks, so every GPU sees the whole loading range."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    return [ChainId(g, g % n_points, g // n_points) for g in range(n_points * replicas) if g % world == rank]


def chains_per_rank(n_points, replicas, world):
    """largest local chain count over ranks (ranks differ by at most one chain)"""
    n = n_points * replicas
    return (n + world - 1) // world


class BlockAverages:
    """Running per-chain averages sampled once per block, like the reference's per-block `number_*.dat` /
    `energy.dat` rows (`write_utils.f90:13-188`): <N> of the active residue, <E_total>, acceptance counters."""

    def __init__(self, n_chains):
        self.n = 0
        self.sum_n = np.zeros(n_chains)
        self.sum_e = np.zeros(n_chains)
        self.sum_n2 = np.zeros(n_chains)
        self.counters = np.zeros((n_chains, N_COUNTERS), dtype=np.int64)

    def add(self, counts, energies, counters=None):
        c = np.asarray(counts, dtype=np.float64)
        self.n += 1
        self.sum_n += c
        self.sum_n2 += c * c
        self.sum_e += np.asarray(energies, dtype=np.float64)
        if counters is not None:
            self.counters = np.asarray(counters, dtype=np.int64).copy()

    def rows(self):
        """[n_chains, 4 + N_COUNTERS]: mean N, var N, mean E_total, blocks, then the move counters"""
        m = max(self.n, 1)
        mean_n = self.sum_n / m
        var_n = np.maximum(self.sum_n2 / m - mean_n ** 2, 0.0)
        return np.concatenate([np.stack([mean_n, var_n, self.sum_e / m, np.full_like(mean_n, self.n)], axis=1),
                               self.counters.astype(np.float64)], axis=1)


def gather_table(ids, rows, n_points, replicas, world=1):
    """All ranks contribute `rows[i]` for chain `ids[i]`; returns the [replicas, n_points, ncol] table on every rank.
    Uses torch.distributed when world > 1 (NCCL on the GPU box, gloo in the CPU tests): ONE all-gather of
    chains_per_rank x (1 + ncol) doubles -- the only collective of the sweep."""
    rows = np.asarray(rows, dtype=np.float64).reshape(len(ids), -1)
    ncol = rows.shape[1]
    table = np.full((replicas, n_points, ncol), np.nan)
    if world == 1:
        for cid, r in zip(ids, rows):
            table[cid.replica, cid.point] = r
        return table
    import torch
    import torch.distributed as dist
    cap = chains_per_rank(n_points, replicas, world)
    buf = np.full((cap, 1 + ncol), -1.0)
    for i, (cid, r) in enumerate(zip(ids, rows)):
        buf[i, 0] = cid.index
        buf[i, 1:] = r
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.from_numpy(buf).to(dev)
    parts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(parts, t)
    out = torch.stack(parts).cpu().numpy().reshape(world * cap, 1 + ncol)
    for row in out:
        g = int(row[0])
        if g >= 0:
            table[g // n_points, g % n_points] = row[1:]
    return table


def isotherm(table):
    """replica average and standard error of <N> per pressure point: [n_points, 2]"""
    mean_n = table[:, :, 0]
    r = mean_n.shape[0]
    se = mean_n.std(axis=0, ddof=1) / np.sqrt(r) if r > 1 else np.zeros(mean_n.shape[1])
    return np.stack([mean_n.mean(axis=0), se], axis=1)


class IsothermSweep:
    """One rank's share of an isotherm sweep on one GPU."""

    def __init__(self, system, n_points=64, replicas=1, lo_atm=1.0e-2, hi_atm=1.0e4, world=1, rank=0, device=0,
                 active_residue=None, recalibrate=None):
        from . import builders
        from .lib import ManiacGPU   # raises if the CUDA library is missing: no CPU fallback
        self.sys, self.n_points, self.replicas, self.world, self.rank = system, n_points, replicas, world, rank
        # input%recalibrate_moves (monte_carlo.f90:78): AdjustMoveStepSizes after every block; default = what the input says
        self.recalibrate = bool(getattr(system, "recalibrate_moves", False)) if recalibrate is None else bool(recalibrate)
        self.ids = deal_chains(n_points, replicas, world, rank)
        self.f_atm = fugacity_points(n_points, lo_atm, hi_atm)
        self.res = int(np.flatnonzero(system.is_active)[0]) if active_residue is None else active_residue
        self.gpu = ManiacGPU(system, n_chains=max(1, len(self.ids)), device=device)
        for c, cid in enumerate(self.ids):
            f = system.fugacity.copy()
            f[self.res] = builders.fugacity_from_atm(self.f_atm[cid.point], system.temperature)
            self.gpu.set_chain_params(c, f, system.temperature)
            self.gpu.set_chain_id(c, cid.seed)     # global Philox stream: the sweep does not depend on the dealing
        self.gpu.system_energy_all()               # one launch set for all chains; the framework's share is computed once
        self.avg = BlockAverages(len(self.ids))
        self.step = 0
        g_mols, g_atoms = self.gpu.chain_capacity()
        C = self.gpu.n_chains
        self._cnt = np.zeros((C, system.nres), dtype=np.int32)
        self._e = np.zeros((C, 6)); self._ctr = np.zeros((C, N_COUNTERS), dtype=np.int64)

    def run_block(self, n_steps, seed=0):
        """one block of `n_steps` MC steps on every local chain, then one sample of N, E and the counters
        (`monte_carlo.f90:42-84`: inner step loop, then the per-block output)."""
        self.gpu.run_batch_async(n_steps, seed=seed)
        if self.recalibrate:
            self.gpu.adjust_step_sizes()
        # raises ManiacGpuError if a creation hit the slot capacity during the block (the reference aborts there)
        self.gpu.download_chains(self._cnt, None, None, self._e, self._ctr)
        n = len(self.ids)
        self.avg.add(self._cnt[:n, self.res], self._e[:n, 5], self._ctr[:n])
        self.step += n_steps

    def run_blocks(self, n_blocks, n_steps, seed=0, output_root=None, trajectory=False, literal=True):
        """`n_blocks` blocks with the per-block exchange off the critical path (`mmc_snapshot_begin/_wait`): block b+1 is
        queued before the host reads block b's snapshot.  With `output_root` every chain also gets the reference's
        per-block files (`energy.dat`, `number_<RES>.dat`, `moves.dat`, optionally `trajectory.lammpstrj`) in
        `<output_root>/p<point>_r<replica>/` (`writers.py`)."""
        from . import writers as W
        ws = [None] * self.gpu.n_chains
        if output_root is not None:
            for c, cid in enumerate(self.ids):
                ws[c] = W.ChainWriter(W.chain_directory(output_root, cid.point, cid.replica), self.sys, literal=literal,
                                      nb_block=n_blocks, trajectory=trajectory)
        n = len(self.ids)

        def sample(block, snap):
            cnt, _, _, e, ctr = snap
            if block > 0:
                self.avg.add(cnt[:n, self.res], e[:n, 5], ctr[:n])
        pipe = W.BlockPipeline(self.gpu, ws, on_block=sample, recalibrate=self.recalibrate)
        if output_root is not None and self.step == 0:
            pipe.write_initial()
        pipe.block = self.step // max(n_steps, 1)
        pipe.run(n_blocks, n_steps, seed=seed)
        self.step += n_blocks * n_steps

    def table(self):
        return gather_table(self.ids, self.avg.rows(), self.n_points, self.replicas, self.world)

    def close(self):
        self.gpu.close()
