"""Host-side logic of the isotherm sweep: chains = pressure points x replicas dealt to ranks, no
data-path collective, one all-gather of the per-chain averages.  The N > 1 path runs here on CPU with gloo,
world_size 2 (the GPU box uses NCCL through the same code)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from maniac_mc_b200 import isotherm as iso


def test_fugacity_points_are_log_spaced():
    f = iso.fugacity_points(64, 1e-2, 1e4)
    assert len(f) == 64 and np.isclose(f[0], 1e-2) and np.isclose(f[-1], 1e4)
    assert np.allclose(np.diff(np.log(f)), np.log(1e6) / 63)
    assert np.allclose(iso.fugacity_points(1, 3.0, 9.0), [3.0])


@pytest.mark.parametrize("n_points,replicas,world", [(64, 8, 1), (64, 8, 2), (64, 8, 8), (64, 1, 4), (5, 3, 4), (3, 1, 8)])
def test_deal_chains_is_a_partition(n_points, replicas, world):
    seen = []
    sizes = []
    for r in range(world):
        ids = iso.deal_chains(n_points, replicas, world, r)
        sizes.append(len(ids))
        for c in ids:
            assert c.index % world == r and c.index == c.replica * n_points + c.point
            assert c.seed == 1000 * c.point + c.replica
        seen += [c.index for c in ids]
    assert sorted(seen) == list(range(n_points * replicas))           # every chain exactly once
    assert max(sizes) - min(sizes) <= 1 and max(sizes) == iso.chains_per_rank(n_points, replicas, world)
    if world <= n_points and n_points % world == 0:                   # every rank sees the whole pressure range
        for r in range(world):
            pts = {c.point for c in iso.deal_chains(n_points, replicas, world, r)}
            assert min(pts) < world and max(pts) >= n_points - world
    with pytest.raises(ValueError):
        iso.deal_chains(n_points, replicas, world, world)


def test_block_averages():
    a = iso.BlockAverages(3)
    a.add([1, 2, 3], [10.0, 20.0, 30.0], np.arange(30).reshape(3, 10))
    a.add([3, 2, 1], [30.0, 20.0, 10.0], np.arange(30).reshape(3, 10) + 1)
    r = a.rows()
    assert r.shape == (3, 14)
    assert np.allclose(r[:, 0], [2, 2, 2]) and np.allclose(r[:, 1], [1, 0, 1]) and np.allclose(r[:, 2], [20, 20, 20])
    assert np.all(r[:, 3] == 2) and np.array_equal(r[:, 4:], np.arange(30).reshape(3, 10) + 1)


def fake_rows(ids):
    """per-chain rows that depend only on the global chain identity"""
    return np.array([[c.point + 0.25 * c.replica, c.index, c.seed] for c in ids], dtype=np.float64).reshape(len(ids), 3)


def test_gather_table_single_rank():
    ids = iso.deal_chains(6, 2)
    t = iso.gather_table(ids, fake_rows(ids), 6, 2, world=1)
    assert t.shape == (2, 6, 3) and not np.isnan(t).any()
    assert np.allclose(t[1, 4], [4.25, 10, 4001])
    m = iso.isotherm(t)
    assert np.allclose(m[:, 0], np.arange(6) + 0.125) and np.allclose(m[:, 1], 0.25 / np.sqrt(2) / np.sqrt(2))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_points, replicas, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ids = iso.deal_chains(n_points, replicas, world, rank)
        t = iso.gather_table(ids, fake_rows(ids), n_points, replicas, world)
        np.save(os.path.join(out, f"table_{rank}.npy"), t)
        # the bench's timing reduction: MAX over ranks of the device-timed durations
        ms = torch.tensor([10.0 + rank], dtype=torch.float64)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        assert float(ms) == 10.0 + world - 1
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_points,replicas", [(8, 2), (5, 3)])
def test_gather_table_world2_gloo(tmp_path, n_points, replicas):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n_points, replicas, str(tmp_path)), nprocs=world, join=True)
    all_ids = iso.deal_chains(n_points, replicas)
    want = iso.gather_table(all_ids, fake_rows(all_ids), n_points, replicas, world=1)
    for r in range(world):
        got = np.load(tmp_path / f"table_{r}.npy")
        assert np.array_equal(got, want)      # every rank holds the complete table, no chain lost or duplicated
