"""CPU, world_size 2 over gloo: the multi-process sharding and gather logic of bench.py's N > 1 path
(the data path itself has no collective)."""
import os
import socket

import numpy as np
import pytest

from support import GOLDEN

SEED = 20261104


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_total, out_path):
    import torch
    import torch.distributed as dist
    from circuitsimulator_b200 import shard
    from circuitsimulator_b200.netlist import Netlist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        nl = Netlist(str(np.load(os.path.join(GOLDEN, "buffer_nominal.npz"))["netlist"]))
        P = shard.shard_params(nl, n_total, rank, world, SEED)  # [P, count]
        first, count = shard.shard_range(n_total, rank, world)
        assert P.shape == (nl.n_params, count)
        # stand-in for the per-instance results a rank produces: [count, probes, points] + per-instance ints
        wave = torch.from_numpy(np.stack([P[:4].T, P[4:8].T], 1).copy())  # [count, 2, 4]
        ids = torch.arange(first, first + count, dtype=torch.int64)
        g_wave = shard.gather_instances(wave, dst=0, dim=0)
        g_ids = shard.gather_instances(ids, dst=0)
        g_cols = shard.gather_instances(torch.from_numpy(P.copy()), dst=0, dim=1)  # instance-minor arrays: dim 1
        if rank == 0:
            np.savez(out_path, wave=g_wave.numpy(), ids=g_ids.numpy(), cols=g_cols.numpy())
        else:
            assert g_wave is None and g_ids is None and g_cols is None
    finally:
        dist.destroy_process_group()


def test_shard_ranges_cover_the_batch_exactly():
    from circuitsimulator_b200 import shard
    for n, w in ((65536, 8), (10, 3), (7, 8), (1, 1)):
        got = [shard.shard_range(n, r, w) for r in range(w)]
        assert got[0][0] == 0 and sum(c for _, c in got) == n
        for (f0, c0), (f1, _) in zip(got, got[1:]):
            assert f1 == f0 + c0
        assert max(c for _, c in got) - min(c for _, c in got) <= 1
    with pytest.raises(ValueError):
        shard.shard_range(4, 4, 4)


@pytest.mark.parametrize("n_total", [64, 37])
def test_two_rank_shards_gather_to_the_single_process_batch(tmp_path, n_total):
    import torch.multiprocessing as mp
    from circuitsimulator_b200.netlist import Netlist
    out = str(tmp_path / "gathered.npz")
    mp.spawn(_worker, args=(2, _free_port(), n_total, out), nprocs=2, join=True)
    g = np.load(out)
    nl = Netlist(str(np.load(os.path.join(GOLDEN, "buffer_nominal.npz"))["netlist"]))
    whole = nl.monte_carlo(n_total, SEED)  # the same global batch drawn by ONE process
    assert np.array_equal(g["ids"], np.arange(n_total))
    assert np.array_equal(g["cols"].view(np.uint64), whole.view(np.uint64))
    assert np.array_equal(g["wave"][:, 0, :].view(np.uint64), np.ascontiguousarray(whole[:4].T).view(np.uint64))
    assert np.array_equal(whole[:, 0], nl.params)  # global instance 0 is the nominal circuit
