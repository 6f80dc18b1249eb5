"""N>1 host logic on CPU: two gloo ranks shard a batch of frame pairs, each computes its slice of
the lookup (with the oracle standing in for the GPU kernels, which need a device), results are
gathered and must equal the single-process result.  No data-path collective is involved."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_pairs, out_path):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
    import torch.distributed as dist
    import oracle
    from raftcorr_b200.shard import gather_pairs, shard_bounds
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    gen = torch.Generator().manual_seed(3)
    f1 = torch.randn(n_pairs, 8, 16, 20, generator=gen)
    f2 = torch.randn(n_pairs, 8, 16, 20, generator=gen)
    coords = torch.rand(n_pairs, 2, 16, 20, generator=gen) * 18
    lo, hi = shard_bounds(n_pairs, rank, world)
    pyr = oracle.build_pyramid(f1[lo:hi], f2[lo:hi], 4)
    local = torch.from_numpy(oracle.lookup(pyr, coords[lo:hi], 4))
    full = gather_pairs(local, n_pairs)
    if rank == 0:
        np.save(out_path, full.numpy())
    dist.destroy_process_group()


def test_shard_bounds_cover_and_balance():
    from raftcorr_b200.shard import all_bounds
    for n in (0, 1, 5, 8, 13):
        for w in (1, 2, 3, 8):
            b = all_bounds(n, w)
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.timeout(300)
def test_two_rank_gloo_sharded_lookup_equals_single_process(tmp_path):
    import oracle
    n_pairs, world = 5, 2   # ragged: 3 + 2
    out = str(tmp_path / "gathered.npy")
    mp.spawn(_worker, args=(world, _free_port(), n_pairs, out), nprocs=world, join=True)
    gen = torch.Generator().manual_seed(3)
    f1 = torch.randn(n_pairs, 8, 16, 20, generator=gen)
    f2 = torch.randn(n_pairs, 8, 16, 20, generator=gen)
    coords = torch.rand(n_pairs, 2, 16, 20, generator=gen) * 18
    ref = oracle.lookup(oracle.build_pyramid(f1, f2, 4), coords, 4)
    assert np.array_equal(np.load(out), ref)
