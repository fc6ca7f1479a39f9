"""CPU, world_size 2 over gloo: the host-side multi-GPU logic (agent / row sharding and the two
collectives).  The arithmetic runs on the GPU only, so the payloads here are synthetic tensors."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from iqlpref_b200 import sharding


def test_shard_bounds_cover_everything_once():
    for n in (0, 1, 7, 64, 1_000_001):
        for world in (1, 2, 3, 8):
            b = [sharding.shard_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1 and sizes == sharding.shard_sizes(n, world)
    with pytest.raises(ValueError):
        sharding.shard_bounds(4, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_agents, n_rows, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = sharding.shard_bounds(n_agents, rank, world)
        losses = torch.stack([torch.arange(lo, hi, dtype=torch.float32) * 10 + k for k in range(3)], dim=1)
        got = sharding.gather_losses(losses, n_agents)
        want = torch.stack([torch.arange(n_agents, dtype=torch.float32) * 10 + k for k in range(3)], dim=1)
        assert torch.equal(got, want)
        rlo, rhi = sharding.shard_bounds(n_rows, rank, world)
        rewards = torch.arange(rlo, rhi, dtype=torch.float32) * 0.5
        full = sharding.gather_rows(rewards, n_rows)
        assert torch.equal(full, torch.arange(n_rows, dtype=torch.float32) * 0.5)
        with pytest.raises(ValueError):
            sharding.gather_rows(rewards[:-1], n_rows)
        torch.save(full, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_agents,n_rows", [(64, 1001), (5, 3)])
def test_gather_over_two_ranks(tmp_path, n_agents, n_rows):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, n_agents, n_rows, str(tmp_path)), nprocs=2, join=True)
    a, b = torch.load(tmp_path / "r0.pt"), torch.load(tmp_path / "r1.pt")
    assert torch.equal(a, b)


def test_single_rank_is_identity():
    x = torch.arange(6.0)
    assert sharding.gather_rows(x, 6) is x
