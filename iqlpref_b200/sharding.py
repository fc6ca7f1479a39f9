"""One process per GPU: how agents and dataset rows are split over ranks, and the only two
collectives the path has.

The reference runs every agent as an independent OS process pinned to a GPU
(`tr_sweeps/launch.sh:82-92`); agents never exchange data.  Here a rank owns a contiguous block of
agents (training) or of dataset rows (relabelling; all ensemble members stay on every rank so the
per-row CVaR selection is local), and `torch.distributed` (NCCL over NVLink on B200, gloo in the CPU
tests) is used only to collect what every rank needs to see:

* `gather_losses`  - the logged `[n_local, 3]` loss rows, every `log_freq` steps (iql.py:1537-1544)
* `gather_rows`    - the relabelled reward slices (4 bytes per dataset row)
"""
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of `n` items owned by `rank`; sizes differ by at most one."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_sizes(n: int, world: int) -> List[int]:
    return [shard_bounds(n, r, world)[1] - shard_bounds(n, r, world)[0] for r in range(world)]


def _world(group=None) -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def gather_rows(local: torch.Tensor, n_total: int, group=None) -> torch.Tensor:
    """All-gather row shards produced with `shard_bounds(n_total, rank, world)` along dim 0.
    Ragged shards are padded to the largest one for the collective and trimmed afterwards."""
    rank, world = _world(group)
    if world == 1:
        if local.shape[0] != n_total:
            raise ValueError("single-rank gather expects the full tensor")
        return local
    sizes = shard_sizes(n_total, world)
    if local.shape[0] != sizes[rank]:
        raise ValueError(f"rank {rank} holds {local.shape[0]} rows, expected {sizes[rank]}")
    width = max(sizes)
    padded = local.new_zeros((width,) + tuple(local.shape[1:]))
    padded[:local.shape[0]] = local
    parts = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(parts, padded, group=group)
    return torch.cat([p[:s] for p, s in zip(parts, sizes)], dim=0)


def gather_losses(local_losses: torch.Tensor, n_agents_total: int, group=None) -> torch.Tensor:
    """[n_local, 3] per rank -> [n_agents_total, 3] on every rank (agent order = global agent id)."""
    return gather_rows(local_losses, n_agents_total, group)


def sharded_ensemble_predict(members, X, reduce: str = "mean", alpha: float = 0.0, activation: str = "relu",
                             device="cuda", group=None) -> torch.Tensor:
    """Row-sharded `ensemble_predict`: every rank relabels its block of rows with ALL members (the CVaR
    selection stays local), then the reward slices are all-gathered."""
    from .relabel import ensemble_predict
    rank, world = _world(group)
    n = X.shape[0]
    lo, hi = shard_bounds(n, rank, world)
    local = ensemble_predict(members, X[lo:hi], reduce, alpha, activation, device)
    return gather_rows(local, n, group)
