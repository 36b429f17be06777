"""Data-parallel plumbing of the decision path: one process per GPU, torch.distributed for the exchange.

Inference shards battles with no collective (battles are independent, mlp_model.py:12: kernel size 1).  The
update needs exactly one exchange: the sum of the flat fp32 gradient vectors.  Each rank
computes d/dparams of (sum of its shard's Huber terms / GLOBAL batch size), so the sum over ranks is the
gradient of the reference's mean loss (trainer.py:187); the [-1, 1] clamp (trainer.py:193-194) is applied to
the reduced gradient, never to the per-rank pieces (clamp does not commute with the sum).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch


def shard_bounds(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of n items for `rank` of `world`; shards are disjoint, ordered and cover
    range(n); sizes differ by at most one."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} not in 0..{world - 1}")
    return (n * rank) // world, (n * (rank + 1)) // world


def active_group():
    """torch.distributed if a process group with more than one rank is initialised, else None."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        return dist
    return None


def allreduce_sum_(grad: torch.Tensor, loss: Optional[torch.Tensor] = None, dist=None) -> None:
    """In-place sum over ranks of the flat gradient (and the loss share).  NCCL on GPUs (one allreduce on one
    contiguous buffer: 175 KB for the default network, latency-bound), gloo in the CPU tests."""
    dist = dist if dist is not None else active_group()
    if dist is None:
        return
    dist.all_reduce(grad, op=dist.ReduceOp.SUM)
    if loss is not None:
        dist.all_reduce(loss, op=dist.ReduceOp.SUM)
