"""Multi-GPU plumbing of the WKV6 path: the recurrence is independent per (batch, head,
direction), so work shards by utterance with NO data-path collective -- rank g owns rows [g*B/G, (g+1)*B/G) like the
reference's one-process-per-GPU inference scripts (go-LF-dev-one-model-paper.sh:91-124).  Training keeps the
reference's DDP gradient all-reduce (train_utils.py:365-366), which this library does not touch."""
from __future__ import annotations

from typing import Tuple

import torch
import torch.distributed as dist


def shard_rows(batch: int, rank: int, world: int) -> Tuple[int, int]:
    """[begin, end) utterance rows of `rank`; remainders go to the lowest ranks, empty shards are legal."""
    base, rem = divmod(batch, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def shard_batch(x: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    b, e = shard_rows(x.shape[0], rank, world)
    return x[b:e].contiguous()


def gather_batch(x: torch.Tensor, batch: int) -> torch.Tensor:
    """Inverse of shard_batch across the default process group (used by tests / evaluation, not by the hot path)."""
    world, rank = dist.get_world_size(), dist.get_rank()
    sizes = [shard_rows(batch, r, world) for r in range(world)]
    pad = max(e - b for b, e in sizes)
    buf = x.new_zeros((pad,) + tuple(x.shape[1:]))
    buf[: x.shape[0]] = x
    outs = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(outs, buf)
    return torch.cat([o[: e - b] for o, (b, e) in zip(outs, sizes)], 0)


def max_over_ranks(value: float, device=None) -> float:
    """Timing convention of bench.py: a multi-GPU number is the MAX over ranks of the device-measured time."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
