"""Multi-GPU plumbing: requests are independent, so a batch is split over the ranks (one process per GPU) with
NO collective on the data path.  torch.distributed is only used to agree on the timing:
a barrier around the timed region, MAX over ranks of the device time, SUM of the bytes.
"""
from __future__ import annotations

import os
from typing import List, Sequence, Tuple

from ._lib import shard_plan


def rank_world() -> Tuple[int, int, int]:
    """(rank, world_size, local_rank) from the torchrun environment (1-process defaults)."""
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def my_shard(msg_bytes: Sequence[int], rank: int, world: int) -> List[int]:
    """Indices of the messages this rank owns: greedy largest-first balancing on bytes (tfs_shard_plan).
    Every rank computes the same plan from the same sizes, so no communication is needed."""
    assign = shard_plan(list(msg_bytes), world)
    return [i for i, a in enumerate(assign) if a == rank]


def aggregate(bytes_local: float, ms_local: float, device=None) -> Tuple[float, float]:
    """(total bytes over all ranks, max milliseconds over all ranks).  No-op without a process group."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(bytes_local), float(ms_local)
    t = torch.tensor([float(bytes_local)], dtype=torch.float64, device=device)
    m = torch.tensor([float(ms_local)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    return float(t.item()), float(m.item())


def throughput_gbps(total_bytes: float, max_ms: float) -> float:
    """Whole-job GB/s: all ranks' bytes over the slowest rank's time."""
    return total_bytes / (max_ms * 1e-3) / 1e9 if max_ms > 0 else 0.0
