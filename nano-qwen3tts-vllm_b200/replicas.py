"""Per-GPU replicas: the path has no exchange step, so N GPUs are N independent
talker+predictor engines with the request streams partitioned across them.  The only cross-rank
traffic is bookkeeping: the barrier around a timed region and the max-over-ranks of its duration.
Works over any ``torch.distributed`` backend (nccl on the box, gloo in the CPU tests)."""
from __future__ import annotations

import torch


def partition_streams(stream_ids: list, world_size: int, rank: int) -> list:
    """Stream ``s`` goes to replica ``s mod world_size`` (the reference pins everything to device 0,
    engine/model_runner/base.py:53; replicas generalise that with one engine group per GPU)."""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    return [s for i, s in enumerate(stream_ids) if i % world_size == rank]


def least_loaded(loads: list[int]) -> int:
    """Router policy for a NEW stream: the replica with the fewest live streams (lowest rank wins ties)."""
    return min(range(len(loads)), key=lambda r: (loads[r], r))


def aggregate(units_local: float, ms_local: float, dist=None, device="cpu") -> tuple[float, float]:
    """(units summed over ranks, duration max over ranks).  ``dist`` None = single replica."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(units_local), float(ms_local)
    u = torch.tensor([float(units_local)], dtype=torch.float64, device=device)
    t = torch.tensor([float(ms_local)], dtype=torch.float64, device=device)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(u.item()), float(t.item())
