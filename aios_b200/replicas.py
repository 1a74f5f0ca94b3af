"""Multi-GPU host logic for the replica path (requests are independent units; every
GPU holds a full model replica + KV pool; there is NO collective on the data path).

The reference never runs two replicas of one model (runtime/src/model_manager.rs:42-56 keeps one
llama-server per *model*); this is the B200 scale-out of its per-request independence.  Only the
bookkeeping (which rank serves which request; whole-job throughput = all units / max-over-ranks
time) touches torch.distributed, and it works on any backend (nccl on GPUs, gloo in CPU tests).
"""
from __future__ import annotations


def assign_requests(n_requests: int, world: int, rank: int) -> list[int]:
    """Round-robin request ids served by `rank` (least-loaded == round-robin for equal-cost units)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    return list(range(rank, n_requests, world))


def aggregate_throughput(local_units: float, local_ms: float, world: int, device=None):
    """(total units, max-over-ranks ms, units/s) — rank-symmetric; uses all_reduce SUM and MAX."""
    if world <= 1:
        return local_units, local_ms, local_units / (local_ms * 1e-3)
    import torch
    import torch.distributed as dist

    u = torch.tensor([float(local_units)], dtype=torch.float64, device=device)
    t = torch.tensor([float(local_ms)], dtype=torch.float64, device=device)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(u.item()), float(t.item()), float(u.item()) / (float(t.item()) * 1e-3)
