"""Multi-GPU host logic for the replica path (requests are independent units; every
GPU holds a full model replica + KV pool; there is NO collective on the data path).

The reference never runs two replicas of one model (runtime/src/model_manager.rs:42-56 keeps one
llama-server per *model*); this is the B200 scale-out of its per-request independence.  Only the
bookkeeping (which rank serves which request; whole-job throughput = all units / max-over-ranks
time) touches torch.distributed, and it works on any backend (nccl on GPUs, gloo in CPU tests).
"""
from __future__ import annotations

import threading


class ReplicaSet:
    """One model replicated on several GPUs of this process (one engine per GPU): the handle ModelManager keeps in
    place of the reference's child process.  Every request goes to the replica with the fewest requests in flight
    ("host dispatcher assigns each request to the least-loaded replica"); ties go round the replicas
    in turn.  The single-replica attributes the inference layer reads (vocab, n_ctx, seq, lock, model) are replica 0's
    — vocabulary and context are identical on every replica."""

    def __init__(self, replicas: list, devices: list[int]):
        if not replicas:
            raise ValueError("no replicas")
        self.replicas, self.devices = list(replicas), list(devices)
        self.inflight = [0] * len(replicas)
        self.served = [0] * len(replicas)
        self._mu = threading.Lock()
        self._turn = 0
        r0 = replicas[0]
        self.name, self.vocab, self.n_ctx = r0.name, r0.vocab, r0.n_ctx
        self.seq, self.lock, self.model = getattr(r0, "seq", None), getattr(r0, "lock", None), getattr(r0, "model", None)

    def pick(self) -> int:
        """Index of the least-loaded replica; the caller owns one in-flight count until release(i)."""
        with self._mu:
            n = len(self.replicas)
            best = min(range(n), key=lambda i: (self.inflight[i], (i - self._turn) % n))
            self._turn = (best + 1) % n
            self.inflight[best] += 1
            self.served[best] += 1
            return best

    def release(self, i: int) -> None:
        with self._mu:
            self.inflight[i] -= 1

    def submit(self, tokens, max_tokens: int, sampling, json_mode: bool = False):
        i = self.pick()
        try:
            req = self.replicas[i].submit(tokens, max_tokens, sampling, json_mode=json_mode)
        except Exception:
            self.release(i)
            raise
        return _Tracked(req, self, i)

    def load(self) -> dict:
        """device -> requests in flight (HealthCheck details)."""
        with self._mu:
            return {d: n for d, n in zip(self.devices, self.inflight)}

    def close(self):
        for r in self.replicas:
            r.close()


class _Tracked:
    """A queued request that gives its replica's in-flight count back when it is freed."""

    def __init__(self, req, owner: ReplicaSet, idx: int):
        self._req, self._owner, self._idx, self.device = req, owner, idx, owner.devices[idx]

    def __getattr__(self, name):
        return getattr(self._req, name)

    def free(self):
        if self._owner is not None:
            self._owner.release(self._idx)
            self._owner = None
        self._req.free()


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
