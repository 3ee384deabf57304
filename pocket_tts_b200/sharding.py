"""Placement of independent utterance streams on the GPUs of one box.

Streams never interact (each owns its KV cache, conv state and noise), weights are replicated, so
there is no collective on the data path: a stream is placed on one rank and stays there.  The only
multi-process traffic is control-plane (who takes which stream, gathering per-rank throughput)."""
from __future__ import annotations

from dataclasses import dataclass, field


def shard_streams(n_streams: int, world_size: int, rank: int) -> list[int]:
    """Static round-robin placement: stream i lives on rank i % world_size (the reference handles
    one utterance per call, so order carries no meaning; round-robin balances ragged lengths)."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank {rank} / world_size {world_size}")
    return list(range(rank, n_streams, world_size))


@dataclass
class LeastLoadedPlacer:
    """Admission-time placement for a serving loop: a new stream goes to the rank with the fewest
    live slots (ties -> lowest rank); closing a stream frees its slot."""
    world_size: int
    capacity: int
    live: list = field(default_factory=list)

    def __post_init__(self):
        self.live = [0] * self.world_size

    def admit(self) -> int:
        rank = min(range(self.world_size), key=lambda r: (self.live[r], r))
        if self.live[rank] >= self.capacity:
            raise RuntimeError("all ranks are at stream capacity")
        self.live[rank] += 1
        return rank

    def release(self, rank: int) -> None:
        if self.live[rank] <= 0:
            raise ValueError(f"rank {rank} has no live stream")
        self.live[rank] -= 1


def aggregate_throughput(local_audio_seconds: float, local_ms: float, group=None) -> tuple[float, float]:
    """Whole-job audio seconds and the slowest rank's time (the job is as slow as its slowest rank);
    works on any torch.distributed backend (nccl on the GPU box, gloo in the CPU tests)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return local_audio_seconds, local_ms
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    t = torch.tensor([local_audio_seconds], dtype=torch.float64, device=dev)
    m = torch.tensor([local_ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(m, op=dist.ReduceOp.MAX, group=group)
    return float(t), float(m)
