"""Stream -> GPU partitioning for the multi-GPU path: streams are independent units (the reference keeps
all per-stream state inside nemo_stream_context, src/nemo-stream.h:177-262, and its server processes sessions
independently, src/nemo-server.cpp:192-271), so N GPUs = N replicas of the weights, each owning the streams whose id
maps to it. There is NO data-path collective; torch.distributed is used only for the benchmark's barrier and the
max-over-ranks / sum-over-ranks reductions of its timing."""
from __future__ import annotations


def owner_of(stream_id: int, world_size: int) -> int:
    """GPU (rank) that owns a stream: stream_id mod world_size."""
    if world_size <= 0:
        raise ValueError("world_size must be positive")
    return stream_id % world_size


def shard_streams(n_streams: int, world_size: int, rank: int) -> list[int]:
    """Global stream ids served by `rank` when n_streams streams are spread over world_size GPUs."""
    if not 0 <= rank < world_size:
        raise ValueError("rank out of range")
    return [s for s in range(n_streams) if owner_of(s, world_size) == rank]


def reduce_timing(local_ms: float, local_units: float, device=None):
    """(max over ranks of local_ms, sum over ranks of local_units). Works with any initialised process group
    (nccl on GPUs, gloo on CPU); with no process group it is the identity."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(local_ms), float(local_units)
    t = torch.tensor([float(local_ms)], dtype=torch.float64, device=device)
    u = torch.tensor([float(local_units)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return float(t.item()), float(u.item())
