"""Sharding of independent stereo streams over ranks (one process per GPU).

The reference parallelises only over stereo pairs that share a read-only constant
(`cpl::jobs::parallel_for(numChannels / 2, ...)`, Source/Spectrum/SpectrumDSP.cpp:83-102); streams
never exchange data, so ranks own contiguous stream ranges, tables are replicated, and the only
communication is an optional final gather of results."""
from typing import Optional, Tuple


def stream_range(rank: int, world: int, n_streams: int) -> Tuple[int, int]:
    """Contiguous, balanced: the first `n_streams % world` ranks get one extra stream."""
    if not (0 <= rank < world) or n_streams < 0:
        raise ValueError("bad rank/world/n_streams")
    base, extra = divmod(n_streams, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_results(local, n_streams: int, dst: int = 0, group=None) -> Optional["torch.Tensor"]:
    """Gather per-rank results [local_streams, ...] into [n_streams, ...] on rank `dst`
    (torch.distributed, any backend; ranks may own different stream counts).  Returns None elsewhere."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    lo, hi = stream_range(rank, world, n_streams)
    assert local.shape[0] == hi - lo, (local.shape, lo, hi)
    counts = [stream_range(r, world, n_streams) for r in range(world)]
    longest = max(b - a for a, b in counts)
    padded = local
    if local.shape[0] < longest:  # pad so every rank contributes the same shape
        pad = torch.zeros((longest - local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        padded = torch.cat([local, pad], dim=0)
    bucket = [torch.empty_like(padded) for _ in range(world)] if rank == dst else None
    dist.gather(padded.contiguous(), bucket, dst=dst, group=group)
    if rank != dst:
        return None
    return torch.cat([bucket[r][: b - a] for r, (a, b) in enumerate(counts)], dim=0)
