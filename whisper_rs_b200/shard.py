"""Segment sharding across the GPUs of one box: independent units, no data-path collective.

A 30 s segment (front-end) and an audio's beam group (decode step) never depend on another unit, so rank g of G
owns the contiguous block of segments [lo, hi) below; beams of one audio stay on one rank.  Results are token ids
and scores, gathered on the host (`gather_results`, torch.distributed object gather -- control plane only).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple


def segment_range(n_segments: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block partition; the first n_segments % world ranks get one extra segment."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("need 0 <= rank < world")
    if n_segments < 0:
        raise ValueError("n_segments < 0")
    base, extra = divmod(n_segments, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def row_range(n_audio: int, beam: int, rank: int, world: int) -> Tuple[int, int]:
    """Rows [lo, hi) of the [n_audio * beam] decode batch owned by `rank`: whole beam groups only."""
    lo, hi = segment_range(n_audio, rank, world)
    return lo * beam, hi * beam


def all_ranges(n_segments: int, world: int) -> List[Tuple[int, int]]:
    return [segment_range(n_segments, r, world) for r in range(world)]


def gather_results(local: Sequence, n_segments: int, group=None) -> List:
    """Collect per-segment results (any picklable objects, in local segment order) on every rank, in global
    segment order.  Uses torch.distributed's object all-gather on whatever backend is initialised (gloo in the
    CPU tests, NCCL's CPU side-channel on a GPU box); single process when torch.distributed is not initialised."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        if len(local) != n_segments:
            raise ValueError("single process must hold every segment")
        return list(local)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lo, hi = segment_range(n_segments, rank, world)
    if len(local) != hi - lo:
        raise ValueError(f"rank {rank} holds {len(local)} results for segments [{lo}, {hi})")
    parts: List = [None] * world
    dist.all_gather_object(parts, list(local), group=group)
    out: List = []
    for p in parts:
        out.extend(p)
    return out
