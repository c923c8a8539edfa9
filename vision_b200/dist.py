"""Multi-GPU partition helpers (host logic only). The path shards by independent units:
pixel-samples (frame indices) for the path tracer, rays for the batch trace API."""
from __future__ import annotations


def frames_for_rank(rank: int, world: int, spp: int):
    """strided sample partition: rank g renders frames g, g+G, g+2G, ... < spp.
    Returns (first_frame, n_frames, stride). The union over ranks is exactly range(spp)."""
    n = (spp - rank + world - 1) // world if rank < spp else 0
    return rank, n, world


def ray_range_for_rank(rank: int, world: int, n_rays: int):
    """contiguous ray chunk [lo, hi) of rank"""
    per = (n_rays + world - 1) // world
    lo = min(n_rays, rank * per)
    return lo, min(n_rays, lo + per)
