"""Frame sharding across ranks: frames are the independent units of the path; the cloud is replicated and
the u32 accumulators are summed once at the end, so any partition of the frames gives the same colours."""
from __future__ import annotations


def frames_of_rank(n_frames: int, rank: int, world: int, mode: str = "interleaved"):
    """Indices of the keyframes rank `rank` processes.  "interleaved": r, r+G, ... (balances the slowly varying
    visible fraction along a path); "blocks": contiguous stretches (pose locality)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    if mode == "interleaved":
        return list(range(rank, n_frames, world))
    if mode == "blocks":
        lo = n_frames * rank // world
        hi = n_frames * (rank + 1) // world
        return list(range(lo, hi))
    raise ValueError(f"unknown sharding mode {mode}")
