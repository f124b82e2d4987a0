"""Frame sharding across ranks: frames are the independent units of the path; the cloud is replicated and
the u32 accumulators are summed once at the end, so any partition of the frames gives the same colours."""
from __future__ import annotations


def frames_of_rank(n_frames: int, rank: int, world: int, mode: str = "interleaved", block: int = 25):
    """Indices of the keyframes rank `rank` processes.  "interleaved": r, r+G, ... (balances the slowly varying
    visible fraction along a path); "blocks": contiguous stretches (pose locality); "dealt": blocks of `block`
    consecutive keyframes dealt round-robin (both: consecutive frames of a rank stay centimetres apart, which the
    seed points of pass 1 rely on, and every rank sees every part of the path)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    if mode == "interleaved":
        return list(range(rank, n_frames, world))
    if mode == "blocks":
        lo = n_frames * rank // world
        hi = n_frames * (rank + 1) // world
        return list(range(lo, hi))
    if mode == "dealt":
        if block < 1:
            raise ValueError("block must be >= 1")
        return [g for b0 in range(0, n_frames, block) if (b0 // block) % world == rank for g in range(b0, min(b0 + block, n_frames))]
    raise ValueError(f"unknown sharding mode {mode}")


def deal_by_cost(costs, world: int, speed=None):
    """Cost-aware static sharding: `costs[b]` is the measured cost of unit b (a block of consecutive keyframes), `speed[r]` the
    relative speed of rank r (default 1).  Longest-processing-time-first: units in decreasing cost go to the rank that would
    finish them earliest.  Returns owner[b].  Deterministic (ties: lower unit index first, lower rank first), so every rank
    computes the same partition from the same gathered numbers; any partition gives the same colours (integer sums)."""
    if world < 1:
        raise ValueError("world must be >= 1")
    n = len(costs)
    speed = [1.0] * world if speed is None else [float(s) for s in speed]
    if len(speed) != world or any(s <= 0 for s in speed):
        raise ValueError("speed: one positive number per rank")
    load = [0.0] * world
    owner = [0] * n
    for b in sorted(range(n), key=lambda i: (-float(costs[i]), i)):
        r = min(range(world), key=lambda q: ((load[q] + float(costs[b])) / speed[q], q))
        owner[b] = r
        load[r] += float(costs[b])
    return owner
