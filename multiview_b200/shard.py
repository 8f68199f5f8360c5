"""Sharding of camera-frames over ranks: every (camera, frame) unit is independent through
a1-a9, so ranks own disjoint cameras and only the per-frame floor histograms are summed across ranks.

Rank g of G owns cameras {g, g+G, g+2G, ...} for all frames (keeps each camera's rectification maps and calibration
resident on one GPU).  With more ranks than cameras the surplus ranks split the frames of a camera instead."""
from __future__ import annotations


def cameras_for_rank(n_cams: int, world: int, rank: int) -> list[int]:
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return list(range(rank, n_cams, world)) if world <= n_cams else [rank % n_cams]


def frames_for_rank(n_frames: int, n_cams: int, world: int, rank: int) -> range:
    """Frames of its camera(s) this rank processes: all of them, unless several ranks share one camera."""
    if world <= n_cams:
        return range(n_frames)
    share = [r for r in range(world) if r % n_cams == rank % n_cams]
    k, idx = len(share), share.index(rank)
    per = -(-n_frames // k)
    return range(min(idx * per, n_frames), min((idx + 1) * per, n_frames))


def units_for_rank(n_cams: int, n_frames: int, world: int, rank: int) -> list[tuple[int, int]]:
    return [(c, f) for c in cameras_for_rank(n_cams, world, rank) for f in frames_for_rank(n_frames, n_cams, world, rank)]
