"""Multi-GPU plan: independent streams dealt to ranks, no data-path collective.

The only communication is the timing reduction: whole-job frames/s = frames decoded by all
ranks / the slowest rank's device time."""
from __future__ import annotations

from typing import List, Tuple


def streams_of_rank(n_streams: int, world: int, rank: int) -> List[int]:
    """stream i -> GPU i mod G"""
    return [i for i in range(n_streams) if i % world == rank]


def split_closed_gops(keyframe_flags: List[bool], world: int) -> List[List[Tuple[int, int]]]:
    """A single long stream is cut at key frames into closed GOPs dealt round-robin;
    returns per rank the list of [first, last) coded-frame ranges."""
    starts = [i for i, k in enumerate(keyframe_flags) if k]
    if not starts or starts[0] != 0:
        raise ValueError("stream must start with a key frame")
    gops = [(s, e) for s, e in zip(starts, starts[1:] + [len(keyframe_flags)])]
    return [[g for j, g in enumerate(gops) if j % world == r] for r in range(world)]


def aggregate_fps(frames_this_rank: int, ms_this_rank: float, dist=None) -> float:
    """total frames over max time; `dist` is torch.distributed (or None for one process)."""
    if dist is None or not dist.is_initialized():
        return frames_this_rank / (ms_this_rank * 1e-3)
    import torch
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([ms_this_rank], dtype=torch.float64, device=dev)
    n = torch.tensor([float(frames_this_rank)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(n, op=dist.ReduceOp.SUM)
    return float(n.item()) / (float(t.item()) * 1e-3)
