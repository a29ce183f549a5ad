"""Stream -> GPU partitioning and the end-of-run statistics gather (the only collective).

Independent camera streams are the unit that shards: stream s runs on rank s % world, with no inter-GPU traffic on the hot
path.  After a run every rank contributes a small statistics vector; torch.distributed (NCCL over NVLink on the GPU box,
gloo in CPU tests) all-gathers it.
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import numpy as np

STAT_KEYS = ("frames", "observations", "seconds", "launches", "h2d_bytes", "d2h_bytes")


def assign_streams(n_streams: int, world: int, rank: int) -> List[int]:
    """Static assignment: stream s -> rank s % world."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    return [s for s in range(n_streams) if s % world == rank]


def gather_stats(dist, local: Dict[str, float], device=None) -> Dict[str, object]:
    """All-gathers one STAT_KEYS vector per rank. Returns per-rank rows plus the job aggregate: frames and bytes add up,
    the job's time is the slowest rank's."""
    import torch
    vec = torch.tensor([float(local.get(k, 0.0)) for k in STAT_KEYS], dtype=torch.float64, device=device)
    if dist is None or not dist.is_initialized():
        rows = [vec.cpu().numpy()]
    else:
        out = [torch.zeros_like(vec) for _ in range(dist.get_world_size())]
        dist.all_gather(out, vec)
        rows = [o.cpu().numpy() for o in out]
    m = np.stack(rows)
    agg = {k: float(m[:, i].sum()) for i, k in enumerate(STAT_KEYS)}
    agg["seconds"] = float(m[:, STAT_KEYS.index("seconds")].max())
    agg["frames_per_sec"] = agg["frames"] / agg["seconds"] if agg["seconds"] > 0 else 0.0
    return {"per_rank": [dict(zip(STAT_KEYS, map(float, r))) for r in rows], "job": agg}
