"""Multi-GPU plumbing for the path: clip sharding (no data-path collective) and the one real exchange — the dataset
mean / std of ssar/supervised/data.py:317-319 as an all-reduce of per-feature (count, sum, sum of squares) in fp64.

One process per GPU (`torchrun`), NCCL over NVLink / NVSwitch on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_clips(durations, world_size: int):
    """Longest-first greedy bin packing of clips onto ranks by duration (work is proportional to samples).
    Returns a list (one entry per rank) of clip indices; deterministic on every rank."""
    order = sorted(range(len(durations)), key=lambda i: (-float(durations[i]), i))
    loads = [0.0] * world_size
    shards = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (loads[k], k))
        shards[r].append(i)
        loads[r] += float(durations[i])
    return shards


def merge_stats(stats: torch.Tensor, group=None) -> torch.Tensor:
    """All-reduce (sum) of a local [3, C] fp64 accumulator (count, sum, sum of squares) across ranks, in place."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats


def mean_std_from_stats(stats: torch.Tensor):
    """(count, sum, sum of squares) -> (mean, population std) per feature: np.mean / np.std over all windows and frames."""
    n, s, ss = stats[0], stats[1], stats[2]
    mean = s / n
    var = torch.clamp(ss / n - mean * mean, min=0.0)
    return mean, torch.sqrt(var)


def dataset_mean_std(local_feature_batches, n_features: int = 59, device=None, group=None):
    """Local feature tensors [..., C] (CUDA) -> global (mean [C], std [C]) over every rank's frames."""
    from . import ops

    stats = None
    for f in local_feature_batches:
        stats = ops.feature_stats(f, stats)
    if stats is None:
        stats = torch.zeros(3, n_features, dtype=torch.float64, device=device)
    return mean_std_from_stats(merge_stats(stats, group))
