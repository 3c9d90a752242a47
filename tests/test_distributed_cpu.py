"""CPU, world_size 2 over gloo: clip sharding and the feature-statistics all-reduce (the path's only collective)."""
import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ssar_b200 import distributed as D
from ssar_b200.synth import synth_features


def test_shard_clips_balanced_and_complete():
    rng = np.random.default_rng(0)
    dur = rng.uniform(120, 360, 1000)
    for world in (1, 2, 4, 8):
        shards = D.shard_clips(dur, world)
        assert sorted(i for s in shards for i in s) == list(range(1000))
        loads = [sum(dur[i] for i in s) for s in shards]
        assert max(loads) - min(loads) <= 360.0  # greedy longest-first: imbalance below one clip
    assert D.shard_clips([], 2) == [[], []]


def _worker(rank, world, port, out):
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    clips = list(range(6))
    mine = D.shard_clips([100 + 10 * c for c in clips], world)[rank]
    stats = torch.zeros(3, 59, dtype=torch.float64)
    for c in mine:  # CPU stand-in of ssar_feature_stats (the CUDA kernel is covered by tests/test_gpu_post.py)
        f = synth_features(200 + 10 * c, 59, c).double()
        stats[0] += f.shape[0]
        stats[1] += f.sum(0)
        stats[2] += (f * f).sum(0)
    mean, std = D.mean_std_from_stats(D.merge_stats(stats))
    if rank == 0:
        out.put((mean.numpy(), std.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_stats_allreduce_world2_matches_numpy():
    import socket

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    mean, std = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    allf = np.concatenate([synth_features(200 + 10 * c, 59, c).double().numpy() for c in range(6)])
    np.testing.assert_allclose(mean, allf.mean(0), rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(std, allf.std(0), rtol=1e-9, atol=1e-9)
