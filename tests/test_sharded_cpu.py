"""CPU tests of the N>1 path (world_size 2, gloo): the all-gather of shard totals and the
routing of global draws to shards, with the device pieces replaced by the oracle.  The
CUDA kernels of the same path are covered on one GPU by test_shard_routing_equals_single_tree
and on several GPUs by tests/multi_gpu_check.py (run under torchrun on the GPU box)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class _OracleShard:
    """Stand-in for ReplayMemory on the CPU: totals and descents come from the oracle."""

    def __init__(self, leaves):
        self.leaves = leaves
        self.device = torch.device("cpu")

    def total_tensor(self, out):
        out[0] = int(np.int64(np.uint64(self.leaves.sum(dtype=np.uint64))))
        return out


def _route_oracle(totals, rank, u):
    from oracle import replay_oracle as ro
    tot = [int(np.uint64(np.int64(t))) for t in totals.tolist()]
    t, shard, local_t = ro.shard_split(tot, u.numpy())
    mine = np.nonzero(shard == rank)[0]
    return local_t[mine], mine, len(mine)


def _worker(rank, world, port, tmp):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import apex_b200
    from oracle import replay_oracle as ro
    n_per, B = 4000, 512
    shards = [np.random.default_rng(100 + r).integers(0, 1 << 28, n_per).astype(np.uint64) for r in range(world)]
    u = torch.from_numpy(np.random.default_rng(7).random(B))            # same draws on every rank
    sampler = apex_b200.ShardedSampler(_OracleShard(shards[rank]), route=_route_oracle)
    assert (sampler.rank, sampler.world) == (rank, world)
    totals = sampler.gather_totals()
    assert totals.tolist() == [int(s.sum(dtype=np.uint64)) for s in shards]
    local_t, pos, cnt = sampler.route(u)
    local_idx = ro.descend(shards[rank], local_t)
    gathered = [None] * world
    dist.all_gather_object(gathered, (pos.tolist(), (rank * n_per + local_idx).tolist()))
    if rank == 0:
        got = np.full(B, -1, np.int64)
        for p, i in gathered:
            got[p] = i
        want = ro.tree_sample(np.concatenate(shards), u.numpy())
        np.save(os.path.join(tmp, "ok.npy"), np.array([int((got == want).all()), sum(len(p) for p, _ in gathered)]))
    dist.destroy_process_group()


def test_sharded_sampling_equals_single_tree_gloo(tmp_path):
    world = 2
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    ok = np.load(tmp_path / "ok.npy")
    assert ok[0] == 1 and ok[1] == 512
