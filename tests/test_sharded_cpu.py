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


def _grad_worker(rank, world, port, tmp):
    """Data-parallel gradient reduction of ShardedLearner (FlatGradients), on gloo: each rank backpropagates
    its rows of a global batch with the loss divided by the GLOBAL batch; the all-reduced flat gradient
    equals the single-process gradient of the whole batch, and identical optimiser steps keep replicas equal."""
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    import apex_b200
    torch.manual_seed(0)                                               # same replica on every rank
    net = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3))
    whole = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3))
    whole.load_state_dict(net.state_dict())
    fg = apex_b200.FlatGradients(net.parameters())
    assert all(p.grad.data_ptr() >= fg.flat.data_ptr() for p in net.parameters())
    opt = torch.optim.RMSprop(net.parameters(), lr=0.01)
    opt_whole = torch.optim.RMSprop(whole.parameters(), lr=0.01)
    g = torch.Generator().manual_seed(3)
    ok = True
    for step in range(3):
        x, y = torch.randn(10, 6, generator=g), torch.randn(10, 3, generator=g)
        rows = slice(0, 7) if rank == 0 else slice(7, 10)              # uneven shares, as shard quotas are
        fg.zero()
        (0.5 * (net(x[rows]) - y[rows]) ** 2).sum().div(10 * 3).backward()
        fg.all_reduce()
        opt_whole.zero_grad()
        (0.5 * (whole(x) - y) ** 2).mean().backward()
        want = torch.cat([p.grad.reshape(-1) for p in whole.parameters()])
        ok &= bool(torch.allclose(fg.flat, want, rtol=1e-5, atol=1e-7))
        opt.step()
        opt_whole.step()
        ok &= all(p.grad.data_ptr() >= fg.flat.data_ptr() and p.grad.data_ptr() < fg.flat.data_ptr() + fg.flat.numel() * 4
                  for p in net.parameters())                           # still views of the flat buffer after a step
        ok &= all(torch.allclose(a, b, rtol=1e-5, atol=1e-7) for a, b in zip(net.parameters(), whole.parameters()))
    np.save(os.path.join(tmp, "grad_ok_%d.npy" % rank), np.array([int(ok)]))
    dist.destroy_process_group()


def test_flat_gradient_allreduce_gloo(tmp_path):
    world = 2
    port = 31500 + os.getpid() % 2000
    mp.spawn(_grad_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    assert all(np.load(tmp_path / ("grad_ok_%d.npy" % r))[0] == 1 for r in range(world))
