#!/usr/bin/env python
"""Multi-GPU parity check, one process per GPU (not collected by pytest):

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tests/multi_gpu_check.py

Every rank owns a shard; both sharded paths — NCCL all-gather of shard totals + K8 routing + local
K3/K4/K5/K2, and the default peer-memory exchange + fused sharded kernel (+ push from the TD/update
kernel) — must return, over all ranks together, exactly the records a single tree over the
concatenated shards returns for the same draws, byte-exact batches, and a loss whose sum over ranks
equals the single-batch mean.  The shards are deliberately uneven (rank 0 carries 8x the priority
mass): every draw a rank owns must arrive, however large its share.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import apex_b200                                   # noqa: E402
from apex_b200 import _C                           # noqa: E402
from oracle import replay_oracle as ro             # noqa: E402
from oracle import td_oracle as to                 # noqa: E402


def shard_data(rank, n, C, F, lanes):
    rng = np.random.default_rng(500 + rank)
    W = C - 1 + 3
    ring = ro.RingOracle(n, C, F, lanes=lanes, fpr=1, lookback=W, alpha=0.6)
    ops = []
    for lane in range(lanes):
        m = n // lanes - W
        frames = rng.integers(0, 256, (m, F), dtype=np.uint8)
        rel = apex_b200.stream_offsets(C, np.full(m, 3), np.minimum(np.arange(m), 500))
        A = rng.integers(0, 4, m).astype(np.int32)
        R = rng.standard_normal(m).astype(np.float32)
        G = np.full(m, 0.97, np.float32)
        # uneven shards: rank 0 carries 32^0.6 = 8x the priority mass of a unit-scale shard, so it owns far
        # more than its fair share of every global batch (several waves of the fused sharded kernel)
        prio = (np.abs(rng.standard_normal(m)) * (32.0 if rank == 0 else 1 + rank) + 1e-3).astype(np.float32)
        ring.insert(lane, frames, rel, A, R, G, prio=prio)
        ops.append((lane, frames, rel, A, R, G, prio))
    return ring, ops


def sharded_learner_check(rank, world, dev):
    """BASELINE configs[3] in small: data-parallel learner over a sharded replay.  The all-reduced
    gradient equals the gradient of one process evaluating the whole global batch (torch autograd on
    the all-gathered rows), every replica holds the same weights after the step, and the ranks'
    losses add up to the global mean."""
    import copy
    tf32 = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    ok = True
    n, C, lanes, B, nA = 4096, 1, 4, 32, 4
    mem = apex_b200.ReplayMemory(n, {"priority_exponent": 0.6, "layout": "stream", "lanes": lanes, "n_step": 3,
                                     "state_shape": (C, 84, 84), "seed": 99})
    g = torch.Generator(device=dev).manual_seed(500 + rank)
    m = n // lanes - 8                                               # one insert call must leave the lookback zone free
    rel = torch.as_tensor(apex_b200.stream_offsets(C, np.full(m, 3), np.minimum(np.arange(m) + 10, 200))).to(dev)
    for lane in range(lanes):
        mem.insert(lane, torch.randint(0, 256, (m, 7056), dtype=torch.uint8, device=dev, generator=g), rel,
                   torch.randint(0, nA, (m,), dtype=torch.int32, device=dev, generator=g),
                   torch.randn(m, device=dev, generator=g), torch.full((m,), 0.97, device=dev),
                   prio=torch.rand(m, device=dev, generator=g) * (32.0 if rank == 0 else 1 + rank) + 0.01)   # rank 0: 8x the mass
    torch.manual_seed(1000 + rank)                                   # replicas start different: setup() must equalise them
    lp = {"remove_old_xp_freq": 100, "q_target_sync_freq": 2500, "min_replay_mem_size": 10, "replay_sample_size": B,
          "load_saved_state": False}
    L = apex_b200.ShardedLearner({"state_shape": [C, 84, 84], "action_dim": nA}, lp, {}, mem)
    L.Q.batch_coupled_mean = L.Q_double.batch_coupled_mean = False   # per-row duelling mean: rows are independent (Q19)
    L.setup()
    Bg, rows = B * world, L.max_rows

    def checksum(net):
        s = torch.stack([p.double().sum() for p in net.state_dict().values()]).sum().reshape(1)
        both = [torch.zeros_like(s) for _ in range(world)]
        dist.all_gather(both, s)
        return [float(x) for x in both]
    equal_start = len(set(checksum(L.Q))) == 1 and len(set(checksum(L.Q_double))) == 1
    if not equal_start:
        print("sharded_learner_check rank %d: replicas differ after setup()" % rank, flush=True)
    ok &= equal_start
    for step in range(2):
        q_before, qd_before = copy.deepcopy(L.Q), copy.deepcopy(L.Q_double)
        loss = L.step_sharded()
        o = mem._outputs(rows, False, L._f32())                      # the learner's batch (float32 straight from the gather)
        cnt = o["count"].clone()
        parts = []
        for name in ("S_t", "S_tpn", "A", "R", "Gamma"):
            mine = o[name].clone()
            every = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(every, mine)
            parts.append(every)
        counts = [torch.zeros_like(cnt) for _ in range(world)]
        dist.all_gather(counts, cnt)
        counts = [int(c.item()) for c in counts]
        report = {"counts": sum(counts) == Bg}
        S_t, S_tpn, A, R, G = (torch.cat([e[r][:counts[r]] for r in range(world)]) for e in parts)
        S_t, S_tpn = S_t.view(Bg, C, 84, 84).float(), S_tpn.view(Bg, C, 84, 84).float()
        with torch.no_grad():
            q_tg, q_on_n = qd_before(S_tpn)[2], q_before(S_tpn)[2]
        q_on = q_before(S_t)[2]
        target = R + G * q_tg.gather(1, q_on_n.argmax(1, keepdim=True)).squeeze(1)
        delta = target - q_on.gather(1, A.long().view(-1, 1)).squeeze(1)
        ref_loss = (0.5 * delta * delta).mean()
        ref_loss.backward()
        ref_grad = torch.cat([p.grad.reshape(-1) for p in q_before.parameters()])
        got = L.grads.flat
        scale = float(ref_grad.abs().max())
        report["grad"] = bool(((got - ref_grad).abs().max() <= 2e-3 * scale + 1e-7).item())
        total = loss.clone().reshape(1)
        dist.all_reduce(total)
        ref_loss = float(ref_loss.detach())
        report["loss"] = abs(float(total) - ref_loss) <= 1e-4 * abs(ref_loss) + 1e-6
        report["replicas"] = len(set(checksum(L.Q))) == 1
        if not all(report.values()):
            print("sharded_learner_check rank %d step %d: %s  grad err %.3e of %.3e, loss %.6f vs %.6f" %
                  (rank, step, report, float((got - ref_grad).abs().max()), scale, float(total), ref_loss), flush=True)
        ok &= all(report.values())
    mem.check_status()
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32
    return bool(ok)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    n, C, F, lanes, Bg, nA = 1 << 14, 2, 256, 4, 256 * world, 4
    rings = [shard_data(r, n, C, F, lanes) for r in range(world)]            # every rank can rebuild every shard
    ring, ops = rings[rank]
    mem = apex_b200.ReplayMemory(n, {"priority_exponent": 0.6, "layout": "stream", "lanes": lanes, "n_step": 3,
                                     "state_shape": (C, 1, F)})
    for op in ops:
        mem.insert(*op[:6], prio=op[6])
    # identical leaves on both sides: take the device's (the fp32 pow is a separate, toleranced item)
    leaves_local = mem.buffer(_C.BUF_LEAVES).cpu().numpy().view(np.uint64).copy()
    all_leaves = [torch.zeros(n, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(all_leaves, torch.as_tensor(leaves_local.view(np.int64)).to(dev))
    all_leaves = [t.cpu().numpy().view(np.uint64) for t in all_leaves]
    for r in range(world):
        rings[r][0].leaves[:] = all_leaves[r]
    sampler = apex_b200.ShardedSampler(mem)
    ok = True
    rng = np.random.default_rng(9)                                           # same draws on every rank
    for step in range(4):
        u = ro.stratified_uniforms(Bg, rng) if step % 2 else rng.random(Bg)
        batch, pos, cnt = sampler.sample(torch.as_tensor(u).to(dev))
        cnt_h = int(cnt.item())
        whole = ro.tree_sample(np.concatenate([rings[r][0].leaves for r in range(world)]), u)
        mine = np.nonzero(whole // n == rank)[0]
        ok &= cnt_h == len(mine) and (pos.cpu().numpy()[:cnt_h] == mine).all()
        local_idx = batch.idx.cpu().numpy()[:cnt_h]
        ok &= (local_idx == whole[mine] % n).all()
        g = ring.gather(local_idx)
        ok &= (batch.S_t.cpu().numpy()[:cnt_h] == g["S_t"]).all() and (batch.S_tpn.cpu().numpy()[:cnt_h] == g["S_tpn"]).all()
        ok &= (batch.A.cpu().numpy()[:cnt_h] == g["A"]).all()
        # K5+K2 on the local rows; loss_mean divided by the GLOBAL batch so that the sum over ranks is the mean
        qrng = np.random.default_rng(1000 + step)
        qg = [qrng.standard_normal((Bg, nA)).astype(np.float32) for _ in range(3)]          # one Q row per global draw
        q = [torch.as_tensor(np.concatenate([x[mine], np.zeros((Bg - cnt_h, nA), np.float32)])).to(dev) for x in qg]
        delta, prio, loss, mean = mem.td_update(batch.idx, q[0], q[1], q[2], batch.A, batch.R, batch.Gamma,
                                                count_dev=cnt, mean_divisor=float(Bg))
        total_loss = mean.clone()
        dist.all_reduce(total_loss)
        # oracle: the whole global batch at once
        gl = [rings[int(i // n)][0] for i in whole]
        A_g = np.array([r_.A[i % n] for r_, i in zip(gl, whole)])
        R_g = np.array([r_.R[i % n] for r_, i in zip(gl, whole)])
        G_g = np.array([r_.G[i % n] for r_, i in zip(gl, whole)])
        o = to.td_forward(qg[0], qg[1], qg[2], A_g, R_g, G_g, alpha=0.6)
        ok &= abs(total_loss.item() - o["loss_mean"]) <= 1e-5 * abs(o["loss_mean"]) + 1e-7
        ok &= np.allclose(prio.cpu().numpy()[:cnt_h], o["prio"][mine], rtol=1e-5, atol=1e-7)
        # write the same priorities into the oracle shard (last writer wins) and compare shard totals
        ring.set_leaves(local_idx, ro.priority_to_leaf(prio.cpu().numpy()[:cnt_h]))
        ok &= mem.total() == ring.total()
        new_leaves = [torch.zeros(n, dtype=torch.int64, device=dev) for _ in range(world)]
        dist.all_gather(new_leaves, torch.as_tensor(ring.leaves.view(np.int64)).to(dev))
        for r in range(world):
            rings[r][0].leaves[:] = new_leaves[r].cpu().numpy().view(np.uint64)
    # ---- peer-memory exchange + fused routing/search/gather (no NCCL on the step path)
    sampler.connect_peers()
    max_rows = Bg                        # the outputs hold the whole global batch: no draw can be dropped
    totals = torch.zeros(world, dtype=torch.int64, device=dev)
    for rnd in range(1, 6):
        sampler.exchange(totals)
        want_tot = [int(rings[r][0].leaves.sum(dtype=np.uint64)) for r in range(world)]
        ok &= [int(x) for x in totals.cpu().numpy().view(np.uint64)] == want_tot
        u_out = torch.zeros(max_rows, dtype=torch.float64, device=dev)
        batch, pos, cnt = sampler.sample_fused(Bg, max_rows, u_out=u_out)
        cnt_h = int(cnt.item())
        # reassemble the global draw vector from what every rank reports for its own rows
        u_glob = torch.zeros(Bg, dtype=torch.float64, device=dev)
        owner = torch.zeros(Bg, dtype=torch.int64, device=dev)
        u_glob[pos[:cnt_h].long()] = u_out[:cnt_h]
        owner[pos[:cnt_h].long()] = 1
        dist.all_reduce(u_glob)
        dist.all_reduce(owner)
        ok &= bool((owner == 1).all().item())                                # every draw owned by exactly one rank
        u = u_glob.cpu().numpy()
        ok &= bool((u == ro.kernel_stratified_uniforms(mem.seed, rnd, rnd - 1, Bg)).all())
        whole = ro.tree_sample(np.concatenate([rings[r][0].leaves for r in range(world)]), u)
        mine = np.nonzero(whole // n == rank)[0]
        ok &= cnt_h == len(mine) and (pos.cpu().numpy()[:cnt_h] == mine).all()
        local_idx = batch.idx.cpu().numpy()[:cnt_h]
        ok &= (local_idx == whole[mine] % n).all()
        g = ring.gather(local_idx)
        ok &= (batch.S_t.cpu().numpy()[:cnt_h] == g["S_t"]).all() and (batch.S_tpn.cpu().numpy()[:cnt_h] == g["S_tpn"]).all()
        q = [torch.randn(max_rows, nA, device=dev) for _ in range(3)]
        delta, prio, loss, mean = mem.td_update(batch.idx, q[0], q[1], q[2], batch.A, batch.R, batch.Gamma,
                                                count_dev=cnt, mean_divisor=float(Bg))
        ring.set_leaves(local_idx, ro.priority_to_leaf(prio.cpu().numpy()[:cnt_h]))
        ok &= mem.total() == ring.total()
        new_leaves = [torch.zeros(n, dtype=torch.int64, device=dev) for _ in range(world)]
        dist.all_gather(new_leaves, torch.as_tensor(ring.leaves.view(np.int64)).to(dev))
        for r in range(world):
            rings[r][0].leaves[:] = new_leaves[r].cpu().numpy().view(np.uint64)
    # ---- push-on-update: the TD/update kernel publishes the new total itself, no exchange() per step
    sampler.push_on_update(True)
    sampler.exchange()
    rnd, call = 6, 5
    for step in range(6):
        u_out = torch.zeros(max_rows, dtype=torch.float64, device=dev)
        batch, pos, cnt = sampler.sample_fused(Bg, max_rows, u_out=u_out)
        cnt_h = int(cnt.item())
        u_glob = torch.zeros(Bg, dtype=torch.float64, device=dev)
        u_glob[pos[:cnt_h].long()] = u_out[:cnt_h]
        dist.all_reduce(u_glob)
        u = u_glob.cpu().numpy()
        ok &= bool((u == ro.kernel_stratified_uniforms(mem.seed, rnd, call, Bg)).all())
        whole = ro.tree_sample(np.concatenate([rings[r][0].leaves for r in range(world)]), u)
        mine = np.nonzero(whole // n == rank)[0]
        ok &= cnt_h == len(mine) and (pos.cpu().numpy()[:cnt_h] == mine).all()
        local_idx = batch.idx.cpu().numpy()[:cnt_h]
        ok &= (local_idx == whole[mine] % n).all()
        q = [torch.randn(max_rows, nA, device=dev) for _ in range(3)]
        delta, prio, loss, mean = mem.td_update(batch.idx, q[0], q[1], q[2], batch.A, batch.R, batch.Gamma,
                                                count_dev=cnt, mean_divisor=float(Bg))       # pushes round rnd+1
        ring.set_leaves(local_idx, ro.priority_to_leaf(prio.cpu().numpy()[:cnt_h]))
        ok &= mem.total() == ring.total()
        new_leaves = [torch.zeros(n, dtype=torch.int64, device=dev) for _ in range(world)]
        dist.all_gather(new_leaves, torch.as_tensor(ring.leaves.view(np.int64)).to(dev))
        for r in range(world):
            rings[r][0].leaves[:] = new_leaves[r].cpu().numpy().view(np.uint64)
        rnd, call = rnd + 1, call + 1
        if step == 2:                                        # an insert changes one shard's total: re-exchange
            m = 40
            ins = (rank % lanes, np.random.default_rng(77 + rank).integers(0, 256, (m, F), dtype=np.uint8),
                   apex_b200.stream_offsets(C, np.full(m, 3), np.full(m, 50)), np.zeros(m, np.int32),
                   np.zeros(m, np.float32), np.ones(m, np.float32))
            pr = np.full(m, 1.0 + rank, np.float32)
            mem.insert(*ins, prio=pr)
            ring.insert(*ins, prio=pr)
            ring.leaves[:] = mem.buffer(_C.BUF_LEAVES).cpu().numpy().view(np.uint64)
            dist.all_gather(new_leaves, torch.as_tensor(ring.leaves.view(np.int64)).to(dev))
            for r in range(world):
                rings[r][0].leaves[:] = new_leaves[r].cpu().numpy().view(np.uint64)
            sampler.exchange()
            rnd += 1
    mem.check_status()
    if not ok:
        print("multi_gpu_check rank %d: replay sections failed" % rank, flush=True)
    ok &= sharded_learner_check(rank, world, dev)
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("multi_gpu_check world=%d: %s" % (world, "PASS" if flag.item() == 1 else "FAIL"))
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
