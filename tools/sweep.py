#!/usr/bin/env python
"""BASELINE configs[4]: capacity x batch sweep of the sample+gather+TD+update step at 1/2/4/8 GPUs, with the
reference's CPU replay path (numpy/torch port and the C/OpenMP port of the oracle) timed beside it.

    python tools/sweep.py [--capacities 65536,...] [--batches 32,...] [--cpu]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/sweep.py ...

One JSON line per point: us per step (CUDA graph of 8 steps, max over ranks), whole-job transitions/s (capacity and
batch are PER GPU: weak scaling), the fused kernel's time and algorithmic GB/s.  With --cpu (rank 0, capacities up to
2M records: the frame store must fit host RAM) the same step on the host cores: records/s of both ports."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import apex_b200  # noqa: E402
import bench  # noqa: E402


_HOST = {}


def host_shard(N, C=4, nA=4):
    """Host-resident synthetic shard of N records (built once per capacity)."""
    if N in _HOST:
        return _HOST[N]
    _HOST.clear()                                        # one frame store at a time in host RAM
    from oracle import replay_oracle as ro
    rng = np.random.default_rng(1234)
    block = rng.integers(0, 256, (1 << 12, bench.F_BYTES), dtype=np.uint8)
    frames = np.empty((N, bench.F_BYTES), np.uint8)
    for lo in range(0, N, 1 << 12):
        frames[lo:min(N, lo + (1 << 12))] = block[: min(N, lo + (1 << 12)) - lo]
    lane_cap = N // 32
    pos = np.arange(N) % lane_cap
    rel = np.concatenate([np.arange(-3 - C + 1, -2), np.arange(-C + 1, 1)])[None, :]
    fidx = ((np.arange(N) // lane_cap * lane_cap)[:, None] + (pos[:, None] + rel) % lane_cap).astype(np.int32)
    A = rng.integers(0, nA, N).astype(np.int32)
    R = np.clip(rng.standard_normal(N), -1, 1).astype(np.float32)
    G = np.where(rng.random(N) < 0.01, 0, bench.GAMMA_N).astype(np.float32)
    leaves = ro.priority_to_leaf(ro.pow_priority(np.abs(rng.standard_normal(N)) + 1e-3, bench.ALPHA))
    _HOST[N] = (leaves, frames, fidx, A, R, G)
    return _HOST[N]


def cpu_point(N, B, C=4, nA=4, seconds=1.5):
    from oracle.c_oracle import CpuReplay
    leaves, frames, fidx, A, R, G = host_shard(N, C, nA)
    rng = np.random.default_rng(7)
    out = {}
    for name, obj in (("numpy_torch_port", bench.NumpyReplay(leaves.copy(), frames, fidx, A, R, G, C, bench.ALPHA)),
                      ("c_openmp_port", CpuReplay(leaves.copy(), frames, fidx, A, R, G, C, bench.F_BYTES, bench.ALPHA))):
        n, dt = bench.time_cpu(obj, B, nA, seconds, 2000, rng)
        out[name] = {"transitions_per_s": round(n * B / dt, 1), "threads": obj.threads}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--capacities", default="65536,262144,1048576,2097152,4194304,16777216")
    ap.add_argument("--batches", default="32,128,512,2048,4096")
    ap.add_argument("--cpu", action="store_true")
    a = ap.parse_args()
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    apex_b200._C.call("apexb200_set_pdl", int(os.environ.get("APEX_PDL", "1")))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_graph(fn, R=8, reps=20):
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                for _ in range(R):
                    fn()
        torch.cuda.current_stream().wait_stream(side)
        for _ in range(3):
            g.replay()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        barrier()
        us = e0.elapsed_time(e1) * 1e3 / (R * reps)
        if world > 1:
            t = torch.tensor([us], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            us = float(t.item())
        return us

    for N in [int(x) for x in a.capacities.split(",")]:
        args = argparse.Namespace(capacity=N, lanes=32, frames_per_state=4, actions=4, batch=512)
        mem = apex_b200.ReplayMemory(N, {"priority_exponent": 0.6, "layout": "stream", "lanes": 32, "n_step": 3,
                                         "state_shape": (4, 84, 84), "seed": 1})
        bench.fill_shard(torch, apex_b200, mem, args, dev, 1234 + rank)
        sampler = None
        if world > 1:
            sampler = apex_b200.ShardedSampler(mem, rank=rank, world_size=world)
            sampler.connect_peers()
            sampler.push_on_update(True)
            sampler.exchange()
        for B in [int(x) for x in a.batches.split(",")]:
            Bg = B * world
            rows = B if world == 1 else Bg
            q = [torch.randn(rows, 4, device=dev) for _ in range(3)]
            out = tuple(torch.empty(rows, device=dev) for _ in range(3)) + (torch.empty(1, device=dev),)
            mem.workspace(rows)

            if world == 1:
                def step():
                    b = mem.sample_tensors(B)
                    mem.td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma, out=out)

                def fused():
                    mem.sample_tensors(B)
            else:
                def step():
                    b, _pos, cnt = sampler.sample_fused(Bg)
                    mem.td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma, out=out, count_dev=cnt, mean_divisor=float(Bg))

                def fused():
                    sampler.sample_fused(Bg)
            step()
            t_step = timed_graph(step)
            if world > 1:
                sampler.push_on_update(False)
            t_fused = timed_graph(fused)
            if world > 1:
                sampler.push_on_update(True)
                sampler.exchange()
            line = {"n_gpus": world, "capacity_per_gpu": N, "batch_per_gpu": B, "us_per_step": round(t_step, 2),
                    "transitions_per_s": round(Bg / t_step * 1e6), "fused_kernel_us": round(t_fused, 2),
                    "fused_algorithmic_GBps_per_gpu": round(B * (4 * 4 * 7056 + 32) / t_fused / 1e3, 1)}
            if a.cpu and rank == 0 and N <= (1 << 21):
                t0 = time.time()
                line["cpu"] = cpu_point(N, B)
                line["cpu"]["seconds_spent"] = round(time.time() - t0, 1)
            if a.cpu and world > 1:
                dist.barrier()
            if rank == 0:
                print(json.dumps(line), flush=True)
        mem.check_status()
        mem.close()
        del mem, sampler
        torch.cuda.empty_cache()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
