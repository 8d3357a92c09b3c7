#!/usr/bin/env python
"""Per-kernel device time on the bench workload, without per-launch event overhead:
a CUDA graph of R back-to-back launches of ONE kernel is replayed and timed as a whole
(time / R = kernel + one inter-node gap).  Usage: python tools/kernel_times.py [--batch 512]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import apex_b200  # noqa: E402
import bench  # noqa: E402


def timed_graph(fn, R=20, reps=30):
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
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (R * reps)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--capacity", type=int, default=1 << 21)
    ap.add_argument("--pdl", type=int, default=1)
    a = ap.parse_args()
    args = argparse.Namespace(capacity=a.capacity, lanes=32, frames_per_state=4, actions=4, batch=a.batch)
    torch.cuda.set_device(0)
    dev = torch.device("cuda:0")
    apex_b200._C.call("apexb200_set_pdl", a.pdl)
    mem = apex_b200.ReplayMemory(args.capacity, {"priority_exponent": 0.6, "layout": "stream", "lanes": 32, "n_step": 3,
                                                 "state_shape": (4, 84, 84), "seed": 1})
    bench.fill_shard(torch, apex_b200, mem, args, dev, 1234)
    B = a.batch
    o = mem._outputs(B, False)
    q = [torch.randn(B, 4, device=dev) for _ in range(3)]
    out = tuple(torch.empty(B, device=dev) for _ in range(3)) + (torch.empty(1, device=dev),)
    mem.workspace(B)
    mem.sample_tensors(B)
    res = {}
    res["sample_gather (fused K3+K4)"] = timed_graph(lambda: mem.sample_tensors(B))
    res["tree_sample (K3)"] = timed_graph(lambda: mem.sample_stratified(B, idx_out=o["idx"], fsel_out=o["fsel"]))
    for v in (1, 0):
        apex_b200._C.call("apexb200_set_gather_variant", v)
        res["gather variant %d (K4)" % v] = timed_graph(lambda: mem.gather(o["idx"], fsel=o["fsel"], out=o))
    apex_b200._C.call("apexb200_set_gather_variant", 1)
    res["td_small (K5+K2)"] = timed_graph(lambda: mem.td_update(o["idx"], q[0], q[1], q[2], o["A"], o["R"], o["Gamma"], out=out))
    res["tree_update (K2 only)"] = timed_graph(lambda: mem.update_priorities(o["idx"], out[0]))
    tot = torch.zeros(1, dtype=torch.int64, device=dev)
    res["tree_total"] = timed_graph(lambda: mem.total_tensor(tot))

    def step():
        b = mem.sample_tensors(B)
        mem.td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma, out=out)
    res["full step (fused + td)"] = timed_graph(step)
    # the sharded form of the two kernels with a group of one (self exchange): what the multi-GPU step runs per rank
    sampler = apex_b200.ShardedSampler(mem, rank=0, world_size=1)
    sampler.connect_peers()
    sampler.exchange()
    res["sample_gather_sharded (world 1)"] = timed_graph(lambda: sampler.sample_fused(B))
    sampler.push_on_update(True)

    def sharded_step():
        b, _pos, cnt = sampler.sample_fused(B)
        mem.td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma, out=out, count_dev=cnt, mean_divisor=float(B))
    res["full sharded step (world 1)"] = timed_graph(sharded_step)
    sampler.push_on_update(False)
    for k, v in res.items():
        print("%-32s %8.2f us" % (k, v))
    mem.check_status()


if __name__ == "__main__":
    main()
