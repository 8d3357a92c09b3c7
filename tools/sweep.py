#!/usr/bin/env python
"""BASELINE configs[4]: capacity x batch sweep of the sample+gather+TD+update step on one GPU.
Prints one JSON line per point: us per step (CUDA graph of 8 steps), transitions/s, and the
fused kernel's algorithmic GB/s.   python tools/sweep.py [--capacities 65536,...] [--batches 32,...]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import apex_b200  # noqa: E402
import bench  # noqa: E402
from tools.kernel_times import timed_graph  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--capacities", default="65536,262144,1048576,2097152,4194304,16777216")
    ap.add_argument("--batches", default="32,128,512,2048,4096")
    a = ap.parse_args()
    torch.cuda.set_device(0)
    dev = torch.device("cuda:0")
    apex_b200._C.call("apexb200_set_pdl", int(os.environ.get("APEX_PDL", "1")))
    for N in [int(x) for x in a.capacities.split(",")]:
        args = argparse.Namespace(capacity=N, lanes=32, frames_per_state=4, actions=4, batch=512)
        mem = apex_b200.ReplayMemory(N, {"priority_exponent": 0.6, "layout": "stream", "lanes": 32, "n_step": 3,
                                         "state_shape": (4, 84, 84), "seed": 1})
        bench.fill_shard(torch, apex_b200, mem, args, dev, 1234)
        for B in [int(x) for x in a.batches.split(",")]:
            q = [torch.randn(B, 4, device=dev) for _ in range(3)]
            out = tuple(torch.empty(B, device=dev) for _ in range(3)) + (torch.empty(1, device=dev),)
            mem.workspace(B)
            mem.sample_tensors(B)

            def step():
                b = mem.sample_tensors(B)
                mem.td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma, out=out)
            t_step = timed_graph(step, R=8, reps=20)
            t_fused = timed_graph(lambda: mem.sample_tensors(B), R=8, reps=20)
            print(json.dumps({"capacity": N, "batch": B, "us_per_step": round(t_step, 2),
                              "transitions_per_s": round(B / t_step * 1e6), "fused_kernel_us": round(t_fused, 2),
                              "fused_algorithmic_GBps": round(B * (4 * 4 * 7056 + 32) / t_fused / 1e3, 1)}), flush=True)
        mem.check_status()
        mem.close()
        del mem
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
