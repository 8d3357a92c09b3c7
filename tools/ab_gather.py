#!/usr/bin/env python
"""Same-box A/B of fused sample+gather variants (apexb200_set_gather_variant): per-launch time of the fused
kernel and of the full step from CUDA graphs, plus a byte-for-byte comparison of the variants' batches for the
same draws.  Usage: python tools/ab_gather.py [--variants 1,2,4] [--batch 512]"""
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
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--capacity", type=int, default=1 << 21)
    ap.add_argument("--variants", default="1,2,4")
    ap.add_argument("--rounds", type=int, default=2)
    a = ap.parse_args()
    args = argparse.Namespace(capacity=a.capacity, lanes=32, frames_per_state=4, actions=4, batch=a.batch)
    torch.cuda.set_device(0)
    dev = torch.device("cuda:0")
    mem = apex_b200.ReplayMemory(args.capacity, {"priority_exponent": 0.6, "layout": "stream", "lanes": 32, "n_step": 3,
                                                 "state_shape": (4, 84, 84), "seed": 1})
    bench.fill_shard(torch, apex_b200, mem, args, dev, 1234)
    B = a.batch
    q = [torch.randn(B, 4, device=dev) for _ in range(3)]
    out = tuple(torch.empty(B, device=dev) for _ in range(3)) + (torch.empty(1, device=dev),)
    mem.workspace(B)
    variants = [int(v) for v in a.variants.split(",")]
    u = (torch.arange(B, device=dev, dtype=torch.float64) + torch.rand(B, device=dev, dtype=torch.float64)) / B
    ref = None
    for v in variants:
        apex_b200._C.call("apexb200_set_gather_variant", v)
        b = mem.sample_tensors(B, u=u)
        cur = [x.clone() for x in (b.idx, b.S_t, b.S_tpn, b.A, b.R, b.Gamma)]
        if ref is None:
            ref = cur
        else:
            print("variant %d equals variant %d byte for byte: %s" % (v, variants[0], all(torch.equal(x, y) for x, y in zip(ref, cur))))

    def step():
        b = mem.sample_tensors(B)
        mem.td_update(b.idx, q[0], q[1], q[2], b.A, b.R, b.Gamma, out=out)
    for r in range(a.rounds):
        for v in variants:
            apex_b200._C.call("apexb200_set_gather_variant", v)
            k_us = timed_graph(lambda: mem.sample_tensors(B))
            s_us = timed_graph(step)
            print(json.dumps({"round": r, "variant": v, "fused_kernel_us": round(k_us, 3), "step_us": round(s_us, 3)}))
    mem.check_status()


if __name__ == "__main__":
    main()
