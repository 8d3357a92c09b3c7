#!/usr/bin/env python
"""Where the end-to-end step (HostStepPipeline) spends its host time: submit (fill pinned inputs + one C call)
vs. blocked in result().  python tools/e2e_breakdown.py [--capacity N]"""
import argparse
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import apex_b200                                   # noqa: E402
import bench                                       # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--capacity", type=int, default=1 << 21)
    ap.add_argument("--steps", type=int, default=3000)
    ap.add_argument("--depth", type=int, default=3)
    a = ap.parse_args()
    args = argparse.Namespace(capacity=a.capacity, lanes=32, frames_per_state=4, actions=4, batch=512)
    dev = torch.device("cuda:0")
    torch.cuda.set_device(0)
    mem = apex_b200.ReplayMemory(args.capacity, {"priority_exponent": 0.6, "layout": "stream", "lanes": 32, "n_step": 3,
                                                 "state_shape": (4, 84, 84), "seed": 4321})
    bench.fill_shard(torch, apex_b200, mem, args, dev, 1234)
    B, nA, P = 512, 4, 4
    rng = np.random.default_rng(1)
    pipe = apex_b200.HostStepPipeline(mem, B, nA, depth=a.depth)
    q_host = rng.standard_normal((P, 3, B, nA)).astype(np.float32)
    u_host = (np.arange(B) + rng.random((64, B))) / B
    for k in range(16):
        pipe.result(pipe.submit(u_host[k % 64], q_host[k % P]))
    torch.cuda.synchronize()
    pc = time.perf_counter
    t_submit = t_result = 0.0
    t0 = pc()
    import collections
    pending = collections.deque()
    for k in range(a.steps):
        s0 = pc()
        pending.append(pipe.submit(u_host[k % 64], q_host[k % P]))
        s1 = pc()
        if len(pending) == pipe.DEPTH:
            pipe.result(pending.popleft())
        s2 = pc()
        t_submit += s1 - s0
        t_result += s2 - s1
    while pending:
        pipe.result(pending.popleft())
    dt = pc() - t0
    # host-only cost of the C call: same call with the GPU idle in between (sync each step)
    c_call = 0.0
    s = pipe.slots[0]
    for k in range(300):
        torch.cuda.synchronize()
        s0 = pc()
        rc = pipe._host_step(*s["args"])
        c_call += pc() - s0
        assert rc == 0
    torch.cuda.synchronize()
    fill = 0.0
    for k in range(2000):
        s0 = pc()
        s["u_host"][:] = u_host[k % 64]
        s["q_host"][:] = q_host[k % P]
        fill += pc() - s0
    # device-side view: an event on the kernel stream before and after every submit
    n = 400
    before = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
    after = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
    for k in range(n):
        before[k].record()
        pending.append(pipe.submit(u_host[k % 64], q_host[k % P]))
        after[k].record()
        if len(pending) == pipe.DEPTH:
            pipe.result(pending.popleft())
    while pending:
        pipe.result(pending.popleft())
    torch.cuda.synchronize()
    period = np.array([after[k].elapsed_time(after[k + 1]) for k in range(50, n - 1)]) * 1e3
    span = np.array([before[k].elapsed_time(after[k]) for k in range(50, n)]) * 1e3
    print("device: period between step ends %.2f us (median %.2f) | span from previous step's end to this step's end %.2f us "
          "(median %.2f)" % (period.mean(), np.median(period), span.mean(), np.median(span)))
    print("per step: total %.2f us | submit %.2f us (fill %.2f us, C call with idle GPU %.2f us) | blocked+result %.2f us" %
          (dt / a.steps * 1e6, t_submit / a.steps * 1e6, fill / 2000 * 1e6, c_call / 300 * 1e6, t_result / a.steps * 1e6))


if __name__ == "__main__":
    main()
