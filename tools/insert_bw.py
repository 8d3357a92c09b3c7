#!/usr/bin/env python
"""K6 insert bandwidth: records/s and GB/s (2 x frame bytes per record) for several chunk sizes."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import apex_b200  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda:0")
N, C, F, lanes = 1 << 21, 4, 7056, 32
mem = apex_b200.ReplayMemory(N, {"priority_exponent": 0.6, "layout": "stream", "lanes": lanes, "n_step": 3, "state_shape": (C, 84, 84)})
mem._ensure()
for m in (5, 160, 1024, 8192, 32768):
    frames = torch.randint(0, 256, (m, F), dtype=torch.uint8, device=dev)
    rel = torch.as_tensor(apex_b200.stream_offsets(C, np.full(m, 3), np.full(m, 100))).to(dev)
    A = torch.zeros(m, dtype=torch.int32, device=dev)
    R = torch.zeros(m, device=dev)
    G = torch.ones(m, device=dev)
    prio = torch.rand(m, device=dev) + 0.01
    for _ in range(3):
        mem.insert(0, frames, rel, A, R, G, prio=prio)
    torch.cuda.synchronize()
    reps = 40
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(reps):
        mem.insert(k % lanes, frames, rel, A, R, G, prio=prio)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    print("insert %6d records: %8.2f us per call  %7.2f M records/s  %7.1f GB/s (read+write of frames)"
          % (m, us, m / us, 2 * m * F / us / 1e3))
mem.check_status()
