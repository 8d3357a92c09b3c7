#!/usr/bin/env python
"""K3 latency vs tree depth: graph of back-to-back tree_sample launches (tiny frames, no gather)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import apex_b200  # noqa: E402
from tools.kernel_times import timed_graph  # noqa: E402

torch.cuda.set_device(0)
for N in (32, 1024, 32768, 1 << 20, 1 << 21, 1 << 24):
    mem = apex_b200.ReplayMemory(N, {"priority_exponent": 0.6, "state_shape": (1, 1, 16)})
    mem._ensure()
    leaves = torch.randint(1, 1 << 28, (N,), dtype=torch.int64, device="cuda")
    apex_b200._C.call("apexb200_tree_load_leaves", mem._h, apex_b200._C.ptr(leaves), 0, N, apex_b200._C.stream_ptr())
    mem.update_sample_probabilities()
    row = []
    for B in (32, 512, 4096):
        idx = torch.empty(B, dtype=torch.int32, device="cuda")
        u = torch.rand(B, dtype=torch.float64, device="cuda")
        t_strat = timed_graph(lambda: mem.sample_stratified(B, idx_out=idx))
        t_u = timed_graph(lambda: mem.sample_indices(u, idx_out=idx))
        row.append("B=%d strat %.2f given-u %.2f" % (B, t_strat, t_u))
    print("N=%-9d levels=%d  " % (N, len(__import__("oracle").replay_oracle.level_sizes(N))), " | ".join(row))
    del mem
