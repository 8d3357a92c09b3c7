#!/usr/bin/env python
"""Small driver for an ncu capture of K3 alone (2M-leaf tree, 512 draws per launch)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import apex_b200  # noqa: E402

torch.cuda.set_device(0)
N, B = 1 << 21, 512
mem = apex_b200.ReplayMemory(N, {"priority_exponent": 0.6, "state_shape": (1, 1, 16)})
mem._ensure()
leaves = torch.randint(1, 1 << 28, (N,), dtype=torch.int64, device="cuda")
apex_b200._C.call("apexb200_tree_load_leaves", mem._h, apex_b200._C.ptr(leaves), 0, N, apex_b200._C.stream_ptr())
mem.update_sample_probabilities()
idx = torch.empty(B, dtype=torch.int32, device="cuda")
for k in range(24):
    u = torch.rand(B, dtype=torch.float64, device="cuda")
    mem.sample_indices(u, idx_out=idx)
torch.cuda.synchronize()
mem.check_status()
print("ok")
