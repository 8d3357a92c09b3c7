#!/usr/bin/env python
"""Priority kernels against HBM bandwidth at sizes where bytes, not latency, dominate:
K1 tree build over 16M leaves, K2 point update and K5(+K2) TD/priority kernels over large batches.
Algorithmic bytes are stated per kernel; times are CUDA events over back-to-back launches."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import apex_b200  # noqa: E402
import bench  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda:0")
peak, peak_src = bench.peaks()


def timed(fn, R=8, reps=10):
    """Device time per call: a CUDA graph of R back-to-back calls, replayed (no host launch cost in the number)."""
    fn()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for _ in range(R):
                fn()
    torch.cuda.current_stream().wait_stream(side)
    for _ in range(2):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (R * reps)


out = []
N, nA = 1 << 24, 4
# tiny frames: this tool is about the tree and the TD arithmetic, not the ring
mem = apex_b200.ReplayMemory(N, {"priority_exponent": 0.6, "layout": "stream", "lanes": 1, "n_step": 3, "state_shape": (1, 1, 16)})
mem._ensure()
g = torch.Generator(device=dev).manual_seed(1)
leaves = torch.randint(1, 1 << 30, (N,), dtype=torch.int64, device=dev, generator=g)
apex_b200._C.call("apexb200_tree_load_leaves", mem._h, apex_b200._C.ptr(leaves), 0, N, apex_b200._C.stream_ptr())
us = timed(lambda: apex_b200._C.call("apexb200_tree_build", mem._h, apex_b200._C.stream_ptr()))
nbytes = N * 8 * (1 + 1 / 32 + 1 / 1024) + N * 8 * (1 / 32 + 1 / 1024 + 1 / 32768)      # every level read once, every parent written once
out.append(dict(kernel="K1 tree_build", units="16M leaves (4 launches, one per level)", us=round(us, 2),
                alg_bytes=int(nbytes), GBps=round(nbytes / us / 1e3, 1), frac=round(nbytes / us / 1e3 / peak, 3)))
for B in (1 << 16, 1 << 18, 1 << 20):
    idx = torch.randint(0, N, (B,), dtype=torch.int32, device=dev, generator=g)
    prio = torch.rand(B, device=dev, generator=g) + 0.01
    us = timed(lambda: mem.update_priorities(idx, prio))
    # per row: idx 4 + prio 4 read, stamp 4 atomic + 4 read/clear, leaf 8 read + 8 write, 4 ancestor atomics of 8 B
    nbytes = B * (4 + 4 + 8 + 16 + 32)
    out.append(dict(kernel="K2 tree_update (stamp + apply)", units="%d random rows of 16M" % B, us=round(us, 2), alg_bytes=nbytes,
                    GBps=round(nbytes / us / 1e3, 1), frac=round(nbytes / us / 1e3 / peak, 3),
                    Mrows_per_s=round(B / us, 1)))
    q = [torch.randn(B, nA, device=dev, generator=g) for _ in range(3)]
    A = torch.randint(0, nA, (B,), dtype=torch.int32, device=dev, generator=g)
    R, G = torch.randn(B, device=dev, generator=g), torch.full((B,), 0.97, device=dev)
    ws = apex_b200.ops.td_workspace(B, dev)
    o4 = tuple(torch.empty(B, device=dev) for _ in range(3)) + (torch.empty(1, device=dev),)
    P_ = apex_b200._C.ptr
    us = timed(lambda: apex_b200._C.call("apexb200_td_fwd", P_(q[0]), P_(q[1]), P_(q[2]), P_(A), P_(R), P_(G), B, nA, 0.6, 0.0, 0, 1.0,
                                         None, P_(o4[0]), P_(o4[1]), P_(o4[2]), P_(o4[3]), P_(ws), int(ws.numel()),
                                         apex_b200._C.stream_ptr()))
    nbytes = B * (3 * nA * 4 + 12 + 12)                 # three Q rows + A, R, Gamma in; delta, prio, loss out
    out.append(dict(kernel="K5 td_multi (no tree)", units="%d rows, %d actions" % (B, nA), us=round(us, 2), alg_bytes=nbytes,
                    GBps=round(nbytes / us / 1e3, 1), frac=round(nbytes / us / 1e3 / peak, 3), Mrows_per_s=round(B / us, 1)))
    td_out = tuple(torch.empty(B, device=dev) for _ in range(3)) + (torch.empty(1, device=dev),)
    mem.workspace(B)
    us = timed(lambda: mem.td_update(idx, q[0], q[1], q[2], A, R, G, out=td_out))
    nbytes = B * (3 * nA * 4 + 12 + 12 + 4 + 8 + 16 + 32)
    out.append(dict(kernel="K5+K2 td_multi + stamp/apply", units="%d rows into 16M leaves" % B, us=round(us, 2), alg_bytes=nbytes,
                    GBps=round(nbytes / us / 1e3, 1), frac=round(nbytes / us / 1e3 / peak, 3), Mrows_per_s=round(B / us, 1)))
mem.check_status()
for r in out:
    print(json.dumps(r))
print(json.dumps({"peak_GBps": peak, "peak_source": peak_src}))
